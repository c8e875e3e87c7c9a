"""Filesystem helpers — drop-in for KGE/utils.py (:6-9, :18-26).  `ns_with_same_type`
(:11-16) has no counterpart here: typed sampling is the CUDA sampler (csrc/common.cuh)."""
import os
import shutil


def rmtree(top):
    shutil.rmtree(top, ignore_errors=True)


def check_path_exist_and_create(path):
    """Start from an empty directory (an existing one is removed, like the reference)."""
    if os.path.exists(path):
        rmtree(path)
    os.makedirs(path)
