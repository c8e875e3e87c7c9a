#!/usr/bin/env python
"""Pack the reference's bundled, already indexed triple files into small .npz fixtures
(tests/golden/data_*.npz) so the GPU box — which has no /root/reference — can train and rank
on the REAL graphs (C1 = fb15k_237 train/valid, C2 rank parity = fb15k valid).

    python tests/golden/make_datasets.py          # in the build container

Only data (integer id triples) is packed, no reference source.  fb15k's train file is listed in
/root/reference/.MISSING_LARGE_BLOBS, so only its valid split exists."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/data"
FILES = {"fb15k237_train": "fb15k_237/train_indexed/train.csv",
         "fb15k237_valid": "fb15k_237/valid_indexed/valid.csv",
         "fb15k_valid": "fb15k/valid_indexed/valid.csv"}

if __name__ == "__main__":
    out = {}
    for name, rel in FILES.items():
        a = np.loadtxt(os.path.join(REF, rel), delimiter=",", dtype=np.int64)
        assert a.ndim == 2 and a.shape[1] == 3 and a.min() >= 0 and a.max() < 65536
        out[name] = a.astype(np.uint16)
        print(name, a.shape, "max ids", a.max(0))
    np.savez_compressed(os.path.join(HERE, "data_bundled.npz"), **out)
    print(os.path.getsize(os.path.join(HERE, "data_bundled.npz")), "bytes")
