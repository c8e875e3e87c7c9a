"""The C-ABI library loads on a CPU-only box and exports every symbol include/kge_abi.h
declares; compute entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "kge_abi.h")).read()
    return sorted(set(re.findall(r"^(?:int|const char\*) (kge_[a-z_]+)\(", src, re.M)))


def test_library_exports_every_declared_symbol():
    from KGE._b200 import lib
    L = lib.load()
    names = _declared()
    assert set(names) == set(lib.SYMBOLS)
    for n in names:
        assert getattr(L, n) is not None
    assert L.kge_abi_version() == lib.KGE_ABI_VERSION


def test_binding_argument_counts_match_header():
    """Every entry point's ctypes argtypes has as many entries as the header's declaration has parameters (a
    parameter added on one side only would otherwise read garbage off the stack instead of failing)."""
    from KGE._b200 import lib
    L = lib.load()
    src = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "kge_abi.h")).read(), flags=re.S)
    seen = 0
    for m in re.finditer(r"^(?:int|const char\*) (kge_[a-z_]+)\((.*?)\);", src, re.M | re.S):
        name, args = m.group(1), " ".join(m.group(2).split())
        n = 0 if args in ("", "void") else args.count(",") + 1
        at = getattr(L, name).argtypes
        if n == 0:
            assert not at, name
        else:
            assert at is not None and len(at) == n, (name, n, None if at is None else len(at))
        seen += 1
    assert seen == len(lib.SYMBOLS)


def test_struct_sizes_match_header():
    from KGE._b200 import lib
    assert C.sizeof(lib.Table) == 56 and C.sizeof(lib.Tables) == 4 * 56 + 16 + 8
    assert C.sizeof(lib.Cfg) == 12 * 4 + 10 * 4 + 2 * 8
    assert C.sizeof(lib.Sampler) == 56 and C.sizeof(lib.Filter) == 24
    assert C.sizeof(lib.Dp) == 24
    assert C.sizeof(lib.Grads) == 760   # kge_grads (gcc sizeof): dense, n_peers, peer slots, owner fields, abort_word, multicast


def test_struct_layouts_match_the_header_as_gcc_sees_it(tmp_path):
    """sizeof and the offset of every field of the nine ABI structs, from a C program compiled against
    include/kge_abi.h, against the ctypes mirrors (same names, same order)."""
    import subprocess
    from KGE._b200 import lib
    pairs = [("kge_table", lib.Table), ("kge_tables", lib.Tables), ("kge_cfg", lib.Cfg), ("kge_sampler", lib.Sampler),
             ("kge_filter", lib.Filter), ("kge_dp", lib.Dp), ("kge_grads", lib.Grads), ("kge_peer_flags", lib.PeerFlags),
             ("kge_shard", lib.Shard)]
    lines = ["#include <stdio.h>", "#include <stddef.h>", '#include "kge_abi.h"', "int main(void) {"]
    for cname, cls in pairs:
        lines.append('  printf("%s %%zu", sizeof(%s));' % (cname, cname))
        for f in cls._fields_:
            lines.append('  printf(" %%zu", offsetof(%s, %s));' % (cname, f[0]))
        lines.append('  printf("\\n");')
    lines += ["  return 0;", "}"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.strip().splitlines()
    for (cname, cls), line in zip(pairs, out):
        got = [int(x) for x in line.split()[1:]]
        want = [C.sizeof(cls)] + [getattr(cls, f[0]).offset for f in cls._fields_]
        assert got == want, (cname, got, want)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from KGE._b200 import lib
    L = lib.load()
    h = C.c_void_p()
    rc = L.kge_ctx_create(0, C.byref(h))
    assert rc == -3 and b"no CPU fallback" in L.kge_last_error()
    from KGE._b200 import engine
    with pytest.raises(RuntimeError):
        engine.Engine({"model": 2, "n_ent": 4, "n_rel": 2, "d_ent": 4, "d_rel": 4}, {})


def test_bad_cfg_is_rejected_without_gpu():
    from KGE._b200 import lib
    L = lib.load()
    cfg = lib.Cfg(abi_version=99)
    n = C.c_size_t()
    assert L.kge_workspace_bytes(C.byref(cfg), 1, 0, C.byref(n)) == -1
    assert b"abi_version" in L.kge_last_error()
