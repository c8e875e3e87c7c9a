"""SASS evidence of the built library, no GPU needed: per-kernel register / stack / shared-memory use
(cuobjdump -res-usage) and per-kernel counts of the Blackwell-specific mnemonics (cuobjdump -sass):
tcgen05 = UTCHMMA / UTCBAR / LDTM, TMA tensor = UTMALDG, TMA bulk = UBLKCP, mbarrier = SYNCS, cluster barrier = UCGABAR.
usage: python scripts/sass_evidence.py <tag>   ->  profiles/<tag>_res_usage.txt, profiles/<tag>_sass_blackwell_mnemonics.txt"""
import collections, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "knowledge-graph-embedding_b200", "lib", "libkge_b200.so")
tag = sys.argv[1] if len(sys.argv) > 1 else "final"


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
rows, cur = [], None
for line in res.splitlines():
    m = re.match(r"\s*Function (\S+):", line)
    if m:
        cur = m.group(1)
    elif cur and "REG:" in line:
        rows.append((cur, " ".join(line.split())))
        cur = None
dm = demangle([r[0] for r in rows])
with open(os.path.join(ROOT, "profiles", tag + "_res_usage.txt"), "w") as fh:
    fh.write("# cuobjdump -res-usage knowledge-graph-embedding_b200/lib/libkge_b200.so (sm_100a): registers per thread, per-thread\n"
             "# stack (call frames AND register spills) and static shared memory per kernel.  Hundreds of bytes of stack at\n"
             "# REG:255 = spilling: the NV = 8 instantiations of k_step / k_rank_generic (rows of 513 .. 1024 floats), none of\n"
             "# which a BASELINE.json configuration runs (C2: NV = 2, 8 bytes; C3's wide rows go through k_step_rot)\n")
    for name, usage in sorted(rows, key=lambda r: dm[r[0]]):
        fh.write("%s\n    %s\n" % (dm[name], usage))

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
pat = re.compile(r"\b(UTCHMMA[\w.]*|UTCBAR[\w.]*|LDTM[\w.]*|UTMALDG[\w.]*|UBLKCP[\w.]*|SYNCS[\w.]*|UCGABAR[\w.]*|UTCCP[\w.]*)")
counts, cur = collections.OrderedDict(), None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    if cur:
        m = pat.search(line)
        if m:
            counts.setdefault(cur, collections.Counter())[m.group(1)] += 1
dm = demangle(list(counts))
with open(os.path.join(ROOT, "profiles", tag + "_sass_blackwell_mnemonics.txt"), "w") as fh:
    fh.write("# cuobjdump -sass knowledge-graph-embedding_b200/lib/libkge_b200.so (sm_100a): counts of Blackwell-specific mnemonics per kernel\n"
             "# (tcgen05 = UTCHMMA/UTCBAR/LDTM, TMA tensor = UTMALDG, TMA bulk = UBLKCP, mbarrier = SYNCS, cluster barrier = UCGABAR).\n"
             "# multimem.st (NVSwitch multicast stores of the owner exchange) assembles to a plain STG on a multicast address.\n\n")
    for name, c in counts.items():
        fh.write(dm[name] + "\n")
        for k, v in sorted(c.items()):
            fh.write("%9d  %s\n" % (v, k))
        fh.write("\n")
print(len(rows), "kernels;", len(counts), "with Blackwell-specific instructions")
