"""Per-kernel counts of the SASS mnemonics that prove the tcgen05 / TMA / mma.sync paths (B200_PROFILING.md):
    python tools/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections, glob, os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MNEM = ["UTCIMMA", "UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTCBAR", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "IMMA", "HMMA",
        "LDSM", "HFMA2", "HMUL2", "FFMA2", "FFMA", "LDS", "STS", "BAR", "ELECT", "USETMAXREG"]
print("# cuobjdump -sass of the sm_100a objects in smoothquant_mixedprecision_b200/build (instruction counts per kernel)")
for obj in sorted(glob.glob(os.path.join(ROOT, "smoothquant_mixedprecision_b200", "build", "*.o"))):
    out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    name, counts, total = None, None, 0
    def flush():
        if name and total:
            short = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
            short = short.replace("(anonymous namespace)::", "").replace("void ", "").split("(")[0]
            hits = " ".join(f"{m}={counts[m]}" for m in MNEM if counts[m])
            print(f"{os.path.basename(obj):16s} {short[:70]:70s} insts={total:6d} {hits}")
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            flush()
            name, counts, total = m.group(1), collections.Counter(), 0
            continue
        m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_]+)", line)
        if m and name:
            total += 1
            op = m.group(1)
            for k in MNEM:
                if op == k or op.startswith(k + "."):
                    counts[k] += 1
    flush()
