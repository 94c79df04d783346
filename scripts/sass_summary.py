"""Static SASS instruction counts per kernel of libtsplat.so (cuobjdump -sass): the mnemonics that show which hardware
paths a kernel uses.  python scripts/sass_summary.py [lib] > profiles/<tag>_sass_summary.txt   (no GPU needed)"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "transplat_b200", "lib", "libtsplat.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
COLS = [("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS"), ("LDGSTS", r"\bLDGSTS"), ("REDG.F32x4/x2", r"\bREDG\.E\.ADD\.F32x[24]"),
        ("REDG.F32", r"\bREDG\.E\.ADD\.F32\."), ("..of which .SYS", r"\bREDG\.E\.ADD\.F32.*STRONG\.SYS"), ("LDGMC", r"\bLDGMC"),
        ("ATOMG/REDG.int", r"\bATOMG|\bREDG\.E\.ADD\.(STRONG|64)"), ("ATOMS", r"\bATOMS"), ("MUFU.EX2", r"MUFU\.EX2"),
        ("MUFU.RCP", r"MUFU\.RCP"), ("SHFL", r"\bSHFL"), ("VOTE", r"\bVOTE"), ("REDUX", r"\bREDUX|\bCREDUX"), ("BAR", r"\bBAR\.SYNC"),
        ("HMMA/UTC*MMA", r"\bHMMA|\bUTC.*MMA|\bIMMA|\bDMMA"), ("total", r"^\s+/\*[0-9a-f]{4}\*/")]
counts, name = collections.OrderedDict(), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(.*", "", name).replace("void ", "").replace("tsplat::", "")
        counts[name] = collections.Counter()
        continue
    if name is None or "/*" not in line:
        continue
    for col, pat in COLS:
        if re.search(pat, line):
            counts[name][col] += 1
print(f"# SASS summary of {os.path.relpath(lib, ROOT)} (cuobjdump -sass, sm_100a): static instruction counts per kernel")
print("# UBLKCP = 1-D TMA bulk copy, SYNCS = mbarrier, LDGSTS = cp.async gather, REDG = global reduction (F32x4/x2 = vector; .SYS scope = "
      "multimem.red through the NVLS multicast alias), LDGMC = multimem.ld_reduce, no tensor-core instruction anywhere (nothing on this path is a contraction)")
print(f"{'kernel':58s}" + "".join(f"{c:>14s}" for c, _ in COLS))
for k, c in counts.items():
    print(f"{k[:57]:58s}" + "".join(f"{c[col]:14d}" for col, _ in COLS))
