"""SASS instruction summary of libb200belos.so: per kernel, how many TMA tensor copies (UTMALDG), bulk copies
(UBLKCP), FP64 tensor-core instructions (DMMA), mbarrier operations (SYNCS), FP64 FMAs (DFMA) and 128/256-bit global
loads the binary really contains.  usage: python tools/sass_summary.py > profiles/r2_sass_summary.txt"""
import collections
import re
import subprocess
import sys

SO = sys.argv[1] if len(sys.argv) > 1 else "trilinos_b200/lib/libb200belos.so"
PATS = [("UTMALDG", r"\bUTMALDG"), ("UBLKCP", r"\bUBLKCP"), ("UBLKPF", r"\bUBLKPF"), ("DMMA", r"\bDMMA"), ("SYNCS", r"\bSYNCS"),
        ("DFMA", r"\bDFMA"), ("LDG.128/256", r"\bLDG\.E\.(128|ENL2\.256|256)|\bLDG\.E\.[A-Z0-9.]*(128|256)"),
        ("LDS", r"\bLDS"), ("ST.*SYS", r"\bST\.E[A-Z0-9.]*SYS|\bSTG\.E[A-Z0-9.]*SYS")]
out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
counts, cur = collections.OrderedDict(), None
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    for name, pat in PATS:
        if re.search(pat, ln):
            counts[cur][name] += 1
dem = subprocess.run(["cu++filt"] + list(counts), capture_output=True, text=True).stdout.splitlines()
print(f"# {SO}: SASS instruction counts per kernel (cuobjdump -sass, sm_100a)")
print("# " + " ".join(f"{n:>11s}" for n, _ in PATS) + "  kernel")
tot = collections.Counter()
for (fn, c), d in zip(counts.items(), dem):
    d = re.sub(r"\(anonymous namespace\)::|b200::|void ", "", d)
    d = re.sub(r"\((int|bool|unsigned int)\)", "", d)
    d = re.sub(r"\(.*", "", d)
    print("  " + " ".join(f"{c[n]:11d}" for n, _ in PATS) + "  " + d)
    tot.update(c)
print("# " + " ".join(f"{tot[n]:11d}" for n, _ in PATS) + "  TOTAL")
