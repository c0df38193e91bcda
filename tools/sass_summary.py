"""profiles/sass_summary.txt: opcode histogram per kernel of the in-tree library (cuobjdump -sass), the mnemonics that prove
tcgen05 / TMEM / TMA / cp.async use (B200_PROFILING.md): UTCHMMA[.2CTA] (tcgen05.mma), LDTM / STTM (tcgen05.ld / st),
UBLKCP (cp.async.bulk), UTMALDG / UTMASTG (tensor-map TMA), LDGSTS (cp.async), HMMA (legacy mma.sync), SYNCS (mbarrier)."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "xtts_b200", "libttts_mel.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
keys = ["UTCHMMA.2CTA", "UTCHMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTMALDG", "UTMASTG", "LDGSTS", "HMMA", "SYNCS", "LDG", "STG", "LDS", "STS", "FFMA", "MUFU"]
cur, per = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        per[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        per[cur]["total"] += 1
        for k in keys:
            if op == k or op.startswith(k + ".") or (k == "UTCHMMA" and op.startswith("UTCHMMA") and ".2CTA" not in op and False):
                per[cur][k] += 1
        if op.startswith("UTCHMMA") and ".2CTA" in op:
            per[cur]["UTCHMMA.2CTA"] += 1
print(f"# {os.path.relpath(lib, ROOT)}: SASS opcode counts per kernel (cuobjdump -sass; static instruction counts)")
print(f"{'kernel':70s} {'total':>6s} " + " ".join(f"{k:>8s}" for k in keys))
tot = collections.Counter()
for name, c in per.items():
    print(f"{name[:70]:70s} {c['total']:6d} " + " ".join(f"{c[k]:8d}" for k in keys))
    tot.update(c)
print(f"{'ALL':70s} {tot['total']:6d} " + " ".join(f"{tot[k]:8d}" for k in keys))
