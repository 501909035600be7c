"""Per-kernel counts of the Blackwell-specific SASS instructions in flownet2-tf_b200/libfn2ops.so
(cuobjdump -sass): UTC*MMA (tcgen05.mma), LDTM/STTM (tcgen05.ld/st), UTMALDG/UTMASTG (TMA tensor load/store),
UBLKCP/UBLKRED (bulk copy / bulk reduce), UTCBAR (tcgen05.commit), SYNCS (mbarrier), ATOMS/RED/ATOMG.
    python tools/sass_summary.py > profiles/sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "flownet2-tf_b200", "libfn2ops.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
KEYS = ["UTCHMMA", "UTCQMMA", "UTCOMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP",
        "UBLKRED", "SYNCS", "ATOMS", "ATOMG", "REDG", "RED.", "STG", "LDG", "ELECT"]
counts, order, cur = collections.defaultdict(collections.Counter), [], None
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(anonymous namespace\)::", "", cur)
        cur = re.sub(r"\(.*$", "", cur).replace("void ", "")
        order.append(cur)
        continue
    if cur is None:
        continue
    m = re.search(r"/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if not m:
        continue
    op = m.group(1)
    counts[cur]["total"] += 1
    for k in KEYS:
        if op.startswith(k):
            counts[cur][k] += 1
            break
print("SASS summary of %s (sm_100a), instruction counts per kernel" % os.path.relpath(so, ROOT))
cols = ["total"] + [k for k in KEYS if any(counts[f][k] for f in order)]
print("%-58s" % "kernel" + "".join("%9s" % c.rstrip(".") for c in cols))
for f in order:
    print("%-58s" % f[:57] + "".join("%9d" % counts[f][c] for c in cols))
