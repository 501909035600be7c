"""Print the SASS neighbourhood (with sample counts) of the hottest instructions of an .ncu-rep.
python tools/ncu_hot.py rep [ntop] [context]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 3; ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 8
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hh, rows = rows[1], rows[2:]
iS, iSrc, iI = hh.index("# Samples"), hh.index("Source"), hh.index("Instructions Executed")
order = sorted(range(len(rows)), key=lambda i: -int(rows[i][iS] or 0))[:ntop]
for o in order:
    print("---- around SASS row", o)
    for i in range(max(0, o - ctx), min(len(rows), o + ctx + 1)):
        print("%s%6s %8s  %s" % (">" if i == o else " ", rows[i][iS], rows[i][iI], rows[i][iSrc][:90]))
