"""Per CUDA source line: warp instructions executed and stall samples, from an .ncu-rep captured with
--import-source on.  python tools/ncu_lines.py rep [ntop]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
out, fname, hh = [], "", None
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hh = r
        iS, iI = hh.index("# Samples"), hh.index("Instructions Executed")
    elif hh and r and r[0] not in ("", "Function Name"):
        try:
            out.append((fname, int(r[0]), r[1].strip(), int(r[iS] or 0), int(r[iI] or 0)))
        except ValueError:
            pass
ti, ts = sum(o[4] for o in out), sum(o[3] for o in out)
print("warp instructions", ti, "samples", ts)
key = 3 if len(sys.argv) > 3 and sys.argv[3] == "smp" else 4
for o in sorted(out, key=lambda o: -o[key])[:ntop]:
    print("%5.1f%% inst %5.1f%% smp  %s:%d  %s" % (100.0 * o[4] / max(ti, 1), 100.0 * o[3] / max(ts, 1), o[0], o[1], o[2][:100]))
