"""DRAM traffic per launch of every kernel of the hot path, from the `ncu --set full` captures in gpurun_out/
(tools/gpu_s24z.sh), written to profiles/traffic.json for bench.py's `roofline.traffic`, plus one text summary per
capture under profiles/.   python tools/make_traffic.py <round tag, e.g. r01_g>"""
import csv, io, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
CAPS = {  # precision -> op -> captures (one kernel each) whose launches make up the op
    "tf32x3": {"corr_fwd": ["x3_fwd"], "corr_bwd": ["x3_bwd"], "warp_fwd": ["warp_fwd"], "warp_bwd": ["warp_bwd"],
               "downsample": ["ds_tables", "ds_band", "ds_combine"]},
    "tf32": {"corr_fwd": ["corr_fwd"], "corr_bwd": ["corr_bwd"], "warp_fwd": ["warp_fwd"], "warp_bwd": ["warp_bwd"],
             "downsample": ["ds_tables", "ds_band", "ds_combine"]}}
out = {"tf32x3": {}, "tf32": {}, "kernels": {}, "how": "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu "
       "--set full --clock-control none, batch 32 (tools/gpu_r2_evidence.sh); per op = sum over the op's kernels (the "
       "memset of warp_bwd's image gradient is not captured)"}
done = set()
for prec, ops in CAPS.items():
    for op, caps in ops.items():
        total = 0
        for c in caps:
            rep = os.path.join(ROOT, "gpurun_out", "prof_%s.ncu-rep" % c)
            if not os.path.exists(rep):
                total = None
                break
            if c not in out["kernels"]:
                raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
                r = list(csv.reader(io.StringIO(raw)))
                h, u, v = r[0], r[1], r[2]
                def get(k):
                    i = h.index(k)
                    x = float(v[i].replace(",", ""))
                    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u[i], 1)
                rd, wr = get("dram__bytes_read.sum"), get("dram__bytes_write.sum")
                out["kernels"][c] = {"kernel": v[h.index("Kernel Name")], "dram_read": rd, "dram_write": wr,
                                     "time_us": float(v[h.index("gpu__time_duration.sum")].replace(",", ""))}
                txt = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep, "12"],
                                     capture_output=True, text=True).stdout
                open(os.path.join(ROOT, "profiles", "%s_ncu_%s.txt" % (tag, c)), "w").write(txt)
            total += out["kernels"][c]["dram_read"] + out["kernels"][c]["dram_write"]
        if total is not None:
            out[prec][op] = total
json.dump(out, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print(json.dumps({k: out[k] for k in ("tf32x3", "tf32")}))
