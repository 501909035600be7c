#!/usr/bin/env python
"""bench.py -- throughput of the flownet2-tf hot path on B200 (see DESIGN.md "Measurement").

One *step* = one pass of the hot path over one batch of B synthetic 512x384 image pairs (per GPU):
    correlation forward + backward at the FlowNetC conv3 shape  [B,48,64,256] -> [B,48,64,441]  (flownet_c.py:40)
    flow_warp forward + backward on                             image [B,384,512,3], flow [B,384,512,2]
    downsample of the ground-truth flow [B,384,512,2] to the five loss scales (flownet_c.py:135-160)
metric = image pairs / s (whole job, all GPUs; batch-sharded, no collective on the data path => weak scaling).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision fp32|tf32|bf16|tf32x3]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0.  `--impl reference` times the CPU restatement of the reference ops (the reference
has no CPU implementation of these ops and its TF GPU ops cannot be built here) on all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DS_SIZES = [(6, 8), (12, 16), (24, 32), (48, 64), (96, 128)]
CORR_ARGS = (1, 20, 1, 2, 20)
HC, WC, CC, D = 48, 64, 256, 441
HI, WI = 384, 512


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("FN2_BENCH_PRECISION", "auto"))
    ap.add_argument("--eager", action="store_true", help="time eager launches instead of CUDA-graph replays")
    ap.add_argument("--batch", type=int, default=32, help="image pairs per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-ref-gpu", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------- bytes per pair
def algorithmic_bytes(batch):
    """Algorithmic HBM bytes per launch group, SURVEY.md section 8(d): every tensor read or written once."""
    f = 4
    act, vol = HC * WC * CC * f, HC * WC * D * f
    img, flo = HI * WI * 3 * f, HI * WI * 2 * f
    ds_out = sum(h * w * 2 * f for h, w in DS_SIZES)
    return {
        "corr_fwd": batch * (2 * act + vol),
        "corr_bwd": batch * (vol + 4 * act),
        "warp_fwd": batch * (2 * img + flo),
        "warp_bwd": batch * (2 * img + flo + 2 * img + flo),   # read image, flow, grad; zero+accumulate dimage; write dflow
        "downsample": batch * (flo + ds_out),   # one fused pass (fn2.downsample_multi): the flow is read once
    }


def corr_flops(batch):
    return {"corr_fwd": batch * 2.0 * HC * WC * D * CC, "corr_bwd": batch * 4.0 * HC * WC * D * CC}


# ------------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for (t, r) in self.rows if t0 - 0.05 <= t <= t1 + 0.05 and len(r) >= 8] or \
               [r for (_, r) in self.rows if len(r) >= 8]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for j, n in enumerate(names) if any(r[4 + j].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "samples": len(rows),
                "power_w_max": max(float(r[2]) for r in rows), "reasons": reasons}


# ------------------------------------------------------------------------------------------- CPU restatement
def cpu_pass(o, data, n):
    """One pass of the hot path over the first n pairs with the C restatement (all host threads)."""
    a, b, g, image, flow, gimg, gt = [x[:n] for x in data]
    o.c_correlation(a, b, *CORR_ARGS)
    o.c_correlation_grad(g, a, b, *CORR_ARGS)
    o.c_flow_warp(image, flow)
    o.c_flow_warp_grad(image, flow, gimg)
    for s in DS_SIZES:
        o.c_downsample(gt, s)


def host_data(batch, seed=0):
    import numpy as np
    import torch
    gen = torch.Generator().manual_seed(1234 + seed)
    lr = torch.nn.functional.leaky_relu
    a = lr(torch.randn(batch, HC, WC, CC, generator=gen), 0.1)
    b = lr(torch.randn(batch, HC, WC, CC, generator=gen), 0.1)
    g = torch.randn(batch, HC, WC, D, generator=gen)
    image = torch.rand(batch, HI, WI, 3, generator=gen)
    flow = torch.randn(batch, HI, WI, 2, generator=gen) * 8
    far = torch.rand(batch, HI, WI, generator=gen) < 0.05
    flow[far] = 2.0 * WI
    gimg = torch.randn(batch, HI, WI, 3, generator=gen)
    gt = torch.randn(batch, HI, WI, 2, generator=gen) * 5
    return [a, b, g, image, flow, gimg, gt]


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1; the CPU arm is meant to use every host thread.  Returns the thread count
    OpenMP will actually use (the number reported as `cores`)."""
    import ctypes
    n = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(n)
    try:
        gomp = ctypes.CDLL("libgomp.so.1")
        gomp.omp_set_num_threads(n)
        return int(gomp.omp_get_max_threads())
    except OSError:
        return n


def run_reference(args, rank):
    """--impl reference: the CPU restatement of the reference ops on all host threads, bounded sample per step."""
    if rank != 0:
        return
    from oracle import oracle as o
    o.lib()
    cores = use_all_host_threads()
    sample = 4
    data = [x.numpy() for x in host_data(sample)]
    for _ in range(min(args.warmup, 1)):
        cpu_pass(o, data, sample)
    t0 = time.time()
    for _ in range(args.steps):
        cpu_pass(o, data, sample)
    dt = time.time() - t0
    val = sample * args.steps / dt
    line = {
        "impl": "reference", "metric": "image_pairs_per_s", "value": val, "unit": "pairs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.batch, "fp32"),
        "cpu_baseline": {"value": val, "unit": "pairs/s", "cores": cores, "kind": "port",
                         "sample": "%d pairs per step (of the %d-pair batch), C restatement oracle/fn2_oracle.c, "
                                   "OpenMP on %d threads" % (sample, args.batch, cores)},
        "e2e": {"value": val, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(batch, precision):
    return {"workload": "hot path over %d synthetic 512x384 pairs per GPU: correlation fwd+bwd @ FlowNetC conv3 "
                        "[%d,48,64,256]->[%d,48,64,441] (k=1,d=20,s1=1,s2=2,pad=20) + flow_warp fwd+bwd "
                        "[%d,384,512,3] + downsample of GT flow to 5 loss scales" % (batch, batch, batch, batch),
            "pairs_per_gpu": batch, "correlation_precision": precision, "parallelism": "batch-sharded replicas",
            "l2": "inputs larger than L2 (1.6 GB per step vs 126 MB)",
            "launch": "CUDA graph replay per op, 5 per step (eager launches with --eager)"}


# ------------------------------------------------------------------------------------------------------ ours
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    import flownet2_tf_b200 as fn2
    from flownet2_tf_b200 import _lib

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()

    precision = args.precision
    if precision == "auto":
        # the tcgen05 path is the one BASELINE.json's target is stated for; its tolerance is in tests/test_gpu_tc.py
        precision = "tf32"
    B = args.batch
    host = host_data(B, seed=rank)
    a, b, g, image, flow, gimg, gt = [x.to(dev) for x in host]
    ops = ["corr_fwd", "corr_bwd", "warp_fwd", "warp_bwd", "downsample"]

    def step(ev=None):
        def mark(i):
            if ev is not None:
                ev[i].record()
        mark(0)
        out = fn2.correlation(a, b, *CORR_ARGS, precision=precision)
        mark(1)
        ga, gb = fn2.correlation_grad(g, a, b, *CORR_ARGS, precision=precision)
        mark(2)
        warped = fn2.flow_warp(image, flow)
        mark(3)
        gi, gf = fn2.flow_warp_grad(image, flow, gimg)
        mark(4)
        ds = fn2.downsample_multi(gt, DS_SIZES)
        mark(5)
        return out, ga, gb, warped, gi, gf, ds

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    warm = max(args.warmup, 3)
    for _ in range(warm):
        step()
    barrier()

    # Each op of the step is captured as its own CUDA graph (the C ABI is capturable: it only enqueues on the given
    # stream); a step = five graph replays with an event between them, so the per-op device times come from INSIDE
    # the timed region and the region is free of Python launch latency.  --eager times plain launches instead.
    side = torch.cuda.Stream(device=dev)
    op_fns = [
        lambda: fn2.correlation(a, b, *CORR_ARGS, precision=precision),
        lambda: fn2.correlation_grad(g, a, b, *CORR_ARGS, precision=precision),
        lambda: fn2.flow_warp(image, flow),
        lambda: fn2.flow_warp_grad(image, flow, gimg),
        lambda: fn2.downsample_multi(gt, DS_SIZES),
    ]
    graphs, keep, launches_per_step = [], [], 0
    if not args.eager:
        with torch.cuda.stream(side):
            for fn in op_fns:
                fn()
                torch.cuda.synchronize()
                gr = torch.cuda.CUDAGraph()
                l0 = _lib.launch_count()
                with torch.cuda.graph(gr, stream=side):
                    keep.append(fn())
                launches_per_step += _lib.launch_count() - l0
                graphs.append(gr)
            for _ in range(warm):
                for gr in graphs:
                    gr.replay()
        barrier()

    sampler = ClockSampler(local) if rank == 0 else None
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(args.steps)]
    launches0 = _lib.launch_count()
    barrier()
    t0 = time.time()
    if graphs:
        with torch.cuda.stream(side):
            for k in range(args.steps):
                for i, gr in enumerate(graphs):
                    evs[k][i].record()
                    gr.replay()
            evs[args.steps - 1][5].record()
    else:
        for k in range(args.steps):
            step(evs[k])
    barrier()
    t1 = time.time()
    launches = launches_per_step * args.steps if graphs else _lib.launch_count() - launches0
    last = 5
    elapsed_ms = evs[0][0].elapsed_time(evs[args.steps - 1][last])
    clocks = sampler.stop(t0, t1) if sampler else None
    if world > 1:
        t = torch.tensor([elapsed_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms = float(t.item())

    def op_end(k, i):   # the event that closes op i of step k
        if i < 4 or not graphs:
            return evs[k][i + 1]
        return evs[k + 1][0] if k + 1 < args.steps else evs[k][5]
    per_op_ms = {name: sum(evs[k][i].elapsed_time(op_end(k, i)) for k in range(args.steps)) / args.steps
                 for i, name in enumerate(ops)}

    # ---- end to end through the public API with HOST buffers (pinned), copies inside the timed region --------
    # Every step uploads ALL of its inputs from pinned host memory and downloads ALL of its results.  Three streams
    # (upload / compute / download) with double-buffered device inputs and host outputs let step k+1's upload run
    # under step k's compute and download (PCIe is full duplex); nothing is skipped or reused between steps.
    e2e = None
    if not args.no_e2e:
        pinned = [x.pin_memory() for x in host]
        h2d = sum(x.numel() * 4 for x in pinned)
        e2e_steps = max(4, min(args.steps, 20))
        s_up, s_dn = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        s_cp = torch.cuda.current_stream(dev)
        dev_in = [[torch.empty_like(x, device=dev) for x in pinned] for _ in range(2)]
        host_out = [None, None]
        up_done = [torch.cuda.Event() for _ in range(2)]
        cp_done = [torch.cuda.Event() for _ in range(2)]
        dn_done = [torch.cuda.Event() for _ in range(2)]
        live = [None, None]

        def e2e_step(k):
            sl = k & 1
            with torch.cuda.stream(s_up):
                s_up.wait_event(cp_done[sl])          # step k-2 has finished reading this input set
                for d, h in zip(dev_in[sl], pinned):
                    d.copy_(h, non_blocking=True)
                up_done[sl].record(s_up)
            s_cp.wait_event(up_done[sl])
            s_cp.wait_event(dn_done[sl])              # step k-2's results have left the device
            da, db, dg, dimg, dflow, dgimg, dgt = dev_in[sl]
            res = [fn2.correlation(da, db, *CORR_ARGS, precision=precision)]
            res += list(fn2.correlation_grad(dg, da, db, *CORR_ARGS, precision=precision))
            res.append(fn2.flow_warp(dimg, dflow))
            res += list(fn2.flow_warp_grad(dimg, dflow, dgimg))
            res += fn2.downsample_multi(dgt, DS_SIZES)
            cp_done[sl].record(s_cp)
            live[sl] = res                            # keep the device results alive until their download ran
            if host_out[sl] is None:
                host_out[sl] = [torch.empty(r.shape, dtype=r.dtype).pin_memory() for r in res]
            with torch.cuda.stream(s_dn):
                s_dn.wait_event(cp_done[sl])
                for h, r in zip(host_out[sl], res):
                    h.copy_(r, non_blocking=True)
                    r.record_stream(s_dn)
                dn_done[sl].record(s_dn)

        for k in range(4):
            e2e_step(k)
        barrier()
        s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s2.record(s_cp)
        s_up.wait_event(s2)
        for k in range(e2e_steps):
            e2e_step(k)
        s_cp.wait_event(dn_done[0])
        s_cp.wait_event(dn_done[1])
        e2.record(s_cp)
        barrier()
        ms = s2.elapsed_time(e2)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        d2h = sum(h.numel() * 4 for h in host_out[0])
        e2e = {"value": world * B * e2e_steps / (ms / 1e3), "unit": "pairs/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "steps": e2e_steps, "ms_per_step": ms / e2e_steps,
               "how": "public Python API, pinned host inputs and outputs, upload/compute/download streams"}
        del live, dev_in

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ------------------------------------------------------------------
    peaks, peak_src = {}, "fallback"
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peak_src = "measured"
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    abytes, flops = algorithmic_bytes(B), corr_flops(B)
    dominant = max(per_op_ms, key=per_op_ms.get)
    per_op = {}
    for name in ops:
        gbs = abytes[name] / (per_op_ms[name] * 1e-3) / 1e9
        per_op[name] = {"ms": round(per_op_ms[name], 4), "algorithmic_mb": round(abytes[name] / 1e6, 2),
                        "gb_s": round(gbs, 1), "hbm_frac": round(gbs / hbm_peak, 4)}
        if name in flops:
            per_op[name]["useful_tflop_s"] = round(flops[name] / (per_op_ms[name] * 1e-3) / 1e12, 2)
    achieved = per_op[dominant]["gb_s"]
    roofline = {"kernel": dominant, "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": round(achieved / hbm_peak, 4), "traffic": None, "peak_source": peak_src,
                "correlation_fwd_bwd_hbm_frac": round(
                    (abytes["corr_fwd"] + abytes["corr_bwd"]) / ((per_op_ms["corr_fwd"] + per_op_ms["corr_bwd"]) * 1e-3)
                    / 1e9 / hbm_peak, 4)}
    # DRAM bytes per launch (read + write) from the committed `ncu --set full` captures of these kernels at batch 32
    # (profiles/traffic.json, written by tools/make_traffic.py); only valid for the batch it was captured at
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof) and B == 32:
        try:
            tr = json.load(open(prof)).get(precision, {})
            roofline["traffic"] = tr.get(dominant)
            for name in ops:
                if name in tr:
                    per_op[name]["dram_mb_ncu"] = round(tr[name] / 1e6, 2)
        except Exception:
            pass

    # ---- CPU baseline: the restatement of the reference ops on the host cores, bounded sample ---------------
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        from oracle import oracle as o
        o.lib()
        ncores = use_all_host_threads()
        sample = min(B, 16)
        data = [x.numpy() for x in host]
        cpu_pass(o, data, 1)
        tc = time.time()
        cpu_pass(o, data, sample)
        dt = time.time() - tc
        cpu = {"value": sample / dt, "unit": "pairs/s", "cores": ncores, "kind": "port",
               "sample": "%d of the %d pairs, one pass, C restatement of the reference kernels (oracle/fn2_oracle.c), "
                         "OpenMP on %d threads, %.1f s" % (sample, B, ncores, dt)}

    # ---- the reference's own CUDA kernels on this GPU (informational; they are not a CPU path) --------------
    ref_gpu = None
    if world == 1 and not args.no_ref_gpu:
        try:
            from oracle import ref
            if ref.available():
                temps = (torch.empty((B, HC + 40, WC + 40, CC), device=dev), torch.empty((B, HC + 40, WC + 40, CC), device=dev))
                def ref_step():
                    ref.correlation(a, b, *CORR_ARGS, temps=temps)
                    ref.correlation_grad(g, a, b, *CORR_ARGS, temps=temps)
                    ref.flow_warp(image, flow)
                    for s in DS_SIZES:
                        ref.downsample(gt, s)
                ref_step()
                torch.cuda.synchronize()
                tr = time.time()
                ref_step()
                torch.cuda.synchronize()
                dt = time.time() - tr
                ref_gpu = {"value": B / dt, "unit": "pairs/s", "ms_per_step": dt * 1e3,
                           "what": "reference src/ops/**.cu.cc kernels compiled unmodified for sm_100a "
                                   "(oracle/_ref), same batch, without flow_warp backward (broken in the reference)"}
        except Exception as exc:  # pragma: no cover
            ref_gpu = {"unavailable": str(exc)[:200]}

    value = world * B * args.steps / (elapsed_ms / 1e3)
    line = {
        "metric": "image_pairs_per_s", "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": {"fp32": "f32", "tf32": "tf32", "bf16": "bf16",
                                                          "tf32x3": "tf32x3"}[precision],
        "data": "synthetic", "config": workload_config(B, precision), "clocks": clocks, "e2e": e2e,
        "gpu_launches": int(launches), "roofline": roofline, "per_op": per_op, "cpu_baseline": cpu,
        "reference_gpu_kernels": ref_gpu,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
