#!/usr/bin/env python
"""bench.py -- throughput of the flownet2-tf hot path on B200 (see DESIGN.md "Measurement").

One *step* = one pass of the hot path over one batch of B synthetic 512x384 image pairs (per GPU):
    correlation forward + backward at the FlowNetC conv3 shape  [B,48,64,256] -> [B,48,64,441]  (flownet_c.py:40)
    flow_warp forward + backward on                             image [B,384,512,3], flow [B,384,512,2]
    downsample of the ground-truth flow [B,384,512,2] to the five loss scales (flownet_c.py:135-160)
metric = image pairs / s (whole job, all GPUs; batch-sharded, no collective on the data path => weak scaling).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision auto|fp32|tf32x3|tf32]
                    [--config hotpath|strips4k|flownetc_train|flownet2_infer]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0.  The headline runs the correlation at the drop-in default precision ("auto" =
split-tf32, which meets the reference's float32 tolerance); `per_precision` carries the other modes.  Extra keys:
`cfg1` (one pair, warm and cold L2), `cfg2` (batch-8 warp + downsample), `reference_gpu_kernels` (the reference's own
CUDA kernels on this GPU, per op), `e2e_copy_ceiling`, `cfg3` / `cfg4` (FlowNetC training step and FlowNet2 stacked
inference through flownet2_tf_b200.graphs, 8 pairs per GPU; also `--config flownetc_train | flownet2_infer`), and
`strips4k` (BASELINE config 5: the 4K correlation cut into N row strips, NCCL halo exchange inside the timed region;
N = 1 is the whole pair on one GPU, the strong-scaling denominator; also `--config strips4k` on its own).
`--impl reference` times the CPU restatement of the reference ops (the reference has no CPU implementation of these
ops and its TF GPU ops cannot be built here) on all host threads.
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
H4K, W4K = 272, 480          # conv3 of a 3840x2160 pair padded to 3840x2176 (net.py:301-309)
REF_SAMPLE = 16              # pairs per step of the CPU arm


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("FN2_BENCH_PRECISION", "auto"))
    ap.add_argument("--config", default="hotpath", choices=["hotpath", "strips4k", "flownetc_train", "flownet2_infer"])
    ap.add_argument("--eager", action="store_true", help="time eager launches instead of CUDA-graph replays")
    ap.add_argument("--batch", type=int, default=32, help="image pairs per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-ref-gpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip per_precision / cfg1 / cfg2 / strips4k extras")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------- bytes per pair
def algorithmic_bytes(batch, hc=HC, wc=WC):
    """Algorithmic HBM bytes per launch group, SURVEY.md section 8(d): every tensor read or written once."""
    f = 4
    act, vol = hc * wc * CC * f, hc * wc * D * f
    img, flo = HI * WI * 3 * f, HI * WI * 2 * f
    ds_out = sum(h * w * 2 * f for h, w in DS_SIZES)
    return {
        "corr_fwd": batch * (2 * act + vol),
        "corr_bwd": batch * (vol + 4 * act),
        "warp_fwd": batch * (2 * img + flo),
        "warp_bwd": batch * (2 * img + flo + 2 * img + flo),   # read image, flow, grad; zero+accumulate dimage; write dflow
        "downsample": batch * (flo + ds_out),   # one fused pass (fn2.downsample_multi): the flow is read once
    }


def corr_flops(batch, hc=HC, wc=WC):
    return {"corr_fwd": batch * 2.0 * hc * wc * D * CC, "corr_bwd": batch * 4.0 * hc * wc * D * CC}


def executed_mma_flops(batch, precision, hc=HC, wc=WC):
    """FLOPs the tensor-core kernels actually issue (band tiles include pairs outside the +-10 window), from their
    tilings: corr_tc.cu (8-row pair tiles, 4 b-rows x 32 cols per block, M256 N128), corr_tc_x3.cu (2-row tiles,
    M128 N256 with the hi/lo split folded in), corr_tc_bwd.cu (8-row tiles, one item per source row, M128 N256 K32 per
    128 channels; three MMAs per step in the split mode).  Parity classes: (hc/2) x (wc/2) pixels each."""
    if precision not in ("tf32", "tf32x3"):
        return None
    hcl, wcl = (hc + 1) // 2, (wc + 1) // 2
    tx = (wcl + 31) // 32
    nhx = 2 if tx > 1 else 1
    # forward tilings: tf32 narrow = CTA pair, 8 a-rows, blocks of 4 b-rows (M256 N128); tf32 wide = 4 a-rows (M128
    # N128); tf32x3 narrow = CTA pair, 4 a-rows, blocks of 4 b-rows with the hi/lo split folded in (M256 N256);
    # tf32x3 wide = 2 a-rows, blocks of 2 b-rows (M128 N128)
    if precision == "tf32":
        rows_f, brows, m, n = (8, 4, 256, 128) if tx == 1 else (4, 4, 128, 128)
    else:
        rows_f, brows, m, n = (4, 4, 256, 256) if tx == 1 else (2, 2, 128, 128)
    fwd = 0.0
    for y0 in range(0, hcl, rows_f):
        lo, hi = max(0, y0 - 10), min(hcl - 1, y0 + rows_f - 1 + 10)
        blocks = (hi - lo + brows) // brows * nhx
        fwd += blocks * 2.0 * m * n * CC
    bwd = 0.0
    for r0 in range(0, hcl, 8):
        lo, hi = max(0, r0 - 10), min(hcl - 1, r0 + 7 + 10)
        items = (hi - lo + 1) * nhx
        bwd += items * 2.0 * CC * 256 * 32 * (3 if precision == "tf32x3" else 1)
    per_image = 4 * tx
    return {"corr_fwd": batch * per_image * fwd, "corr_bwd": batch * per_image * bwd * 2}


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
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(n)
    try:
        gomp = ctypes.CDLL("libgomp.so.1")
        gomp.omp_set_num_threads(n)
        return int(gomp.omp_get_max_threads())
    except OSError:
        return n


def workload_config(batch):
    """The same dict in both arms (ours and --impl reference): the workload, not the implementation."""
    return {"workload": "hot path over %d synthetic 512x384 pairs per GPU: correlation fwd+bwd @ FlowNetC conv3 "
                        "[%d,48,64,256]->[%d,48,64,441] (k=1,d=20,s1=1,s2=2,pad=20) + flow_warp fwd+bwd "
                        "[%d,384,512,3] + downsample of GT flow to 5 loss scales" % (batch, batch, batch, batch),
            "pairs_per_gpu": batch, "parallelism": "batch-sharded replicas",
            "l2": "inputs larger than L2 (1.6 GB per step vs 126 MB)"}


def run_reference(args, rank):
    """--impl reference: the CPU restatement of the reference ops on all host threads; every step is a bounded sample
    (REF_SAMPLE pairs) of the batch, with the warm-up the caller asked for."""
    if rank != 0:
        return
    from oracle import oracle as o
    o.lib()
    cores = use_all_host_threads()
    sample = min(REF_SAMPLE, args.batch)
    data = [x.numpy() for x in host_data(sample)]
    for _ in range(args.warmup):
        cpu_pass(o, data, sample)
    t0 = time.time()
    for _ in range(args.steps):
        cpu_pass(o, data, sample)
    dt = time.time() - t0
    val = sample * args.steps / dt
    line = {
        "impl": "reference", "metric": "image_pairs_per_s", "value": val, "unit": "pairs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.batch),
        "cpu_baseline": {"value": val, "unit": "pairs/s", "cores": cores, "kind": "port",
                         "sample": "%d pairs per step (of the %d-pair batch), C restatement oracle/fn2_oracle.c, "
                                   "OpenMP on %d threads" % (sample, args.batch, cores)},
        "e2e": {"value": val, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------- timing helpers
def graph_time(torch, fns, reps, stream, rounds=3):
    """Median device time (ms) of ONE pass over the list of callables `fns`, each a (rotating) variant of the same op:
    they are captured into one CUDA graph so that consecutive launches never find their inputs in L2, and the graph
    replay is timed with CUDA events on `stream`."""
    with torch.cuda.stream(stream):
        for fn in fns:
            fn()
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=stream):
            keep = [fn() for fn in fns]
        for _ in range(2):
            gr.replay()
        torch.cuda.synchronize()
        ts = []
        for _ in range(rounds):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(reps):
                gr.replay()
            e1.record(stream)
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) / (reps * len(fns)))
    del keep
    ts.sort()
    return ts[len(ts) // 2]


def op_entry(ms, nbytes, hbm_peak, flops=None, mma=None, tf32_peak=None):
    gbs = nbytes / (ms * 1e-3) / 1e9
    e = {"ms": round(ms, 4), "algorithmic_mb": round(nbytes / 1e6, 2), "gb_s": round(gbs, 1),
         "hbm_frac": round(gbs / hbm_peak, 4)}
    if flops:
        e["useful_tflop_s"] = round(flops / (ms * 1e-3) / 1e12, 2)
    if mma and tf32_peak:
        e["mma_tflop_s"] = round(mma / (ms * 1e-3) / 1e12, 1)
        e["tensor_frac"] = round(mma / (ms * 1e-3) / 1e12 / tf32_peak, 4)
    return e


def load_peaks():
    peaks, src = {}, "fallback"
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        src = "measured"
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    # kind::tf32 runs at half the bf16 rate (K = 8 against 16 per instruction); no tf32 GEMM is measured by the driver
    tf32 = float(peaks.get("bf16_tflops", 1590.0)) / 2.0
    return hbm, tf32, src


# ----------------------------------------------------------------------------------------- strips (config 5)
def strips_bench(torch, dist, fn2, dev, rank, world, precision, steps, warm, hbm_peak):
    """BASELINE config 5: ONE 4K pair's conv3 correlation [1,272,480,256], forward + backward, cut into `world` row
    strips (strong scaling); the halo rows of input_b and the mirror halo of backprops_b travel by NCCL send/recv inside
    the timed region.  Strips of >= 100 rows (2 GPUs: 136) overlap the exchange with their interior rows; the 68- and
    34-row strips of 4 and 8 GPUs exchange first and run one kernel per pass (flownet2_tf_b200.strips.overlap_used; the
    line says which under collective.overlap)."""
    from flownet2_tf_b200 import strips
    if H4K % world != 0 or (H4K // world) % 2 != 0 or H4K // world < 20:
        return {"unavailable": "272 rows do not split into %d even strips of >= 20 rows" % world}
    rows = H4K // world
    gen = torch.Generator(device=dev).manual_seed(77 + rank)
    lr = torch.nn.functional.leaky_relu
    a = lr(torch.randn(1, rows, W4K, CC, device=dev, generator=gen), 0.1).requires_grad_()
    if world > 1:
        b_ext, b_view = strips.alloc_b_ext(1, rows, W4K, CC, dev)
        b_view.copy_(lr(torch.randn(1, rows, W4K, CC, device=dev, generator=gen), 0.1))
        b = b_view.detach().requires_grad_()
    else:
        b_ext = None
        b = lr(torch.randn(1, rows, W4K, CC, device=dev, generator=gen), 0.1).requires_grad_()
    go = torch.randn(1, rows, W4K, D, device=dev, generator=gen)

    def step():
        a.grad = None
        b.grad = None
        if world > 1:
            out = strips.correlation_strips(a, b, *CORR_ARGS, precision=precision, b_ext=b_ext)
        else:
            out = fn2.correlation(a, b, *CORR_ARGS, precision=precision)
        out.backward(go)
        return out

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(warm, 3)):
        step()
    sync()
    # the exchange alone (what the limiting collective costs): 20 rows of b each way per boundary, ncclSend/Recv
    comm_ms = None
    if world > 1:
        h = strips.halo_rows(CORR_ARGS[1], CORR_ARGS[3])
        up, down = strips._neighbours(None)
        own = b_ext[:, (h if up is not None else 0):][:, :rows]
        su, sd = own[:, :h], own[:, rows - h:]
        ru = torch.empty_like(su) if up is not None else None
        rd = torch.empty_like(sd) if down is not None else None
        for _ in range(3):
            strips._finish(strips._start_exchange(su, ru, sd, rd, up, down, None))
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            strips._finish(strips._start_exchange(su, ru, sd, rd, up, down, None))
        e1.record()
        sync()
        t = torch.tensor([e0.elapsed_time(e1) / 20], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        comm_ms = float(t.item())
    def timed_steps(fn):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        sync()
        ms_ = e0.elapsed_time(e1) / steps
        if world > 1:
            t_ = torch.tensor([ms_], device=dev)
            dist.all_reduce(t_, op=dist.ReduceOp.MAX)
            ms_ = float(t_.item())
        return ms_

    eager_ms = timed_steps(step)
    # The whole step -- kernels, the two ncclSend/Recv exchanges, the adds -- captured in ONE CUDA graph: every call only
    # enqueues work, so the step replays without Python between its ~12 launches (at 4K the eager step is bound by the
    # host: ~0.9 ms of launch calls for ~0.3 ms of device work per rank).  Every rank decides the same way.
    graph_ms, graph_note = None, None
    if os.environ.get("FN2_STRIPS_GRAPH", "1") != "0":
        try:
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(3):
                    step()
                a.grad = None
                b.grad = None
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=side):
                    step()
                for _ in range(3):
                    gr.replay()
            torch.cuda.current_stream(dev).wait_stream(side)
            sync()
            graph_ms = timed_steps(gr.replay)
        except Exception as exc:  # pragma: no cover
            graph_note = "capture failed: %s" % str(exc)[:160]
    ms = graph_ms if graph_ms is not None else eager_ms
    ab, fl = algorithmic_bytes(1, H4K, W4K), corr_flops(1, H4K, W4K)
    nbytes = ab["corr_fwd"] + ab["corr_bwd"]
    return {
        "workload": "one 4K pair: correlation fwd+bwd at conv3 [1,272,480,256]->[1,272,480,441], %d row strips of %d"
                    % (world, rows),
        "precision": precision, "scaling": "strong", "n_gpus": world, "steps": steps, "ms_per_step": round(ms, 4),
        "pairs_per_s": round(1e3 / ms, 1), "gb_s_aggregate": round(nbytes / (ms * 1e-3) / 1e9, 1),
        "hbm_frac_of_n_gpus": round(nbytes / (ms * 1e-3) / 1e9 / (hbm_peak * world), 4),
        "useful_tflop_s": round((fl["corr_fwd"] + fl["corr_bwd"]) / (ms * 1e-3) / 1e12, 1),
        "collective": None if world == 1 else {
            "what": "ncclSend/ncclRecv (torch.distributed batch_isend_irecv), 20 rows of input_b per direction per "
                    "boundary in forward, 20 rows of backprops_b partial sums in backward",
            "bytes_per_direction_per_boundary": 20 * W4K * CC * 4, "comm_nranks": world,
            "exchange_alone_ms": round(comm_ms, 4),
            "overlap": bool(strips.overlap_used(1, rows, W4K, CC, 20, 2, rank, world)),
            "overlap_note": "true: the exchange runs under the strip's interior rows; false: exchange first, then "
                            "one kernel over the strip (strips under %d rows)" % strips.OVERLAP_MIN_ROWS},
        "launch": "one CUDA graph per step (kernels + NCCL send/recv captured)" if graph_ms is not None
                  else "eager autograd calls (correlation_strips forward + backward)",
        "ms_per_step_eager": round(eager_ms, 4), "graph_note": graph_note,
    }


# ------------------------------------------------------------------------- caller graphs (configs 3 and 4)
def graphs_bench(torch, dist, dev, rank, world, precision, config, steps, warm, batch):
    """BASELINE configs 3 and 4, end to end through flownet2_tf_b200.graphs (the three ops are this library's kernels,
    the convolutions are cuDNN library calls on channels_last tensors, TF32 allowed as torch's default):
      flownetc_train : FlowNetC forward + multi-scale loss + backward + Adam step, `batch` 512x384 pairs per GPU,
                       random-init weights, synthetic pairs; N > 1 = DistributedDataParallel (NCCL gradient all-reduce)
      flownet2_infer : FlowNet2 (CSS + SD + fusion) stacked inference, `batch` pairs per GPU, no gradients."""
    from flownet2_tf_b200 import graphs as G
    torch.manual_seed(1234)
    gen = torch.Generator(device=dev).manual_seed(99 + rank)
    inputs = {"input_a": torch.rand(batch, HI, WI, 3, device=dev, generator=gen),
              "input_b": torch.rand(batch, HI, WI, 3, device=dev, generator=gen)}
    if config == "flownetc_train":
        net = G.FlowNetC(precision=precision).to(dev)
        ops = net.ops
        model = net
        if world > 1:
            model = torch.nn.parallel.DistributedDataParallel(net, device_ids=[dev.index], gradient_as_bucket_view=True)
        opt = torch.optim.Adam(net.parameters(), lr=1e-4)     # training_schedules.py: Adam, 1e-4
        gt = torch.randn(batch, HI, WI, 2, device=dev, generator=gen) * 5

        def step():
            opt.zero_grad(set_to_none=True)
            out = model(inputs)
            loss, _ = G.multiscale_loss(gt, out, ops)
            loss = loss + G.l2_regularization(net, 4e-4)      # flownet_c.py:27: weight decay 0.0004
            loss.backward()
            opt.step()
            return loss
        what = "FlowNetC forward + multi-scale EPE loss + backward + Adam step"
    else:
        net = G.FlowNet2(precision=precision).to(dev).eval()

        def step():
            with torch.no_grad():
                return net(inputs)["flow"]
        what = "FlowNet2 (CSS + SD + fusion) stacked inference"

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(warm, 3)):
        step()
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        res = step()
    e1.record()
    sync()
    ms = e0.elapsed_time(e1) / steps
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ok = bool(torch.isfinite(res.detach() if hasattr(res, "detach") else res).all())
    return {"metric": "image_pairs_per_s", "value": world * batch / (ms * 1e-3), "unit": "pairs/s", "n_gpus": world,
            "steps": steps, "warmup": max(warm, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": precision, "data": "synthetic",
            "config": {"workload": "%s, %d synthetic 512x384 pairs per GPU, random-init weights" % (what, batch),
                       "pairs_per_gpu": batch,
                       "parallelism": ("DistributedDataParallel, NCCL gradient all-reduce over %d ranks" % world)
                       if (config == "flownetc_train" and world > 1) else "batch-sharded replicas",
                       "convolutions": "cuDNN (library), channels_last, TF32 allowed; correlation / flow_warp / "
                                       "downsample: this library's kernels, correlation precision %s" % precision},
            "finite": ok, "launch": "eager PyTorch"}


# ------------------------------------------------------------------------------------------------------ ours
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    # With several ranks on one host, give each its own slice of the cores (its pinned buffers are then first-touched
    # and its copy threads run next to each other instead of all ranks sharing every core).
    affinity = None
    if world > 1 and hasattr(os, "sched_getaffinity"):
        try:
            cores = sorted(os.sched_getaffinity(0))
            per = len(cores) // world
            if per >= 1:
                os.sched_setaffinity(0, cores[local * per:(local + 1) * per])
                affinity = "%d cores per rank (%d-%d for rank 0)" % (per, cores[0], cores[per - 1])
        except OSError:
            pass

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
    hbm_peak, tf32_peak, peak_src = load_peaks()

    precision = args.precision
    if precision == "auto":
        precision = "tf32x3"   # what fn2.correlation() runs by default for this call: meets the fp32 tolerance
    B = args.batch

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.config in ("flownetc_train", "flownet2_infer"):
        gb = args.batch if args.batch != 32 or args.config == "flownetc_train" else 8   # cfg4: 64 pairs over 8 GPUs
        line = graphs_bench(torch, dist, dev, rank, world, precision, args.config, args.steps, args.warmup, gb)
        if rank == 0:
            print(json.dumps(line), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    if args.config == "strips4k":
        res = strips_bench(torch, dist, fn2, dev, rank, world, precision, args.steps, args.warmup, hbm_peak)
        if rank == 0:
            line = {"metric": "image_pairs_per_s", "value": res.get("pairs_per_s"), "unit": "pairs/s (4K pairs)",
                    "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                    "ms_per_step": res.get("ms_per_step"), "higher_is_better": True, "scaling": "strong",
                    "vs_baseline": None, "dtype": precision, "data": "synthetic",
                    "config": {"workload": res.get("workload"), "parallelism": "row strips + NCCL halo exchange"},
                    "strips4k": res}
            print(json.dumps(line), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    host = host_data(B, seed=rank)
    a, b, g, image, flow, gimg, gt = [x.to(dev) for x in host]
    ops = ["corr_fwd", "corr_bwd", "warp_fwd", "warp_bwd", "downsample"]

    op_fns = [
        lambda: fn2.correlation(a, b, *CORR_ARGS, precision=precision),
        lambda: fn2.correlation_grad(g, a, b, *CORR_ARGS, precision=precision),
        lambda: fn2.flow_warp(image, flow),
        lambda: fn2.flow_warp_grad(image, flow, gimg),
        lambda: fn2.downsample_multi(gt, DS_SIZES),
    ]

    def step(ev=None):
        res = []
        for i, fn in enumerate(op_fns):
            if ev is not None:
                ev[i].record()
            res.append(fn())
        if ev is not None:
            ev[5].record()
        return res

    warm = max(args.warmup, 3)
    for _ in range(warm):
        step()
    barrier()

    # Each op of the step is captured as its own CUDA graph (the C ABI is capturable: it only enqueues on the given
    # stream); a step = five graph replays with an event between them, so the per-op device times come from INSIDE
    # the timed region and the region is free of Python launch latency.  --eager times plain launches instead.
    side = torch.cuda.Stream(device=dev)
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
    elapsed_ms = evs[0][0].elapsed_time(evs[args.steps - 1][5])
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
    # `e2e_copy_ceiling` is the same loop with the compute left out: what the box's host memory / PCIe path allows.
    e2e = ceiling = None
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

        def e2e_step(k, compute=True):
            sl = k & 1
            with torch.cuda.stream(s_up):
                s_up.wait_event(cp_done[sl])          # step k-2 has finished reading this input set
                for d, h in zip(dev_in[sl], pinned):
                    d.copy_(h, non_blocking=True)
                up_done[sl].record(s_up)
            s_cp.wait_event(up_done[sl])
            s_cp.wait_event(dn_done[sl])              # step k-2's results have left the device
            if compute:
                da, db, dg, dimg, dflow, dgimg, dgt = dev_in[sl]
                res = [fn2.correlation(da, db, *CORR_ARGS, precision=precision)]
                res += list(fn2.correlation_grad(dg, da, db, *CORR_ARGS, precision=precision))
                res.append(fn2.flow_warp(dimg, dflow))
                res += list(fn2.flow_warp_grad(dimg, dflow, dgimg))
                res += fn2.downsample_multi(dgt, DS_SIZES)
                live[sl] = res                        # keep the device results alive until their download ran
            else:
                res = live[sl]
            cp_done[sl].record(s_cp)
            if host_out[sl] is None:
                host_out[sl] = [torch.empty(r.shape, dtype=r.dtype).pin_memory() for r in res]
            with torch.cuda.stream(s_dn):
                s_dn.wait_event(cp_done[sl])
                for h, r in zip(host_out[sl], res):
                    h.copy_(r, non_blocking=True)
                    r.record_stream(s_dn)
                dn_done[sl].record(s_dn)

        def e2e_run(compute):
            barrier()
            s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s2.record(s_cp)
            s_up.wait_event(s2)
            for k in range(e2e_steps):
                e2e_step(k, compute)
            s_cp.wait_event(dn_done[0])
            s_cp.wait_event(dn_done[1])
            e2.record(s_cp)
            barrier()
            ms = s2.elapsed_time(e2)
            if world > 1:
                t = torch.tensor([ms], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            return ms

        for k in range(4):
            e2e_step(k)
        ms = e2e_run(True)
        d2h = sum(h.numel() * 4 for h in host_out[0])
        e2e = {"value": world * B * e2e_steps / (ms / 1e3), "unit": "pairs/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "steps": e2e_steps, "ms_per_step": ms / e2e_steps,
               "how": "public Python API, pinned host inputs and outputs, upload/compute/download streams"}
        ms = e2e_run(False)
        ceiling = {"value": world * B * e2e_steps / (ms / 1e3), "unit": "pairs/s", "ms_per_step": ms / e2e_steps,
                   "gb_s_per_gpu_each_way": [round(h2d / (ms / e2e_steps * 1e-3) / 1e9, 1),
                                             round(d2h / (ms / e2e_steps * 1e-3) / 1e9, 1)],
                   "what": "the same loop with the compute left out: pinned H2D of every input and D2H of every result "
                           "of a step on all %d ranks at once (one cudaMemcpyAsync per tensor); the part of the e2e "
                           "loss that belongs to the box" % world,
                   "cpu_affinity": affinity}
        del live, dev_in

    # ---- config 5 inside the default multi-GPU run, so that the scaling run records a data-path collective ----
    # (N = 1 gives the strong-scaling denominator: the whole 4K pair on one GPU, no exchange)
    strips4k = None
    if not args.no_extras:
        try:
            strips4k = strips_bench(torch, dist, fn2, dev, rank, world, precision, 20, 3, hbm_peak)
        except Exception as exc:  # pragma: no cover
            strips4k = {"unavailable": str(exc)[:300]}

    # ---- configs 3 and 4 (the callers of the path) on every N, after the hot path's own timed region ------------
    cfg3 = cfg4 = None
    if not args.no_extras:
        for name in ("flownetc_train", "flownet2_infer"):
            try:
                res = graphs_bench(torch, dist, dev, rank, world, precision, name, 10, 3, 8)
                res = {k: res[k] for k in ("value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "config",
                                           "finite", "launch")}
            except Exception as exc:  # pragma: no cover
                res = {"unavailable": str(exc)[:300]}
            if name == "flownetc_train":
                cfg3 = res
            else:
                cfg4 = res

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ------------------------------------------------------------------
    abytes, flops = algorithmic_bytes(B), corr_flops(B)
    mma = executed_mma_flops(B, precision) or {}
    dominant = max(per_op_ms, key=per_op_ms.get)
    per_op = {name: op_entry(per_op_ms[name], abytes[name], hbm_peak, flops.get(name), mma.get(name), tf32_peak)
              for name in ops}
    achieved = per_op[dominant]["gb_s"]
    roofline = {"kernel": dominant, "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": round(achieved / hbm_peak, 4), "traffic": None, "peak_source": peak_src,
                "correlation_fwd_bwd_hbm_frac": round(
                    (abytes["corr_fwd"] + abytes["corr_bwd"]) / ((per_op_ms["corr_fwd"] + per_op_ms["corr_bwd"]) * 1e-3)
                    / 1e9 / hbm_peak, 4),
                "tf32_peak_tflops": tf32_peak, "tf32_peak_source": "bf16 peak (%s) / 2" % peak_src}
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

    extras = world == 1 and not args.no_extras
    tstream = torch.cuda.Stream(device=dev)

    # ---- the correlation at every precision mode, same batch, same timing method ------------------------------
    per_precision = None
    if extras:
        per_precision = {}
        for p in ("tf32x3", "tf32", "fp32"):
            reps = 1 if p == "fp32" else 10
            try:
                f_ms = graph_time(torch, [lambda p=p: fn2.correlation(a, b, *CORR_ARGS, precision=p)], reps, tstream)
                b_ms = graph_time(torch, [lambda p=p: fn2.correlation_grad(g, a, b, *CORR_ARGS, precision=p)], reps,
                                  tstream)
            except Exception as exc:  # pragma: no cover
                per_precision[p] = {"unavailable": str(exc)[:200]}
                continue
            m = executed_mma_flops(B, p) or {}
            per_precision[p] = {
                "corr_fwd": op_entry(f_ms, abytes["corr_fwd"], hbm_peak, flops["corr_fwd"], m.get("corr_fwd"), tf32_peak),
                "corr_bwd": op_entry(b_ms, abytes["corr_bwd"], hbm_peak, flops["corr_bwd"], m.get("corr_bwd"), tf32_peak),
                "fwd_bwd_hbm_frac": round((abytes["corr_fwd"] + abytes["corr_bwd"]) / ((f_ms + b_ms) * 1e-3) / 1e9
                                          / hbm_peak, 4),
                "tolerance": {"tf32x3": "1e-5 rel + 1e-6 abs (the op's fp32 tolerance)",
                              "tf32": "2.5e-3 of the tensor's max", "fp32": "1e-5 rel + 1e-6 abs"}[p],
                "what": {"tf32x3": "tcgen05 split-tf32 (the drop-in default)", "tf32": "tcgen05 kind::tf32",
                         "fp32": "CUDA cores, any geometry"}[p]}

    # ---- CPU baseline: the restatement of the reference ops on the host cores, bounded sample ---------------
    cpu = None
    o = None
    if world == 1 and not args.no_cpu_baseline:
        from oracle import oracle as o
        o.lib()
        ncores = use_all_host_threads()
        sample = min(B, REF_SAMPLE)
        data = [x.numpy() for x in host]
        cpu_pass(o, data, 1)
        tc = time.time()
        cpu_pass(o, data, sample)
        dt = time.time() - tc
        cpu = {"value": sample / dt, "unit": "pairs/s", "cores": ncores, "kind": "port",
               "sample": "%d of the %d pairs, one pass, C restatement of the reference kernels (oracle/fn2_oracle.c), "
                         "OpenMP on %d threads, %.1f s" % (sample, B, ncores, dt)}

    # ---- BASELINE config 1: ONE pair's correlation forward, warm and cold L2, beside the CPU port ---------------
    cfg1 = cfg2 = None
    if extras:
        a1, b1 = a[:1].contiguous(), b[:1].contiguous()
        scrub = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
        nb1 = algorithmic_bytes(1)["corr_fwd"]
        cfg1 = {"workload": "correlation forward, one 512x384 pair at conv3 [1,48,64,256] (11.7 MB: fits the L2)"}
        for p in ("tf32x3", "tf32"):
            warm_ms = graph_time(torch, [lambda p=p: fn2.correlation(a1, b1, *CORR_ARGS, precision=p)] * 8, 10, tstream)
            ts = []
            with torch.cuda.stream(tstream):
                for _ in range(12):
                    scrub.zero_()                      # 256 MB > 126 MB of L2: the pair comes from HBM again
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(tstream)
                    fn2.correlation(a1, b1, *CORR_ARGS, precision=p)
                    e1.record(tstream)
                    torch.cuda.synchronize()
                    ts.append(e0.elapsed_time(e1))
            ts.sort()
            cold_ms = ts[len(ts) // 2]
            cfg1[p] = {"warm_l2_ms": round(warm_ms, 4), "cold_l2_ms": round(cold_ms, 4),
                       "warm_pairs_per_s": round(1e3 / warm_ms, 1), "cold_pairs_per_s": round(1e3 / cold_ms, 1),
                       "cold_hbm_frac": round(nb1 / (cold_ms * 1e-3) / 1e9 / hbm_peak, 4)}
        if o is not None:
            an, bn = host[0][:1].numpy(), host[1][:1].numpy()
            o.c_correlation(an, bn, *CORR_ARGS)
            tc = time.time()
            for _ in range(5):
                o.c_correlation(an, bn, *CORR_ARGS)
            cfg1["cpu_port_ms"] = round((time.time() - tc) / 5 * 1e3, 2)
        del scrub
        # ---- BASELINE config 2: flow_warp + downsample on batch 8, rotating inputs so nothing is found in L2 ----
        ROT = 6
        gen = torch.Generator(device=dev).manual_seed(5)
        IMG = [torch.rand(8, HI, WI, 3, device=dev, generator=gen) for _ in range(ROT)]
        FLO = [torch.randn(8, HI, WI, 2, device=dev, generator=gen) * 8 for _ in range(ROT)]
        GI = [torch.randn(8, HI, WI, 3, device=dev, generator=gen) for _ in range(ROT)]
        ab8 = algorithmic_bytes(8)
        cfg2 = {"workload": "flow_warp fwd/bwd + downsample on batch-8 512x384 image/flow pairs (%d rotating input "
                            "sets, 300 MB > L2)" % ROT}
        for name, mk in (("warp_fwd", lambda i: (lambda: fn2.flow_warp(IMG[i], FLO[i]))),
                         ("warp_bwd", lambda i: (lambda: fn2.flow_warp_grad(IMG[i], FLO[i], GI[i]))),
                         ("downsample", lambda i: (lambda: fn2.downsample_multi(FLO[i], DS_SIZES)))):
            ms = graph_time(torch, [mk(i) for i in range(ROT)], 5, tstream)
            cfg2[name] = op_entry(ms, ab8[name], hbm_peak)
        cfg2["pairs_per_s"] = round(8e3 / sum(cfg2[n]["ms"] for n in ("warp_fwd", "warp_bwd", "downsample")), 1)
        del IMG, FLO, GI
        # ---- the same two flow_warp kernels on the reference's REAL flow (data/samples/0flow.flo as packed in
        # tests/golden/sample0.npz): smooth, mean 2 px, down to -109 px -- the other regime from the iid N(0, 8 px) ----
        fix = os.path.join(ROOT, "tests", "golden", "sample0.npz")
        if os.path.exists(fix):
            import numpy as np
            z = np.load(fix)
            img1 = torch.from_numpy(z["image"].astype(np.float32)).to(dev)
            flo1 = torch.from_numpy(z["flow"]).to(dev)
            IMG = [torch.stack([img1.roll(7 * (i * B + j), 0) for j in range(B)]).contiguous() for i in range(2)]
            FLO = [torch.stack([flo1.roll(7 * (i * B + j), 0) for j in range(B)]).contiguous() for i in range(2)]
            GI = [torch.randn(B, HI, WI, 3, device=dev, generator=gen) for _ in range(2)]
            cfg2["on_the_reference_sample_flow"] = {
                "workload": "batch %d of data/samples pair 0 (rows rolled per image), real ground-truth flow" % B,
                "warp_fwd": op_entry(graph_time(torch, [(lambda i=i: fn2.flow_warp(IMG[i], FLO[i])) for i in range(2)],
                                                5, tstream), abytes["warp_fwd"], hbm_peak),
                "warp_bwd": op_entry(graph_time(torch, [(lambda i=i: fn2.flow_warp_grad(IMG[i], FLO[i], GI[i]))
                                                        for i in range(2)], 5, tstream), abytes["warp_bwd"], hbm_peak)}
            del IMG, FLO, GI

    # ---- the reference's own CUDA kernels on this GPU, per op (they are the reference's only implementation) ----
    ref_gpu = None
    if world == 1 and not args.no_ref_gpu:
        try:
            from oracle import ref
            if ref.available():
                temps = (torch.empty((B, HC + 40, WC + 40, CC), device=dev),
                         torch.empty((B, HC + 40, WC + 40, CC), device=dev))
                ref_ops = [
                    ("corr_fwd", lambda: ref.correlation(a, b, *CORR_ARGS, temps=temps)),
                    ("corr_bwd", lambda: ref.correlation_grad(g, a, b, *CORR_ARGS, temps=temps)),
                    ("warp_fwd", lambda: ref.flow_warp(image, flow)),
                    ("warp_bwd", lambda: ref.flow_warp_grad(image, flow, gimg)),
                    ("downsample", lambda: [ref.downsample(gt, s) for s in DS_SIZES]),
                ]
                reps = 5
                times = {n: [] for n, _ in ref_ops}
                for r in range(reps + 1):              # first round = warm-up
                    for n, fn in ref_ops:
                        torch.cuda.synchronize()
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()                    # the reference launches on the legacy default stream,
                        fn()                           # which is torch's current stream here
                        e1.record()
                        torch.cuda.synchronize()
                        if r > 0:
                            times[n].append(e0.elapsed_time(e1))
                med = {n: sorted(v)[len(v) // 2] for n, v in times.items()}
                total = sum(med.values())
                ref_gpu = {"value": B / (total * 1e-3), "unit": "pairs/s", "ms_per_step": round(total, 3),
                           "per_op_ms": {n: round(v, 3) for n, v in med.items()},
                           "speedup_of_this_library": {n: round(med[n] / per_op_ms[n], 1) for n in med},
                           "speedup_of_this_library_step": round(total / sum(per_op_ms.values()), 1),
                           "reps": reps,
                           "what": "reference src/ops/**.cu.cc kernels compiled unmodified for sm_100a (oracle/_ref), "
                                   "same batch, CUDA events, median of %d; flow_warp backward through its launcher "
                                   "(the reference's OpKernel cannot run it: flow_warp_grad.cc:36-37)" % reps}
        except Exception as exc:  # pragma: no cover
            ref_gpu = {"unavailable": str(exc)[:200]}

    value = world * B * args.steps / (elapsed_ms / 1e3)
    cfg = workload_config(B)
    line = {
        "metric": "image_pairs_per_s", "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": {"fp32": "f32", "tf32": "tf32", "tf32x3": "tf32x3"}[precision],
        "data": "synthetic", "config": cfg, "correlation_precision": precision,
        "launch": "CUDA graph replay per op, 5 per step" if graphs else "eager launches",
        "clocks": clocks, "e2e": e2e, "e2e_copy_ceiling": ceiling,
        "gpu_launches": int(launches), "roofline": roofline, "per_op": per_op, "per_precision": per_precision,
        "cpu_baseline": cpu, "cfg1": cfg1, "cfg2": cfg2, "cfg3": cfg3, "cfg4": cfg4,
        "strips4k": strips4k, "reference_gpu_kernels": ref_gpu,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
