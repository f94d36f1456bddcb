#!/usr/bin/env python
"""bench.py -- headline benchmark of the roi_align / nms hot path (see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]

One "step" = one MultiScaleRoIAlign forward + backward over one batch (BASELINE.json configs[2]:
4 FPN levels of an 800x1088 image, batch 16 per GPU, 512 RoIs/img, 14x14 output, sampling_ratio 2,
fp32, channels_last features) through the C-ABI library.  `value` = RoIs/s (fwd+bwd), whole job.
Every rank works on its own 16 images (shard by image, no data-path collective) -> "weak" scaling.
Inputs (1.18 GB features + 1.64 GB upstream gradient per rank) are far larger than the 126 MB L2, so
no L2 flush is needed between iterations.

Extra objects on the JSON line: roofline (dominant kernel vs the measured HBM peak), e2e (same step
through the *_host C-ABI with pinned host buffers, copies inside the timed region), cpu_baseline
(the reference's own CPU kernels, oracle/_ref, on a bounded sample), clocks, gpu_launches, and
`secondary` (cfg1 roi_align, cfg3 bf16, cfg2 batched_nms boxes/s).

--impl reference times the reference's CPU implementation (oracle/_ref/libtvref.so, all host
threads) on a bounded sample of the same workload and prints the same line with impl=reference.
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

SHAPES = [(200, 272), (100, 136), (50, 68), (25, 34)]
SCALES = [0.25, 0.125, 0.0625, 0.03125]
IMAGE = (800, 1088)
C = 256


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)  # ~0.3 s timed region at 1.6 ms per step
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--workload", default="cfg3_fp32", choices=["cfg3_fp32", "cfg3_bf16"])
    p.add_argument("--no-extras", action="store_true", help="skip e2e / cpu_baseline / secondary legs")
    p.add_argument("--gather", action="store_true", help="(kept for compatibility: the cfg4 exchange now always runs at N > 1)")
    return p.parse_args()


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# --------------------------------------------------------------------------------------- workload
def make_rois(gen, batch, per_img, lo, hi, device):
    """per_img boxes per image, sqrt(area) log-uniform in [lo, hi], clipped to the image, grouped by image."""
    k = batch * per_img
    h, w = IMAGE
    u = lambda: torch.rand(k, generator=gen, device=device)
    side = torch.exp(u() * (torch.log(torch.tensor(hi / lo))) + torch.log(torch.tensor(lo)))
    aspect = torch.exp((u() - 0.5) * 1.0)
    bw, bh = (side * aspect).clamp(max=w), (side / aspect).clamp(max=h)
    cx, cy = u() * w, u() * h
    x1, y1 = (cx - bw / 2).clamp(0, w - 2), (cy - bh / 2).clamp(0, h - 2)
    x2, y2 = torch.maximum((cx + bw / 2).clamp(max=w), x1 + 1), torch.maximum((cy + bh / 2).clamp(max=h), y1 + 1)
    b = torch.arange(batch, device=device).repeat_interleave(per_img).float()
    return torch.stack([b, x1, y1, x2, y2], 1).contiguous()


def make_cfg3(device, dtype, seed, batch=16, per_img=512, ph=14):
    gen = torch.Generator(device=device).manual_seed(seed)
    feats = [torch.randn(batch, C, h, w, generator=gen, device=device, dtype=torch.float32).to(dtype)
             .contiguous(memory_format=torch.channels_last) for h, w in SHAPES]
    rois = make_rois(gen, batch, per_img, 16.0, 800.0, device)
    grad = torch.randn(batch * per_img, C, ph, ph, generator=gen, device=device, dtype=torch.float32).to(dtype) \
        .contiguous(memory_format=torch.channels_last)
    return feats, rois, grad


def bind_to_gpu_numa_node(index):
    """Pins this process to the CPUs next to its GPU (NVML's ideal affinity) before any host buffer is
    allocated, so the pinned staging memory of the e2e leg is first-touched on the GPU's NUMA node.  With 8
    ranks on one box the host<->device copies otherwise cross the socket interconnect."""
    if os.environ.get("B200DET_BENCH_NO_AFFINITY"):
        return
    try:
        import pynvml as nv

        nv.nvmlInit()
        nv.nvmlDeviceSetCpuAffinity(nv.nvmlDeviceGetHandleByIndex(index))
    except Exception:
        pass


class ClockSampler:
    """Samples SM clock / throttle reasons with NVML every 10 ms while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.ok, self._stop = [], set(), False, threading.Event()
        try:
            import pynvml as nv

            nv.nvmlInit()
            self.nv, self.h = nv, nv.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # pragma: no cover
            self.err = repr(e)
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.01)

    def start(self):
        if self.ok:
            self.t.start()

    def stop(self):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"]}
        self._stop.set()
        self.t.join()
        s = self.samples or [0]
        return {"sm_mhz": statistics.median(s), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def physical_gpu_index(local):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local])
        except Exception:
            return local
    return local


# --------------------------------------------------------------------------------------- reference arm
def cpu_reference_sample(feats_cpu, rois_cpu, grad_cpu, ph, stride, threads, repeats=1):
    """The reference's CPU roi_align (oracle/_ref, unmodified kernels) on every `stride`-th RoI of the
    workload, forward + backward, level by level like its Python MultiScaleRoIAlign does
    (torchvision/ops/poolers.py:200-222).  Returns (seconds per pass, RoIs per pass, kind)."""
    import oracle

    kind = "reference" if oracle.have_ref() else "port"
    impl = "ref" if kind == "reference" else "oracle"
    sel = torch.arange(0, rois_cpu.shape[0], stride)
    r = rois_cpu[sel].contiguous()
    # grad_cpu is either the full [K,...] gradient or already the selected rows
    g = (grad_cpu if grad_cpu.shape[0] == sel.numel() else grad_cpu[sel]).contiguous()
    levels = oracle.fpn_level_map(r, 2, 5)
    per_level = []
    for l in range(4):
        idx = torch.nonzero(levels == l).flatten()
        per_level.append((r[idx].contiguous(), g[idx].contiguous()))
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        for l, (rl, gl) in enumerate(per_level):
            n, c, h, w = feats_cpu[l].shape
            oracle.roi_align_forward(feats_cpu[l], rl, ph, ph, SCALES[l], 2, False, impl=impl, threads=threads)
            oracle.roi_align_backward(gl, rl, SCALES[l], ph, ph, n, c, h, w, 2, False, impl=impl, threads=threads)
        best = min(best, time.perf_counter() - t0)
    return best, int(sel.numel()), kind


def run_reference(args, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    dtype = torch.float32
    gen = torch.Generator().manual_seed(1234 + 3)
    batch, per_img, ph = 16, 512, 14
    feats = [torch.randn(batch, C, h, w, generator=gen) for h, w in SHAPES]
    rois = make_rois(gen, batch, per_img, 16.0, 800.0, "cpu")
    # the whole batch per step (a sample would spread the reference's per-call cost -- allocating and zero-filling
    # the 1.18 GB gradient like its at::zeros -- over fewer RoIs and under-count the CPU); ~3 s of CPU work per step
    stride = 1
    torch.set_num_threads(threads)  # torchrun exports OMP_NUM_THREADS=1: the zero fills would run on one core
    try:
        import oracle

        oracle.lib("ref").tvref_set_num_threads(threads)
    except Exception:
        pass
    k_s = len(range(0, rois.shape[0], stride))
    gsel = torch.randn(k_s, C, ph, ph, generator=gen)  # upstream gradient of the sampled RoIs only

    times = []
    for i in range(args.warmup + args.steps):
        sec, n, kind = cpu_reference_sample(feats, rois, gsel, ph, stride, threads)
        if i >= args.warmup:
            times.append(sec)
        if sum(times) > 240:  # keep the whole run within a few minutes
            break
    ms = 1e3 * sum(times) / len(times)
    val = n / (ms / 1e3)
    sample = (f"all {n} RoIs of the cfg3 batch per step, fwd+bwd, 4 levels, {threads} host threads (forward sharded "
              "over RoIs, backward over images: <= 16 threads)")
    line = {
        "impl": "reference", "metric": "roi_align_fwd_bwd_rois_per_s", "value": val, "unit": "RoIs/s",
        "n_gpus": args.gpus, "steps": len(times), "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config("cfg3_fp32"),
        "cpu_baseline": {"value": val, "unit": "RoIs/s", "cores": threads, "kind": kind, "sample": sample,
                         "nms": cpu_nms_leg()},
        "e2e": {"value": val, "unit": "RoIs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def port_evidence(key):
    """The SM -> L2 request port of the backward in the committed ncu capture of `key` (profiles/traffic.json): the unit
    that bounds the scatter; None when there is no capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get(key, {})
        return {"bwd": t.get("bwd_sm_l2_port"), "dram_bytes_bwd": t.get("bwd"), "source": t.get("source")}
    except Exception:
        return None


def workload_config(name):
    return {
        "workload": f"{name}: MultiScaleRoIAlign fwd+bwd, 4 FPN levels of 800x1088 (200x272..25x34), C=256, "
                    "batch 16 per GPU, 512 RoIs/img, 14x14 output, sampling_ratio 2, aligned False "
                    "(BASELINE.json configs[2])",
        "layout": "channels_last features / outputs / gradients",
        "rois_per_step_per_gpu": 8192,
        "sharding": "by image, 16 images per rank, no data-path collective",
        "l2": "inputs (1.18 GB features + 1.64 GB grad per rank in fp32) exceed the 126 MB L2; no flush needed",
    }


# --------------------------------------------------------------------------------------- b200 arm
def time_events(fn, steps, warmup, sync):
    """Device time of `steps` calls of fn() in ms (CUDA events on the current stream)."""
    for _ in range(warmup):
        fn()
    sync()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    b.synchronize()
    return a.elapsed_time(b)


def run_b200(args, rank, world, local):
    import torch.distributed as dist

    from cpu_vision_b200 import capi

    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    bind_to_gpu_numa_node(local)
    dtype = torch.bfloat16 if args.workload == "cfg3_bf16" else torch.float32
    es = 2 if dtype == torch.bfloat16 else 4
    batch, per_img, ph = 16, 512, 14
    feats, rois, grad = make_cfg3(dev, dtype, 1234 + 3 + rank, batch, per_img, ph)
    K = rois.shape[0]
    shapes = [(batch, C, h, w) for h, w in SHAPES]
    out = torch.empty(K, C, ph, ph, device=dev, dtype=dtype).contiguous(memory_format=torch.channels_last)
    levels = torch.empty(K, dtype=torch.int32, device=dev)
    gfeats = [torch.empty_like(f) for f in feats]

    def fwd():
        capi.msroi_align_forward(feats, rois, SCALES, ph, ph, 2, False, out=out)

    lv_holder = {}

    def fwd_keep_levels():
        _, lv = capi.msroi_align_forward(feats, rois, SCALES, ph, ph, 2, False, out=out)
        lv_holder["lv"] = lv

    fwd_keep_levels()
    levels = lv_holder["lv"]

    def bwd():
        capi.msroi_align_backward(grad, rois, levels, SCALES, shapes, ph, ph, 2, False, layout=capi.NHWC, outs=gfeats)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        fwd()
        bwd()
    barrier()
    sampler = ClockSampler(physical_gpu_index(local))
    sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * args.steps + 1)]
    ev[0].record()
    for i in range(args.steps):
        fwd()
        ev[2 * i + 1].record()
        bwd()
        ev[2 * i + 2].record()
    torch.cuda.synchronize()
    barrier()
    clocks = sampler.stop()
    total_ms = ev[0].elapsed_time(ev[-1])
    fwd_ms = sum(ev[2 * i].elapsed_time(ev[2 * i + 1]) for i in range(args.steps)) / args.steps
    bwd_ms = sum(ev[2 * i + 1].elapsed_time(ev[2 * i + 2]) for i in range(args.steps)) / args.steps
    if world > 1:
        t = torch.tensor([total_ms, fwd_ms, bwd_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, fwd_ms, bwd_ms = t.tolist()
    ms_per_step = total_ms / args.steps
    value = world * K / (ms_per_step / 1e3)

    feat_bytes = sum(n * c * h * w for n, c, h, w in shapes) * es
    out_bytes = K * C * ph * ph * es
    algo = {"fwd": feat_bytes + out_bytes + K * 20, "bwd": out_bytes + feat_bytes + K * 20}
    peak, peak_src = measured_peak()
    # DRAM bytes of the dominant kernel from the committed ncu --set full capture (profiles/traffic.json, written by
    # tools/ncu_summary.py --traffic from the same .ncu-rep as the text summary); null when no capture matches
    traffic = {}
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(args.workload, {})
    except Exception:
        pass
    dom = "bwd" if bwd_ms >= fwd_ms else "fwd"
    dom_ms = bwd_ms if dom == "bwd" else fwd_ms
    achieved = algo[dom] / (dom_ms / 1e3) / 1e9
    roofline = {
        "bound": "hbm", "kernel": f"msroi_align_{dom}", "achieved": achieved, "peak": peak, "unit": "GB/s",
        "frac": achieved / peak, "traffic": traffic.get(dom), "traffic_source": traffic.get("source"),
        # the busiest unit of the dominant kernel in the same ncu capture: the SM -> L2 request port
        "sm_l2_port": traffic.get(dom + "_sm_l2_port"),
        "peak_source": peak_src, "algorithmic_bytes_per_launch": algo[dom],
        "fwd": {"ms": fwd_ms, "GBps": algo["fwd"] / (fwd_ms / 1e3) / 1e9, "frac": algo["fwd"] / (fwd_ms / 1e3) / 1e9 / peak},
        "bwd": {"ms": bwd_ms, "GBps": algo["bwd"] / (bwd_ms / 1e3) / 1e9, "frac": algo["bwd"] / (bwd_ms / 1e3) / 1e9 / peak},
        "fwd_bwd_frac": (algo["fwd"] + algo["bwd"]) / (ms_per_step / 1e3) / 1e9 / peak,
    }
    line = {
        "metric": "roi_align_fwd_bwd_rois_per_s", "value": value, "unit": "RoIs/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16" if dtype == torch.bfloat16 else "f32",
        "data": "synthetic", "config": workload_config(args.workload), "roofline": roofline, "clocks": clocks,
        # kernels of ours per step: fpn_level_map + msroi forward + msroi backward (memsets not counted)
        "gpu_launches": 3 * args.steps,
    }
    if not args.no_extras:
        line["e2e"] = e2e_leg(capi, feats, rois, grad, levels, shapes, ph, dtype, local, world, K, dist)
        if rank == 0 and world == 1:
            line["cpu_baseline"] = cpu_leg(feats, rois, grad, ph)
        del feats, grad, gfeats, out
        torch.cuda.empty_cache()
        cfg4 = cfg4_leg(capi, dev, rank, world, dist, peak)
        if world > 1:
            line["cfg4"] = cfg4["cfg4"]
            line["allgather"] = cfg4["allgather"]
        torch.cuda.empty_cache()
        line["cfg5"] = cfg5_leg(capi, dev, rank, world, dist)
        if rank == 0 and world == 1:
            line["secondary"] = secondary_legs(capi, dev, peak)
            line["secondary"]["cfg4_64img_x_2000"] = cfg4["cfg4"]
            line["secondary"]["stock_cuda_kernels"] = subprocess_leg("stock_cuda_bench.py", "tvref_cuda.so")
            line["secondary"]["cfg5_model"] = subprocess_leg("cfg5_model_bench.py", os.path.join("torchvision", "_C.so"))
    if rank == 0:
        print(json.dumps(line), flush=True)


def e2e_leg(capi, feats, rois, grad, levels, shapes, ph, dtype, local, world, K, dist):
    """Same step through the host-buffer C-ABI: pinned host inputs, H2D + kernels + D2H all timed."""
    hf = [f.cpu().contiguous(memory_format=torch.channels_last).pin_memory() for f in feats]
    hr = rois.cpu().pin_memory()
    hg = grad.cpu().contiguous(memory_format=torch.channels_last).pin_memory()
    hout = torch.empty_like(hg).pin_memory()
    hlv = torch.empty(K, dtype=torch.int32).pin_memory()
    hgf = [torch.empty_like(f).pin_memory() for f in hf]
    hf = [f.contiguous(memory_format=torch.channels_last) for f in hf]
    steps, warm = 5, 2

    def step():
        capi.msroi_align_forward_host(hf, hr, SCALES, ph, ph, 2, False, device=local, out=hout, levels=hlv)
        capi.msroi_align_backward_host(hg, hr, hlv, SCALES, shapes, ph, ph, 2, False, layout=capi.NHWC,
                                       device=local, outs=hgf)

    for _ in range(warm):
        step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    sec = (time.perf_counter() - t0) / steps  # the host calls return only after the D2H copies completed
    if world > 1:
        t = torch.tensor([sec], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sec = t.item()
    capi.host_release()
    nb = lambda ts: sum(t.numel() * t.element_size() for t in ts)
    return {"value": world * K / sec, "unit": "RoIs/s", "ms_per_step": sec * 1e3,
            "h2d_bytes_per_step": nb(hf) + 2 * nb([hr]) + nb([hg]) + nb([hlv]),
            "d2h_bytes_per_step": nb([hout]) + nb([hlv]) + nb(hgf), "steps": steps,
            "api": "b200det_msroi_align_forward_host + b200det_msroi_align_backward_host (pinned host buffers)"}


def cpu_nms_leg():
    """cfg2 on the host: the reference's CPU nms kernel (cpu/nms_kernel.cpp:9-77, one thread by construction)
    through its batched_nms strategy rule for a CPU tensor (boxes.py:72: numel > 4000 -> one nms per class,
    boxes.py:98-112), and plain nms of the same 5000 boxes."""
    import oracle

    impl = "ref" if oracle.have_ref() else "oracle"
    gen = torch.Generator().manual_seed(1234 + 2)
    n = 5000
    xy = torch.rand(n, 2, generator=gen) * 1000
    boxes = torch.cat([xy, xy + torch.rand(n, 2, generator=gen) * 200 + 10], 1)
    scores = torch.randperm(n, generator=gen).float() / n
    idxs = torch.randint(0, 80, (n,), generator=gen)

    def vanilla():
        keep = torch.zeros(n, dtype=torch.bool)
        for c in torch.unique(idxs):
            sel = torch.where(idxs == c)[0]
            keep[sel[oracle.nms(boxes[sel], scores[sel], 0.5, impl=impl)]] = True
        k = torch.where(keep)[0]
        return k[scores[k].sort(descending=True)[1]]

    def best(fn, reps):
        b = float("inf")
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            b = min(b, time.perf_counter() - t0)
        return b

    tb, tn = best(vanilla, 5), best(lambda: oracle.nms(boxes, scores, 0.5, impl=impl), 5)
    return {"batched_nms_5000x80_ms": tb * 1e3, "batched_nms_boxes_per_s": n / tb, "nms_5000_ms": tn * 1e3,
            "nms_boxes_per_s": n / tn, "cores": 1, "kind": "reference" if impl == "ref" else "port",
            "sample": "cfg2 whole (5000 boxes x 80 classes, IoU 0.5), best of 5"}


def cpu_leg(feats, rois, grad, ph):
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    fc = [f.float().cpu().contiguous() for f in feats]
    sec, n, kind = cpu_reference_sample(fc, rois.cpu(), grad.float().cpu().contiguous(), ph, 1, threads, repeats=2)
    return {"value": n / sec, "unit": "RoIs/s", "cores": threads, "kind": kind, "seconds": sec,
            "sample": f"all {n} RoIs of the same batch, fwd+bwd, best of 2; forward sharded over RoIs, backward "
                      "over images (<=16 threads)",
            "nms": cpu_nms_leg()}


def secondary_legs(capi, dev, peak):
    """Other BASELINE.json configs, device-timed with an L2 flush between iterations where the
    working set is smaller than L2."""
    res = {}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    sync = torch.cuda.synchronize

    def timed(fn, iters=20, flush_l2=True):
        for _ in range(3):
            fn()
        sync()
        tot = 0.0
        for _ in range(iters):
            if flush_l2:
                flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            tot += a.elapsed_time(b)
        return tot / iters

    # cfg1: single level 2x256x200x272, 1000 RoIs/img, 7x7, sr=2, aligned
    gen = torch.Generator(device=dev).manual_seed(1234 + 1)
    x = torch.randn(2, C, 200, 272, generator=gen, device=dev).contiguous(memory_format=torch.channels_last)
    r = make_rois(gen, 2, 1000, 16.0, 512.0, dev)
    g = torch.randn(2000, C, 7, 7, generator=gen, device=dev).contiguous(memory_format=torch.channels_last)
    gi = torch.empty_like(x)
    f_ms = timed(lambda: capi.roi_align_forward(x, r, 7, 7, 0.25, 2, True))
    b_ms = timed(lambda: capi.roi_align_backward(g, r, 0.25, 7, 7, 2, C, 200, 272, 2, True, layout=capi.NHWC, out=gi))
    fb = x.numel() * 4 + g.numel() * 4 + 2000 * 20
    res["cfg1_roi_align_fp32"] = {
        "fwd_ms": f_ms, "bwd_ms": b_ms, "fwd_rois_per_s": 2000 / (f_ms / 1e3),
        "fwd_bwd_rois_per_s": 2000 / ((f_ms + b_ms) / 1e3), "fwd_frac_of_hbm": fb / (f_ms / 1e3) / 1e9 / peak,
        "bwd_frac_of_hbm": fb / (b_ms / 1e3) / 1e9 / peak, "l2": "flushed between iterations",
        "sm_l2_port": port_evidence("cfg1")}
    del x, g, gi
    # cfg2: batched_nms 5000 boxes x 80 classes, IoU 0.5 (coordinate-trick branch on CUDA, boxes.py:72)
    n = 5000
    xy = torch.rand(n, 2, generator=gen, device=dev) * 1000
    boxes = torch.cat([xy, xy + torch.rand(n, 2, generator=gen, device=dev) * 200 + 10], 1)
    scores = (torch.randperm(n, generator=gen, device=dev).float() / n)
    idxs = torch.randint(0, 80, (n,), generator=gen, device=dev)
    ws = capi.NmsWorkspace(n, dev)
    ms = timed(lambda: capi.batched_nms(boxes, scores, idxs, 0.5, capi.BNMS_AUTO, workspace=ws, sync=False),
               flush_l2=False)
    ms_sync = timed(lambda: capi.batched_nms(boxes, scores, idxs, 0.5, capi.BNMS_AUTO, workspace=ws, sync=True),
                    flush_l2=False)
    kept = capi.batched_nms(boxes, scores, idxs, 0.5, capi.BNMS_AUTO, workspace=ws).numel()
    res["cfg2_batched_nms_5000x80"] = {
        "ms": ms, "boxes_per_s": n / (ms / 1e3), "iou_pairs_per_s": n * (n - 1) / 2 / (ms / 1e3), "kept": kept,
        "ms_with_count_readback": ms_sync,
        "frac_of_bandwidth_bound": ((n * 20 + kept * 8) / (peak * 1e9)) / (ms / 1e3)}
    ms = timed(lambda: capi.nms(boxes, scores, 0.5, workspace=ws, sync=False), flush_l2=False)
    res["nms_5000"] = {"ms": ms, "boxes_per_s": n / (ms / 1e3)}
    # cfg5 (RPN side): filter_proposals for 64 images of 800x1088, 5 FPN levels x 3 anchors,
    # rpn_pre_nms_top_n = rpn_post_nms_top_n = 2000 (8663 candidates per image): the batched path
    # (clip/filter kernel + ONE segmented NMS) next to the reference's shape of the same work -- a Python
    # loop of per-image batched_nms calls -- both on our kernels, both ending with the boxes on the device
    from cpu_vision_b200 import detection

    B = 64
    napl = [200 * 272 * 3, 100 * 136 * 3, 50 * 68 * 3, 25 * 34 * 3, 13 * 17 * 3]
    A = sum(napl)
    ctr = torch.rand(B, A, 2, generator=gen, device=dev) * torch.tensor([1088.0, 800.0], device=dev)
    wh = torch.exp(torch.rand(B, A, 2, generator=gen, device=dev) * 3.5 + 2.5)
    props = torch.cat([ctr - wh / 2, ctr + wh / 2], 2)
    obj = torch.randn(B * A, 1, generator=gen, device=dev)
    shapes = [(800, 1088)] * B
    f1_ms = timed(lambda: detection.filter_proposals(props, obj, shapes, napl, 2000, 2000, 0.7), iters=5,
                  flush_l2=False)
    o2 = obj.reshape(B, -1)
    idx = detection._get_top_n_idx(o2, napl, 2000)
    bi = torch.arange(B, device=dev)[:, None]
    pb, ps = props[bi, idx], torch.sigmoid(o2[bi, idx])
    lv = torch.cat([torch.full((n_,), i_, dtype=torch.int64, device=dev) for i_, n_ in enumerate(napl)])[idx]
    per = pb.shape[1]
    off = [i * per for i in range(B + 1)]
    pbf, psf, lvf = pb.reshape(-1, 4).contiguous(), ps.reshape(-1).contiguous(), lv.reshape(-1).contiguous()
    seg_ms = timed(lambda: capi.nms_segmented(pbf, psf, lvf, off, 0.7, top_n=2000, sync=False), iters=10,
                   flush_l2=False)

    def loop():
        return [capi.batched_nms(pb[i], ps[i], lv[i], 0.7)[:2000] for i in range(B)]

    loop_ms = timed(loop, iters=3, flush_l2=False)
    del props, obj, ctr, wh, pb, ps, lv, pbf, psf, lvf
    # cfg5 (RoIHeads side): postprocess_detections for 64 images x 1000 proposals x 91 classes (roi_heads defaults:
    # score 0.05, NMS 0.5, 100 detections per image): softmax / box decoding (library element-wise ops), clip +
    # filters, compaction of the survivors, ONE segmented class-aware NMS, top-100
    P_, C_ = 1000, 91
    rp = []
    for _ in range(B):
        c_ = torch.rand(P_, 2, generator=gen, device=dev) * torch.tensor([1088.0, 800.0], device=dev)
        w_ = torch.exp(torch.rand(P_, 2, generator=gen, device=dev) * 3.0 + 2.5)
        rp.append(torch.cat([c_ - w_ / 2, c_ + w_ / 2], 1))
    lg = torch.randn(B * P_, C_, generator=gen, device=dev) * 2.5
    rg = torch.randn(B * P_, C_ * 4, generator=gen, device=dev) * 0.5
    f1b_ms = timed(lambda: detection.postprocess_detections(lg, rg, rp, shapes, 0.05, 0.5, 100), iters=5,
                   flush_l2=False)
    det_n = sum(b_.shape[0] for b_ in detection.postprocess_detections(lg, rg, rp, shapes, 0.05, 0.5, 100)[0])
    del rp, lg, rg
    # cfg3 in bf16 (BASELINE.json configs[2] names fp32 and bf16): same kernels, 2-byte elements, packed
    # red.global.add.v4.bf16x2 in the backward; the working set (1.41 GB) exceeds L2, no flush
    fb_, rb_, gb_ = make_cfg3(dev, torch.bfloat16, 1239)
    Kb = rb_.shape[0]
    ob_ = torch.empty(Kb, C, 14, 14, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    gfb_ = [torch.empty_like(f) for f in fb_]
    _, lvb_ = capi.msroi_align_forward(fb_, rb_, SCALES, 14, 14, 2, False, out=ob_)
    shp_ = [(16, C, h, w) for h, w in SHAPES]
    bf_f = timed(lambda: capi.msroi_align_forward(fb_, rb_, SCALES, 14, 14, 2, False, out=ob_), iters=10, flush_l2=False)
    bf_b = timed(lambda: capi.msroi_align_backward(gb_, rb_, lvb_, SCALES, shp_, 14, 14, 2, False, layout=capi.NHWC,
                                                   outs=gfb_), iters=10, flush_l2=False)
    bytes_b = sum(f.numel() for f in fb_) * 2 + gb_.numel() * 2 + Kb * 24
    res["cfg3_bf16"] = {
        "fwd_ms": bf_f, "bwd_ms": bf_b, "fwd_bwd_rois_per_s": Kb / ((bf_f + bf_b) / 1e3),
        "algorithmic_bytes_per_launch": bytes_b, "fwd_frac_of_hbm": bytes_b / (bf_f / 1e3) / 1e9 / peak,
        "bwd_frac_of_hbm": bytes_b / (bf_b / 1e3) / 1e9 / peak,
        "note": "parity within 1e-2 of the fp32 oracle (tests/test_roi_align_gpu.py); also selectable as --workload cfg3_bf16"}
    del fb_, rb_, gb_, ob_, gfb_
    # cfg3 as an NCHW caller sees it (the reference's models hand NCHW-contiguous features to roi_align):
    # the dispatcher shim re-lays out -> channels_last kernels -> re-lays out; results are NCHW-contiguous
    import cpu_vision_b200  # noqa: F401

    feats, rois, grad = make_cfg3(dev, torch.float32, 1237)
    # cfg3 with the op's DEFAULT sampling_ratio (<= 0: adaptive grids, what torchvision.ops.roi_align(x, boxes, 14)
    # and the reference's test_ops.py use): roi_align_anygrid_sm100.cu against the generic one-thread-per-element
    # kernels (the reference's parallelisation, B200DET_ANYGRID=0) on the same channels_last tensors
    lvl = capi.fpn_level_map(rois, 2, 5)
    oa_ = torch.empty(rois.shape[0], C, 14, 14, device=dev).contiguous(memory_format=torch.channels_last)
    ga_ = [torch.empty_like(f) for f in feats]
    shp3 = [(16, C, h, w) for h, w in SHAPES]
    ad = {}
    for sw in ("1", "0"):
        os.environ["B200DET_ANYGRID"] = sw
        ad[sw] = (timed(lambda: capi.msroi_align_forward(feats, rois, SCALES, 14, 14, 0, False, levels=lvl, out=oa_),
                        iters=5, flush_l2=False),
                  timed(lambda: capi.msroi_align_backward(grad, rois, lvl, SCALES, shp3, 14, 14, 0, False,
                                                          layout=capi.NHWC, outs=ga_), iters=5, flush_l2=False))
    os.environ.pop("B200DET_ANYGRID", None)
    res["cfg3_fp32_adaptive_sampling"] = {
        "fwd_ms": ad["1"][0], "bwd_ms": ad["1"][1], "fwd_bwd_rois_per_s": rois.shape[0] / (sum(ad["1"]) / 1e3),
        "generic_kernels_fwd_ms": ad["0"][0], "generic_kernels_bwd_ms": ad["0"][1],
        "note": "sampling_ratio 0 (grid = ceil(roi / bins): 1-2 samples per bin and axis on these levels); bwd includes "
                "the zero fill of the 1.18 GB gradient"}
    del oa_, ga_
    fn, gn = [f.contiguous() for f in feats], grad.contiguous()
    del feats, grad
    hs_, ws_ = [h for h, _ in SHAPES], [w for _, w in SHAPES]
    nf = timed(lambda: torch.ops.b200det.msroi_align(fn, rois, SCALES, 14, 14, 2, False, 2, 5, 224, 4, 1e-6), iters=5,
               flush_l2=False)
    nb = timed(lambda: torch.ops.b200det.msroi_align_backward(gn, rois, lvl, SCALES, hs_, ws_, 16, C, 14, 14, 2, False,
                                                              False), iters=5, flush_l2=False)
    res["cfg3_fp32_nchw_caller"] = {
        "fwd_ms": nf, "bwd_ms": nb, "fwd_bwd_rois_per_s": rois.shape[0] / ((nf + nb) / 1e3),
        "path": "b200det_relayout NCHW->NHWC, channels_last kernels, b200det_relayout back (inside the timed region)",
        "generic_strided_kernels_ms": "fwd 9.3 / bwd 11.4 (tools/kbench_nchw.py, B200DET_NCHW_RELAYOUT=0)"}
    del fn, gn
    res["cfg5_rpn_postprocess_64img"] = {
        "candidates_per_image": per, "filter_proposals_ms": f1_ms,
        "filter_proposals_boxes_per_s": B * per / (f1_ms / 1e3),
        "nms_segmented_ms": seg_ms, "nms_segmented_boxes_per_s": B * per / (seg_ms / 1e3),
        "per_image_loop_of_batched_nms_ms": loop_ms, "speedup_vs_loop": loop_ms / seg_ms}
    res["cfg5_roi_heads_postprocess_64img"] = {
        "instances_per_image": P_ * (C_ - 1), "postprocess_detections_ms": f1b_ms,
        "instances_per_s": B * P_ * (C_ - 1) / (f1b_ms / 1e3), "detections": det_n}
    return res


def subprocess_leg(script, needs):
    """Legs that need the reference's package (baseline/_ref) run in a process of their own: it must not be
    imported next to the stand-alone product (both define torchvision:: schemas)."""
    import subprocess

    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref, needs)):
        return {"unavailable": f"baseline/_ref/{needs} not built (python baseline/build_ref.py --cuda)"}
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([ref, ROOT, env.get("PYTHONPATH", "")])
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    try:
        p = subprocess.run([sys.executable, os.path.join(ROOT, "baseline", script)], env=env, capture_output=True,
                           text=True, timeout=900)
        if p.returncode != 0:
            return {"unavailable": (p.stderr or p.stdout)[-300:]}
        return json.loads(p.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        return {"unavailable": repr(e)[:300]}


def cfg4_leg(capi, dev, rank, world, dist, peak):
    """BASELINE.json configs[3]: 64 images x 2000 RoIs (7x7, 4 FPN levels, fp32), images split 64 / N per rank
    (STRONG scaling).  Forward written straight into this rank's slice of the gather buffer, backward, then the
    exchange north_star names: an NCCL all-gather of the pooled features [128000, 256, 7, 7] (6.42 GB) and of the
    kept indices of a per-rank batched_nms.  The all-gather is also pipelined behind the forward (image chunks:
    chunk c is pooled while chunk c - 1 is on the wire; the buffer is then ordered chunk, rank, RoI)."""
    from cpu_vision_b200 import sharding

    B, per_img, ph = 64, 2000, 7
    lo, hi = sharding.image_block(B, rank, world)
    nb = hi - lo
    gen = torch.Generator(device=dev).manual_seed(1234 + 4 + rank)
    feats = [torch.randn(nb, C, h, w, generator=gen, device=dev).contiguous(memory_format=torch.channels_last)
             for h, w in SHAPES]
    rois = make_rois(gen, nb, per_img, 16.0, 800.0, dev)
    K = nb * per_img
    grad = torch.randn(K, C, ph, ph, generator=gen, device=dev).contiguous(memory_format=torch.channels_last)
    gfeats = [torch.empty_like(f) for f in feats]
    shapes = [(nb, C, h, w) for h, w in SHAPES]
    full, mine = sharding.gather_buffer(K, (C, ph, ph), torch.float32, dev, rank, world, torch.channels_last)
    levels = capi.fpn_level_map(rois, 2, 5)
    sync = torch.cuda.synchronize

    def timed(fn, iters=5, warm=2):
        for _ in range(warm):
            fn()
        if world > 1:
            dist.barrier()
        sync()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        b.synchronize()
        ms = a.elapsed_time(b) / iters
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms

    fwd = lambda: capi.msroi_align_forward(feats, rois, SCALES, ph, ph, 2, False, levels=levels, out=mine)
    bwd = lambda: capi.msroi_align_backward(grad, rois, levels, SCALES, shapes, ph, ph, 2, False, layout=capi.NHWC,
                                            outs=gfeats)
    f_ms, b_ms = timed(fwd), timed(bwd)
    fb = sum(f.numel() for f in feats) * 4
    ob = K * C * ph * ph * 4
    res = {"cfg4": {
        "images": B, "rois": B * per_img, "images_per_rank": nb, "scaling": "strong", "fwd_ms": f_ms, "bwd_ms": b_ms,
        "bwd_rois_per_s": B * per_img / (b_ms / 1e3), "fwd_bwd_rois_per_s": B * per_img / ((f_ms + b_ms) / 1e3),
        "bwd_frac_of_hbm": (fb + ob + K * 20) / (b_ms / 1e3) / 1e9 / peak,
        "fwd_frac_of_hbm": (fb + ob + K * 20) / (f_ms / 1e3) / 1e9 / peak,
        "sm_l2_port_16_image_slice": port_evidence("cfg4_slice")}}
    if world == 1:
        return res
    # --- the exchange
    ag_ms = timed(lambda: dist.all_gather_into_tensor(full, mine), iters=3, warm=1)
    del grad, gfeats
    torch.cuda.empty_cache()
    tmp_in = torch.empty_like(mine)
    tmp_out = torch.empty_like(full)
    plain_ms = timed(lambda: dist.all_gather_into_tensor(tmp_out, tmp_in), iters=3, warm=1)
    del tmp_in, tmp_out
    recv = (world - 1) * mine.numel() * 4
    # kept indices: one batched_nms per rank (cfg2 shape), gathered as (count, padded int64[capacity])
    n = 5000
    xy = torch.rand(n, 2, generator=gen, device=dev) * 1000
    boxes = torch.cat([xy, xy + torch.rand(n, 2, generator=gen, device=dev) * 200 + 10], 1)
    scores = torch.randperm(n, generator=gen, device=dev).float() / n
    idxs = torch.randint(0, 80, (n,), generator=gen, device=dev)
    keep = capi.batched_nms(boxes, scores, idxs, 0.5)
    cnt = torch.tensor([keep.numel()], device=dev)
    kept_ms = timed(lambda: sharding.all_gather_keep(keep, cnt, n), iters=3, warm=1)
    # pipelined: chunk c pooled while chunk c - 1 is gathered (NCCL's stream waits for the kernels enqueued before
    # the call; async_op returns at once, so the next chunk's kernels overlap the transfer)
    chunks = min(4, nb)
    ipc = nb // chunks
    chunks = nb // ipc
    Kc = ipc * per_img
    piped = torch.empty((chunks * world * Kc, C, ph, ph), dtype=torch.float32, device=dev,
                        memory_format=torch.channels_last)
    parts = []
    for c in range(chunks):
        r = rois[c * Kc:(c + 1) * Kc].clone()
        r[:, 0] -= c * ipc
        parts.append(([f[c * ipc:(c + 1) * ipc] for f in feats], r, levels[c * Kc:(c + 1) * Kc].contiguous(),
                      piped[(c * world) * Kc:((c + 1) * world) * Kc], piped[(c * world + rank) * Kc:(c * world + rank + 1) * Kc]))

    def pipelined():
        hs = []
        for fs, r, lv, dst, my in parts:
            capi.msroi_align_forward(fs, r, SCALES, ph, ph, 2, False, levels=lv, out=my)
            hs.append(dist.all_gather_into_tensor(dst, my, async_op=True))
        for h in hs:
            h.wait()

    def sequential():
        for fs, r, lv, dst, my in parts:
            capi.msroi_align_forward(fs, r, SCALES, ph, ph, 2, False, levels=lv, out=my)
        torch.cuda.current_stream().synchronize()
        for fs, r, lv, dst, my in parts:
            dist.all_gather_into_tensor(dst, my)

    pipe_ms = timed(pipelined, iters=3, warm=1)
    seq_ms = timed(sequential, iters=3, warm=1)
    del piped, parts
    torch.cuda.empty_cache()
    # fused: ONE kernel pools the rows and stores them into every rank's buffer over NVLink peer memory
    # (b200det_msroi_align_forward_push; the two 4-byte all-reduces that order the ranks are inside the timed region)
    push_ms, push_same, push_err = float("nan"), False, None
    try:
        pg = sharding.PushGather(K, C, ph, ph, torch.float32, dev)
        push_ms = timed(lambda: pg.forward(feats, rois, SCALES, ph, ph, 2, False, levels=levels), iters=3, warm=1)
        fwd()
        dist.all_gather_into_tensor(full, mine)
        sync()
        same = torch.tensor([int(torch.equal(pg.full, full))], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        push_same = bool(same.item())
        pg.close()
    except Exception as e:  # peer memory cannot be mapped on this box: the NCCL numbers above stand
        push_err = f"{type(e).__name__}: {e}"[:300]
    res["allgather"] = {
        "what": "pooled features of cfg4, [128000, 256, 7, 7] fp32 = 6.42 GB, all_gather_into_tensor in place "
                "(the forward wrote this rank's slice of the buffer)",
        "ms": ag_ms, "bytes_received_per_gpu": recv, "recv_GBps_per_gpu": recv / (ag_ms / 1e3) / 1e9,
        "nccl_plain_allgather_same_size_ms": plain_ms, "nccl_plain_recv_GBps_per_gpu": recv / (plain_ms / 1e3) / 1e9,
        "frac_of_nccl_plain": plain_ms / ag_ms,
        "kept_indices_allgather_ms": kept_ms, "kept_indices_bytes_per_rank": (n + 1) * 8,
        "forward_then_gather_ms": seq_ms, "pipelined_forward_gather_ms": pipe_ms, "chunks": chunks,
        "overlap_frac": (seq_ms - pipe_ms) / max(min(f_ms, ag_ms), 1e-9),
        "push_fused_forward_gather_ms": push_ms, "push_recv_GBps_per_gpu": recv / (push_ms / 1e3) / 1e9,
        "push_vs_pipelined_nccl": pipe_ms / push_ms, "push_equals_nccl_result_on_every_rank": push_same,
        "push_error": push_err,
        "push_what": "b200det_msroi_align_forward_push: the forward kernel stores every pooled row into all ranks' "
                     "buffers (CUDA IPC peer memory over NVLink), no collective pass; includes the two 4-byte "
                     "all-reduces that order the ranks",
        "limiter": "the all-gather: every GPU receives (N-1)/N of 6.42 GB over NVLink while the local kernels "
                   "take fwd_ms + bwd_ms"}
    return res


def cfg5_leg(capi, dev, rank, world, dist):
    """BASELINE.json configs[4] on its synthetic shapes at N GPUs: the post-backbone chain of
    fasterrcnn_resnet50_fpn for 64 images of 800x1088 split 64 / N per rank (strong scaling, sharded by image, no
    exchange): filter_proposals (5 levels x 3 anchors, 2000 / 2000, IoU 0.7) -> MultiScaleRoIAlign of the kept
    proposals (7x7, 4 levels) -> postprocess_detections (91 classes, 0.05 / 0.5 / 100).  Device-timed, max over ranks."""
    from cpu_vision_b200 import detection, sharding

    B = 64
    lo, hi = sharding.image_block(B, rank, world)
    nb = hi - lo
    gen = torch.Generator(device=dev).manual_seed(1234 + 5 + rank)
    napl = [200 * 272 * 3, 100 * 136 * 3, 50 * 68 * 3, 25 * 34 * 3, 13 * 17 * 3]
    A = sum(napl)
    ctr = torch.rand(nb, A, 2, generator=gen, device=dev) * torch.tensor([1088.0, 800.0], device=dev)
    wh = torch.exp(torch.rand(nb, A, 2, generator=gen, device=dev) * 3.5 + 2.5)
    props = torch.cat([ctr - wh / 2, ctr + wh / 2], 2)
    del ctr, wh
    obj = torch.randn(nb * A, 1, generator=gen, device=dev)
    shapes = [(800, 1088)] * nb
    feats = [torch.randn(nb, C, h, w, generator=gen, device=dev).contiguous(memory_format=torch.channels_last)
             for h, w in SHAPES]
    P_, C_ = 1000, 91
    lg = torch.randn(nb * P_, C_, generator=gen, device=dev) * 2.5
    rg = torch.randn(nb * P_, C_ * 4, generator=gen, device=dev) * 0.5
    state = {}

    def rpn():
        state["boxes"], _ = detection.filter_proposals(props, obj, shapes, napl, 2000, 2000, 0.7)

    def pool():
        bx = state["boxes"]
        ids = torch.cat([torch.full((b.shape[0], 1), float(i), device=dev) for i, b in enumerate(bx)])
        rois = torch.cat([ids, torch.cat(bx)], 1)
        state["rois"] = rois.shape[0]
        capi.msroi_align_forward(feats, rois, SCALES, 7, 7, 2, False)

    def post():
        # the box head runs between pooling and this step in the model; its outputs are synthetic here
        detection.postprocess_detections(lg, rg, [b[:P_] if b.shape[0] >= P_ else torch.cat([b, b[:1].expand(P_ - b.shape[0], 4)])
                                                  for b in state["boxes"]], shapes, 0.05, 0.5, 100)

    def timed(fn, iters=5, warm=2):
        for _ in range(warm):
            fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        b.synchronize()
        ms = a.elapsed_time(b) / iters
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms

    r_ms, p_ms, d_ms = timed(rpn), timed(pool), timed(post)
    return {"images": B, "images_per_rank": nb, "scaling": "strong", "filter_proposals_ms": r_ms,
            "box_roi_pool_ms": p_ms, "rois_pooled_per_rank": state.get("rois"), "postprocess_detections_ms": d_ms,
            "chain_ms": r_ms + p_ms + d_ms, "images_per_s": B / ((r_ms + p_ms + d_ms) / 1e3),
            "note": "synthetic RPN / box-head outputs of the shapes fasterrcnn_resnet50_fpn produces for 800x1088 "
                    "images; the reference's own model drives the same kernels at N = 1 in secondary.cfg5_model"}


def gather_leg(out, world, dist):
    """NCCL all-gather of the pooled features (the optional exchange step of SURVEY.md 8e)."""
    full = torch.empty((world * out.shape[0],) + tuple(out.shape[1:]), dtype=out.dtype, device=out.device,
                       memory_format=torch.channels_last)
    for _ in range(2):
        dist.all_gather_into_tensor(full, out)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        dist.all_gather_into_tensor(full, out)
    b.record()
    b.synchronize()
    ms = a.elapsed_time(b) / 5
    t = torch.tensor([ms], device=out.device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    recv = (world - 1) * out.numel() * out.element_size()
    return {"ms": ms, "recv_GBps_per_gpu": recv / (ms / 1e3) / 1e9, "bytes_received_per_gpu": recv}


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU fallback")
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    try:
        run_b200(args, rank, world, local)
    finally:
        if world > 1:
            import torch.distributed as dist

            dist.destroy_process_group()


if __name__ == "__main__":
    main()
