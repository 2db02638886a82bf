#!/usr/bin/env python
"""Benchmark of the curve-layer hot path (BASELINE.json metric: curve-layer fwd+bwd Mpixels/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (config.workload): BASELINE.json configs[3] -- synthetic 4K (3840x2160) RGB fp32, batch 64
per GPU, one step = one fused forward + one fused backward (with image gradient) of the layer over
the whole batch = 530.8 Mpx.  At N > 1 (torchrun) every rank runs the same per-GPU batch on its own
GPU (batch sharding, no collective on the data path): weak scaling, value = N * units / max-rank time.

Printed JSON (rank 0, one line):
  value     Mpx/s with inputs resident in HBM, CUDA-event timed, max over ranks
  e2e       the same step driven from pinned HOST buffers through the public module API: per step the
            image + mask are copied host->device, forward+backward run, and the regulariser and knot
            gradients are read back device->host
  roofline  the dominant kernel (fused backward, 37 algorithmic B/px) against the measured HBM peak;
            roofline_fwd / roofline_step / roofline_bwd_no_gimg (25 B/px: the GCURLNet case, img is data) beside it
  fwd_only  forward-only (inference) legs at every N: resident-input rate, and host round trips with the
            output copied back (fp32: 13 B/px up + 12 down; 8-bit: 4 up + 3 down)
  e2e_u8    the training step fed from 8-bit host data (4 B/px up instead of 13), at every N
  e2e.h2d_peak_gbs / frac_of_h2d_peak   a plain pinned host->device copy of the same chunks on all ranks at once
  gpu_eager_baseline  the same op sequence run eagerly by PyTorch on this GPU (the reference's own
            way of using a GPU), one 1080p tile
  gpu_compile_baseline  the same op sequence under torch.compile (inductor) on this GPU, 1080p x 4
            (dev/compile_baseline.py in a subprocess): the strongest "reference on B200" there is
  extras    non-headline measurements: BASELINE configs[1] (1080p x16 forward), configs[2] / configs[4]
            (GCURLNet optimiser step at 512x512x32, under DDP when N > 1), the 8-bit paths, TriSpace, CURLLoss
  cpu_baseline  the CPU oracle (the reference's ATen op sequence, restated; kind "port") on this
            box's host cores over a bounded sample of the same workload
`--impl reference` times only that CPU path (the reference has no native code and /root/reference
does not travel to the GPU box, so the timed code is oracle/curl_oracle.py, bit-identical to it).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "curve_layer_fwd_bwd_mpixels_per_s"
UNIT = "Mpx/s"
H4K, W4K = 2160, 3840
FWD_BYTES_PER_PX = 25      # 12 img + 12 out + 1 bool mask            
BWD_BYTES_PER_PX = 37      # 12 img + 12 gout + 12 gimg + 1 bool mask


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = [r for (t, r) in self.rows if t0 - 0.05 <= t <= t1 + 0.15] or [r for (_, r) in self.rows]
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def make_inputs(B, H, W, dev, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    img = torch.rand(B, 3, H, W, generator=g, device=dev)
    mask = torch.rand(B, 1, H, W, generator=g, device=dev) > 0.1
    L = 0.1 * torch.randn(B, 48, generator=g, device=dev)
    R = 0.1 * torch.randn(B, 48, generator=g, device=dev)
    Hk = 0.1 * torch.randn(B, 64, generator=g, device=dev)
    return img, mask, L, R, Hk


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (reference op sequence) on the host cores, bounded sample
# ------------------------------------------------------------------------------------------------
def cpu_reference_step(sample):
    from oracle import curl_oracle as O
    img, mask, L, R, Hk, gout, greg = sample
    return O.curl_layer_fwd_bwd(img, mask, L, R, Hk, gout, greg)


def cpu_sample(h=1080, w=1920):
    g = torch.Generator().manual_seed(1234)
    img = torch.rand(1, 3, h, w, generator=g)
    mask = torch.rand(1, 1, h, w, generator=g) > 0.1
    L, R, Hk = (0.1 * torch.randn(1, n, generator=g) for n in (48, 48, 64))
    gout = torch.randn(1, 3, h, w, generator=g)
    return img, mask, L, R, Hk, gout, torch.ones(1)


def time_cpu(steps, warmup, h=1080, w=1920):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sample = cpu_sample(h, w)
    for _ in range(warmup):
        cpu_reference_step(sample)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_reference_step(sample)
    dt = (time.perf_counter() - t0) / steps
    mpx = h * w / 1e6
    return {"value": mpx / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "one %dx%d tile of a 4K image, batch 1, fwd+bwd with image gradient, fp32 (%.2f Mpx/step, "
                      "%d steps, %.2f s/step); oracle/curl_oracle.py = the reference's ATen op sequence "
                      "(bit-identical to the patched reference on CPU)" % (h, w, mpx, steps, dt)}, dt


def time_cpu_real_images(reps=2):
    """the reference's CPU path on the nine real photographs it ships (512 x 340/341,
    8-bit; tests/golden/real.npz carries them to the GPU box), one image at a time, batch 1 -- forward under
    no_grad (what infer.py would do with this layer) and forward + backward."""
    import numpy as np
    from oracle import curl_oracle as O
    path = os.path.join(ROOT, "tests", "golden", "real.npz")
    if not os.path.exists(path):
        return {"unavailable": "tests/golden/real.npz missing"}
    g = np.load(path)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    items = []
    gen = torch.Generator().manual_seed(4)
    for i in range(int(g["n"])):
        x = O.u8_to_unit(torch.from_numpy(np.ascontiguousarray(g["%d.u8" % i])).permute(2, 0, 1)[None]).contiguous()
        L, R, Hk = (torch.from_numpy(g["%d.%s" % (i, k)]) for k in "LRH")
        m = torch.ones(1, 1, x.shape[2], x.shape[3], dtype=torch.bool)
        items.append((x, m, L, R, Hk, torch.randn(x.shape, generator=gen), torch.ones(1)))
    mpx = sum(it[0].shape[2] * it[0].shape[3] for it in items) / 1e6
    with torch.no_grad():
        for it in items[:2]:
            O.curl_layer(*it[:5])
        t0 = time.perf_counter()
        for _ in range(reps):
            for it in items:
                O.curl_layer(*it[:5])
        dt_f = (time.perf_counter() - t0) / reps
    O.curl_layer_fwd_bwd(*items[0])
    t0 = time.perf_counter()
    for _ in range(reps):
        for it in items:
            O.curl_layer_fwd_bwd(*it)
    dt_fb = (time.perf_counter() - t0) / reps
    return {"fwd_mpx_per_s": mpx / dt_f, "fwd_bwd_mpx_per_s": mpx / dt_fb, "cores": cores, "images": len(items),
            "mpx_total": mpx, "fwd_ms_per_image": dt_f / len(items) * 1e3, "fwd_bwd_ms_per_image": dt_fb / len(items) * 1e3,
            "what": "the 9 example photographs of the reference (adobe5k_dpe/curl_example_test_input, 512x340/341, "
                    "u8/255), one at a time, batch 1, mask == 1, knots 0.1*randn seed 1234; oracle on the host cores"}


def time_gpu_compile(timeout_s=420):
    """dev/compile_baseline.py in a subprocess (a failed or hung compile must not take the bench line with it)."""
    cmd = [sys.executable, os.path.join(ROOT, "dev", "compile_baseline.py"), "4", "1080", "1920"]
    try:
        proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=timeout_s, cwd=ROOT)
    except subprocess.TimeoutExpired:
        return {"unavailable": "timed out after %d s" % timeout_s}
    for ln in reversed(proc.stdout.strip().splitlines()):
        if ln.startswith("{"):
            try:
                return json.loads(ln)
            except ValueError:
                break
    return {"unavailable": "rc %d: %s" % (proc.returncode, proc.stderr.strip()[-300:])}


def time_gpu_eager(dev, steps=3, warmup=2, h=1080, w=1920):
    """The reference's own way of running this path on a GPU: the same ATen op sequence, eagerly, on
    the B200.  ~650 launches per forward and every intermediate kept for autograd
    (~3.6 kB/px), which is why the sample is one 1080p tile: 4K x 64 would need ~1.9 TB."""
    sample = tuple(t.to(dev) for t in cpu_sample(h, w))
    for _ in range(warmup):
        cpu_reference_step(sample)
    torch.cuda.synchronize(dev)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        cpu_reference_step(sample)
    e1.record()
    torch.cuda.synchronize(dev)
    dt = e0.elapsed_time(e1) / 1e3 / steps
    mpx = h * w / 1e6
    peak_gb = torch.cuda.max_memory_allocated(dev) / 1e9
    return {"value": mpx / dt, "unit": UNIT, "kind": "port, eager PyTorch on this GPU",
            "sample": "one %dx%d tile, batch 1, fwd+bwd with image gradient, fp32 (%.2f Mpx/step, %d steps, "
                      "%.1f ms/step); oracle/curl_oracle.py on cuda" % (h, w, mpx, steps, dt * 1e3),
            "peak_mem_gb": peak_gb}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 5))
    warmup = max(1, min(args.warmup, 1))
    base, dt = time_cpu(steps, warmup)
    base["real_images"] = time_cpu_real_images()
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "CPU arm runs a bounded sample of the workload (see cpu_baseline.sample)"},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


WORKLOAD = "4K (3840x2160) synthetic RGB fp32, batch 64 per GPU, curve-layer fwd+bwd (BASELINE.json configs[3])"


# ------------------------------------------------------------------------------------------------
# GPU arm
def net_training_step_extra(dev, world, local, timed):
    """BASELINE.json configs[2] (N = 1) and the training half of configs[4] (N > 1): one optimiser step of
    the layer's caller, GCURLNet (reference model.py:179-203), at 512 x 512 x 32 per GPU.  The backbone is
    library code on PyTorch/cuDNN (north_star); the reference's timm EfficientNetV2 is absent from this
    image, so torchvision's efficientnet_v2_s with a 160-wide head stands in.  Loss =
    L1 + 1e-6 * regulariser, Adam as in main.py:236-237.  At N > 1 the net is wrapped in
    DistributedDataParallel exactly as main.py:224-225 does: the only collective is DDP's NCCL gradient
    all-reduce of the backbone weights; the curve layer adds none.  Runs on every rank."""
    import torch.nn as nn
    import torch.nn.functional as F
    import curl_b200
    from curl_b200.parallel import max_over_ranks
    try:
        from torchvision.models import efficientnet_v2_s
    except Exception as e:
        return {"unavailable": "torchvision: %s" % e}
    Bn, Hn, Wn = 32, 512, 512
    torch.manual_seed(99)                                    # same initial weights on every rank
    net = curl_b200.GCURLNet(backbone=efficientnet_v2_s(weights=None, num_classes=160)).to(dev)
    net.train()
    model = nn.parallel.DistributedDataParallel(net, device_ids=[local], output_device=local) if world > 1 else net
    opt = torch.optim.Adam(net.parameters(), lr=5e-7, betas=(0.5, 0.999))
    g = torch.Generator(device=dev).manual_seed(555 + local)
    x = torch.rand(Bn, 3, Hn, Wn, generator=g, device=dev)
    y = (x * 0.9 + 0.05).contiguous()
    m = torch.rand(Bn, 1, Hn, Wn, generator=g, device=dev) > 0.1

    def step():
        opt.zero_grad(set_to_none=True)
        out, reg = model(x, m)
        loss = F.l1_loss(out, y) + 1e-6 * reg.mean()
        loss.backward()
        opt.step()
    ms = max_over_ranks(timed(step, 10), dev)
    # the curve layer's own share of that step: forward + backward without image gradient (25 + 25 B/px)
    layer = net.curllayer
    k = [(0.1 * torch.randn(Bn, n, generator=g, device=dev)).requires_grad_(True) for n in (48, 48, 64)]
    go = torch.randn(Bn, 3, Hn, Wn, generator=g, device=dev)
    gr = torch.ones(Bn, device=dev)

    def layer_step():
        o, r = layer(x, m, *k)
        torch.autograd.backward([o, r], [go, gr])
    ms_layer = max_over_ranks(timed(layer_step, 20), dev)
    mpx = world * Bn * Hn * Wn / 1e6
    res = {"step_ms": ms, "images_per_s": world * Bn / (ms / 1e3), "mpx_per_s": mpx / (ms / 1e3),
           "curve_layer_fwd_bwd_ms": ms_layer, "curve_layer_share": ms_layer / ms,
           "backbone": "torchvision efficientnet_v2_s (random init, fp32, cuDNN), %.1f M parameters"
                       % (sum(p.numel() for p in net.backbone.parameters()) / 1e6),
           "what": "GCURLNet optimiser step, 512x512 batch 32 per GPU: backbone -> [L|R|H] -> CURLLayer -> "
                   "L1 + 1e-6*reg -> backward -> Adam; " +
                   ("DistributedDataParallel over NCCL (gradient all-reduce of the backbone only), max over ranks"
                    if world > 1 else "single GPU")}
    del model, net, opt, x, y, m, go
    return res


# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch.distributed as dist
    import curl_b200
    from curl_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the curve layer has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout to the one JSON line
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))
    _lib.load()

    B, H, W = args.batch, args.height, args.width
    n_px = B * H * W
    layer = curl_b200.CURLLayer().to(dev)
    img, mask, L, R, Hk = make_inputs(B, H, W, dev, 1234 + rank)
    img.requires_grad_(True)
    L.requires_grad_(True); R.requires_grad_(True); Hk.requires_grad_(True)
    gout = torch.randn(B, 3, H, W, device=dev, generator=torch.Generator(device=dev).manual_seed(99 + rank))
    greg = torch.ones(B, device=dev)

    def step():
        for t in (img, L, R, Hk):
            t.grad = None
        out, reg = layer(img, mask, L, R, Hk)
        torch.autograd.backward([out, reg], [gout, greg])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()

    # ---- timed region: resident inputs, CUDA events on the launching (current) stream -------------
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)
    launches0 = _lib.launches
    barrier()
    t_wall0 = time.time()
    start = torch.cuda.Event(enable_timing=True); end = torch.cuda.Event(enable_timing=True)
    start.record()
    for k in range(args.steps):
        for t in (img, L, R, Hk):
            t.grad = None
        ev[k][0].record()
        out, reg = layer(img, mask, L, R, Hk)
        ev[k][1].record()
        torch.autograd.backward([out, reg], [gout, greg])
        ev[k][2].record()
    end.record()
    barrier()
    t_wall1 = time.time()
    launches = _lib.launches - launches0
    clocks = sampler.stop(t_wall0, t_wall1)
    ms_total = start.elapsed_time(end)
    fwd_ms = statistics.mean(e[0].elapsed_time(e[1]) for e in ev)
    bwd_ms = statistics.mean(e[1].elapsed_time(e[2]) for e in ev)
    del out, reg

    def max_ranks(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    ms_step = max_ranks(ms_total) / args.steps
    value = world * n_px / 1e6 / (ms_step / 1e3)

    def timed_all(fn, reps, warm=2):
        """ms per call of fn on every rank at once: barrier + sync on both sides, CUDA events, max over ranks."""
        for _ in range(warm):
            fn()
        barrier()
        a = torch.cuda.Event(enable_timing=True); b_ = torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b_.record()
        barrier()
        return max_ranks(a.elapsed_time(b_)) / reps

    # ---- forward only (inference; north_star states the multi-GPU target on batch-sharded inference) -------
    launches_before_legs = _lib.launches

    def fwd_only():
        with torch.no_grad():
            return layer(img, mask, L, R, Hk)
    fwd_only_ms = timed_all(fwd_only, args.steps)
    fwd_only_value = world * n_px / 1e6 / (fwd_only_ms / 1e3)

    # ---- backward without the image gradient (the GCURLNet case: img is data; 25 B/px) ----------------------
    img_data = img.detach()
    ev2 = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps)]

    def step_no_gimg(k=None):
        for t in (L, R, Hk):
            t.grad = None
        o, r = layer(img_data, mask, L, R, Hk)
        if k is not None:
            ev2[k][0].record()
        torch.autograd.backward([o, r], [gout, greg])
        if k is not None:
            ev2[k][1].record()
    for _ in range(2):
        step_no_gimg()
    barrier()
    for k in range(args.steps):
        step_no_gimg(k)
    barrier()
    bwd_nog_ms = max_ranks(statistics.mean(e[0].elapsed_time(e[1]) for e in ev2))
    for t in (L, R, Hk):
        t.grad = None

    # ---- e2e: pinned host buffers -> H2D -> fwd+bwd -> D2H of the step's results -------------------
    # The public module API is driven from HOST data the way the reference's training loop does
    # (main.py:279-283: batch.to(device, non_blocking=True) then net(...)): the batch is fed in
    # chunks of images, chunk i+1's copy (copy stream) overlapping chunk i's kernels (compute stream).
    for t in (img, L, R, Hk):
        t.grad = None
    h_img = torch.empty(img.shape, dtype=torch.float32, pin_memory=True)
    h_mask = torch.empty(mask.shape, dtype=torch.bool, pin_memory=True)
    h_img.copy_(img.detach()); h_mask.copy_(mask)
    h_res = torch.empty(B, 1 + 160, dtype=torch.float32, pin_memory=True)
    d_img = img.detach()          # reuse the resident buffers as the H2D destination
    d_mask = mask
    Ld, Rd, Hd = L.detach(), R.detach(), Hk.detach()
    e2e_steps = max(1, min(args.steps, 5))
    chunk = max(1, min(B, args.e2e_chunk))
    spans = [(lo, min(B, lo + chunk)) for lo in range(0, B, chunk)]
    copy_stream = torch.cuda.Stream(device=dev)
    main_stream = torch.cuda.current_stream(dev)
    d_res = torch.empty(B, 1 + 160, dtype=torch.float32, device=dev)

    def e2e_step():
        ready = []
        copy_stream.wait_stream(main_stream)          # previous step's kernels are done with the buffers
        with torch.cuda.stream(copy_stream):
            for lo, hi in spans:
                d_img[lo:hi].copy_(h_img[lo:hi], non_blocking=True)
                d_mask[lo:hi].copy_(h_mask[lo:hi], non_blocking=True)
                ev_c = torch.cuda.Event()
                ev_c.record(copy_stream)
                ready.append(ev_c)
        for (lo, hi), ev_c in zip(spans, ready):
            main_stream.wait_event(ev_c)
            x = d_img[lo:hi].requires_grad_(True)
            l, r, h = (t[lo:hi].clone().requires_grad_(True) for t in (Ld, Rd, Hd))
            o, rg = layer(x, d_mask[lo:hi], l, r, h)
            torch.autograd.backward([o, rg], [gout[lo:hi], greg[lo:hi]])
            d_res[lo:hi, 0] = rg.detach()
            d_res[lo:hi, 1:49] = l.grad; d_res[lo:hi, 49:97] = r.grad; d_res[lo:hi, 97:161] = h.grad
        h_res.copy_(d_res, non_blocking=True)

    e2e_step()
    barrier()
    launches_e0 = _lib.launches
    s2 = torch.cuda.Event(enable_timing=True); e2 = torch.cuda.Event(enable_timing=True)
    s2.record()
    for _ in range(e2e_steps):
        e2e_step()
    e2.record()
    barrier()
    t2 = torch.tensor([s2.elapsed_time(e2)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    e2e_ms = t2.item() / e2e_steps
    e2e_value = world * n_px / 1e6 / (e2e_ms / 1e3)
    h2d = h_img.numel() * 4 + h_mask.numel()
    d2h = h_res.numel() * 4

    # ---- what the host->device path can do: the same pinned chunks, plain copies, all ranks at once ---------
    def h2d_probe():
        with torch.cuda.stream(copy_stream):
            for lo, hi in spans:
                d_img[lo:hi].copy_(h_img[lo:hi], non_blocking=True)
                d_mask[lo:hi].copy_(h_mask[lo:hi], non_blocking=True)
        main_stream.wait_stream(copy_stream)
    h2d_ms = timed_all(h2d_probe, 3, warm=1)
    h2d_peak_gbs = h2d / (h2d_ms / 1e3) / 1e9                 # per rank, measured with every rank copying

    # ---- forward-only host round trips (inference e2e), fp32 and 8-bit, at every N ---------------------------
    out_stream = torch.cuda.Stream(device=dev)
    # every chunk's output is copied back (same bytes, same D2H queue) into a two-slot pinned ring instead of a
    # whole-batch host buffer: 8 ranks would otherwise pin 8 x 6.4 GB more host memory for nothing
    h_out = torch.empty((2, chunk) + tuple(img.shape[1:]), dtype=torch.float32, pin_memory=True)
    d_out = torch.empty_like(d_img)

    def infer_f32():
        ready = []
        copy_stream.wait_stream(main_stream)
        with torch.cuda.stream(copy_stream):
            for lo, hi in spans:
                d_img[lo:hi].copy_(h_img[lo:hi], non_blocking=True)
                d_mask[lo:hi].copy_(h_mask[lo:hi], non_blocking=True)
                ev_c = torch.cuda.Event(); ev_c.record(copy_stream); ready.append(ev_c)
        with torch.no_grad():
            for (lo, hi), ev_c in zip(spans, ready):
                main_stream.wait_event(ev_c)
                o, _ = layer(d_img[lo:hi], d_mask[lo:hi], Ld[lo:hi], Rd[lo:hi], Hd[lo:hi])
                d_out[lo:hi] = o
                ev_k = torch.cuda.Event(); ev_k.record(main_stream)
                out_stream.wait_event(ev_k)
                with torch.cuda.stream(out_stream):
                    h_out[(lo // chunk) % 2, :hi - lo].copy_(d_out[lo:hi], non_blocking=True)
        main_stream.wait_stream(out_stream)
    infer_f32_ms = timed_all(infer_f32, 3, warm=1)
    del h_out, d_out

    d_img8 = torch.empty(img.shape, dtype=torch.uint8, device=dev)
    d_out8 = torch.empty(img.shape, dtype=torch.uint8, device=dev)
    h_img8 = torch.empty(img.shape, dtype=torch.uint8, pin_memory=True)
    h_img8.copy_((img.detach() * 255).to(torch.uint8))
    h_out8 = torch.empty((2, chunk) + tuple(img.shape[1:]), dtype=torch.uint8, pin_memory=True)

    def infer_u8():
        ready = []
        copy_stream.wait_stream(main_stream)
        with torch.cuda.stream(copy_stream):
            for lo, hi in spans:
                d_img8[lo:hi].copy_(h_img8[lo:hi], non_blocking=True)
                d_mask[lo:hi].copy_(h_mask[lo:hi], non_blocking=True)
                ev_c = torch.cuda.Event(); ev_c.record(copy_stream); ready.append(ev_c)
        with torch.no_grad():
            for (lo, hi), ev_c in zip(spans, ready):
                main_stream.wait_event(ev_c)
                o8, _ = layer.forward_u8(d_img8[lo:hi], d_mask[lo:hi], Ld[lo:hi], Rd[lo:hi], Hd[lo:hi])
                d_out8[lo:hi] = o8
                ev_k = torch.cuda.Event(); ev_k.record(main_stream)
                out_stream.wait_event(ev_k)
                with torch.cuda.stream(out_stream):
                    h_out8[(lo // chunk) % 2, :hi - lo].copy_(d_out8[lo:hi], non_blocking=True)
        main_stream.wait_stream(out_stream)
    infer_u8_ms = timed_all(infer_u8, 3, warm=1)
    with torch.no_grad():
        fwd_u8_ms = timed_all(lambda: layer.forward_u8(d_img8, d_mask, Ld, Rd, Hd), 5, warm=2)
    del d_out8, h_out8

    # ---- the training step fed from 8-bit host data (what data.py holds before TF.to_tensor), at every N -----
    from curl_b200.functional import to_tensor_u8

    def u8_train_step():
        ready = []
        copy_stream.wait_stream(main_stream)
        with torch.cuda.stream(copy_stream):
            for lo, hi in spans:
                d_img8[lo:hi].copy_(h_img8[lo:hi], non_blocking=True)
                d_mask[lo:hi].copy_(h_mask[lo:hi], non_blocking=True)
                ev_c = torch.cuda.Event()
                ev_c.record(copy_stream)
                ready.append(ev_c)
        for (lo, hi), ev_c in zip(spans, ready):
            main_stream.wait_event(ev_c)
            x = to_tensor_u8(d_img8[lo:hi]).requires_grad_(True)
            l, r, h = (t[lo:hi].clone().requires_grad_(True) for t in (Ld, Rd, Hd))
            o, rg = layer(x, d_mask[lo:hi], l, r, h)
            torch.autograd.backward([o, rg], [gout[lo:hi], greg[lo:hi]])
            d_res[lo:hi, 0] = rg.detach()
            d_res[lo:hi, 1:49] = l.grad; d_res[lo:hi, 49:97] = r.grad; d_res[lo:hi, 97:161] = h.grad
        h_res.copy_(d_res, non_blocking=True)
    e2e_u8_ms = timed_all(u8_train_step, 3, warm=1)
    h2d_u8 = h_img8.numel() + h_mask.numel()
    del d_img8, h_img8
    launches_legs = _lib.launches - launches_before_legs

    # ---- extras (reported beside the headline, never part of it) -----------------------------------
    def timed(fn, reps=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(dev)
        a = torch.cuda.Event(enable_timing=True); b_ = torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b_.record()
        torch.cuda.synchronize(dev)
        return a.elapsed_time(b_) / reps

    extras = {}
    if rank == 0 and world == 1 and not args.no_extras:
        with torch.no_grad():
            # BASELINE.json configs[1]: forward only, 1080p, batch 16 (398 MB per tensor > L2)
            gi, gm, gl, gr, gh = make_inputs(16, 1080, 1920, dev, 4321)
            ms = timed(lambda: layer(gi, gm, gl, gr, gh), 20)
            extras["config1_fwd_1080p_b16"] = {"ms": ms, "mpx_per_s": 16 * 1080 * 1920 / 1e6 / (ms / 1e3),
                                               "gbs": 16 * 1080 * 1920 * FWD_BYTES_PER_PX / (ms / 1e3) / 1e9}
            del gi, gm, gl, gr, gh
        # Second input distribution: the reference's nine example photographs (tests/golden/real.npz),
        # tiled to 4K, batch 9, 8-bit values / 255 -- smooth regions, real saturation statistics, dark areas
        rpath = os.path.join(ROOT, "tests", "golden", "real.npz")
        if os.path.exists(rpath):
            import numpy as np
            rg = np.load(rpath)
            tiles = []
            for i in range(int(rg["n"])):
                a = torch.from_numpy(np.ascontiguousarray(rg["%d.u8" % i])).permute(2, 0, 1)          # [3, h, w] u8
                reps = (1, -(-H4K // a.shape[1]), -(-W4K // a.shape[2]))
                tiles.append(a.repeat(*reps)[:, :H4K, :W4K])
            r_img = (torch.stack(tiles).to(dev).float() / 255.0).contiguous().requires_grad_(True)
            Br = r_img.shape[0]
            r_mask = torch.ones(Br, 1, H4K, W4K, dtype=torch.bool, device=dev)
            r_k = [torch.from_numpy(np.concatenate([rg["%d.%s" % (i, k)] for i in range(Br)])).to(dev).requires_grad_(True) for k in "LRH"]
            r_go = gout[:Br]
            r_gr = greg[:Br]
            with torch.no_grad():
                ms_rf = timed(lambda: layer(r_img, r_mask, *r_k), 10)

            def real_step():
                for t in [r_img] + r_k:
                    t.grad = None
                o, r = layer(r_img, r_mask, *r_k)
                torch.autograd.backward([o, r], [r_go, r_gr])
            ms_rs = timed(real_step, 10)
            rpx = Br * H4K * W4K
            extras["real_images_tiled_4k_b%d" % Br] = {
                "fwd_ms": ms_rf, "fwd_mpx_per_s": rpx / 1e6 / (ms_rf / 1e3), "fwd_bwd_ms": ms_rs,
                "fwd_bwd_mpx_per_s": rpx / 1e6 / (ms_rs / 1e3),
                "headline_fwd_bwd_mpx_per_s_same_run": value,
                "what": "the reference's 9 example photographs tiled to 3840x2160, batch 9, u8/255, mask == 1, knots 0.1*randn: "
                        "the kernels are arithmetic-bound and branch-free, so the rate should not depend on the image content"}
            del r_img, r_mask, r_k, tiles
        # TriSpace polynomial residual (TriSpaceRegNet's per-pixel path), spatial=True (126 coefficients),
        # BASELINE configs[2]'s shape: 512 x 512, batch 32; forward incl. generate_image, and backward
        import torch.nn as nn
        tri = curl_b200.TriSpaceRegNet(spatial=True, backbone=nn.Identity()).to(dev)
        gt = torch.Generator(device=dev).manual_seed(7)
        ti = torch.rand(32, 3, 512, 512, generator=gt, device=dev)
        tg = torch.randn(32, 3, 512, 512, generator=gt, device=dev)
        tc = [(0.3 * torch.randn(32, 3, 126, generator=gt, device=dev)).requires_grad_(True) for _ in range(3)]
        with torch.no_grad():
            ms_f = timed(lambda: tri._fused(ti, tc[0], tc[1], tc[2], True), 10)

        def tri_step():
            for c in tc:
                c.grad = None
            tri._fused(ti, tc[0], tc[1], tc[2], True).backward(tg)
        ms_s = timed(tri_step, 10)
        tpx = 32 * 512 * 512
        extras["trispace_512_b32_spatial"] = {
            "fwd_ms": ms_f, "fwd_mpx_per_s": tpx / 1e6 / (ms_f / 1e3), "fwd_bwd_ms": ms_s,
            "fwd_bwd_mpx_per_s": tpx / 1e6 / (ms_s / 1e3),
            "fwd_tflops_fp32": tpx * 3 * 2 * 3 * 125 / (ms_f / 1e3) / 1e12,
            "what": "fused generate_residual + generate_image, degree 4, 5 variables; FMA-bound (nested evaluation, 2250 flop/px forward); backward from the saved sigmoid outputs"}
        del ti, tg, tc
        # CURLLoss (model.py:36-118): fused colour terms (one forward + one backward kernel) and the whole
        # criterion step incl. library MS-SSIM, on the headline batch (pred = the layer's output)
        from curl_b200.functional import loss_terms
        crit = curl_b200.CURLLoss().to(dev)
        with torch.no_grad():
            l_target = (img.detach() * 0.9 + 0.05)
            l_pred = layer(img.detach(), mask, Ld, Rd, Hd)[0]
            ms_lf = timed(lambda: loss_terms(l_pred, l_target, mask), 10)
        l_pred.requires_grad_(True)
        l_w = torch.full((B, 5), 1e-7, device=dev)
        l_gl = torch.full((B, 1) + tuple(img.shape[2:]), 1e-7, device=dev)

        def terms_step():
            l_pred.grad = None
            s, lp, _ = loss_terms(l_pred, l_target, mask)
            torch.autograd.backward([s, lp], [l_w, l_gl])
        ms_ls = timed(terms_step, 5)

        def crit_step():
            l_pred.grad = None
            crit(l_pred, l_target, mask).backward()
        ms_lc = timed(crit_step, 3)
        extras["curlloss_4k_b%d" % B] = {
            "terms_fwd_ms": ms_lf, "terms_fwd_mpx_per_s": n_px / 1e6 / (ms_lf / 1e3),
            "terms_fwd_gbs": n_px * 33 / (ms_lf / 1e3) / 1e9,
            "terms_fwd_bwd_ms": ms_ls, "terms_fwd_bwd_mpx_per_s": n_px / 1e6 / (ms_ls / 1e3),
            "criterion_fwd_bwd_ms": ms_lc,
            "what": "fused colour terms: 33 B/px forward (pred, target, mask in; 2 L planes out), 41 B/px backward; "
                    "criterion = terms + MS-SSIM of the L planes on the library convolution path"}
        del l_pred, l_target, l_gl, crit

    if not args.no_extras:
        try:
            extras["gcurlnet_train_512_b32" if world == 1 else "gcurlnet_ddp_train_512_b32"] = \
                net_training_step_extra(dev, world, local, timed)
        except Exception as e:   # an extra never fails the headline line (every rank reaches the barrier below)
            extras["gcurlnet_train_512_b32"] = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:200])}

    if rank == 0:
        peak, peak_src = measured_peak()
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):       # DRAM bytes per pixel measured by ncu (committed), scaled to this launch
            with open(tpath) as f:
                traffic = json.load(f)["bwd_bytes_per_px"] * n_px / 1e9
        bwd_gbs = n_px * BWD_BYTES_PER_PX / (bwd_ms / 1e3) / 1e9
        fwd_gbs = n_px * FWD_BYTES_PER_PX / (fwd_ms / 1e3) / 1e9
        step_gbs = n_px * (FWD_BYTES_PER_PX + BWD_BYTES_PER_PX) / ((fwd_ms + bwd_ms) / 1e3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD if (B, H, W) == (64, H4K, W4K) else
                       "synthetic RGB fp32 %dx%dx%d per GPU, curve-layer fwd+bwd" % (B, H, W),
                       "batch_per_gpu": B, "height": H, "width": W, "mask": "bool, ~10% masked",
                       "knots": "16 per curve, sigma 0.1", "image_gradient": True,
                       "l2": "inputs exceed L2 (%.1f GB per tensor), no flush needed" % (n_px * 12 / 1e9),
                       "parallelism": "batch-sharded, no collective" if world > 1 else "single GPU"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms, "steps": e2e_steps, "chunk_images": chunk,
                    "h2d_gbs_achieved": h2d / (e2e_ms / 1e3) / 1e9, "h2d_peak_gbs": h2d_peak_gbs,
                    "frac_of_h2d_peak": (h2d / (e2e_ms / 1e3) / 1e9) / h2d_peak_gbs,
                    "h2d_peak_what": "the same pinned chunks copied host->device with nothing else running, on all %d "
                                     "rank(s) at once; GB/s per rank, slowest rank" % world,
                    "what": "pinned host img+mask -> H2D (copy stream, chunked) -> CURLLayer fwd+bwd -> D2H reg + knot grads"},
            "e2e_u8": {"value": world * n_px / 1e6 / (e2e_u8_ms / 1e3), "unit": UNIT, "ms_per_step": e2e_u8_ms,
                       "h2d_bytes_per_step": h2d_u8, "d2h_bytes_per_step": d2h,
                       "h2d_gbs_achieved": h2d_u8 / (e2e_u8_ms / 1e3) / 1e9,
                       "what": "the same training step fed from 8-bit host data: pinned uint8 image + bool mask -> H2D "
                               "(4 B/px instead of 13) -> curl_u8_to_unit_f32 (device-side TF.to_tensor, exact) -> "
                               "CURLLayer fwd+bwd with image gradient -> D2H reg + knot grads"},
            "fwd_only": {"value": fwd_only_value, "unit": UNIT, "ms_per_step": fwd_only_ms,
                         "achieved_gbs_per_gpu": n_px * FWD_BYTES_PER_PX / (fwd_only_ms / 1e3) / 1e9,
                         "frac": n_px * FWD_BYTES_PER_PX / (fwd_only_ms / 1e3) / 1e9 / peak,
                         "u8_kernel_ms": fwd_u8_ms, "u8_kernel_value": world * n_px / 1e6 / (fwd_u8_ms / 1e3),
                         "e2e_f32": {"value": world * n_px / 1e6 / (infer_f32_ms / 1e3), "ms_per_step": infer_f32_ms,
                                     "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": n_px * 12,
                                     "what": "pinned host fp32 img + mask -> H2D -> forward -> fp32 output copied back "
                                             "(three streams, chunked)"},
                         "e2e_u8": {"value": world * n_px / 1e6 / (infer_u8_ms / 1e3), "ms_per_step": infer_u8_ms,
                                    "h2d_bytes_per_step": h2d_u8, "d2h_bytes_per_step": n_px * 3,
                                    "what": "pinned host uint8 img + mask -> H2D -> curl_layer_fwd_u8 -> uint8 output "
                                            "copied back (7 B/px over PCIe)"},
                         "what": "forward only (inference), whole batch per launch, resident inputs; max over ranks"},
            "gpu_launches": launches,
            "gpu_launches_other_legs": launches_legs,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": "bwd_tma_kernel<OpLayer> (fused recompute backward, 37 B/px)",
                         "achieved": bwd_gbs, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                         "frac": bwd_gbs / peak, "frac_of_8TBps": bwd_gbs / 8000.0, "traffic": traffic,
                         "traffic_unit": "GB per launch",
                         "traffic_source": "NOT measured in this run: the committed ncu constant of profiles/traffic.json "
                                           "(dram read+write bytes per pixel of one --set full capture) x pixels per launch",
                         "algorithmic_gb_per_launch": n_px * BWD_BYTES_PER_PX / 1e9,
                         "ms_per_launch": bwd_ms},
            "roofline_fwd": {"kernel": "fwd_kernel<OpLayer> (fused forward, 25 B/px)", "achieved": fwd_gbs,
                             "frac": fwd_gbs / peak, "ms_per_launch": fwd_ms},
            "roofline_bwd_no_gimg": {"kernel": "bwd_tma_kernel<OpLayer, .., false> (no image gradient: the GCURLNet "
                                               "case, 25 B/px)",
                                     "achieved": n_px * FWD_BYTES_PER_PX / (bwd_nog_ms / 1e3) / 1e9,
                                     "frac": n_px * FWD_BYTES_PER_PX / (bwd_nog_ms / 1e3) / 1e9 / peak,
                                     "ms_per_launch": bwd_nog_ms},
            "roofline_step": {"bytes_per_px": FWD_BYTES_PER_PX + BWD_BYTES_PER_PX, "achieved": step_gbs,
                              "frac": step_gbs / peak, "frac_of_8TBps": step_gbs / 8000.0},
        }
        if extras:
            line["extras"] = extras
        if world == 1 and not args.no_cpu:
            base, _ = time_cpu(3, 1)
            base["real_images"] = time_cpu_real_images()
            line["cpu_baseline"] = base
            try:
                del img, gout, h_img, d_img
                torch.cuda.empty_cache()
                torch.cuda.reset_peak_memory_stats(dev)
                line["gpu_eager_baseline"] = time_gpu_eager(dev)
            except Exception as e:   # reported baseline only: never fail the bench line for it
                line["gpu_eager_baseline"] = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:200])}
            if not args.no_compile:
                torch.cuda.empty_cache()
                line["gpu_compile_baseline"] = time_gpu_compile()
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--height", type=int, default=H4K)
    ap.add_argument("--width", type=int, default=W4K)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline and gpu_eager_baseline legs")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra (non-headline) measurements")
    ap.add_argument("--no-compile", action="store_true", help="skip the torch.compile baseline (a subprocess, ~1-2 min)")
    ap.add_argument("--e2e-chunk", type=int, default=8, help="images per host->device chunk in the e2e leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
