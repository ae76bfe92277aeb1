#!/usr/bin/env python
"""Benchmark of the GenSeg hot path on B200: U-Net (1 -> 2 classes, bilinear) 256x256 bf16 training step
(forward + CE+Dice loss + backward + RMSprop), global batch 64, data-parallel over N GPUs.

    python bench.py --gpus N --steps K --warmup W            # our arm (one JSON line on rank 0)
    python bench.py --impl reference --steps K --warmup W    # the CPU oracle port on the host cores
    python bench.py --workload pix2pix|genseg|eval ...       # BASELINE configs 3 / 4 / 5 (same JSON contract, own metric)

See the module docstrings of semantic_segmentation_b200 and DESIGN.md for what is timed.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

GLOBAL_BATCH = 64
H = W = 256
N_CLASSES = 2
# algorithmic FLOPs per image: U-Net(1,2,bilinear) fwd 79.818 GF, training step 239.38 GF
TRAIN_GF_PER_IMG = 239.38
METRIC = "U-Net train images/sec @256^2 (1->2 classes, bilinear, bf16, global batch 64)"


def synthetic_batch(n, seed=4321):
    """JSRT-shaped synthetic batch: image in [0,1), mask = two filled ellipses (~27% foreground)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(n, 1, H, W, generator=g)
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    m = (((xx - 80) / 48) ** 2 + ((yy - 128) / 90) ** 2 < 1) | (((xx - 176) / 48) ** 2 + ((yy - 128) / 90) ** 2 < 1)
    y = m.long()[None].repeat(n, 1, 1).contiguous()
    return x, y


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [t.strip() for t in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops", 1590.0), d.get("bf16_tflops_sustained", 1400.0), d.get("hbm_gbs", 6650.0), "measured"
    return 1590.0, 1400.0, 6650.0, "fallback"


def cpu_oracle_step_time(batch: int, steps: int, warmup: int):
    """Time the CPU oracle (port of the reference's path) on a bounded sample: fwd + CE+Dice + bwd."""
    from oracle import genseg_oracle as O
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    st = O.seeded_state(O.unet_state_spec(1, N_CLASSES, True), 1234)
    params = {k: v.clone().requires_grad_(True) for k, v in st.items() if v.dtype.is_floating_point and "running" not in k}
    work = dict(st)
    work.update(params)
    x, y = synthetic_batch(batch)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        logits = O.unet_forward(work, x, True, training=True, update_running=True)
        loss, _, _ = O.seg_loss(logits, y)
        grads = torch.autograd.grad(loss, list(params.values()))
        del grads
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    times.sort()
    return times[len(times) // 2], threads


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = 4
    steps = max(1, min(args.steps, 5))
    warm = max(1, min(args.warmup, 2))
    t, threads = cpu_oracle_step_time(batch, steps, warm)
    v = batch / t
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "U-Net(1,2,bilinear) 256x256 fwd+CE+Dice+bwd, CPU oracle port of the reference path",
                       "global_batch": GLOBAL_BATCH, "parallelism": "host cores"},
            "cpu_baseline": {"value": v, "unit": "images/s", "cores": threads, "kind": "port",
                             "sample": "batch %d x %d steps of the 64-image workload (median)" % (batch, steps)},
            "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def time_kernel(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e-3


def _conv_pass_tflops(b, cin, cout, hw):
    """fprop / dgrad / wgrad of one 3x3 layer timed alone with CUDA events on the launching stream (TFLOP/s each)."""
    from semantic_segmentation_b200 import ops
    from semantic_segmentation_b200._lib import MODE_CONV, MODE_CONVT, view_nhwc
    dev = torch.device("cuda")
    x = torch.randn(b, hw, hw, cin, device=dev, dtype=torch.bfloat16)
    w = torch.randn(cout, cin, 3, 3, device=dev) * 0.03
    wp, wpb = ops.pack_weight(w, True, torch.bfloat16), ops.pack_weight(w, False, torch.bfloat16)
    y = torch.empty(b, hw, hw, cout, device=dev, dtype=torch.bfloat16)
    dx = torch.empty_like(x)
    xv, yv, dxv = view_nhwc(x), view_nhwc(y), view_nhwc(dx)
    nsl = ops.conv_wgrad_plan(xv, yv, MODE_CONV, 3, 1, 1, use_tc=1)     # partial-sum slices (no atomics), see DESIGN.md
    dws = torch.empty(nsl, 9, cout, cin, device=dev, dtype=torch.float32)
    flops = 2.0 * b * hw * hw * 9 * cin * cout
    out = {}
    out["fprop"] = flops / time_kernel(lambda: ops.conv_gather(xv, wp, None, yv, MODE_CONV, 3, 1, 1, use_tc=1, want_stats=True))
    out["dgrad"] = flops / time_kernel(lambda: ops.conv_gather(yv, wpb, None, dxv, MODE_CONVT, 3, 1, 1, use_tc=1))
    out["wgrad"] = flops / time_kernel(lambda: ops.conv_wgrad_slices(xv, yv, dws, nsl, MODE_CONV, 3, 1, 1, use_tc=1))
    return {k: v / 1e12 for k, v in out.items()}, flops


def dominant_kernel_roofline(per_gpu_batch: int):
    """The kernel with the largest share of the step (profiles/*launch_summary*) is conv_tc_halo_kernel<256,2>; its
    heaviest launch is up2.conv[0] (512 -> 256 ch, 3x3, 64x64; 9.66 GF per image and pass).  The 64- and 128-column
    variants are reported beside it on up4.conv[0] (128 -> 64 @256^2) and its data gradient (64 -> 128)."""
    main, flops = _conv_pass_tflops(per_gpu_batch, 512, 256, 64)
    small, _ = _conv_pass_tflops(per_gpu_batch, 128, 64, 256)
    return main, small, flops


def committed_traffic(per_gpu_batch: int):
    """dram__bytes_read.sum + dram__bytes_write.sum of the roofline kernel from the committed `ncu --set full` capture
    (profiles/roofline_traffic.json), valid for the batch it was captured at."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        d = json.load(open(p))
        return d["dram_bytes_per_launch"] if int(d.get("per_gpu_batch", -1)) == per_gpu_batch else None
    except Exception:
        return None


def run_secondary(args):
    """BASELINE configs 3 (pix2pix), 4 (genseg), 5 (eval) through the same contract: W warm-up steps, K timed steps between
    barriers, CUDA events, max over ranks, clocks sampled during the timed region, end-to-end leg with the host -> device
    copy of every step's inputs and a device -> host read of its result.  Weak scaling: 8 images per GPU."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_workloads as BW
    from semantic_segmentation_b200 import _lib, parallel
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        parallel.init_data_parallel(sync_bn=not args.no_sync_bn, overlap_grad_allreduce=not args.no_overlap,
                                    grad_comm_dtype=args.grad_comm_dtype)
    rank = parallel.rank()
    lib = _lib.load()
    per = args.per_gpu_batch or 8
    lines = []
    cases = [(64, 256), (16, 512), (4, 1024)] if args.workload == "eval" else [None]
    for case in cases:
        if args.workload == "pix2pix":
            wl = BW.pix2pix(dev, per, rank, world)
        elif args.workload == "genseg":
            wl = BW.genseg(dev, per, rank, world)
        else:
            wl = BW.evaluate(dev, case[0], case[1])
        host = wl["host_batch"]
        static = tuple(t.to(dev) for t in host)

        def barrier():
            if world > 1:
                torch.distributed.barrier()
            torch.cuda.synchronize()

        def timed(fn, steps):
            barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(steps):
                fn()
            e.record()
            barrier()
            t = torch.tensor([s.elapsed_time(e) * 1e-3], device=dev, dtype=torch.float64)
            if world > 1:
                torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            return float(t.item())

        for _ in range(max(args.warmup, 3)):
            wl["step"](static)
        # Workloads whose step holds its per-step state on the device (capturable Adam, device-side dropout call count) are
        # captured into ONE CUDA graph and replayed: the eager Pix2Pix step issues ~425 launches through ctypes and is
        # host-bound on a slow host (13.0 against 8.4 ms of kernels, profiles/r02r_run.log).
        graph, static_out, launches_per_step = None, None, None
        if wl.get("graphable") and not args.no_graph:
            try:
                barrier()
                graph = torch.cuda.CUDAGraph()
                c0 = lib.gsb_launch_count()
                with torch.cuda.graph(graph):
                    static_out = wl["step"](static)
                launches_per_step = int(lib.gsb_launch_count() - c0)
                graph.replay()
                torch.cuda.synchronize()
            except Exception as exc:  # pragma: no cover - reported on stderr, the run goes on eagerly
                graph = None
                sys.stderr.write('bench.py: CUDA graph capture failed, running eagerly: %r\n' % (exc,))
                torch.cuda.synchronize()

        def run_step():
            if graph is not None:
                graph.replay()
                return static_out
            return wl["step"](static)

        for _ in range(3):
            run_step()
        sampler = ClockSampler(local_rank)
        if rank == 0:
            sampler.start()
        l0 = lib.gsb_launch_count()
        torch.cuda.profiler.start()
        t_dev = timed(run_step, args.steps)
        torch.cuda.profiler.stop()
        launches = launches_per_step * args.steps if graph is not None else int(lib.gsb_launch_count() - l0)
        clocks = sampler.stop() if rank == 0 else None

        def e2e_step():
            for d, h in zip(static, host):
                d.copy_(h, non_blocking=True)             # this step's inputs from pinned host memory
            return float(run_step().item())               # and its result back to the host

        e2e_step()
        t_e2e = timed(e2e_step, args.steps)
        units = wl["units"] * world * args.steps
        if rank != 0:
            continue
        burst, sustained, hbm, src = peaks()
        line = {"metric": wl["metric"], "value": units / t_dev, "unit": "images/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": t_dev / args.steps * 1e3, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": wl["workload"], "global_batch": wl["units"] * world, "per_gpu_batch": wl["units"],
                           "parallelism": "dp%d" % world if args.workload != "eval" else "replicas x%d" % world,
                           "sync_bn": bool(world > 1 and not args.no_sync_bn and args.workload != "eval"),
                           "grad_comm_dtype": args.grad_comm_dtype if world > 1 and args.workload != "eval" else None,
                           "cuda_graph": graph is not None, "l2": "per-step working set exceeds the 126 MB L2; no flush needed"},
                "e2e": {"value": units / t_e2e, "unit": "images/s",
                        "h2d_bytes_per_step": int(sum(t.numel() * t.element_size() for t in host)) * world, "d2h_bytes_per_step": 4 * world},
                "gpu_launches": launches, "clocks": clocks}
        if wl.get("flops_per_unit"):
            tf = wl["flops_per_unit"] * wl["units"] * args.steps / t_dev / 1e12
            line["roofline"] = {"bound": "tensor" if args.workload == "eval" else "hbm+tensor (weight streaming at the deep levels)",
                                "achieved": tf, "peak": sustained, "unit": "TFLOP/s", "frac": tf / sustained, "traffic": None,
                                "peak_source": src + " (sustained: whole step)", "flops_note": wl["flops_note"],
                                "weight_bytes_note": wl.get("weight_bytes_note")}
        lines.append(line)
        del wl, static
        torch.cuda.empty_cache()
    for line in lines:
        print(json.dumps(line), flush=True)


def run_ours(args):
    import semantic_segmentation_b200 as S
    from semantic_segmentation_b200 import _lib, parallel
    from semantic_segmentation_b200.optim import RMSprop
    from semantic_segmentation_b200.unet import UNet
    from semantic_segmentation_b200.util.dice_score import seg_loss

    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        parallel.init_data_parallel(sync_bn=not args.no_sync_bn, overlap_grad_allreduce=not args.no_overlap,
                                    grad_comm_dtype=args.grad_comm_dtype)
    rank = parallel.rank()
    lib = _lib.load()
    S.set_precision("bf16")
    if GLOBAL_BATCH % world != 0:
        raise SystemExit("global batch 64 must be divisible by the number of GPUs")
    per = args.per_gpu_batch or GLOBAL_BATCH // world
    torch.manual_seed(1234)              # the reference's default initialisation (SURVEY 8c Protocol A), same on every rank
    net = UNet(1, N_CLASSES, True)
    net = net.to(dev).train()
    params = list(net.parameters())
    opt = RMSprop(params, lr=1e-4, weight_decay=1e-8, momentum=0.9)
    # seg_loss is a mean over the GLOBAL batch (CE count and Dice sums are all-reduced), so the SUM of the ranks' gradients
    # is already the full-batch gradient: no 1/world here (that is for per-rank-mean losses such as GANLoss / L1)
    opt.grad_scale = 1.0
    xg, yg = synthetic_batch(per * world)
    lo, hi = rank * per, (rank + 1) * per
    x_host, y_host = xg[lo:hi].contiguous().pin_memory(), yg[lo:hi].contiguous().pin_memory()
    x_dev, y_dev = x_host.to(dev), y_host.to(dev)

    def step(x, y):
        opt.zero_grad(set_to_none=True)
        loss = seg_loss(net(x), y)
        loss.backward()
        parallel.allreduce_gradients(params)
        opt.step()
        return loss

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(steps):
            fn()
        e.record()
        barrier()
        t = torch.tensor([s.elapsed_time(e) * 1e-3], device=dev, dtype=torch.float64)
        if world > 1:
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return float(t.item())

    for _ in range(max(args.warmup, 3)):
        step(x_dev, y_dev)
    # The whole step (forward, loss, backward, NCCL all-reduces, RMSprop) is captured once into a CUDA graph and
    # replayed: at 8 images per GPU the ~330 kernel launches per step are otherwise host-bound.
    static_x, static_y = x_dev.clone(), y_dev.clone()
    graph, static_loss, launches_per_step = None, None, None
    if not args.no_graph:
        try:
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                step(static_x, static_y)
            torch.cuda.current_stream(dev).wait_stream(side)
            barrier()
            graph = torch.cuda.CUDAGraph()
            c0 = lib.gsb_launch_count()
            with torch.cuda.graph(graph):
                static_loss = step(static_x, static_y)
            launches_per_step = int(lib.gsb_launch_count() - c0)
            graph.replay()
            torch.cuda.synchronize()
        except Exception as exc:  # pragma: no cover - reported in the JSON line
            graph = None
            sys.stderr.write('bench.py: CUDA graph capture failed, running eagerly: %r\n' % (exc,))
            torch.cuda.synchronize()

    def run_step():
        if graph is not None:
            graph.replay()
            return static_loss
        return step(static_x, static_y)

    for _ in range(3):
        run_step()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = lib.gsb_launch_count()
    torch.cuda.profiler.start()      # ncu --profile-from-start off captures exactly the timed steps
    t_dev = timed(run_step, args.steps)
    torch.cuda.profiler.stop()
    launches = launches_per_step * args.steps if graph is not None else int(lib.gsb_launch_count() - l0)
    clocks = sampler.stop() if rank == 0 else None

    # End to end: every step's inputs come from pinned host memory and its loss is read back to the host.  The copy of
    # step i+1 is enqueued on a copy stream into the other staging buffer BEFORE the host blocks on the loss of step i
    # (what a prefetching input pipeline does), so it runs under step i's kernels; the first copy is exposed.  Exactly
    # `steps` host->device copies and `steps` device->host reads happen inside the timed region.
    copy_stream = torch.cuda.Stream(device=dev)
    stage = [(torch.empty_like(static_x), torch.empty_like(static_y)) for _ in range(2)]
    h2d_done = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    state = {"i": 0, "total": 0}

    def enqueue_h2d(slot):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[slot])          # the staging buffer has been read by its previous user
            stage[slot][0].copy_(x_host, non_blocking=True)
            stage[slot][1].copy_(y_host, non_blocking=True)
            h2d_done[slot].record(copy_stream)

    def e2e_reset(total):
        state["i"], state["total"] = 0, total
        cur = torch.cuda.current_stream(dev)
        for ev in consumed:
            ev.record(cur)

    def e2e_step():
        i, cur = state["i"], torch.cuda.current_stream(dev)
        slot = i & 1
        if i == 0:
            enqueue_h2d(slot)                                # nothing to hide the first copy behind
        cur.wait_event(h2d_done[slot])
        static_x.copy_(stage[slot][0], non_blocking=True)    # staging -> the graph's static inputs (device copy)
        static_y.copy_(stage[slot][1], non_blocking=True)
        consumed[slot].record(cur)
        loss = run_step()
        if i + 1 < state["total"]:
            enqueue_h2d(slot ^ 1)                            # next step's inputs, under this step's kernels
        state["i"] = i + 1
        return float(loss.item())                            # device -> host read of the step's loss

    e2e_reset(2)
    e2e_step(); e2e_step()
    e2e_reset(args.steps)
    t_e2e = timed(e2e_step, args.steps)
    images = per * world * args.steps
    value, e2e = images / t_dev, images / t_e2e
    if rank != 0:
        return
    burst, sustained, hbm, src = peaks()
    kern, small, flops = dominant_kernel_roofline(per)
    step_tf = TRAIN_GF_PER_IMG * 1e9 * per * args.steps / t_dev / 1e12
    line = {"metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": t_dev / args.steps * 1e3, "higher_is_better": True,
            "scaling": "strong" if not args.per_gpu_batch else "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": "U-Net(n_channels=1,n_classes=2,bilinear) 256x256 train step: fwd + CE+Dice + bwd + RMSprop",
                       "global_batch": per * world, "per_gpu_batch": per, "parallelism": "dp%d" % world,
                       "sync_bn": bool(world > 1 and not args.no_sync_bn), "cuda_graph": graph is not None, "e2e_inputs": "pinned host -> device every step, double-buffered prefetch on a copy stream",
                       "grad_allreduce": ("overlapped_in_backward" if not args.no_overlap else "after_backward") if world > 1 else None,
                       "grad_comm_dtype": args.grad_comm_dtype if world > 1 else None,
                       "l2": "per-step working set (activations+gradients, >1 GB per GPU) exceeds the 126 MB L2; no flush needed"},
            "e2e": {"value": e2e, "unit": "images/s", "h2d_bytes_per_step": int(x_host.numel() * 4 + y_host.numel() * 8) * world,
                    "d2h_bytes_per_step": 4 * world},
            "gpu_launches": launches, "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": "conv_tc_halo_kernel<256,2> fprop (up2.conv[0] 512->256 3x3 @64^2)",
                         "achieved": kern["fprop"], "peak": burst, "unit": "TFLOP/s", "frac": kern["fprop"] / burst,
                         "traffic": committed_traffic(per), "peak_source": src + " (burst: kernel timed alone)",
                         "flops_per_launch": flops,
                         "other_kernels_tflops": {"up2.conv0 dgrad<256>": kern["dgrad"], "up2.conv0 wgrad<128>": kern["wgrad"],
                                                  "up4.conv0 fprop<64>": small["fprop"], "up4.conv0 dgrad<128>": small["dgrad"],
                                                  "up4.conv0 wgrad<64>": small["wgrad"]},
                         "whole_step_tflops_per_gpu": step_tf, "whole_step_frac_of_sustained": step_tf / sustained}}
    if not args.no_hbm_roofline:
        # the non-convolution kernels against the measured copy bandwidth (target >= 0.70), each timed alone
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_hbm
        del static_x, static_y
        torch.cuda.empty_cache()
        rh = bench_hbm.hbm_rooflines(per, hbm)
        line["roofline_hbm"] = {"peak": hbm, "unit": "GB/s", "peak_source": src, "per_gpu_batch": per,
                                "kernels": {k: {"gbs": round(v["gbs"], 1), "frac": round(v["frac"], 3), "us": round(v["us"], 1),
                                                "bytes": v["bytes"], "shape": v["shape"]} for k, v in rh.items()}}
    if world == 1 and not args.no_cpu_baseline:
        t, threads = cpu_oracle_step_time(4, 3, 1)
        line["cpu_baseline"] = {"value": 4 / t, "unit": "images/s", "cores": threads, "kind": "port",
                                "sample": "batch 4 x 3 steps (median) of the same U-Net step on the host cores"}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--per-gpu-batch", type=int, default=0, help="override 64/N (weak scaling experiments)")
    ap.add_argument("--no-sync-bn", action="store_true")
    ap.add_argument("--no-overlap", action="store_true", help="gradient all-reduce after backward instead of inside it")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel eagerly instead of replaying a CUDA graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-hbm-roofline", action="store_true", help="skip the per-kernel HBM roofline block")
    ap.add_argument("--workload", default="unet", choices=["unet", "pix2pix", "genseg", "eval"],
                    help="unet = the headline (BASELINE config 2); pix2pix / genseg / eval = configs 3 / 4 / 5")
    ap.add_argument("--grad-comm-dtype", default="bf16", choices=["fp32", "bf16"],
                    help="dtype of the gradient buckets on the wire (N > 1); fp32 makes N ranks equal 1 rank to fp32 rounding")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback (use --impl reference)")
    if args.workload != "unet":
        run_secondary(args)
        return
    run_ours(args)


if __name__ == "__main__":
    main()
