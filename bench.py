#!/usr/bin/env python
"""bench.py -- rays/s of the volume-rendering hot path on N B200s (no data-path collective: images are
independent).

    python bench.py --gpus N --steps K --warmup W                  # this repo's CUDA path, config 2
    python bench.py --config {1,2,3s,3c,4,5} ...                    # the other BASELINE.json configs
    python bench.py --impl reference --steps K --warmup W           # the reference's own CPU path

A step = one pass of the hot path over one batch of synthetic inputs of the chosen BASELINE.json config
(default 2: 16 images, 128x128 rays each, 64 coarse + 64 fine samples per ray, perspective cameras, render
forward + l1 loss (run.py:2236-2238) + backward to planes, attention values, cam2world and focal).
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

METRIC = "rays/sec (fwd+bwd)"
UNIT = "rays/s"
PLANE_BYTES = 3 * 32 * 256 * 256 * 4          # P, bytes per image
S = 64                                         # depth_samples_per_ray (run.py:512)

# BASELINE.json configs.  B = images per GPU ("weak") or in total over all ranks ("strong").
CONFIGS = {
    "1": dict(label="config1: p3d_car forward render, 1 image 64x64 rays, orthographic, deterministic depths",
              B=1, res=64, camera="ortho", scene_range=1.4, white=False, backward=False, randomize=False,
              grads=(), scaling="weak"),
    "2": dict(label="config2: p3d_car inversion render fwd+bwd, perspective, grads to planes/attention/cam/focal",
              B=16, res=128, camera="persp", scene_range=1.4, white=False, backward=True, randomize=True,
              grads=("planes", "palette", "cam", "focal"), scaling="weak"),
    "3s": dict(label="config3 (ShapeNet chairs): inversion fwd+bwd, batch 32 IN TOTAL sharded over the ranks, perspective, "
                     "scene_range 0.55, white background, one focal value for all images, no pose gradient "
                     "(data/loaders.py:120-123, 426-435)",
               B=32, res=128, camera="persp_shared_focal", scene_range=0.55, white=True, backward=True,
               randomize=True, grads=("planes", "palette"), scaling="strong"),
    "3c": dict(label="config3 (CUB): inversion fwd+bwd, batch 32 IN TOTAL sharded over the ranks, orthographic cameras with "
                     "T[3,3] != 1 (data/datasets.py:556-567), scene_range 2.0, grads to planes/attention/cam",
               B=32, res=128, camera="cub", scene_range=2.0, white=False, backward=True, randomize=True,
               grads=("planes", "palette", "cam"), scaling="strong"),
    "4": dict(label="config4: perspective eval path (eval_kitti_persp.py:706, 948-950, 221-228): 1 image, 0-dim focal, "
                    "principal point, z-buffer depth, scene_range 3.0, white background, fwd+bwd incl. pose",
              B=1, res=128, camera="persp_eval", scene_range=3.0, white=True, backward=True, randomize=True,
              grads=("planes", "palette", "cam", "focal"), zbuffer=True, scaling="weak"),
    "5": dict(label="config5: generator-training gradient set (planes, attention, W0, b0, W1, b1, beta, alpha; run.py:980-986), "
                    "4 images per GPU (batch 32 at 8 GPUs) + all-reduce of the generator's 32 164 056 and the discriminator's "
                    "28 921 344 fp32 gradients over NCCL when more than one rank runs",
              B=4, res=128, camera="persp", scene_range=1.4, white=False, backward=True, randomize=True,
              grads=("planes", "palette", "W0", "b0", "W1", "b1", "beta", "alpha"), scaling="weak",
              allreduce=(32164056, 28921344)),
}
WKEYS = ("W0", "b0", "W1", "b1", "beta", "alpha")


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return p, "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc is not None:
            self.proc.kill()          # the exact PID we started
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ncu --set full capture of one field_bwd launch at config 2 (dram__bytes_read.sum + dram__bytes_write.sum)
NCU_BWD_DRAM = {"bytes": None, "source": None}
try:
    with open(os.path.join(ROOT, "profiles", "r02_bwd_dram_bytes.json")) as _f:
        NCU_BWD_DRAM = json.load(_f)
except Exception:
    NCU_BWD_DRAM = {"bytes": 1204762000 + 596777728, "source": "profiles/r01z_field_ws_ncu_full_summary.txt"}


def make_inputs(cfg, B, seed_shift=0, res=None):
    """Synthetic inputs of one config on the CPU: smooth 256x256 planes, decoder calibrated so
    that the mask is non-trivial, random object-centred cameras of the config's model, l1 target.
    Returns a dict of CPU tensors: planes, palette, cam, focal (None / [B] / 0-dim), center (None / [B,2]),
    target, and the decoder dict `field`."""
    from nfi_b200 import synth
    H = W = res or cfg["res"]
    planes = synth.make_planes(B, res=256, seed=1 + seed_shift)
    fd = synth.make_field(planes, seed=3)
    pal = synth.make_palette(B, 10, seed=4 + seed_shift)
    kind = cfg["camera"]
    center = None
    if kind == "cub":
        cam, focal = synth.make_ortho_cub_cameras(B, seed=7 + seed_shift), None
    elif kind == "ortho":
        cam, focal = synth.make_cameras(B, seed=5 + seed_shift, perspective=False)
    else:
        cam, focal = synth.make_cameras(B, seed=2 + seed_shift, perspective=True)
        cam = cam.clone()
        cam[:, :3, 3] *= cfg["scene_range"] / 1.4            # keep the box at the same apparent size
        if kind == "persp_shared_focal":                      # data/loaders.py:426-435: one focal, expanded
            focal = focal[:1].expand(B).contiguous()
        elif kind == "persp_eval":                            # eval_kitti_persp.py:948-950
            focal = focal[0].clone()                          # 0-dim
            center = 0.5 + (torch.rand(B, 2, generator=torch.Generator().manual_seed(11 + seed_shift)) - 0.5) * 0.1
    target = torch.rand(B, H, W, 3, generator=torch.Generator().manual_seed(7 + seed_shift)) * 2 - 1
    return dict(planes=planes, palette=pal, cam=cam, focal=focal, center=center, target=target, field=fd)


def bind_to_gpu_numa_node(index):
    """Multi-rank runs: pin this process to the cores of the NUMA node its GPU hangs off, BEFORE the pinned host
    buffers are allocated (first touch then places them on that node).  torchrun starts every rank on the same
    cores by default; with 8 ranks copying 0.8 GB per step each way, host memory on the wrong socket halves the
    copy bandwidth per GPU (profiles/r01w_pcie_probe_8gpus_concurrent.txt).  Returns the node or None."""
    try:
        bdf = torch.cuda.get_device_properties(index).pci_bus_id if hasattr(torch.cuda.get_device_properties(index), "pci_bus_id") else None
        if bdf is None:
            q = subprocess.run(["nvidia-smi", "-i", str(index), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                               capture_output=True, text=True, timeout=20).stdout.strip()
            bdf = q
        bdf = bdf.lower()
        if len(bdf.split(":")[0]) == 8:                  # nvidia-smi prints an 8-digit PCI domain, sysfs uses 4
            bdf = bdf[4:]
        with open("/sys/bus/pci/devices/%s/numa_node" % bdf) as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            cpus = set()
            for part in f.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if allowed:
            os.sched_setaffinity(0, allowed)
            return node
    except Exception:
        pass
    return None


def auto_chunks(B):
    """Image chunks of the pipelined end-to-end leg: short first / last chunks (their copy-in / copy-out overlaps
    nothing), 4-image chunks between (sweep on one B200, ms per step at B=16: 8x2 23.5 | 1,1,4,4,4,1,1 22.6 |
    1,3,4,4,3,1 22.4).  Small batches go as one chunk."""
    if B < 8:
        return [B]
    mid, n_chunks = B - 2, [1]
    while mid > 0:
        n_chunks.append(min(4, mid) if mid not in (5, 6, 7) else 3)
        mid -= n_chunks[-1]
    left, right = [], []
    for i, v in enumerate(sorted(n_chunks[1:])):        # smaller chunks towards the two ends
        (left if i % 2 == 0 else right).append(v)
    return [1] + left + right[::-1] + [1]


# ---------------------------------------------------------------------------------------------
# this repo's arm
# ---------------------------------------------------------------------------------------------
def run_ours(args):
    import torch.distributed as dist
    from nfi_b200 import _lib, decoder, render
    from nfi_b200.dist import shard_range
    cfg = CONFIGS[args.config]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py (impl=ours) needs a CUDA device; there is no CPU fallback"
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa_node(local) if world > 1 and not args.no_numa_bind else None
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()
    B_cfg = args.batch or cfg["B"]
    if cfg["scaling"] == "strong":                 # fixed total batch, sharded by image (dist.shard_range)
        lo, hi = shard_range(B_cfg, rank, world)
        B, B_total = hi - lo, B_cfg
    else:
        B, B_total = B_cfg, B_cfg * world
    assert B > 0, "more ranks than images"
    H = W = args.res or cfg["res"]
    R = B * H * W
    inp = make_inputs(cfg, B, seed_shift=rank, res=H)
    fd = inp["field"]
    want = cfg["grads"]
    backward = cfg["backward"]
    pdt = torch.bfloat16 if args.plane_dtype == "bf16" else torch.float32
    wdev = {k: fd[k].to(dev).requires_grad_(k in want) for k in WKEYS}
    fld = decoder.Field(*[wdev[k] for k in WKEYS], cfg["scene_range"], 10, True)
    no_cam = backward and "cam" not in want
    names = [k for k in ("planes", "palette", "cam", "focal", "center", "target") if inp[k] is not None]
    leaf_in = [k for k in names if k in want]                      # per-step inputs that get a gradient
    leaf_w = [k for k in WKEYS if k in want]                       # resident parameters that get one (config 5)

    def make_step(check_hits, chunked):
        def step(*tensors):
            t = dict(zip(names, tensors))
            for k in leaf_in:
                t[k] = t[k].requires_grad_()
            nb = t["cam"].shape[0]
            rgb, depth, mask = render.render_planes(
                t["planes"], t["palette"], fld, H, W, t["cam"], t.get("focal"), t.get("center"), None, S,
                randomize=cfg["randomize"], white_background=cfg["white"], force_no_cam_grad=no_cam,
                zbuffer_depth=cfg.get("zbuffer", False), check_hits=check_hits, plane_dtype=pdt)
            if not backward:
                return (rgb.sum() * 0, rgb, depth, mask)
            diff = (rgb - t["target"]).abs()                       # run.py:2236-2238 (l1)
            loss = diff.sum() * (1.0 / (H * W * 3)) if chunked else diff.mean() * nb
            grads = torch.autograd.grad(loss, [t[k] for k in leaf_in] + [wdev[k] for k in leaf_w])
            return (loss, rgb, depth, mask, *grads)
        return step

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- config 5: the training step's gradient all-reduce (run.py:636-644, 980-986, 1064-1073) ------
    ar_bufs, ar_ms = [], []
    if cfg.get("allreduce") and world > 1:
        ar_bufs = [torch.zeros(n, dtype=torch.float32, device=dev) for n in cfg["allreduce"]]

    def allreduce_payloads(timed, outs=None):
        if world > 1 and leaf_w and outs is not None:
            # the decoder / beta / alpha gradients of this path: one flat bucket over NCCL (dist.allreduce_grads,
            # the sum nn.DataParallel performs at run.py:636-644)
            from nfi_b200.dist import allreduce_grads
            allreduce_grads(list(outs[-len(leaf_w):]), params=[wdev[k] for k in leaf_w])
        if not ar_bufs:
            return
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record()
        for b in ar_bufs:
            dist.all_reduce(b)
        ev[1].record()
        if timed:
            ar_ms.append(ev)

    # ---- `value`: inputs resident in HBM ------------------------------------------------------
    # Small batches (<= 4 images: configs 4, 5 and config 3 at 8 GPUs) are in the launch-latency regime -- a step is ~40
    # launches of 5-500 us -- so the step is captured once and replayed as a CUDA graph (nfi_b200.graph.GraphedStep, the
    # same front end the inversion loop would use: its shapes are static); large batches run eagerly.
    step = make_step(True, False)
    resident = [inp[k].to(dev) for k in names]
    torch.manual_seed(1234 + rank)
    n_warm = 1 if args.e2e_only else args.warmup
    res_mode = args.resident_mode if args.resident_mode != "auto" else ("graph" if B <= 4 else "eager")
    gstep = None
    if res_mode == "graph":
        from nfi_b200.graph import GraphedStep
        gstep = GraphedStep(make_step("defer", False), [t.detach() for t in resident], warmup=max(3, n_warm))
        run_step = lambda: gstep()
    else:
        run_step = lambda: step(*[t.detach() for t in resident])
    for _ in range(n_warm):
        allreduce_payloads(False, run_step())
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    timed = ("nfi_field_fwd", "nfi_field_bwd", "nfi_field_bwd_train")
    _lib.PROFILE.start(timed)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        allreduce_payloads(True, run_step())
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = _lib.PROFILE.launches
    if gstep is not None:
        # kernels inside a graph replay cannot be bracketed with events: the per-launch times of the roofline come
        # from three eager steps AFTER the timed region; the launch count is the captured one times the replays
        gstep.check_hits()
        launches = gstep.launches_per_replay * args.steps
        for _ in range(3):
            step(*[t.detach() for t in resident])
        torch.cuda.synchronize()
    _lib.PROFILE.stop()
    durs = _lib.PROFILE.durations_ms()
    clocks = sampler.finish() if rank == 0 else None
    ar_time = sum(a.elapsed_time(b) for a, b in ar_ms) / max(1, len(ar_ms))
    del resident

    # ---- `e2e`: host buffers in, host buffers out, through the public API ---------------------
    # pinned host inputs -> (chunks of images over copy-in / compute / copy-out streams, chunk step replayed as a
    # CUDA graph) -> pinned host outputs: loss, rgb, depth, mask and every gradient.  One pipeline, one code path,
    # at every GPU count; a CUDA fault here fails the run (no retry, no fallback).
    pin = lambda t: t.contiguous().pin_memory()
    host_in = [pin(inp[k]) for k in names]
    shapes_out = [(), (B, H, W, 3), (B, H, W), (B, H, W)]
    if backward:
        shapes_out += [tuple(inp[k].shape) for k in leaf_in] + [tuple(fd[k].shape) for k in leaf_w]
    inp = None
    import gc
    gc.collect()
    host_out = [pin(torch.empty(s)) for s in shapes_out]
    h2d = sum(t.numel() * 4 for t in host_in)
    d2h = sum(t.numel() * 4 for t in host_out)
    if args.e2e_chunks == "auto":
        sizes = auto_chunks(B) if not leaf_w else [B]           # shared-parameter gradients: one chunk
    else:
        sizes = [int(v) for v in args.e2e_chunks.split(",")]
        if len(sizes) == 1:
            sizes = [B // sizes[0]] * sizes[0]
        assert sum(sizes) == B, "--e2e-chunks does not partition the batch"
    chunk_step = make_step("defer", True)
    from nfi_b200.host_pipeline import GraphedHostStepPipeline, HostStepPipeline
    if args.e2e_mode == "graph":
        pipe = GraphedHostStepPipeline(chunk_step, B, sizes, dev, host_in, depth=args.e2e_depth)
    else:
        pipe = HostStepPipeline(chunk_step, B, sizes, dev)
    n_e2e = max(3, min(args.steps, 10))
    for _ in range(max(1, min(args.warmup, 3))):
        pipe.run(host_in, host_out)
        allreduce_payloads(False)
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(n_e2e):
        pipe.run(host_in, host_out)                         # returns when the host results are complete
        allreduce_payloads(False)
        _ = float(host_out[0])
    f1.record()
    barrier()
    e2e_ms = f0.elapsed_time(f1)
    if args.e2e_only:
        print(json.dumps({"e2e_ms": e2e_ms, "n_e2e": n_e2e, "sizes": sizes}))
        return

    # max over ranks
    if world > 1:
        t = torch.tensor([ms_total, e2e_ms, ar_time], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, e2e_ms, ar_time = float(t[0]), float(t[1]), float(t[2])
    if rank != 0:
        dist.destroy_process_group()
        return
    args.resident_mode = res_mode
    report(args, cfg, world, B, B_total, H, W, ms_total, e2e_ms, n_e2e, sizes, h2d, d2h, launches, durs, clocks,
           ar_time, dev, numa)
    if world > 1:
        dist.destroy_process_group()


def launches_per_step(times, args):
    """Timed launches of one entry point per step: they were recorded over `steps` eager steps, or over the three
    eager steps that follow a graph-replayed timed region."""
    n = 3 if args.resident_mode == "graph" else args.steps
    return len(times) / max(1, n)


def report(args, cfg, world, B, B_total, H, W, ms_total, e2e_ms, n_e2e, sizes, h2d, d2h, launches, durs, clocks,
           ar_time, dev, numa=None):
    R = B * H * W
    rays_job = (B_total if cfg["scaling"] == "strong" else B * world) * H * W
    ms_step = ms_total / args.steps
    value = rays_job / (ms_step * 1e-3)
    e2e_value = rays_job / (e2e_ms / n_e2e * 1e-3)
    pk, pk_src = peaks()
    samples = R * S
    bwd = durs.get("nfi_field_bwd", []) + durs.get("nfi_field_bwd_train", [])
    fwd = durs.get("nfi_field_fwd", [])
    bwd_ms = sum(bwd) / max(1, len(bwd))
    fwd_ms = sum(fwd) / max(1, len(fwd))
    pose = "cam" in cfg["grads"]
    if cfg["backward"]:
        # dominant kernel: the field backward (one launch = one 64-sample pass over every ray).  Algorithmic bytes of
        # ONE launch: planes read once + plane gradient written once + per sample (upstream float4 16
        # + depth 4 [+ d/d point 12 with pose gradients]) + rays 24 B
        train = "W0" in cfg["grads"]
        kernel, k_ms, k_all = ("field_bwd_ws_kernel" if not train else
                               "field_bwd_ws_kernel<STASH> + wgrad_reduce_kernel (one nfi_field_bwd_train call)"), bwd_ms, bwd
        a_bytes = 2 * B * PLANE_BYTES + samples * (16 + 4 + (12 if pose else 0)) + R * 24
        if train:      # + the per-sample rows F, H, dHpre, dY (704 B) written by the backward and read by the reduction
            a_bytes += samples * 2 * 705
    else:
        # forward only: planes read once + per sample (depth 4 in, float4 16 out) + rays 24 B
        kernel, k_ms, k_all = "field_fwd_ws_kernel", fwd_ms, fwd
        a_bytes = B * PLANE_BYTES + samples * 20 + R * 24
    ach = a_bytes / (k_ms * 1e-3) / 1e9 if k_ms > 0 else None
    dtype = "f32" if args.plane_dtype == "f32" else "f32 (bf16 plane reads)"
    out = {
        "metric": METRIC if cfg["backward"] else "rays/sec (fwd)", "value": value, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": cfg["scaling"], "vs_baseline": None, "dtype": dtype, "data": "synthetic",
        "config": {"workload": "%s; B=%d images/GPU (%d over %d GPU%s), %dx%d rays, 64+64 samples/ray"
                               % (cfg["label"], B, B_total, world, "" if world == 1 else "s", H, W),
                   "id": args.config, "rays_per_step_per_gpu": R,
                   "l2": ("inputs (planes %d MB + gradients) exceed the 126 MB L2; no flush" % (B * PLANE_BYTES >> 20))
                   if B * PLANE_BYTES * (2 if cfg["backward"] else 1) > 126 << 20 else
                   "inputs fit the 126 MB L2: a 256 MB buffer is not written between steps -- L2-warm number, stated",
                   "rng": "torch.rand uniforms drawn per step (parity mode)" if cfg["randomize"] else "deterministic depths",
                   "resident": "step replayed as a CUDA graph (launch-latency regime, <= 4 images)"
                   if args.resident_mode == "graph" else "eager launches",
                   "e2e": "pinned host buffers; chunks of %s images pipelined over copy-in / compute / copy-out streams%s"
                   % ("+".join(str(v) for v in sizes),
                      ", chunk step replayed as a CUDA graph" if args.e2e_mode == "graph" else "")},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": n_e2e, "ms_per_step": e2e_ms / n_e2e,
                **({"numa_node_rank0": numa} if numa is not None else {})},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": kernel, "achieved": ach, "peak": pk["hbm_gbs"],
                     "unit": "GB/s", "frac": (ach / pk["hbm_gbs"]) if ach else None,
                     # dram__bytes_read.sum + dram__bytes_write.sum of one launch at config 2, from the committed
                     # ncu --set full capture named in traffic_source
                     "traffic": NCU_BWD_DRAM["bytes"] if (args.config == "2" and B == 16 and H == 128) else None,
                     "traffic_source": NCU_BWD_DRAM["source"],
                     "peak_source": pk_src, "algorithmic_bytes_per_launch": a_bytes,
                     "launch_ms": k_ms, "launches_timed": len(k_all),
                     "share_of_step": (k_ms * launches_per_step(k_all, args)) / ms_step if ms_step else None,
                     "mlp_tflops_bwd_data": samples * 5504 / (bwd_ms * 1e-3) / 1e12 if bwd_ms > 0 else None,
                     # what ncu says actually limits this launch: not HBM, not the tensor pipe
                     "measured_limiter": "instruction streams and hand-offs of the gather / MLP-chain / scatter roles (5.8 ms "
                                         "with all memory traffic removed); the plane-gradient reductions alone need 4.3 ms "
                                         "at the measured L2 reduction cap (216 M red lines per launch at 49.6 G lines/s) "
                                         "-- DESIGN.md 4.2",
                     "field_fwd_launch_ms": fwd_ms,
                     "field_fwd_share_of_step": (fwd_ms * launches_per_step(fwd, args)) / ms_step if ms_step else None},
    }
    if cfg.get("allreduce"):
        nbytes = sum(cfg["allreduce"]) * 4
        out["allreduce"] = {"payload_bytes": nbytes, "ranks": world, "ms_per_step": ar_time if world > 1 else None,
                            "bus_gbs": (2 * (world - 1) / world * nbytes / (ar_time * 1e-3) / 1e9)
                            if (world > 1 and ar_time > 0) else None,
                            "note": "two flat fp32 buffers (generator 32 164 056, discriminator 28 921 344 elements) "
                                    "all-reduced with NCCL inside every timed step; single rank: no collective"}
    if not args.no_cpu_baseline and world == 1:
        try:                                  # reported baseline only: never fails the bench line
            out["gpu_eager_baseline"] = gpu_eager_baseline(dev, cfg, B=min(B, args.eager_batch), res=H)
        except Exception as e:                # noqa: BLE001
            out["gpu_eager_baseline"] = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:200])}
        torch.cuda.empty_cache()
        out["cpu_baseline"] = cpu_baseline(cfg, sample_res=min(args.cpu_res, H))
    elif world > 1:
        out["cpu_baseline"] = None
    print(json.dumps(out))


# ---------------------------------------------------------------------------------------------
# Reference arms: the reference's OWN code for the path -- `render` lifted from run.py / eval_kitti_persp.py,
# its `sampler` closure, `TriplanarDecoder.forward` and lib/nerf_utils.py, unmodified, from the files staged in
# git-ignored oracle/_ref/ (oracle/build_ref.py; /root/reference in the build container) -- on the same synthetic
# workload (the out-of-scope StyleGAN2 synthesis is replaced by a stub that returns the planes).  kind =
# "reference".  Only when those files are absent does the oracle port (oracle/nfi_oracle.py) stand in (kind "port").
# ---------------------------------------------------------------------------------------------
def reference_step_fn(cfg, B, res, device):
    """-> (step() -> loss value, rays per step, kind)"""
    import ref_loader
    H = W = res
    inp = make_inputs(cfg, B, res=res)
    fd = inp["field"]
    want = cfg["grads"]
    mv = lambda t: None if t is None else t.to(device)
    base = {k: mv(inp[k]) for k in ("planes", "palette", "cam", "focal", "center", "target")}
    no_cam = cfg["backward"] and "cam" not in want
    train_w = any(k in want for k in WKEYS)
    if ref_loader.available():
        import ref_model
        script = "eval_kitti_persp.py" if cfg.get("zbuffer") else "run.py"
        render = ref_model.reference_render_fn(script, scene_range=cfg["scene_range"], white_background=cfg["white"])
        G = ref_model.reference_generator(base["planes"], fd, cfg["scene_range"], trainable=train_w).to(device)
        wparams = {"W0": G.decoder.net[0].weight, "b0": G.decoder.net[0].bias, "W1": G.decoder.net[2].weight,
                   "b1": G.decoder.net[2].bias, "beta": G.beta, "alpha": G.alpha}
        ws = torch.zeros(B, 15, 512, device=device)

        def step():
            lv = {k: base[k].clone().requires_grad_() for k in ("planes", "palette", "cam", "focal")
                  if k in want and base[k] is not None}
            t = {**base, **lv}
            G.synthesis_network.planes = t["planes"]
            out = render(G, H, W, t["cam"], t["focal"], t["center"], None, ws, S, randomize=cfg["randomize"],
                         extra_model_inputs={"attention_values": t["palette"]}, force_no_cam_grad=no_cam)
            if not cfg["backward"]:
                return float(out[0].sum())
            loss = (out[0] - t["target"]).abs().mean() * B
            torch.autograd.grad(loss, list(lv.values()) + [wparams[k] for k in WKEYS if k in want])
            return float(loss.detach())
        return step, B * H * W, "reference"

    import nfi_oracle as O
    w = {k: fd[k].to(device).requires_grad_(k in want) for k in WKEYS}
    fp = O.FieldParams(w["W0"], w["b0"], w["W1"], w["b1"], cfg["scene_range"], True, 10, w["beta"], w["alpha"])
    gen = torch.Generator(device=device).manual_seed(1234)

    def step():
        lv = {k: base[k].clone().requires_grad_() for k in ("planes", "palette", "cam", "focal")
              if k in want and base[k] is not None}
        t = {**base, **lv}
        u_s = torch.rand(B, H, W, S, generator=gen, device=device)
        u_p = torch.rand(B * H * W, S, generator=gen, device=device)
        rgb, depth, mask = O.render(t["planes"].view(B, 3, 32, 256, 256), t["palette"], fp, H, W, t["cam"], t["focal"],
                                    t["center"], None, S, cfg["white"], u_s, u_p, cfg["randomize"], True, no_cam,
                                    cfg.get("zbuffer", False))
        if not cfg["backward"]:
            return float(rgb.sum())
        loss = (rgb - t["target"]).abs().mean() * B
        torch.autograd.grad(loss, list(lv.values()) + [w[k] for k in WKEYS if k in want])
        return float(loss.detach())
    return step, B * H * W, "port"


def _what(kind):
    return ("the reference's own PyTorch code (run.py::render, sampler closure, TriplanarDecoder, lib/nerf_utils.py; "
            "oracle/_ref)" if kind == "reference" else "oracle port of the reference PyTorch path")


def gpu_eager_baseline(dev, cfg, B, res, steps=3):
    """The reference's eager PyTorch path ON THE SAME GPU ("the kernel to beat"), TF32 off as
    run.py:59-60 sets it, at the repo arm's batch when it fits (bounded by --eager-batch)."""
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    torch.cuda.reset_peak_memory_stats(dev)
    step, rays, kind = reference_step_fn(cfg, B, res, dev)
    step()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / steps
    peak_gb = torch.cuda.max_memory_allocated(dev) / 2**30
    return {"value": rays / (ms * 1e-3), "unit": UNIT, "kind": kind, "ms_per_step": ms,
            "sample": "%s run eagerly on this GPU (TF32 off): %d image(s) %dx%d rays (64+64 samples), %d timed steps after 1 "
                      "warm-up; peak %.1f GiB allocated" % (_what(kind), B, res, res, steps, peak_gb)}


def cpu_baseline(cfg, sample_res=128, budget_s=20.0):
    """Bounded CPU sample on the box's host cores: B = 1 at sample_res x sample_res."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    step, rays, kind = reference_step_fn(cfg, 1, sample_res, "cpu")
    t0 = time.perf_counter()
    step()                                   # warm-up
    warm = time.perf_counter() - t0
    n, t0 = 0, time.perf_counter()
    while True:
        step()
        n += 1
        dt = time.perf_counter() - t0
        if dt + warm > budget_s or n >= 8:
            break
    return {"value": rays * n / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": "%s, 1 image %dx%d rays (64+64 samples) of this config, %d timed steps after 1 warm-up, torch CPU "
                      "with %d threads" % (_what(kind), sample_res, sample_res, n, cores)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = CONFIGS[args.config]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    res = min(args.cpu_res, args.res or cfg["res"])
    step, rays, kind = reference_step_fn(cfg, 1, res, "cpu")
    t0 = time.perf_counter()
    step()
    first = time.perf_counter() - t0
    if first * (args.steps + args.warmup) > 300.0 and res > 64:      # keep the whole run within a few minutes
        res = 64
        step, rays, kind = reference_step_fn(cfg, 1, res, "cpu")
    for _ in range(max(0, args.warmup - 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = rays * args.steps / dt
    sample = ("each step = bounded sample of this config: 1 image %dx%d rays (64+64 samples/ray), %s, %d torch threads"
              % (res, res, _what(kind), cores))
    print(json.dumps({
        "impl": "reference", "metric": METRIC if cfg["backward"] else "rays/sec (fwd)", "value": value, "unit": UNIT,
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": cfg["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": cfg["label"] + " (CPU arm: " + sample + ")", "id": args.config},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="2", choices=sorted(CONFIGS), help="BASELINE.json config (default 2)")
    ap.add_argument("--batch", type=int, default=0, help="override the config's batch (images per GPU, or in total "
                                                           "for the strong-scaling configs 3s / 3c)")
    ap.add_argument("--res", type=int, default=0, help="override the config's image size")
    ap.add_argument("--cpu-res", type=int, default=128, help="image size of the bounded CPU sample (BASELINE.md 4: 128)")
    ap.add_argument("--eager-batch", type=int, default=16, help="largest batch of the eager-PyTorch-on-GPU baseline")
    ap.add_argument("--e2e-chunks", default="auto",
                    help="image chunks of the pipelined end-to-end leg: a count, or chunk sizes like 1,1,4,4,4,1,1")
    ap.add_argument("--e2e-mode", choices=("graph", "eager"), default="graph",
                    help="chunk step of the end-to-end leg: CUDA-graph replay (default, at every GPU count) or eager launches")
    ap.add_argument("--e2e-depth", type=int, default=2, help="graph instances used round-robin (graph mode)")
    ap.add_argument("--resident-mode", choices=("auto", "eager", "graph"), default="auto",
                    help="device-resident leg: eager launches or CUDA-graph replay; auto = graph for <= 4 images per GPU")
    ap.add_argument("--plane-dtype", choices=("f32", "bf16"), default="f32",
                    help="bf16: the gather reads a bfloat16 copy of the planes (the 1e-2 mode); default f32 (1e-4)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-numa-bind", action="store_true", help="multi-rank runs: do not pin the rank to its GPU's NUMA node")
    ap.add_argument("--e2e-only", action="store_true", help=argparse.SUPPRESS)   # developer switch
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else max(args.warmup, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
