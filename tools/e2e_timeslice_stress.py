"""Stress run behind DESIGN.md 5.1: TWO processes time-slicing one GPU (all three round-1 sightings of the launch failure
had a second CUDA context alive), each looping the graphed end-to-end pipeline of bench.py -- one on config 2 (almost no idle
tiles), one on config 3s (small scene box: many idle tiles) -- for SECONDS seconds, checking every step's host outputs.

    python tools/e2e_timeslice_stress.py [SECONDS]        # parent: starts the two workers, prints their summaries
"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(config, batch, seconds):
    import torch
    import bench
    from nfi_b200 import decoder, render
    from nfi_b200.host_pipeline import GraphedHostStepPipeline
    cfg = bench.CONFIGS[config]
    dev = torch.device("cuda", 0)
    B, H, W, S = batch, cfg["res"], cfg["res"], 64
    inp = bench.make_inputs(cfg, B)
    fd = inp["field"]
    fld = decoder.Field(*[fd[k].to(dev) for k in bench.WKEYS], cfg["scene_range"], 10, True)
    names = [k for k in ("planes", "palette", "cam", "focal", "center", "target") if inp[k] is not None]
    leaf = [k for k in names if k in cfg["grads"]]
    no_cam = "cam" not in cfg["grads"]

    def chunk_step(*tensors):
        t = dict(zip(names, tensors))
        for k in leaf:
            t[k] = t[k].requires_grad_()
        rgb, depth, mask = render.render_planes(t["planes"], t["palette"], fld, H, W, t["cam"], t.get("focal"), t.get("center"),
                                                None, S, randomize=True, white_background=cfg["white"],
                                                force_no_cam_grad=no_cam, check_hits="defer")
        loss = (rgb - t["target"]).abs().sum() * (1.0 / (H * W * 3))
        return (loss, rgb, depth, mask, *torch.autograd.grad(loss, [t[k] for k in leaf]))

    pin = lambda x: x.contiguous().pin_memory()
    host_in = [pin(inp[k]) for k in names]
    host_out = [pin(torch.empty(s)) for s in [(), (B, H, W, 3), (B, H, W), (B, H, W)] + [tuple(inp[k].shape) for k in leaf]]
    pipe = GraphedHostStepPipeline(chunk_step, B, bench.auto_chunks(B), dev, host_in)
    n, first, t0 = 0, None, time.time()
    while time.time() - t0 < seconds:
        pipe.run(host_in, host_out)
        assert all(bool(torch.isfinite(x).all()) for x in host_out), "non-finite output in step %d" % n
        loss = float(host_out[0])
        first = loss if first is None else first
        assert abs(loss - first) <= 0.05 * abs(first), "step %d: loss %.6f far from the first %.6f" % (n, loss, first)
        n += 1
    print("config %s B=%d: %d end-to-end steps (%d chunk graphs replayed) in %.0f s, loss %.5f, no fault"
          % (config, B, n, n * len(pipe.sizes), time.time() - t0, first))


if __name__ == "__main__":
    if len(sys.argv) >= 4 and sys.argv[1] == "--worker":
        worker(sys.argv[2], int(sys.argv[3]), float(sys.argv[4]))
    else:
        seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--worker", c, str(b), str(seconds)],
                                  stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
                 for c, b in (("2", 16), ("3s", 16))]
        rc = 0
        for p in procs:
            out = p.communicate(timeout=seconds + 600)[0]
            print(out.strip().splitlines()[-1] if out.strip() else "(no output)")
            if p.returncode:
                print(out[-3000:])
            rc |= p.returncode
        sys.exit(rc)
