import os, sys, types, torch
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests")); sys.path.insert(0, os.path.join(os.getcwd(), "oracle"))
import test_gpu_parity as T
import nfi_b200
from nfi_b200 import ops, decoder, render, synth, nerf_utils
nfi = types.SimpleNamespace(ops=ops, dec=decoder, render=render, nu=nerf_utils, synth=synth)
for seed in (0, 1, 2, 3):
    for (B, HW) in ((1, 32), (1, 64)):
        e, flips, n = T._oracle_vs_cuda(nfi, B, HW, HW, 64, 64, True, False, False, False, seed=seed)
        print(seed, B, HW, "flips", flips, {k: "%.1e" % v for k, v in e.items() if k in ("d_cam", "d_focal")})
