"""Developer probe: pinned host <-> device copy bandwidth on this box, one direction and both at once
(the e2e leg of bench.py moves 406 MB in and 408 MB out per step)."""
import torch

dev = torch.device("cuda", 0)
n = 406 * 1000 * 1000 // 4
h_in = torch.empty(n).pin_memory()
h_out = torch.empty(n).pin_memory()
d_in = torch.empty(n, device=dev)
d_out = torch.empty(n, device=dev)
s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    s1.synchronize(); s2.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def h2d():
    s1.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1)


def d2h():
    s2.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s2)


def both():
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)


gb = n * 4 / 1e9
for name, fn in (("h2d", h2d), ("d2h", d2h), ("both", both)):
    ms = timed(fn)
    print("%s: %.2f ms for %.0f MB each way -> %.1f GB/s per direction" % (name, ms, gb * 1e3, gb / ms * 1e3))
