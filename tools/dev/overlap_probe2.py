import torch, time
dev = torch.device("cuda", 0)
a = torch.randn(8192, 8192, device=dev); b = torch.randn(8192, 8192, device=dev)
big_d = torch.empty(100 * 1024 * 1024 // 4, device=dev); big_h = torch.empty(100 * 1024 * 1024 // 4).pin_memory()
side = torch.cuda.Stream()
def work(kind):
    if kind == "mm":
        for _ in range(4): torch.mm(a, b)
    else:
        for _ in range(40): a.add_(1.0)
def timed(kind, copy):
    for _ in range(2): work(kind)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(8):
        if copy == "d2h":
            with torch.cuda.stream(side): big_h.copy_(big_d, non_blocking=True)
        elif copy == "h2d":
            with torch.cuda.stream(side): big_d.copy_(big_h, non_blocking=True)
        work(kind)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 8
for kind in ("mm", "ew"):
    for c in ("none", "h2d", "d2h"):
        print(kind, c, "ms %.2f" % timed(kind, c))
