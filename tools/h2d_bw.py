"""H2D bandwidth from pinned host memory on this box: buffer left alone vs rewritten by the CPU right before each copy (dirty in the
CPU caches), the situation of the library's staging block."""
import torch, time, numpy as np
n = 15_360_000
x = torch.empty(n, dtype=torch.uint8).pin_memory()
src = torch.randint(0, 255, (n,), dtype=torch.uint8)
y = torch.empty_like(x, device="cuda")
def run(rewrite, threads):
    torch.set_num_threads(threads)
    ts = []
    for it in range(12):
        if rewrite: x.copy_(src)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); y.copy_(x, non_blocking=True); e1.record(); torch.cuda.synchronize()
        if it >= 2: ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    print(f"H2D 15.36 MB, rewritten={rewrite}, cpu threads={threads}: {ms:.3f} ms = {n/ms/1e6:.1f} GB/s")
run(False, 1); run(True, 1); run(True, 8)
