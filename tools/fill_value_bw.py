import torch
dev = torch.device("cuda", 0)
def timeit(fn, n=50):
    for _ in range(5): fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e-3
n = int(62.72e6/4)
bufs = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(4)]
for v in (0, -1, 7, 0, -1, 0x55555555):
    t = timeit(lambda i: bufs[i % 4].fill_(v))
    print(f"fill {v:#x}: {t*1e6:.1f} us  {62.72e-3/t:.0f} GB/s")
