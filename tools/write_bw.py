"""Development aid: sustained write bandwidth of plain fills / copies on this GPU, for the roofline discussion."""
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

for mb, nbuf in ((62.72, 4), (62.72, 1), (250.88, 4), (1024, 2)):
    n = int(mb * 1e6 / 4)
    bufs = [torch.empty(n, dtype=torch.int32, device=dev) for _ in range(nbuf)]
    src = torch.empty(n, dtype=torch.int32, device=dev)
    t = timeit(lambda i: bufs[i % nbuf].fill_(7))
    t2 = timeit(lambda i: bufs[i % nbuf].copy_(src))
    t3 = timeit(lambda i: torch.cuda.current_stream().synchronize() or bufs[i % nbuf].zero_())
    print(f"{mb:8.2f} MB x{nbuf}: fill {t*1e6:7.1f} us = {mb/1e3/t:7.1f} GB/s written | copy {t2*1e6:7.1f} us = {2*mb/1e3/t2:7.1f} GB/s (r+w)")
