import torch, time
x = torch.empty(64*480*752, dtype=torch.uint8).pin_memory()
d = torch.empty_like(x, device='cuda')
y = torch.empty(4*1024*1024, dtype=torch.uint8).pin_memory(); dy = torch.empty_like(y, device='cuda')
for name, src, dst in (("h2d 23MB", x, d), ("d2h 23MB", d, x), ("h2d 4MB", y, dy), ("d2h 4MB", dy, y)):
    for _ in range(3): dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): dst.copy_(src, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(name, round(ms, 3), "ms", round(src.numel() / ms / 1e6, 1), "GB/s")
# bidirectional
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(20):
    with torch.cuda.stream(s1): d.copy_(x, non_blocking=True)
    with torch.cuda.stream(s2): y.copy_(dy, non_blocking=True)
torch.cuda.synchronize(); ms = (time.perf_counter() - t) / 20 * 1e3
print("bidir 23MB in + 4MB out", round(ms, 3), "ms")
