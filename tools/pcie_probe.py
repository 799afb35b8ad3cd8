"""D2H / H2D bandwidth of pinned copies on this box: one stream vs two concurrent streams, and by chunk size."""
import torch, time
dev = torch.device("cuda:0")
N = 1 << 30
src = torch.empty(2 * N, dtype=torch.uint8, device=dev)
dst = torch.empty(2 * N, dtype=torch.uint8).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); t = time.perf_counter(); fn(); torch.cuda.synchronize(); best = min(best, time.perf_counter() - t)
    return best
def one():
    with torch.cuda.stream(s1): dst.copy_(src, non_blocking=True)
def two():
    with torch.cuda.stream(s1): dst[:N].copy_(src[:N], non_blocking=True)
    with torch.cuda.stream(s2): dst[N:].copy_(src[N:], non_blocking=True)
def chunks(sz):
    def f():
        with torch.cuda.stream(s1):
            for o in range(0, 2 * N, sz): dst[o:o + sz].copy_(src[o:o + sz], non_blocking=True)
    return f
def h2d():
    with torch.cuda.stream(s1): src.copy_(dst, non_blocking=True)
def both():
    with torch.cuda.stream(s1): dst[:N].copy_(src[:N], non_blocking=True)
    with torch.cuda.stream(s2): src[N:].copy_(dst[N:], non_blocking=True)
print("d2h 1 stream  GB/s", 2 * N / run(one) / 1e9)
print("d2h 2 streams GB/s", 2 * N / run(two) / 1e9)
for sz in (1 << 20, 16 << 20, 256 << 20): print("d2h chunks", sz >> 20, "MiB GB/s", 2 * N / run(chunks(sz)) / 1e9)
print("h2d 1 stream  GB/s", 2 * N / run(h2d) / 1e9)
print("d2h+h2d concurrent, each 1 GiB: GB/s per direction", N / run(both) / 1e9)
