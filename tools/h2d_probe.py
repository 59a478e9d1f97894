import torch, time
n = 64*49104*90
h = torch.empty(n, dtype=torch.float32).pin_memory()
d = torch.empty(n, dtype=torch.float32, device="cuda")
s2 = torch.cuda.Stream(); s3 = torch.cuda.Stream(); s4 = torch.cuda.Stream()
def single():
    d.copy_(h, non_blocking=True)
def multi(k):
    streams = [torch.cuda.current_stream(), s2, s3, s4][:k]
    step = n // k
    for i, s in enumerate(streams):
        lo, hi = i*step, (n if i == k-1 else (i+1)*step)
        with torch.cuda.stream(s):
            d[lo:hi].copy_(h[lo:hi], non_blocking=True)
    for s in streams[1:]:
        torch.cuda.current_stream().wait_stream(s)
for name, fn in [("single", single), ("2 streams", lambda: multi(2)), ("4 streams", lambda: multi(4))]:
    for _ in range(2): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5): fn()
    torch.cuda.synchronize(); dt = (time.perf_counter()-t0)/5
    print(name, round(dt*1e3,2), "ms", round(n*4/dt/1e9,1), "GB/s")
