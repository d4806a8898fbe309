"""Do two neighbouring GPUs have independent host links?  H2D from pinned memory to GPU 0 alone, GPU 1 alone, both at once."""
import torch, time
n = 64 << 20
hs = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(2)]
ds = [torch.empty(n, dtype=torch.uint8, device=f"cuda:{i}") for i in range(2)]
ss = [torch.cuda.Stream(device=f"cuda:{i}") for i in range(2)]
def run(which, reps=20):
    for i in which:
        torch.cuda.synchronize(i)
    t0 = time.perf_counter()
    for _ in range(reps):
        for i in which:
            with torch.cuda.stream(ss[i]):
                ds[i].copy_(hs[i], non_blocking=True)
    for i in which:
        torch.cuda.synchronize(i)
    dt = time.perf_counter() - t0
    return len(which) * reps * n / dt / 1e9
run([0]); run([1])
print(f"H2D GPU0 alone {run([0]):.1f} GB/s, GPU1 alone {run([1]):.1f} GB/s, both at once {run([0, 1]):.1f} GB/s (sum)")
# and the peer hop that would bring GPU 1's half over to GPU 0
a = torch.empty(6 << 20, dtype=torch.uint8, device="cuda:1"); b = torch.empty(6 << 20, dtype=torch.uint8, device="cuda:0")
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.device(0):
    b.copy_(a); torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        b.copy_(a, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
print(f"peer copy of 6 MiB GPU1 -> GPU0: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us")
# the host-facing step's copy pattern without any compute: half a batch over each link, then the hop, double-buffered
B, half = 1024, 512
pool = [torch.randn(B, 3, 32, 32).pin_memory() for _ in range(16)]
d0, d1 = torch.device("cuda:0"), torch.device("cuda:1")
xd = [torch.empty(B, 3, 32, 32, device=d0) for _ in range(2)]
xv = [torch.empty(B - half, 3, 32, 32, device=d1) for _ in range(2)]
s_own, s_hop = torch.cuda.Stream(device=d0), torch.cuda.Stream(device=d0)
s_via = torch.cuda.Stream(device=d1)
evs = [tuple(torch.cuda.Event() for _ in range(3)) for _ in range(2)]
def step(i, split=True):
    f = i & 1
    arrived, own, ev = evs[f]
    x = pool[i % 16]
    if not split:
        with torch.cuda.stream(s_own):
            xd[f].copy_(x, non_blocking=True)
        return
    s_via.wait_event(ev); 
    with torch.cuda.stream(s_via):
        xv[f].copy_(x[half:], non_blocking=True); arrived.record(s_via)
    with torch.cuda.stream(s_own):
        xd[f][:half].copy_(x[:half], non_blocking=True); own.record(s_own)
    with torch.cuda.stream(s_hop):
        s_hop.wait_event(arrived)
        xd[f][half:].copy_(xv[f], non_blocking=True)
        s_hop.wait_event(own); ev.record(s_hop)
for split in (False, True):
    for i in range(8): step(i, split)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    t0 = time.perf_counter()
    for i in range(200): step(i, split)
    t1 = time.perf_counter()
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    t2 = time.perf_counter()
    print(f"copy pattern split={split}: {1e6 * (t2 - t0) / 200:.1f} us per 1024-sample batch (host enqueue {1e6 * (t1 - t0) / 200:.1f} us)")
