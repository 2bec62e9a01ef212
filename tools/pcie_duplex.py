"""PCIe copy rates of the box: one direction alone and both directions at once (pinned host memory)."""
import torch, time
n = 512_000_000
h1 = torch.empty(n, dtype=torch.uint8).pin_memory(); h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
d1 = torch.empty(n, dtype=torch.uint8, device="cuda"); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(h2d, d2h, chunks=1):
    torch.cuda.synchronize(); t = time.perf_counter()
    c = n // chunks
    for k in range(chunks):
        if h2d:
            with torch.cuda.stream(s1): d1[k*c:(k+1)*c].copy_(h1[k*c:(k+1)*c], non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2): h2[k*c:(k+1)*c].copy_(d2[k*c:(k+1)*c], non_blocking=True)
    torch.cuda.synchronize(); return (time.perf_counter() - t) * 1e3
for name, a in (("h2d", (1, 0)), ("d2h", (0, 1)), ("both", (1, 1))):
    for ch in (1, 32):
        run(*a, ch); ts = [run(*a, ch) for _ in range(5)]
        print(f"{name} chunks={ch}: {min(ts):.2f} ms  ({n/min(ts)/1e6:.1f} GB/s per direction)")
