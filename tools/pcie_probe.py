"""Tuning aid: PCIe ceilings for the end-to-end path (pinned host memory, 4 GB each way)."""
import torch, time
n = 4 << 30
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True); h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d_in = torch.empty(n, dtype=torch.uint8, device="cuda"); d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(name, fn):
    fn(); torch.cuda.synchronize(); t = time.perf_counter(); fn(); torch.cuda.synchronize(); dt = time.perf_counter() - t
    print(f"{name:28s} {dt * 1e3:7.1f} ms")
    return dt
a = run("H2D 4 GB alone", lambda: d_in.copy_(h_in, non_blocking=True)); print(f"  {n / a / 1e9:.1f} GB/s")
b = run("D2H 4 GB alone", lambda: h_out.copy_(d_out, non_blocking=True)); print(f"  {n / b / 1e9:.1f} GB/s")
def both():
    with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
c = run("H2D + D2H concurrently", both); print(f"  {n / c / 1e9:.1f} GB/s each way")
