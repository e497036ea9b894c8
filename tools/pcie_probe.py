"""PCIe characterisation for the host-buffer path: H2D alone, D2H alone, both at once, per copy size (queued back to back)."""
import torch, time, json
dev = torch.device("cuda:0")
out = {}
def timed(fn, reps):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for kb in (256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 65536):
    n = kb << 10
    h_in = torch.empty(n, dtype=torch.uint8).pin_memory(); h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_in.fill_(1); h_out.fill_(2)
    d_in = torch.empty(n, dtype=torch.uint8, device=dev); d_out = torch.empty(n, dtype=torch.uint8, device=dev)
    def h2d():
        with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
    def d2h():
        with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
    def both():
        h2d(); d2h()
    reps = max(10, min(200, (256 << 20) // n))
    r = []
    for _ in range(3):
        r.append((timed(h2d, reps), timed(d2h, reps), timed(both, reps)))
    a, b, c = (min(x[i] for x in r) for i in range(3))
    out["%dKB" % kb] = "h2d %6.1f us %5.1f GB/s | d2h %6.1f us %5.1f GB/s | both %6.1f us" % (a * 1e6, n / a / 1e9, b * 1e6, n / b / 1e9, c * 1e6)
print(json.dumps(out, indent=1))
