import torch, time
n = 67108864
h = torch.empty(n, dtype=torch.uint8).pin_memory()
h2 = torch.empty(n // 2, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
d2 = torch.empty(n // 2, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=20):
    fn(); torch.cuda.synchronize()
    a = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - a) / reps
ms = t(lambda: d.copy_(h, non_blocking=True)) * 1e3
print("H2D 67 MB alone: %.3f ms = %.1f GB/s" % (ms, n / ms / 1e6))
def both():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
ms = t(both) * 1e3
print("H2D 67 MB || D2H 33.5 MB: %.3f ms per pair (H2D %.1f GB/s)" % (ms, n / ms / 1e6))
def two_halves():
    d[: n // 2].copy_(h[: n // 2], non_blocking=True); d[n // 2:].copy_(h[n // 2:], non_blocking=True)
ms = t(two_halves) * 1e3
print("H2D 2 x 33.5 MB same stream: %.3f ms" % ms)
