"""tools/calib_pcie.py — what the host link of the box gives (development aid): contiguous H2D / D2H of pinned
memory alone and together, and the library's streamed col-major upload / download (sblas_dense_upload_rows_*),
which is what bench.py's e2e leg moves per step."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
K = pkg.kernels
dev = torch.device("cuda:0")
rows, k = 1 << 22, 64                      # 2 GiB fp64 operand
n = rows * k
h = torch.ones(n, dtype=torch.float64, pin_memory=True)
h2 = torch.empty(n, dtype=torch.float64, pin_memory=True)
d = torch.empty(n, dtype=torch.float64, device=dev)
d2 = torch.ones(n, dtype=torch.float64, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
GB = n * 8 / 1e9


def t(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    a = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - a) / reps


ms = t(lambda: d.copy_(h, non_blocking=True))
print("H2D contiguous %.1f GB alone: %.1f GB/s" % (GB, GB / ms))
ms = t(lambda: h2.copy_(d2, non_blocking=True))
print("D2H contiguous alone: %.1f GB/s" % (GB / ms))


def both():
    with torch.cuda.stream(s1):
        d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)


ms = t(both)
print("H2D || D2H contiguous: %.1f GB/s each way (pair in %.1f ms)" % (GB / ms, ms * 1e3))
ms = t(lambda: K.dense_upload_rows(h, rows, 0, rows, k, d, k))
print("streamed col-major upload (2-D DMA chunks + device transposes): %.1f GB/s" % (GB / ms))
ms = t(lambda: K.dense_download_rows(d2, k, rows, k, h2, rows, 0))
print("streamed col-major download: %.1f GB/s" % (GB / ms))


def both_streamed():
    with torch.cuda.stream(s1):
        K.dense_upload_rows(h, rows, 0, rows, k, d, k, stream=s1)
    with torch.cuda.stream(s2):
        K.dense_download_rows(d2, k, rows, k, h2, rows, 0, stream=s2)


ms = t(both_streamed)
print("streamed upload || download: %.1f GB/s each way (pair in %.1f ms)" % (GB / ms, ms * 1e3))
