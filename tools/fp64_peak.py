"""Measures the FP64 denominators MEASURED_PEAKS.json lacks: cuBLAS DGEMM 8192^3 burst /
sustained, and a batched DGEMM at the path's block sizes.  Writes one JSON object to stdout."""
import json
import time

import torch

dev = torch.device("cuda")
out = {"gpu": torch.cuda.get_device_name(0)}


def time_ms(fn, n):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


N = 8192
a = torch.randn(N, N, dtype=torch.float64, device=dev)
b = torch.randn(N, N, dtype=torch.float64, device=dev)
c = torch.empty_like(a)
for _ in range(2):
    torch.matmul(a, b, out=c)
torch.cuda.synchronize()
best = min(time_ms(lambda: torch.matmul(a, b, out=c), 1) for _ in range(10))
out["dgemm_8192_burst_tflops"] = 2 * N**3 / (best * 1e-3) / 1e12
t0 = time.time()
n = 0
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
while time.time() - t0 < 4.0:
    torch.matmul(a, b, out=c)
    n += 1
    if n % 4 == 0:
        torch.cuda.synchronize()
e1.record()
torch.cuda.synchronize()
out["dgemm_8192_sustained_tflops"] = 2 * N**3 * n / (e0.elapsed_time(e1) * 1e-3) / 1e12
del a, b, c
# batched at the path's sizes: (225 x 225) x (225 x 1024) per element, 20 elements
for (m, k, nn, bt) in [(225, 225, 1024, 20), (900, 900, 1024, 40), (70, 70, 70, 4096)]:
    A = torch.randn(bt, m, k, dtype=torch.float64, device=dev)
    B = torch.randn(bt, k, nn, dtype=torch.float64, device=dev)
    C = torch.empty(bt, m, nn, dtype=torch.float64, device=dev)
    for _ in range(3):
        torch.bmm(A, B, out=C)
    torch.cuda.synchronize()
    best = min(time_ms(lambda: torch.bmm(A, B, out=C), 3) for _ in range(5))
    out["dgemm_batched_%dx%dx%d_b%d_tflops" % (m, k, nn, bt)] = 2.0 * m * k * nn * bt / (best * 1e-3) / 1e12
# HBM copy for context
x = torch.empty(1 << 28, dtype=torch.float64, device=dev)
y = torch.empty_like(x)
for _ in range(2):
    y.copy_(x)
best = min(time_ms(lambda: y.copy_(x), 1) for _ in range(10))
out["hbm_copy_gbs"] = 2 * x.numel() * 8 / (best * 1e-3) / 1e9
print(json.dumps(out))
