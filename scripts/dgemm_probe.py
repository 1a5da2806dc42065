"""cuBLAS DGEMM on this box (the Gram kernel's denominator): rate, and -- under ncu -- which pipe it runs on."""
import torch, time
p, n = 2048, 262144
a = torch.randn(n, p, dtype=torch.float64, device="cuda")
for _ in range(2):
    c = a.t() @ a
torch.cuda.synchronize()
best = 1e9
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); c = a.t() @ a; e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print("cuBLAS DGEMM %d x %d x %d: %.3f ms, %.2f TFLOP/s" % (p, p, n, best, 2.0 * p * p * n / best / 1e9))
