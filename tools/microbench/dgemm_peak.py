"""cuBLAS DGEMM peak (8192^3) burst + sustained: the FP64 tensor roofline denominator."""
import json, time, torch
n = 8192
a = torch.randn(n, n, dtype=torch.float64, device="cuda"); b = torch.randn(n, n, dtype=torch.float64, device="cuda")
c = torch.empty_like(a)
for _ in range(3): torch.matmul(a, b, out=c)
torch.cuda.synchronize()
best = 1e9
for _ in range(10):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); torch.matmul(a, b, out=c); e1.record(); e1.synchronize()
    best = min(best, e0.elapsed_time(e1))
burst = 2 * n**3 / best * 1e-9
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
t0 = time.time(); k = 0; e0.record()
while time.time() - t0 < 4.0:
    for _ in range(5): torch.matmul(a, b, out=c); k += 1
    torch.cuda.synchronize()
e1.record(); e1.synchronize()
sus = 2 * n**3 * k / e0.elapsed_time(e1) * 1e-9
print(json.dumps({"bench": "cublas_dgemm_8192", "burst_tflops": round(burst, 2), "sustained_tflops": round(sus, 2), "best_ms": round(best, 3)}))
