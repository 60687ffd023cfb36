"""One-off probe of the GPU box: host cores, TBB presence, FP64 GEMM peak (cuBLAS via torch) for the roofline."""
import glob
import json
import os
import time

import torch

out = {"nproc": os.cpu_count()}
try:
    out["cpu_model"] = [l.split(":")[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")][0]
except Exception:
    pass
out["tbb"] = glob.glob("/usr/lib/**/libtbb*", recursive=True)[:4] + glob.glob("/usr/include/**/tbb.h", recursive=True)[:2]
out["gpu"] = torch.cuda.get_device_name(0)
for n in (4096, 8192):
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    for _ in range(2):
        torch.matmul(a, b)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        torch.matmul(a, b)
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e))
    out[f"dgemm_{n}_tflops"] = 2 * n ** 3 / best / 1e9
print(json.dumps(out))
