"""Dense TF32 and bf16 cuBLAS GEMM rates on this box (BASELINE.md section 2 asks for the TF32 figure to be measured):
torch.matmul 8192^3, best of 10 (burst) and back to back for 3 s (sustained), CUDA events.  One JSON line."""
import json
import time

import torch

N = 8192
out = {}
for name, dt, tf32 in (('tf32', torch.float32, True), ('bf16', torch.bfloat16, False)):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    a = torch.randn(N, N, device='cuda', dtype=dt)
    b = torch.randn(N, N, device='cuda', dtype=dt)
    for _ in range(3):
        a @ b
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        a @ b
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    n, e0, e1 = 0, torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    while time.time() - t0 < 3.0:
        for _ in range(20):
            a @ b
        n += 20
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    out[name + '_tflops'] = 2 * N ** 3 / (best * 1e-3) / 1e12
    out[name + '_tflops_sustained'] = 2 * N ** 3 * n / (e0.elapsed_time(e1) * 1e-3) / 1e12
out['gpu'] = torch.cuda.get_device_name(0)
out['how'] = 'torch.matmul %d^3 (cuBLAS), allow_tf32 for the fp32 case; best of 10 and 3 s back to back' % N
print(json.dumps(out))
