"""cuBLAS FP64 and TF32 GEMM throughput on this box (roofline denominators that MEASURED_PEAKS.json does not hold)."""
import json, torch
def t(dtype, n, tf32=False, reps=5):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    a = torch.randn(n, n, device="cuda", dtype=dtype); b = torch.randn(n, n, device="cuda", dtype=dtype)
    torch.matmul(a, b); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return 2.0 * n ** 3 / best * 1e-9
print(json.dumps({"cublas_fp64_tflops": t(torch.float64, 8192), "cublas_tf32_tflops": t(torch.float32, 8192, True),
                  "cublas_fp32_tflops": t(torch.float32, 8192, False)}))
