#!/usr/bin/env python
"""Dense tf32 / bf16 matmul peaks measured the way MEASURED_PEAKS.json measures bf16 (torch.matmul 8192^3, best of
10 and 4 s back to back), for the tensor-pipe fractions of the BSR path ("tf32 peak to be
measured the same way in the first GPU session").  Writes profiles/r2_tf32_peak.json."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def peak(dtype, tf32):
    n = 8192
    torch.backends.cuda.matmul.allow_tf32 = tf32
    a = torch.randn(n, n, device="cuda", dtype=dtype)
    b = torch.randn(n, n, device="cuda", dtype=dtype)
    for _ in range(3):
        a @ b
    best = 1e9
    for _ in range(10):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        a @ b
        e.record()
        e.synchronize()
        best = min(best, s.elapsed_time(e))
    t0 = time.perf_counter()
    torch.cuda.synchronize()
    iters = 0
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    while time.perf_counter() - t0 < 4.0:
        for _ in range(20):
            a @ b
        iters += 20
        torch.cuda.synchronize()
    e.record()
    e.synchronize()
    return 2.0 * n ** 3 / best / 1e9, 2.0 * n ** 3 * iters / s.elapsed_time(e) / 1e9


def main():
    out = os.path.join(ROOT, sys.argv[1] if len(sys.argv) > 1 else "profiles/r2_tf32_peak.json")
    tf_b, tf_s = peak(torch.float32, True)
    bf_b, bf_s = peak(torch.bfloat16, False)
    d = {"tf32_tflops": tf_b, "tf32_tflops_sustained": tf_s, "bf16_tflops": bf_b, "bf16_tflops_sustained": bf_s,
         "gpu_name": torch.cuda.get_device_name(0), "torch": torch.__version__,
         "how": "torch.matmul 8192^3 (2*N^3): fp32 inputs with allow_tf32 = True, and bf16; best of 10 (burst) and "
                "back to back for 4 s (sustained), CUDA events"}
    json.dump(d, open(out, "w"), indent=1)
    print(json.dumps(d))


if __name__ == "__main__":
    main()
