"""block_n sweep of gtn_gemm_bf16 on the step's NT shapes (dev tool): TFLOP/s per (M, N, K, block_n, out dtype)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gantron_b200 import ops

shapes = [(51200, 2048, 88, torch.float32), (51200, 512, 4096, torch.float32), (51200, 256, 4096, torch.float32),
          (51264, 256, 256, torch.bfloat16), (51264, 256, 80, torch.bfloat16), (9600, 2048, 512, torch.float32),
          (51200, 4096, 256, torch.float32)]
for M, N, K, od in shapes:
    A = torch.randn(M, K, device="cuda").bfloat16()
    B = torch.randn(N, K, device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=od)
    for bn in (64, 128, 256):
        if N % bn:
            continue
        for _ in range(3):
            ops.gemm(A, B, out, block_n=bn)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            ops.gemm(A, B, out, block_n=bn)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"M={M} N={N} K={K} {str(od)[6:]:9s} block_n={bn:3d}: {ms * 1e3:7.1f} us  {2.0 * M * N * K / ms / 1e9:7.0f} TFLOP/s  "
              f"{(M * N * out.element_size()) / ms / 1e6:6.0f} GB/s out")
