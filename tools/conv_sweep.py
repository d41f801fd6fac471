"""block_n sweep of gtn_conv1d_bf16 on the Discriminator's conv shapes (dev tool)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gantron_b200 import ops
from gantron_b200.conv_ops import pack_conv_weight

# (B, P positions, Cin, Cout, k, dil): forward convs of the Discriminator at B = 64, T = 800 (P = 40) and their dgrads
shapes = [(64, 40, 1600, 1024, 5, 1), (64, 40, 1024, 512, 5, 2), (64, 40, 512, 512, 5, 2), (64, 40, 512, 80, 5, 2),
          (64, 40, 1024, 1600, 5, 1), (64, 40, 512, 1024, 5, 2)]
for B, P, cin, cout, k, dil in shapes:
    w = torch.randn(cout, cin, k, device="cuda") * 0.02
    wp, ce, cp = pack_conv_weight(w, 1)
    x = torch.randn(B, P, cin, device="cuda").bfloat16()
    out = torch.empty(B, P, cout, device="cuda", dtype=torch.bfloat16)
    bias = torch.zeros(cout, device="cuda")
    pad = dil * (k - 1) // 2
    for bn in (64, 128, 256):
        try:
            for _ in range(3):
                ops.conv1d(wp, ce, cp, k, dil, pad, x, out, None, bias=bias, act=ops.ACT_TANH, dropout=(0.5, 1, 32), drop_first=True, block_n=bn)
        except Exception as exc:
            print(f"cin={cin} cout={cout} block_n={bn}: {str(exc)[:80]}")
            continue
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            ops.conv1d(wp, ce, cp, k, dil, pad, x, out, None, bias=bias, act=ops.ACT_TANH, dropout=(0.5, 1, 32), drop_first=True, block_n=bn)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"B={B} P={P} cin={cin} cout={cout} k={k} block_n={bn:3d}: {ms * 1e3:7.1f} us  {2.0 * B * P * cin * cout * k / ms / 1e9:6.0f} TFLOP/s")
