"""Per-phase clock64() profile of CTA 0 of the persistent decoder kernels in training (dev tool)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gantron_b200 import decoder_ops  # noqa: E402
from gantron_b200.hparams import HParams  # noqa: E402
from gantron_b200.model import Decoder  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
T = int(sys.argv[2]) if len(sys.argv) > 2 else 32
torch.manual_seed(0)
hp = HParams("use_noise=True,use_labels=False,b200_precision=bf16")
dec = Decoder(hp).cuda().train()
mem = torch.cat((torch.randn(B, 150, 512).cuda() * 0.5, torch.rand(B, 1, 512).repeat_interleave(150, 1).cuda()), -1).requires_grad_(True)
mels = (torch.randn(B, 80, T) * 2 - 5).clamp(-11.5, 2).cuda()
# slots of the pipelined training kernel (gtn_decoder_fwd3.cu, PROF3)
FW = {0: "step start (gsync2 of t-1 done)", 2: "A: acc done", 3: "A: exchange done",
      4: "A: cells + qpart done", 5: "arrived barrier 1", 10: "C(t-1): acc done", 11: "C(t-1): exchange done",
      12: "C(t-1): cells done", 13: "attn prep (im2col, loc MMA) done", 14: "barrier 1 wait done", 15: "attn: q bulk copies issued",
      24: "attn: q partials landed", 16: "attn: q summed", 17: "attn: energies", 18: "attn: pair exchange", 19: "attn: softmax",
      20: "attn done (context)", 21: "arrived barrier 2", 23: "barrier 2 wait done"}
FO = [0, 2, 3, 4, 5, 10, 11, 12, 13, 14, 15, 24, 16, 17, 18, 19, 20, 21, 23]
BW = {0: "step start", 1: "P1: TMA issued", 2: "P1: acc done", 3: "P1: exchange done", 4: "P1: epilogue done", 28: "barrier 1: CTA synced", 29: "barrier 1: release done", 14: "P2 prepare: window loaded", 6: "P2 prepare (in barrier shadow)", 5: "gsync done",
      10: "P2: loads + window", 22: "P2: memory tile landed", 23: "P2: dw done", 24: "P2: dw totals done", 11: "P2: dot (block sum)", 12: "P2: cluster sync 1", 7: "P2: softmax backward", 8: "P2: tanh tile landed", 9: "P2: dth rows done", 13: "P2: dq/dv combined", 30: "P2: G accumulator complete", 15: "P2: G to smem",
      16: "P2: d wcat", 17: "P2: cluster sync 2", 18: "P2 done", 19: "P2: dq in smem", 27: "cells: operands loaded, Wq tile landed", 25: "cells: Wq^T dq product done", 26: "cells: partials combined", 20: "cells done", 21: "gsync done"}
BO = [0, 1, 2, 3, 4, 28, 29, 14, 6, 5, 10, 22, 23, 24, 11, 12, 7, 8, 9, 13, 30, 15, 16, 17, 19, 27, 25, 26, 20, 21]
for rep in range(2):
    pf = torch.zeros(16 * 32 + 4 * 160, dtype=torch.int64, device="cuda")
    pb = torch.zeros(2 * 16 * 32, dtype=torch.int64, device="cuda")
    decoder_ops.PROF_BUFFER, decoder_ops.PROF_BUFFER_BWD = pf, pb
    dec.zero_grad()
    mel, gate, align = decoder_ops.decoder_train_forward(dec, mem, mels, torch.full((B,), 150).cuda(), 5)
    (mel.square().mean() + gate.square().mean()).backward()
    torch.cuda.synchronize()
for name, buf, names, order in (("forward", pf[:512], FW, FO), ("backward CTA 0", pb[:512], BW, BO), ("backward CTA 48", pb[512:], BW, BO)):
    pr = buf.cpu().view(16, 32)
    t = 8
    print(f"--- {name} B={B}: cycles since step start (frame slot {t})")
    prev = 0
    for sl in order:
        if int(pr[t, sl]):
            v = int(pr[t, sl] - pr[t, 0])
            print(f"   {names[sl]:40s} {v:8d}  (+{v - prev})")
            prev = v
    print(f"   next step start          {int(pr[t + 1, 0] - pr[t, 0]):8d}")

# per-CTA skew of the forward kernel at frame 8 (%globaltimer, ns): loop top, barrier-1 arrive, barrier-1 done, barrier-2 arrive
g = pf[512:512 + 4 * 128].cpu().view(128, 4).double()
if float(g.abs().sum()) > 0:
    base = g[:, 0].min()
    names = ["frame start", "arrive barrier 1", "barrier 1 passed", "arrive barrier 2"]
    for k in range(4):
        col = g[:, k] - base
        print(f"   skew {names[k]:20s}: min {col.min():8.0f} ns  median {col.median():8.0f}  max {col.max():8.0f}  (CTA of max: {int(col.argmax())})")
    late = (g[:, 3] - g[:, 3].min())
    print("   arrive-2 lateness by rank (ns):", [round(float(late[r::4].mean())) for r in range(4)])
    print("   arrive-2 lateness by cluster (ns):", [round(float(late[4 * c:4 * c + 4].mean())) for c in range(32)])
    late1 = (g[:, 1] - g[:, 1].min())
    print("   arrive-1 lateness by cluster (ns):", [round(float(late1[4 * c:4 * c + 4].mean())) for c in range(32)])
