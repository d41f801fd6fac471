"""Time the persistent decode kernel (autoregressive, forced step count) for both kernel generations."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gantron_b200 import decoder_ops  # noqa: E402
from gantron_b200.hparams import HParams  # noqa: E402
from gantron_b200.model import Decoder  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
torch.manual_seed(0)
hp = HParams(f"use_noise=True,use_labels=False,b200_precision=bf16,max_decoder_steps={steps}")
dec = Decoder(hp).cuda().eval()
for ver in ("2",):
    os.environ["GANTRON_B200_DECODER"] = ver
    for B in (1, 64, 256):
        mem = torch.cat((torch.randn(B, 150, 512).cuda() * 0.5, torch.rand(B, 1, 512).repeat_interleave(150, 1).cuda()), -1)
        best = None
        for it in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            r = decoder_ops.decoder_inference(dec, mem, max_steps=steps, gate_threshold=0.5, seed=7, prec=1,
                                              const_tail=True, ignore_stop=True)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
        print(f"v{ver} B={B}: {1e3 * best / steps:.1f} us/frame-step  ({B * steps / best * 1e3:.0f} frames/s) "
              f"mel absmax {float(r[0].abs().max()):.3f}", flush=True)

# ---- per-phase clock64() profile of CTA 0 of the batch-1 kernel ----
if os.environ.get("GTN_PROF"):
    names = ["frame start", "prenet done", "att LSTM mat-vec done", "att cells + q partial published", "barrier 1 done",
             "h_att / q gathered", "energies published", "barrier E done", "softmax done", "dec LSTM mat-vec done",
             "dec cells + mel partials published", "barrier 2 done"]
    pf = torch.zeros(16 * 16, dtype=torch.int64, device="cuda")
    decoder_ops.PROF_BUFFER = pf
    mem = torch.cat((torch.randn(1, 150, 512).cuda() * 0.5, torch.rand(1, 1, 512).repeat_interleave(150, 1).cuda()), -1)
    decoder_ops.decoder_inference(dec, mem, max_steps=16, gate_threshold=0.5, seed=7, prec=1, const_tail=True, ignore_stop=True)
    torch.cuda.synchronize()
    pr = pf.cpu().view(16, 16)
    t = 8
    prev = 0
    for i, n in enumerate(names):
        v = int(pr[t, i] - pr[t, 0])
        print(f"   {n:40s} {v:8d}  (+{v - prev})")
        prev = v
    print(f"   next frame start                         {int(pr[t + 1, 0] - pr[t, 0]):8d}")

# ---- per-phase clock64() profile of CTA 0 of the general kernel (decoder_fwd2_kernel, inference) at B = 64 / 256 ----
if os.environ.get("GTN_PROF2"):
    os.environ["GANTRON_B200_INFER1"] = "0"
    order = [(28, "frame start (prenet 1 GEMM)"), (29, "prenet 1 done"), (30, "barrier"), (31, "prenet 2 done"), (0, "barrier; att LSTM start"),
             (4, "att LSTM done"), (5, "barrier"), (24, "attn: q / im2col / MMA issued"), (25, "attn: energies"), (27, "attn: softmax"),
             (6, "attention done"), (8, "barrier"), (13, "dec LSTM done"), (18, "barrier"), (19, "projection + gate done"), (23, "barrier")]
    for B in (64, 256):
        pf = torch.zeros(16 * 32, dtype=torch.int64, device="cuda")
        decoder_ops.PROF_BUFFER = pf
        mem = torch.cat((torch.randn(B, 150, 512).cuda() * 0.5, torch.rand(B, 1, 512).repeat_interleave(150, 1).cuda()), -1)
        decoder_ops.decoder_inference(dec, mem, max_steps=16, gate_threshold=0.5, seed=7, prec=1, const_tail=True, ignore_stop=True)
        torch.cuda.synchronize()
        decoder_ops.PROF_BUFFER = None
        pr = pf.cpu().view(16, 32)
        t = 8
        print(f"--- decoder_fwd2_kernel inference B={B}: cycles since frame start (frame slot {t})")
        prev = 0
        for sl, n in order:
            v = int(pr[t, sl] - pr[t, 28])
            print(f"   {n:40s} {v:8d}  (+{v - prev})")
            prev = v
        print(f"   next frame start                         {int(pr[t + 1, 28] - pr[t, 28]):8d}")
        print("   raw slots:", [int(pr[t, s] - pr[t, 28]) for s in range(32)])
