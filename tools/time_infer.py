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

# ---- per-phase clock64() profile of CTA 0 (v2 only) ----
if os.environ.get("GTN_PROF"):
    os.environ["GANTRON_B200_DECODER"] = "2"
    names = {0: "step start", 1: "att: TMA issued", 2: "att: acc done", 3: "att: exchange done", 4: "att: epilogue done",
             5: "gsync1 done", 24: "attn: mma issued", 25: "attn: energies", 26: "attn: cluster sync", 27: "attn: softmax",
             6: "attn done", 8: "gsync2 done", 9: "dec: TMA issued", 10: "dec: acc done", 11: "dec: exchange done",
             13: "dec done", 28: "frame start (pre1)", 29: "pre1 done", 30: "gsync done", 31: "pre2 done",
             18: "gsync3 done", 14: "proj: TMA issued", 15: "proj: acc done", 16: "proj: exchange done", 19: "proj done", 23: "gsync4 done"}
    order = [28, 29, 30, 31, 0, 1, 2, 3, 4, 5, 24, 25, 26, 27, 6, 8, 9, 10, 11, 13, 18, 14, 15, 16, 19, 23]
    for B, train in ((1, False), (64, False), (64, True)):
        prof = torch.zeros(16 * 32, dtype=torch.int64, device="cuda")
        decoder_ops.PROF_BUFFER = prof
        if train:
            dec.train()
            mem = torch.cat((torch.randn(B, 150, 512).cuda() * 0.5, torch.rand(B, 1, 512).repeat_interleave(150, 1).cuda()), -1)
            mels = (torch.randn(B, 80, 16) * 2 - 5).clamp(-11.5, 2).cuda()
            with torch.no_grad():
                decoder_ops.decoder_train_forward(dec, mem, mels, torch.full((B,), 150).cuda(), 5)
        else:
            dec.eval()
            mem = torch.cat((torch.randn(B, 150, 512).cuda() * 0.5, torch.rand(B, 1, 512).repeat_interleave(150, 1).cuda()), -1)
            decoder_ops.decoder_inference(dec, mem, max_steps=16, gate_threshold=0.5, seed=7, prec=1, const_tail=True,
                                          ignore_stop=True)
        torch.cuda.synchronize()
        pr = prof.cpu().view(16, 32)
        print(f"--- B={B} train={train}: cycles since step start, frame 8 (and frame 9 start)")
        t = 8
        base = int(pr[t, 28]) if int(pr[t, 28]) else int(pr[t, 0])
        prev = base
        for sl in order:
            if int(pr[t, sl]):
                v = int(pr[t, sl])
                print(f"   {names[sl]:22s} {v - base:8d}  (+{v - prev})")
                prev = v
        nxt = int(pr[t + 1, 28]) if int(pr[t + 1, 28]) else int(pr[t + 1, 0])
        print(f"   next frame start       {nxt - base:8d}")
    decoder_ops.PROF_BUFFER = None
