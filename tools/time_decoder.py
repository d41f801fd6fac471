"""Time the persistent decoder kernels at the BASELINE shapes (dev tool, not the bench)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gantron_b200.hparams import HParams
from gantron_b200.model import Decoder
from gantron_b200 import decoder_ops, _lib

def ev():
    return torch.cuda.Event(enable_timing=True)

def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 800
    Tin = 150
    torch.manual_seed(0)
    hp = HParams("use_noise=True,use_labels=False,b200_precision=bf16")
    dec = Decoder(hp).cuda().train()
    enc = torch.randn(B, Tin, 512).cuda() * 0.5
    noise = torch.rand(B, 1, 512).repeat_interleave(Tin, 1).cuda()
    memory = torch.cat((enc, noise), -1).requires_grad_(True)
    mels = (torch.randn(B, 80, T) * 2 - 5).clamp(-11.5, 2).cuda()
    lens = torch.full((B,), Tin).cuda()
    for it in range(3):
        dec.zero_grad()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        mel, gate, align = decoder_ops.decoder_train_forward(dec, memory, mels, lens, 1234 + it)
        e1.record()
        loss = (mel ** 2).mean() + (gate ** 2).mean()
        loss.backward()
        e2.record()
        torch.cuda.synchronize(); t1 = time.perf_counter()
        print(f"iter {it}: fwd {e0.elapsed_time(e1):.2f} ms  bwd {e1.elapsed_time(e2):.2f} ms  wall {1e3*(t1-t0):.1f} ms "
              f"frames/s {B*T/(t1-t0):.0f}  launches {_lib.launch_count()}", flush=True)
    # inference
    dec.eval()
    for Bi in (1, 256):
        mem = torch.cat((torch.randn(Bi, Tin, 512).cuda() * 0.5, torch.rand(Bi, 1, 512).repeat_interleave(Tin, 1).cuda()), -1)
        for it in range(2):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            out = decoder_ops.decoder_inference(dec, mem, max_steps=1000, gate_threshold=0.5, seed=7, prec=1,
                                                const_tail=True, ignore_stop=True)
            torch.cuda.synchronize(); t1 = time.perf_counter()
            print(f"infer B={Bi}: {1e3*(t1-t0):.1f} ms for {out[4]} steps -> {1e6*(t1-t0)/out[4]:.1f} us/step, {Bi*out[4]/(t1-t0):.0f} frames/s", flush=True)

main()
