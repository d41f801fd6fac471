"""Short batched decode (B = 256, T_in = 150, bf16) for ncu: `ncu -k regex:decoder_fwd2 --set full python tools/prof_infer256.py`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gantron_b200 import decoder_ops  # noqa: E402
from gantron_b200.hparams import HParams  # noqa: E402
from gantron_b200.model import Decoder  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 16
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
torch.manual_seed(0)
hp = HParams(f"use_noise=True,use_labels=False,b200_precision=bf16,max_decoder_steps={steps}")
dec = Decoder(hp).cuda().eval()
mem = torch.cat((torch.randn(B, 150, 512).cuda() * 0.5, torch.rand(B, 1, 512).repeat_interleave(150, 1).cuda()), -1)
for _ in range(2):
    r = decoder_ops.decoder_inference(dec, mem, max_steps=steps, gate_threshold=0.5, seed=7, prec=1, const_tail=True, ignore_stop=True)
torch.cuda.synchronize()
print("ok", float(r[0].abs().max()))
