"""One teacher-forced decoder forward + backward (B, T) for Nsight Compute (dev tool): `ncu -k regex:decoder_ ...`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gantron_b200 import decoder_ops  # noqa: E402
from gantron_b200.hparams import HParams  # noqa: E402
from gantron_b200.model import Decoder  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
T = int(sys.argv[2]) if len(sys.argv) > 2 else 100
torch.manual_seed(0)
dec = Decoder(HParams("use_noise=True,use_labels=False,b200_precision=bf16")).cuda().train()
mem = torch.cat((torch.randn(B, 150, 512).cuda() * 0.5, torch.rand(B, 1, 512).repeat_interleave(150, 1).cuda()), -1).requires_grad_(True)
mels = (torch.randn(B, 80, T) * 2 - 5).clamp(-11.5, 2).cuda()
for rep in range(2):
    dec.zero_grad()
    mel, gate, align = decoder_ops.decoder_train_forward(dec, mem, mels, torch.full((B,), 150).cuda(), 5)
    (mel.square().mean() + gate.square().mean()).backward()
    torch.cuda.synchronize()
print("ok", float(mel.sum()))
