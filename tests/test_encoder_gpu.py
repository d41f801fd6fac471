"""Text encoder (reference model.py:154-224) on the sm_100a kernels vs the CPU oracle (stock torch conv / BatchNorm /
packed nn.LSTM, i.e. the reference's own ops): outputs and every gradient, identical Philox dropout masks, ragged lengths.

bf16 mode (conv stack + persistent BiLSTM kernel): stated tolerance 3e-2 relative Frobenius on outputs, 8e-2 on the LSTM's
gradients, 1.5e-1 on the gradients that went back through the three relu + BatchNorm layers (relu decisions of bf16-rounded
pre-activations near zero flip, and each flip moves a whole BatchNorm statistic).
fp32 mode: conv stack in 3-plane fp32-accurate arithmetic (1e-3), LSTM on the stock op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(got, ref, l2=True):
    got, ref = got.detach().float().cpu(), ref.detach().float().cpu()
    if l2:
        return float((got - ref).norm()) / (float(ref.norm()) + 1e-12)
    return float((got - ref).abs().max()) / (float(ref.abs().max()) + 1e-12)


def _setup(prec, B, T, seed):
    from gantron_b200.hparams import HParams
    from gantron_b200.model import Encoder
    torch.manual_seed(seed)
    hp = HParams(f"use_noise=True,use_labels=False,b200_precision={prec}")
    enc = Encoder(hp).cuda()
    with torch.no_grad():
        for seq in enc.convolutions:
            seq[1].weight.uniform_(0.5, 1.5); seq[1].bias.uniform_(-0.3, 0.3)
    g = torch.Generator().manual_seed(seed + 1)
    emb = torch.randn(B, 512, T, generator=g) * 0.3
    lens = torch.tensor([max(T - 5 * i, 2) for i in range(B)])
    emb = emb * (torch.arange(T)[None, None, :] < lens[:, None, None])      # zero-padded text positions
    R = torch.randn(B, T, 512, generator=g)
    return hp, enc, emb, lens, R


def _oracle(enc, emb, lens, R, seed, training):
    from oracle import gantron_oracle as O, philox as PX
    sd = {"encoder." + k: v.detach().float().cpu().clone().requires_grad_(v.dtype.is_floating_point)
          for k, v in enc.state_dict().items()}
    B, C, T = emb.shape
    keeps = [torch.from_numpy(PX.keep_mask(seed, PX.SITE_ENC0 + i, 0, (B, T, C), 0.5)).permute(0, 2, 1).contiguous()
             for i in range(3)] if training else None
    x = emb.clone().requires_grad_(True)
    out = O.encoder_forward(sd, x, lens, keeps, training)
    (out * R).sum().backward()
    return out.detach(), x.grad, sd


@pytest.mark.parametrize("B,T", [(5, 37), (20, 150), (64, 150)])
def test_encoder_bf16_forward_backward(B, T):
    hp, enc, emb, lens, R = _setup("bf16", B, T, seed=B)
    enc.train()
    seed = 555 + B
    enc._forced_seed = seed
    x = emb.cuda().requires_grad_(True)
    out = enc(x, lens.cuda())
    assert out.shape == (B, T, 512)
    (out * R.cuda()).sum().backward()
    oo, ogx, sd = _oracle(enc, emb, lens, R, seed, True)
    pad = torch.arange(T)[None, :] >= lens[:, None]
    assert float(out.detach().cpu()[pad].abs().max()) == 0.0, "outputs beyond the sequence length must be zero (packed LSTM)"
    errs = {k: _rel(p.grad, sd['encoder.' + k].grad) for k, p in enc.named_parameters() if not k.endswith("conv.bias")}
    errs["d(embedded text)"] = _rel(x.grad, ogx)
    errs["out"] = _rel(out, oo)
    msg = "; ".join(f"{k} {e:.2e}" for k, e in sorted(errs.items(), key=lambda kv: -kv[1])[:8])
    assert errs["out"] <= 3e-2, msg
    assert all(e <= 8e-2 for k, e in errs.items() if k.startswith("lstm.")), msg
    assert all(e <= 1.5e-1 for e in errs.values()), msg
    print("ENCODER", B, T, msg)


def test_encoder_inference_eval_mode():
    """Encoder.inference (model.py:216-224): eval-mode BatchNorm, no dropout, no packing."""
    hp, enc, emb, lens, R = _setup("bf16", 1, 43, seed=9)
    enc.eval()
    with torch.no_grad():
        out = enc.inference(emb.cuda())
    from oracle import gantron_oracle as O
    sd = {"encoder." + k: v.detach().float().cpu() for k, v in enc.state_dict().items()}
    with torch.no_grad():
        oo = O.encoder_forward(sd, emb, None, None, False)
    assert _rel(out, oo) <= 3e-2, f"encoder inference {_rel(out, oo):.3e}"


def test_encoder_fp32_mode_conv_stack():
    """fp32 mode: conv stack on the 3-plane kernels (1e-3), LSTM on the stock op."""
    hp, enc, emb, lens, R = _setup("fp32", 4, 29, seed=3)
    enc.train()
    enc._forced_seed = 77
    x = emb.cuda().requires_grad_(True)
    out = enc(x, lens.cuda())
    (out * R.cuda()).sum().backward()
    oo, ogx, sd = _oracle(enc, emb, lens, R, 77, True)
    assert _rel(out, oo, l2=False) <= 1e-3, f"encoder out fp32 mode {_rel(out, oo, l2=False):.3e}"
    assert _rel(x.grad, ogx, l2=False) <= 2e-3, f"d(embedded text) fp32 mode {_rel(x.grad, ogx, l2=False):.3e}"
