"""Postnet, Discriminator and loss kernels vs the CPU oracle: values and every gradient, identical Philox masks.

fp32 mode (3 bf16 planes) must meet the north star's 1e-3 relative bound; bf16 mode has the stated looser bounds."""
import pytest
import torch

from conftest import assert_close

pytestmark = pytest.mark.gpu


def _hp(prec, extra=""):
    from gantron_b200.hparams import HParams
    return HParams(f"use_noise=True,use_labels=False,b200_precision={prec}" + extra)


def _rel(got, ref, l2=False):
    got, ref = got.detach().float().cpu(), ref.detach().float().cpu()
    if l2:
        return float((got - ref).norm()) / (float(ref.norm()) + 1e-12)
    return float((got - ref).abs().max()) / (float(ref.abs().max()) + 1e-12)


def _sd_cpu(mod, prefix=""):
    return {prefix + k: v.detach().float().cpu().clone().requires_grad_(v.dtype.is_floating_point)
            for k, v in mod.state_dict().items()}


def _masks_cl(seed, site0, n, B, T, chans):
    """Keep masks numbered channels-last (b,t,c), returned as (B,C,T) for the oracle."""
    from oracle import philox as PX
    return [torch.from_numpy(PX.keep_mask(seed, site0 + i, 0, (B, T, chans[i]), 0.5)).permute(0, 2, 1).contiguous()
            for i in range(n)]


@pytest.mark.parametrize("prec,B,T,ragged", [("fp32", 3, 37, True), ("fp32", 2, 150, False), ("bf16", 4, 61, True)])
def test_postnet_forward_backward(prec, B, T, ragged):
    from gantron_b200.model import Postnet
    from gantron_b200 import postnet_ops
    from oracle import gantron_oracle as O, philox as PX
    torch.manual_seed(11)
    hp = _hp(prec)
    net = Postnet(hp).cuda().train()
    with torch.no_grad():      # non-trivial BN affine parameters
        for seq in net.convolutions:
            seq[1].weight.uniform_(0.5, 1.5); seq[1].bias.uniform_(-0.3, 0.3)
    g = torch.Generator().manual_seed(5)
    x = (torch.randn(B, 80, T, generator=g) * 2 - 5).clamp(-11.5, 2)
    R = torch.randn(B, 80, T, generator=g)
    xm = None
    if ragged:
        lens = torch.tensor([max(T - 5 * i, 1) for i in range(B)])
        xm = x.masked_fill(torch.arange(T)[None, None, :] >= lens[:, None, None], 0.0)
    seed = 777
    sd = _sd_cpu(net, "postnet.")
    rm0 = net.convolutions[0][1].running_mean.clone()
    xg = x.cuda().requires_grad_(True)
    y = postnet_ops.postnet_forward(net, xg, seed, 1 if prec == "bf16" else 3, wgrad_input=None if xm is None else xm.cuda())
    (y * R.cuda()).sum().backward()

    d = O.Dims.from_hparams(hp)
    chans = [512, 512, 512, 512, 80]
    keeps = _masks_cl(seed, PX.SITE_POSTNET0, 5, B, T, chans)
    xo = x.clone().requires_grad_(True)
    stats = {}
    yo = O.postnet(sd, xo, d, keeps, True, x_masked_for_wgrad=xm, stats_out=stats)
    (yo * R).sum().backward()

    if prec == "fp32":
        assert_close(y, yo, "postnet out", atol=2e-4, rtol=1e-3)
        assert _rel(xg.grad, xo.grad) <= 1e-3, f"d(input) {_rel(xg.grad, xo.grad):.3e}"
        worst = sorted(((_rel(p.grad, sd['postnet.' + k].grad), k) for k, p in net.named_parameters()), reverse=True)
        # conv biases feed straight into BatchNorm: their true gradient is 0 and both sides hold round-off noise
        worst = [(e, k) for e, k in worst if not k.endswith("conv.bias")]
        assert worst[0][0] <= 1e-3, "postnet grads: " + "; ".join(f"{k} {e:.2e}" for e, k in worst[:6])
        for k, p in net.named_parameters():
            if k.endswith("conv.bias"):
                scale = float(net.get_parameter(k.replace("bias", "weight")).grad.abs().max())
                assert float(p.grad.abs().max()) <= 1e-3 * scale * 10 + 1e-4, f"{k} grad should be ~0"
        # running statistics (momentum 0.1, unbiased variance)
        m, v = stats[0]
        assert_close(net.convolutions[0][1].running_mean, 0.9 * rm0.cpu() + 0.1 * m, "running_mean", atol=1e-5, rtol=1e-3)
        assert_close(net.convolutions[0][1].running_var, 0.9 * torch.ones(512) + 0.1 * v, "running_var", atol=1e-5, rtol=1e-3)
        assert int(net.convolutions[0][1].num_batches_tracked) == 1
    else:
        # bf16 operands/activations, fp32 accumulate: stated tolerance 3e-2 relative Frobenius on outputs and grads
        assert _rel(y, yo, l2=True) <= 3e-2, f"postnet out bf16 {_rel(y, yo, l2=True):.3e}"
        assert _rel(xg.grad, xo.grad, l2=True) <= 5e-2
        for k, p in net.named_parameters():
            if not k.endswith("conv.bias"):
                e = _rel(p.grad, sd['postnet.' + k].grad, l2=True)
                assert e <= 5e-2, f"{k}: {e:.3e}"


def test_postnet_eval_mode_uses_running_stats():
    from gantron_b200.model import Postnet
    from oracle import gantron_oracle as O
    torch.manual_seed(3)
    hp = _hp("fp32")
    net = Postnet(hp).cuda().eval()
    with torch.no_grad():
        for seq in net.convolutions:
            seq[1].running_mean.uniform_(-0.2, 0.2); seq[1].running_var.uniform_(0.5, 1.5)
    x = torch.randn(2, 80, 33)
    with torch.no_grad():
        y = net(x.cuda())
    yo = O.postnet(_sd_cpu(net, "postnet."), x, O.Dims.from_hparams(hp), None, False)
    assert_close(y, yo, "postnet eval", atol=2e-4, rtol=1e-3)


def test_postnet_eval_mode_backward():
    """Backward through eval-mode BatchNorm (running statistics are constants): the reference supports it
    (autograd through model.py:146-151 with self.training False)."""
    from gantron_b200.model import Postnet
    from oracle import gantron_oracle as O
    torch.manual_seed(4)
    hp = _hp("fp32")
    net = Postnet(hp).cuda().eval()
    with torch.no_grad():
        for seq in net.convolutions:
            seq[1].running_mean.uniform_(-0.2, 0.2); seq[1].running_var.uniform_(0.5, 1.5)
            seq[1].weight.uniform_(0.5, 1.5); seq[1].bias.uniform_(-0.3, 0.3)
    g = torch.Generator().manual_seed(6)
    x, R = torch.randn(2, 80, 41, generator=g), torch.randn(2, 80, 41, generator=g)
    sd = _sd_cpu(net, "postnet.")
    xg = x.cuda().requires_grad_(True)
    (net(xg) * R.cuda()).sum().backward()
    xo = x.clone().requires_grad_(True)
    (O.postnet(sd, xo, O.Dims.from_hparams(hp), None, False) * R).sum().backward()
    assert _rel(xg.grad, xo.grad) <= 1e-3, f"d(input) {_rel(xg.grad, xo.grad):.3e}"
    worst = sorted(((_rel(p.grad, sd['postnet.' + k].grad), k) for k, p in net.named_parameters()), reverse=True)
    assert worst[0][0] <= 1e-3, "postnet eval grads: " + "; ".join(f"{k} {e:.2e}" for e, k in worst[:6])


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_prenet_module_forward_backward(prec):
    """Stand-alone Prenet.forward (model.py:102-105): honours b200_precision and is differentiable (input + weights)."""
    from gantron_b200.model import Decoder
    from oracle import gantron_oracle as O, philox as PX
    torch.manual_seed(8)
    dec = Decoder(_hp(prec)).cuda()
    pre = dec.prenet
    assert pre._prec == (3 if prec == "fp32" else 1)
    pre._forced_seed = 4242
    g = torch.Generator().manual_seed(9)
    x, R = torch.randn(5, 3, 80, generator=g), torch.randn(5, 3, 256, generator=g)
    xg = x.cuda().requires_grad_(True)
    y = pre(xg)
    (y * R.cuda()).sum().backward()
    w1 = pre.layers[0].linear_layer.weight.detach().cpu().clone().requires_grad_(True)
    w2 = pre.layers[1].linear_layer.weight.detach().cpu().clone().requires_grad_(True)
    k1 = torch.from_numpy(PX.keep_mask(4242, PX.SITE_PRENET1, 0, (15, 256), 0.5)).reshape(5, 3, 256)
    k2 = torch.from_numpy(PX.keep_mask(4242, PX.SITE_PRENET2, 0, (15, 256), 0.5)).reshape(5, 3, 256)
    xo = x.clone().requires_grad_(True)
    yo = O.prenet(xo, w1, w2, k1, k2)
    (yo * R).sum().backward()
    # bf16: relu decisions of bf16-rounded pre-activations near 0 flip (same bound as the decoder tests: 8e-2 Frobenius)
    tol, gtol = (1e-3, 1e-3) if prec == "fp32" else (3e-2, 8e-2)
    l2 = prec != "fp32"
    assert _rel(y, yo, l2) <= tol, f"prenet out {_rel(y, yo, l2):.3e}"
    assert _rel(xg.grad, xo.grad, l2) <= gtol, f"d(input) {_rel(xg.grad, xo.grad, l2):.3e}"
    assert _rel(pre.layers[0].linear_layer.weight.grad, w1.grad, l2) <= gtol
    assert _rel(pre.layers[1].linear_layer.weight.grad, w2.grad, l2) <= gtol


@pytest.mark.parametrize("prec,B,T,train_d", [("fp32", 3, 173, True), ("fp32", 2, 200, False), ("bf16", 4, 415, True)])
def test_discriminator_adversarial_loss(prec, B, T, train_d):
    """adversarial_loss incl. the tail-window cat (T % 20 != 0), the raw reshape, dropout -> tanh order, masked mean."""
    from gantron_b200.model import Discriminator
    from oracle import gantron_oracle as O, philox as PX
    torch.manual_seed(21)
    hp = _hp(prec)
    D = Discriminator(hp).cuda().train()
    for p in D.parameters():
        p.requires_grad_(train_d)
    g = torch.Generator().manual_seed(6)
    mel = (torch.randn(B, 80, T, generator=g) * 2 - 5).clamp(-11.5, 2)
    lens = torch.tensor([max(T - 41 * i, 1) for i in range(B)])
    seed = 4242
    D._forced_seed = seed
    mg = mel.cuda().requires_grad_(True)
    loss = D.adversarial_loss(mg, lens.cuda())
    loss.backward()

    d = O.Dims.from_hparams(hp)
    P = (T - T % 20 + (20 if T % 20 else 0)) // 20
    chans = [c[1] for c in O.disc_layer_dims(d)[:4]]
    keeps = _masks_cl(seed, PX.SITE_DISC0, 4, B, P, chans)
    sd = _sd_cpu(D)
    mo = mel.clone().requires_grad_(True)
    lo = O.adversarial_loss(sd, mo, lens, d, keeps)
    lo.backward()
    if prec == "fp32":
        assert abs(float(loss) - float(lo)) <= 1e-3 * abs(float(lo)) + 1e-5, f"adv loss {float(loss)} vs {float(lo)}"
        assert _rel(mg.grad, mo.grad) <= 1e-3, f"d(mel) {_rel(mg.grad, mo.grad):.3e}"
        if train_d:
            worst = sorted(((_rel(p.grad, sd[k].grad), k) for k, p in D.named_parameters()), reverse=True)
            assert worst[0][0] <= 1e-3, "D grads: " + "; ".join(f"{k} {e:.2e}" for e, k in worst[:6])
        else:
            assert all(p.grad is None for p in D.parameters())
    else:
        assert abs(float(loss) - float(lo)) <= 3e-2 * abs(float(lo)) + 3e-3
        assert _rel(mg.grad, mo.grad, l2=True) <= 6e-2, f"d(mel) bf16 {_rel(mg.grad, mo.grad, l2=True):.3e}"
        for k, p in D.named_parameters():
            e = _rel(p.grad, sd[k].grad, l2=True)
            assert e <= 6e-2, f"{k}: {e:.3e}"


def test_discriminator_forward_shape_and_eval_dropout_off():
    from gantron_b200.model import Discriminator
    from oracle import gantron_oracle as O
    torch.manual_seed(2)
    hp = _hp("fp32")
    D = Discriminator(hp).cuda().eval()      # nn.Dropout modules inactive -> deterministic
    x = torch.randn(2, 100, 80)
    with torch.no_grad():
        out = D(x.cuda())
    assert tuple(out.shape) == (2, 5)
    ref = O.discriminator_forward(_sd_cpu(D), x, O.Dims.from_hparams(hp), None)
    assert_close(out, ref, "D forward eval", atol=2e-4, rtol=1e-3)


@pytest.mark.parametrize("ragged", [False, True])
def test_tacotron2_loss_and_attention_loss(ragged):
    from gantron_b200.loss_function import Tacotron2Loss
    from oracle import gantron_oracle as O
    g = torch.Generator().manual_seed(9)
    B, T, Tin = 3, 50, 23
    buf = torch.randn(B, T, 81, generator=g)                  # channels-last storage like the decoder output
    post = torch.randn(B, 80, T, generator=g)
    tgt = torch.randn(B, 80, T, generator=g)
    gate_t = (torch.rand(B, T, generator=g) > 0.8).float()
    align = torch.softmax(torch.randn(B, T, Tin, generator=g) * 2, -1)
    if ragged:
        in_len, out_len = torch.tensor([23, 20, 9]), torch.tensor([50, 31, 17])
    else:
        in_len, out_len = torch.full((B,), Tin), torch.full((B,), T)
    bufc = buf.cuda().requires_grad_(True)
    postc = post.cuda().requires_grad_(True)
    alignc = align.cuda().requires_grad_(True)
    mel_c, gate_c = bufc[:, :, :80].transpose(1, 2), bufc[:, :, 80]
    ml, gl, al = Tacotron2Loss()([mel_c, postc, gate_c, alignc], (tgt.cuda(), gate_t.cuda()), in_len.cuda(), out_len.cuda())
    (ml * 1.3 + gl * 0.7 + 10 * al).backward()

    bo = buf.clone().requires_grad_(True)
    po = post.clone().requires_grad_(True)
    ao = align.clone().requires_grad_(True)
    mlo, glo = O.tacotron2_losses(bo[:, :, :80].transpose(1, 2), po, bo[:, :, 80], tgt, gate_t)
    alo = O.attention_loss(ao, in_len, out_len)
    (mlo * 1.3 + glo * 0.7 + 10 * alo).backward()
    for a, b, n in ((ml, mlo, "mel_loss"), (gl, glo, "gate_loss"), (al, alo, "attn_loss")):
        assert abs(float(a) - float(b)) <= 1e-5 * abs(float(b)) + 1e-6, f"{n}: {float(a)} vs {float(b)}"
    assert_close(bufc.grad, bo.grad, "d(mel|gate)", atol=1e-7, rtol=1e-4)
    assert_close(postc.grad, po.grad, "d(mel_post)", atol=1e-7, rtol=1e-4)
    assert_close(alignc.grad, ao.grad, "d(align)", atol=1e-7, rtol=1e-4)


def test_gradient_penalty_double_backward():
    """gradient_penalty (train.py:32-69) differentiates twice through the Discriminator (create_graph=True at
    train.py:55-57).  The drop-in routes that one call through a twice-differentiable composite with the kernels' own Philox
    dropout stream; value and D-parameter gradients against the CPU oracle."""
    from gantron_b200.model import Discriminator, get_mask_from_lengths
    from oracle import gantron_oracle as O, philox as PX
    from torch.autograd import grad as torch_grad
    torch.manual_seed(21)
    hp = _hp("fp32", ",gradient_penalty_lambda=10,clipping_value=0")
    D = Discriminator(hp).cuda().train()
    D._forced_seed = 31337
    B, T = 3, 60
    g = torch.Generator().manual_seed(2)
    real = (torch.randn(B, 80, T, generator=g) * 2 - 5).clamp(-11.5, 2)
    fake = (torch.randn(B, 80, T, generator=g) * 2 - 5).clamp(-11.5, 2)
    lens = torch.tensor([60, 47, 40])
    alpha = torch.rand(B, 1, 1, generator=g)

    def penalty(disc_fn, real, fake, lens, alpha, mask_fn):
        # train.py:32-69, equal widths (generated_data has the width of real_data)
        a = alpha.expand_as(fake)
        inter = a * real + (1 - a) * fake
        mask = mask_fn(lens).unsqueeze(1).expand_as(inter)
        inter = inter.masked_fill(mask == False, 0).detach().requires_grad_(True)      # noqa: E712 (as the reference)
        prob = disc_fn(inter.transpose(1, 2))
        grads = torch_grad(outputs=prob, inputs=inter, grad_outputs=torch.ones_like(prob), create_graph=True, retain_graph=True)[0]
        grads = grads.masked_fill(mask == False, 0).reshape(B, -1)                     # noqa: E712
        return ((torch.sqrt(torch.sum(grads ** 2, dim=1) + 1e-12) - 1) ** 2).mean()

    # (the double backward of the convolutions runs outside the forward's no-TF32 scope: switch it off here too)
    with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        gp = penalty(D, real.cuda(), fake.cuda(), lens.cuda(), alpha.cuda(), get_mask_from_lengths)
        gp.backward()
    sd = _sd_cpu(D)
    P = T // 20
    chans = [c[1] for c in O.disc_layer_dims(O.Dims.from_hparams(hp))[:4]]
    keeps = _masks_cl(31337, PX.SITE_DISC0, 4, B, P, chans)
    d = O.Dims.from_hparams(hp)
    gpo = penalty(lambda x: O.discriminator_forward(sd, x, d, keeps), real, fake, lens, alpha,
                  lambda L: torch.arange(T)[None, :] < L[:, None])
    gpo.backward()
    assert abs(float(gp) - float(gpo)) <= 1e-3 * abs(float(gpo)) + 1e-6, f"gradient penalty {float(gp)} vs {float(gpo)}"
    # (the head's bias does not reach the input gradient: its gradient is None on both sides)
    pairs = [(k, p.grad, sd[k].grad) for k, p in D.named_parameters()]
    assert all((a is None) == (b is None) for _, a, b in pairs), [k for k, a, b in pairs if (a is None) != (b is None)]
    worst = sorted(((_rel(a, b), k) for k, a, b in pairs if a is not None), reverse=True)
    assert len(worst) >= 9 and worst[0][0] <= 2e-3, "gradient-penalty grads: " + "; ".join(f"{k} {e:.2e}" for e, k in worst[:5])
