"""Persistent decode kernel (autoregressive inference) vs the CPU oracle, step for step.

fp32 mode (3-plane bf16 split) must match the fp32 oracle to ~1e-4; bf16 mode to a stated looser bound.
Dropout masks are the shared Philox scheme (oracle/philox.py <-> gtn_common.cuh)."""
import numpy as np
import pytest
import torch

from conftest import assert_close

pytestmark = pytest.mark.gpu


def _setup(B, T_in, steps, prec_name, seed=0, style_const=True, hp_extra=""):
    from gantron_b200.hparams import HParams
    from gantron_b200.model import Decoder
    torch.manual_seed(seed)
    hp = HParams(f"use_noise=True,use_labels=False,max_decoder_steps={steps},b200_precision={prec_name}" + hp_extra)
    dec = Decoder(hp).cuda().eval()
    g = torch.Generator().manual_seed(seed + 1)
    enc = torch.randn(B, T_in, hp.encoder_embedding_dim, generator=g) * 0.5
    tail = dec.decoder_in_dim - hp.encoder_embedding_dim
    if style_const:
        style = torch.rand(B, 1, tail, generator=g).repeat_interleave(T_in, dim=1)
    else:
        style = torch.rand(B, T_in, tail, generator=g)
    memory = torch.cat((enc, style), -1)
    return hp, dec, memory


def _oracle(hp, dec, memory, steps, seed, force=True):
    from oracle import gantron_oracle as O, philox as PX
    d = O.Dims.from_hparams(hp)
    d.spk = 0; d.emo = 0
    d.noise = dec.decoder_in_dim - hp.encoder_embedding_dim
    sd = {"decoder." + k: v.detach().float().cpu() for k, v in dec.state_dict().items()}
    B = memory.shape[0]

    def keep_fn(t):
        return (torch.from_numpy(PX.keep_mask(seed, PX.SITE_PRENET1, t, (B, hp.prenet_dim), 0.5)),
                torch.from_numpy(PX.keep_mask(seed, PX.SITE_PRENET2, t, (B, hp.prenet_dim), 0.5)))
    with torch.no_grad():
        return O.decoder_inference(sd, memory, d, keep_fn, max_steps=steps, force_steps=steps if force else None)


@pytest.mark.parametrize("B,T_in,steps", [(1, 17, 6), (3, 23, 5), (9, 40, 4), (70, 12, 3), (130, 11, 3)])
def test_infer_fp32_mode_matches_oracle(B, T_in, steps):
    from gantron_b200 import decoder_ops
    hp, dec, memory = _setup(B, T_in, steps, "fp32")
    seed = 0xABCDEF12345
    mel, gate, align, lengths, n = decoder_ops.decoder_inference(
        dec, memory.cuda(), max_steps=steps, gate_threshold=0.5, seed=seed, prec=3, ignore_stop=True)
    om, og, oa, ol, _ = _oracle(hp, dec, memory, steps, seed)
    assert n == steps
    assert_close(align, oa, "alignments", atol=2e-5, rtol=1e-3)
    assert_close(mel, om, "mel", atol=2e-4, rtol=1e-3)
    assert_close(gate, og, "gate", atol=2e-4, rtol=1e-3)


def test_infer_nonconstant_style_uses_full_memory():
    from gantron_b200 import decoder_ops
    hp, dec, memory = _setup(2, 19, 4, "fp32", style_const=False)
    seed = 77
    mel, gate, align, lengths, n = decoder_ops.decoder_inference(
        dec, memory.cuda(), max_steps=4, gate_threshold=0.5, seed=seed, prec=3, ignore_stop=True)
    om, og, oa, _, _ = _oracle(hp, dec, memory, 4, seed)
    assert_close(align, oa, "alignments", atol=2e-5, rtol=1e-3)
    assert_close(mel, om, "mel", atol=2e-4, rtol=1e-3)


def test_infer_bf16_mode_close_to_oracle():
    """bf16 tolerance: 3e-2 absolute on mel (|mel| ~ 1) over the first frames; alignments 2e-2."""
    from gantron_b200 import decoder_ops
    hp, dec, memory = _setup(4, 31, 5, "bf16")
    seed = 4242
    mel, gate, align, lengths, n = decoder_ops.decoder_inference(
        dec, memory.cuda(), max_steps=5, gate_threshold=0.5, seed=seed, prec=1, ignore_stop=True)
    om, og, oa, _, _ = _oracle(hp, dec, memory, 5, seed)
    assert_close(align, oa, "alignments bf16", atol=2e-2, rtol=0)
    assert_close(mel, om, "mel bf16", atol=3e-2, rtol=3e-2)


@pytest.mark.parametrize("B,T_in,steps", [(6, 21, 4), (70, 12, 4), (130, 11, 3), (256, 9, 3)])
def test_infer_bf16_batched_paths_close_to_oracle(B, T_in, steps):
    """Batched decode in bf16 mode (B > 4: projection + prenet inside the clusters; B > 64: the 128 / 256 column tiles),
    same bf16 tolerance as above."""
    from gantron_b200 import decoder_ops
    hp, dec, memory = _setup(B, T_in, steps, "bf16")
    seed = 991
    mel, gate, align, lengths, n = decoder_ops.decoder_inference(
        dec, memory.cuda(), max_steps=steps, gate_threshold=0.5, seed=seed, prec=1, ignore_stop=True)
    om, og, oa, _, _ = _oracle(hp, dec, memory, steps, seed)
    assert n == steps
    assert_close(align, oa, "alignments bf16", atol=2e-2, rtol=0)
    assert_close(mel, om, "mel bf16", atol=3e-2, rtol=3e-2)
    assert_close(gate, og, "gate bf16", atol=3e-2, rtol=3e-2)


@pytest.mark.parametrize("flag", ["GTN_INFER_NO_TAIL", "GTN_INFER_NO_MULTI_ATTN", "GTN_FWD2_GENERIC"])
@pytest.mark.parametrize("B,T_in", [(37, 33), (100, 1), (130, 64), (256, 150), (256, 211)])
def test_infer_bf16_batched_paths_vs_the_paths_they_replace(B, T_in, flag, monkeypatch):
    """The cluster-local projection / prenet (tail_phase) against the grid-wide GEMM phases, and the all-sequences-at-once
    attention (attention_infer_multi) against the one-at-a-time phase: same operands, different summation order.  Frame 0
    must agree to fp32 rounding; three frames so that one bf16 rounding flip cannot run away."""
    from gantron_b200 import decoder_ops
    hp, dec, memory = _setup(B, T_in, 3, "bf16")
    run = lambda: decoder_ops.decoder_inference(dec, memory.cuda(), max_steps=3, gate_threshold=0.5, seed=31, prec=1,
                                                ignore_stop=True)
    a = run()
    monkeypatch.setenv(flag, "1")
    b = run()
    monkeypatch.delenv(flag)
    assert_close(a[2][:, 0], b[2][:, 0].cpu(), "alignments frame 0", atol=1e-6, rtol=1e-4)
    assert_close(a[0][:, :, 0], b[0][:, :, 0].cpu(), "mel frame 0", atol=2e-3, rtol=2e-3)
    assert_close(a[0], b[0].cpu(), "mel", atol=2e-2, rtol=2e-2)
    assert_close(a[2], b[2].cpu(), "alignments", atol=1e-2, rtol=0)


@pytest.mark.parametrize("B,T_in", [(8, 23), (70, 17), (200, 9)])
def test_infer_bf16_batched_stop_logic_fast_vs_generic(B, T_in, monkeypatch):
    """Per-row stop frames and the end of the loop in the FAST instantiation (gate test inside the clusters' tail, every CTA
    evaluates done[] behind the tail's grid barrier) against the generic instantiation (gate test in the projection GEMM
    phase).  Gate weights scaled so that rows fire at different frames; both sides see the same operands."""
    from gantron_b200 import decoder_ops
    steps = 12
    hp, dec, memory = _setup(B, T_in, steps, "bf16", seed=5)
    with torch.no_grad():
        dec.gate_layer.linear_layer.weight.mul_(40.0)
        dec.gate_layer.linear_layer.bias.zero_()
    # the gate feeds nothing back (model.py:411-417): a forced run gives every row's logit trajectory, and a bias at the 80 %
    # quantile of all logits makes a row fire at the first frame it exceeds it (different frames for different rows)
    g0 = decoder_ops.decoder_inference(dec, memory.cuda(), max_steps=steps, gate_threshold=0.5, seed=123, prec=1,
                                       ignore_stop=True)[1][:, :, 0].float().cpu()
    with torch.no_grad():
        dec.gate_layer.linear_layer.bias.fill_(-float(torch.quantile(g0.flatten(), 0.8)))
    run = lambda: decoder_ops.decoder_inference(dec, memory.cuda(), max_steps=steps, gate_threshold=0.5, seed=123, prec=1)
    mel, gate, align, lengths, n = run()
    monkeypatch.setenv("GTN_FWD2_GENERIC", "1")
    mel_g, gate_g, align_g, lengths_g, n_g = run()
    monkeypatch.delenv("GTN_FWD2_GENERIC")
    assert n == n_g and lengths.cpu().tolist() == lengths_g.cpu().tolist()
    assert len(set(lengths.cpu().tolist())) > 1, "rows should fire at different frames"
    assert int(lengths.max()) == n
    assert_close(mel, mel_g.cpu(), "mel", atol=2e-2, rtol=2e-2)
    assert_close(gate, gate_g.cpu(), "gate", atol=5e-2, rtol=5e-2)


def test_infer_stops_on_gate_like_reference():
    """gate bias pushed high -> sigmoid(gate) > 0.5 on the first frame: output has exactly 1 frame (the frame
    on which the gate fired is kept, model.py:482-487)."""
    from gantron_b200 import decoder_ops
    hp, dec, memory = _setup(1, 17, 8, "fp32")
    dec.gate_layer.linear_layer.bias.data.fill_(30.0)
    mel, gate, align, lengths, n = decoder_ops.decoder_inference(
        dec, memory.cuda(), max_steps=8, gate_threshold=0.5, seed=5, prec=3)
    om, og, oa, ol, _ = _oracle(hp, dec, memory, 8, 5, force=False)
    assert n == om.shape[-1] == 1 and tuple(mel.shape) == tuple(om.shape)
    assert int(lengths[0]) == int(ol[0]) == 1
    assert_close(mel, om, "mel", atol=2e-4, rtol=1e-3)


# ---------------------------------------------------------------------------------------------------------------
# batch 1: the on-chip-weights kernel (gtn_decoder_infer1.cu)
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("T_in,steps", [(17, 8), (150, 12), (211, 6)])
def test_infer1_batch1_kernel_matches_oracle(T_in, steps):
    """bf16 weights / fp32 activations; stated tolerance as the other bf16 inference tests (mel 1e-2 abs on |mel| ~ 1)."""
    from gantron_b200 import decoder_ops
    hp, dec, memory = _setup(1, T_in, steps, "bf16", seed=T_in)
    assert decoder_ops.infer1_supported(dec.packed().dims, 1, T_in, 1), "the batch-1 kernel should cover this shape"
    seed = 0x1B1
    mel, gate, align, lengths, n = decoder_ops.decoder_inference(
        dec, memory.cuda(), max_steps=steps, gate_threshold=0.5, seed=seed, prec=1, ignore_stop=True)
    om, og, oa, ol, _ = _oracle(hp, dec, memory, steps, seed)
    assert n == steps and mel.shape == (1, 80, steps) and gate.shape == (1, steps, 1) and align.shape == (1, steps, T_in)
    assert_close(align, oa, "alignments", atol=1e-3, rtol=0)
    assert_close(mel, om, "mel", atol=1e-2, rtol=1e-2)
    assert_close(gate, og, "gate", atol=1e-2, rtol=1e-2)


def test_infer1_batch1_kernel_vs_clustered_kernel_and_reproducible(monkeypatch):
    """Same call through both decode kernels (bf16): they agree to bf16 accuracy, and the batch-1 kernel is bit-reproducible
    (its cross-CTA sums are fixed-point)."""
    from gantron_b200 import decoder_ops
    hp, dec, memory = _setup(1, 150, 20, "bf16", seed=3)
    kw = dict(max_steps=20, gate_threshold=0.5, seed=99, prec=1, ignore_stop=True)
    a = decoder_ops.decoder_inference(dec, memory.cuda(), **kw)
    b = decoder_ops.decoder_inference(dec, memory.cuda(), **kw)
    assert torch.equal(a[0], b[0]) and torch.equal(a[2], b[2]), "batch-1 kernel is not bit-reproducible"
    monkeypatch.setenv("GANTRON_B200_INFER1", "0")
    c = decoder_ops.decoder_inference(dec, memory.cuda(), **kw)
    assert_close(a[0], c[0], "mel infer1 vs clustered", atol=1e-2, rtol=1e-2)
    assert_close(a[2], c[2], "alignments infer1 vs clustered", atol=1e-3, rtol=0)


def test_infer1_batch1_stops_on_gate():
    from gantron_b200 import decoder_ops
    hp, dec, memory = _setup(1, 33, 9, "bf16", seed=4)
    dec.gate_layer.linear_layer.bias.data.fill_(30.0)
    mel, gate, align, lengths, n = decoder_ops.decoder_inference(
        dec, memory.cuda(), max_steps=9, gate_threshold=0.5, seed=5, prec=1)
    assert n == 1 and tuple(mel.shape) == (1, 80, 1) and int(lengths[0]) == 1
    om, og, oa, ol, _ = _oracle(hp, dec, memory, 9, 5, force=False)
    assert om.shape[-1] == 1
    assert_close(mel, om, "mel", atol=1e-2, rtol=1e-2)
