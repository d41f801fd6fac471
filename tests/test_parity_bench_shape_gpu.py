"""Parity AT THE BENCHMARKED SHAPE AND PRECISION (BASELINE.json configs[1] / configs[2]): T_in = 150 (ragged), T_out = 800,
fp32-accurate mode AND bf16 (the precision bench.py times), the B = 64 kernel template, 1000-step autoregressive decode
and batch > 1 per-row stopping -- CUDA path (through the C ABI) against the CPU oracle with identical Philox masks.

Bounds:
  fp32 mode   outputs and every gradient <= 1e-3 of the tensor's max-abs (the north star's bound), also at frame 800.
  bf16 mode   operands (weights, recurrent states, contexts, encoder memory, saved activations) rounded to bf16, fp32
              accumulation and cell state.  Stated bound at T_out = 800: mel / gate max-abs error <= 1e-2 and mean-abs
              error <= 2e-3 (|mel| ~ 1.3), alignments <= 1e-3 absolute, gradients <= 2e-2 relative Frobenius error per
              tensor, except the two prenet weights (<= 8e-2: relu decisions of bf16-rounded pre-activations near 0 flip).
              Measured on B200 (profiles/r2_parity_bench_shape.md): mel 4.1e-3 / 8e-4, alignments 2.7e-4, gradients
              <= 9e-3 (prenet 5.0e-2); no growth of the error with the frame index in either mode.
Each long case also writes its error-vs-frame curve to gpurun_out/parity_*.json (copied to profiles/ by hand).
"""
import json
import os

import pytest
import torch

from conftest import assert_close
import test_decoder_train_gpu as TD
import test_decoder_infer_gpu as TI

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ORACLE_CACHE = {}


def _dump(name, payload):
    try:
        d = os.path.join(ROOT, "gpurun_out")
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, f"parity_{name}.json"), "w") as f:
            json.dump(payload, f)
    except OSError:
        pass


def _build(B, T_in, T, prec_name, in_len, seed=0):
    hp, dec, memory, mels, _, R = TD._build(B, T_in, T, prec_name, ragged=False, seed=seed)
    return hp, dec, memory, mels, torch.tensor(in_len).long(), R


def _oracle_cached(key, hp, dec, memory, mels, in_len, R, seed):
    if key not in _ORACLE_CACHE:
        _ORACLE_CACHE[key] = TD._oracle(hp, dec, memory, mels, in_len, R, seed)
    return _ORACLE_CACHE[key]


def _frame_curve(got, ref, frame_dim):
    """max-abs error per output frame, as a list."""
    err = (got.detach().float().cpu() - ref.detach().float()).abs()
    dims = [i for i in range(err.dim()) if i != frame_dim]
    return [float(v) for v in err.amax(dim=dims)]


def _grad_errors(dec, sd, mem_grad, o_mem_grad):
    out = {}
    for k, p in dec.named_parameters():
        ref = sd["decoder." + k].grad
        g = p.grad.detach().cpu()
        out[k] = dict(max_rel=float((g - ref).abs().max()) / (float(ref.abs().max()) + 1e-12),
                      fro_rel=float((g - ref).norm()) / (float(ref.norm()) + 1e-12))
    out["memory"] = dict(max_rel=float((mem_grad.cpu() - o_mem_grad).abs().max()) / (float(o_mem_grad.abs().max()) + 1e-12),
                         fro_rel=float((mem_grad.cpu() - o_mem_grad).norm()) / (float(o_mem_grad.norm()) + 1e-12))
    return out


LONG = dict(B=4, T_in=150, T=800, in_len=[150, 143, 129, 101], seed=0x5EED0800)


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_train_t150_t800_values_and_grads(prec):
    """Decoder.forward + BPTT at T_in = 150 ragged / T_out = 800 (the benchmark's sequence shape), B = 4."""
    c = LONG
    hp, dec, memory, mels, in_len, R = _build(c["B"], c["T_in"], c["T"], prec, c["in_len"])
    mel, gate, align, g_mem = TD._run_ours(dec, memory, mels, in_len, R, c["seed"])
    om, og, oa, sd, o_gmem = _oracle_cached("long", hp, dec, memory, mels, in_len, R, c["seed"])
    ge = _grad_errors(dec, sd, g_mem, o_gmem)
    curve = dict(mel=_frame_curve(mel, om, 2), gate=_frame_curve(gate, og, 1), align=_frame_curve(align, oa, 1))
    mel_err = (mel.detach().cpu() - om).abs()
    summary = dict(prec=prec, shape=dict(B=c["B"], T_in=c["T_in"], T_out=c["T"], in_len=c["in_len"]),
                   mel_max=float(mel_err.max()), mel_mean=float(mel_err.mean()), mel_ref_absmax=float(om.abs().max()),
                   gate_max=max(curve["gate"]), align_max=max(curve["align"]),
                   mel_max_first100=max(curve["mel"][:100]), mel_max_last100=max(curve["mel"][-100:]),
                   worst_grad_max_rel=max(v["max_rel"] for v in ge.values()),
                   worst_grad_fro_rel=max(v["fro_rel"] for v in ge.values()))
    _dump(f"train_t800_{prec}", dict(summary=summary, grads=ge, curve_stride8={k: v[::8] for k, v in curve.items()}))
    print("PARITY", json.dumps(summary))
    if prec == "fp32":
        assert_close(align, oa, "alignments", atol=2e-5, rtol=1e-3)
        assert_close(mel, om, "mel", atol=1e-3 * float(om.abs().max()), rtol=0)
        assert_close(gate, og, "gate", atol=1e-3 * max(float(og.abs().max()), 1.0), rtol=0)
        TD._check_grads(dec, sd, g_mem, o_gmem, rel=1e-3)
    else:
        assert_close(align, oa, "alignments bf16", atol=1e-3, rtol=0)
        assert summary["mel_max"] <= 1e-2 and summary["mel_mean"] <= 2e-3, summary
        assert summary["gate_max"] <= 1e-2, summary
        # no drift: the last 100 frames are not worse than twice the first 100
        assert summary["mel_max_last100"] <= 2.0 * summary["mel_max_first100"] + 1e-3, summary
        bad = {k: v["fro_rel"] for k, v in ge.items() if v["fro_rel"] > (8e-2 if k.startswith("prenet.") else 2e-2)}
        assert not bad, f"bf16 gradient error (relative Frobenius) above the stated bound: {bad}"


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_train_b64_template_t150(prec):
    """B = 64 (the <64> kernel template of the benchmark), T_in = 150 ragged (75 positions per CTA of a sequence pair),
    32 frames."""
    B, T_in, T, seed = 64, 150, 32, 0xB64
    in_len = [max(150 - 3 * i, 60) for i in range(B)]
    hp, dec, memory, mels, in_len, R = _build(B, T_in, T, prec, in_len, seed=3)
    mel, gate, align, g_mem = TD._run_ours(dec, memory, mels, in_len, R, seed)
    om, og, oa, sd, o_gmem = _oracle_cached("b64", hp, dec, memory, mels, in_len, R, seed)
    if prec == "fp32":
        assert_close(align, oa, "alignments", atol=2e-5, rtol=1e-3)
        assert_close(mel, om, "mel", atol=3e-4, rtol=1e-3)
        assert_close(gate, og, "gate", atol=3e-4, rtol=1e-3)
        TD._check_grads(dec, sd, g_mem, o_gmem, rel=1e-3)
    else:
        assert_close(align, oa, "alignments bf16", atol=1e-2, rtol=0)
        assert_close(mel, om, "mel bf16", atol=5e-2, rtol=5e-2)
        TD._check_grads(dec, sd, g_mem, o_gmem, rel=5e-2, l2=True)


@pytest.mark.parametrize("T_in", [65, 129, 256])
def test_train_fp32_tin_boundaries(T_in):
    """Attention halves of 33 / 65 / 128 positions (tile-row and warp-trip boundaries), ragged, 32 frames."""
    B, T, seed = 3, 32, 0x7100 + T_in
    hp, dec, memory, mels, in_len, R = _build(B, T_in, T, "fp32", [T_in, T_in - 1, (T_in * 2) // 3], seed=T_in)
    mel, gate, align, g_mem = TD._run_ours(dec, memory, mels, in_len, R, seed)
    om, og, oa, sd, o_gmem = TD._oracle(hp, dec, memory, mels, in_len, R, seed)
    assert_close(align, oa, "alignments", atol=2e-5, rtol=1e-3)
    assert_close(mel, om, "mel", atol=3e-4, rtol=1e-3)
    TD._check_grads(dec, sd, g_mem, o_gmem, rel=1e-3)


# ---------------------------------------------------------------------------------------------------------------
# autoregressive decode
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_infer_b1_1000_forced_steps(prec):
    """BASELINE configs[2]: batch 1, max_decoder_steps = 1000, T_in = 150 -- every frame against the oracle."""
    from gantron_b200 import decoder_ops
    steps, seed = 1000, 0x1000
    hp, dec, memory = TI._setup(1, 150, steps, prec, seed=5)
    mel, gate, align, lengths, n = decoder_ops.decoder_inference(
        dec, memory.cuda(), max_steps=steps, gate_threshold=0.5, seed=seed, prec=3 if prec == "fp32" else 1,
        ignore_stop=True)
    key = "infer1000"
    if key not in _ORACLE_CACHE:
        _ORACLE_CACHE[key] = TI._oracle(hp, dec, memory, steps, seed)
    om, og, oa, ol, _ = _ORACLE_CACHE[key]
    assert n == steps and mel.shape == (1, 80, steps) and gate.shape == (1, steps, 1)
    curve = dict(mel=_frame_curve(mel, om, 2), align=_frame_curve(align, oa, 1))
    summary = dict(prec=prec, steps=steps, mel_max=max(curve["mel"]), align_max=max(curve["align"]),
                   mel_max_first100=max(curve["mel"][:100]), mel_max_last100=max(curve["mel"][-100:]),
                   mel_ref_absmax=float(om.abs().max()))
    _dump(f"infer_b1_1000_{prec}", dict(summary=summary, curve_stride10={k: v[::10] for k, v in curve.items()}))
    print("PARITY", json.dumps(summary))
    if prec == "fp32":
        assert_close(align, oa, "alignments", atol=2e-5, rtol=1e-3)
        assert_close(mel, om, "mel", atol=1e-3 * float(om.abs().max()), rtol=0)
        assert_close(gate, og, "gate", atol=1e-3 * max(float(og.abs().max()), 1.0), rtol=0)
    else:
        # autoregressive feedback of the bf16-rounded mel: stated bound 1e-2 abs on |mel| ~ 1.4 over 1000 frames
        # (measured 2.9e-3, flat over the frames)
        assert_close(align, oa, "alignments bf16", atol=1e-3, rtol=0)
        assert summary["mel_max"] <= 1e-2, summary
        assert summary["mel_max_last100"] <= 2.0 * summary["mel_max_first100"] + 1e-3, summary


def _gate_setup(B, T_in, steps, seed, n_never, gap=0.04):
    """Decoder whose rows fire their gate at DIFFERENT, planned frames (or never).  The gate output feeds nothing back
    (model.py:411-417), so re-weighting the gate layer does not change the trajectories: its weights on the per-sequence
    constant noise columns of the context are solved so that they add a chosen offset to each row's logit (a per-row
    bias), placed halfway below a record high of that row's logit trajectory (margin >= gap / 2 against rounding)."""
    hp, dec, memory = TI._setup(B, T_in, steps, "fp32", seed=seed)
    Hd = hp.decoder_rnn_dim
    with torch.no_grad():
        w = dec.gate_layer.linear_layer.weight
        w *= 30.0
        w[:, Hd + 512:] = 0.0
        dec.gate_layer.linear_layer.bias.zero_()
    g0 = TI._oracle(hp, dec, memory, steps, seed)[1][:, :, 0]                # forced run: logits of every row and frame
    targets = torch.linspace(3, steps - 4, B - n_never).round().tolist()
    level, plan = torch.empty(B), []
    for b in range(B):
        run_max = torch.cummax(g0[b], 0).values
        cands = [f for f in range(2, steps - 1) if float(g0[b, f] - run_max[f - 1]) >= gap]
        if b >= B - n_never or not cands:
            level[b] = float(g0[b].max()) + 1.0
            plan.append(steps)
            continue
        f = min(cands, key=lambda f: abs(f - targets[b]))
        level[b] = float(g0[b, f] + run_max[f - 1]) / 2
        plan.append(f + 1)
    noise = memory[:, 0, 512:].double()
    wc = noise.t() @ torch.linalg.solve(noise @ noise.t(), (-level).double()[:, None])
    with torch.no_grad():
        dec.gate_layer.linear_layer.weight[:, Hd + 512:] = wc.t().float().to(w.device)
    return hp, dec, memory, plan


def test_infer_batch8_rows_stop_independently():
    """Batch extension (the reference raises at model.py:486 for B > 1): per-row `done` flags.  Half of the rows fire
    their gate at different frames, the others never: the loop runs to max_decoder_steps, and every row equals a B = 1
    run of the oracle (= the reference's loop) cut at its own stop frame."""
    from gantron_b200 import decoder_ops
    from oracle import gantron_oracle as O, philox as PX
    B, T_in, steps, seed = 8, 37, 24, 21
    hp, dec, memory, plan = _gate_setup(B, T_in, steps, seed, 3)
    mel, gate, align, lengths, n = decoder_ops.decoder_inference(
        dec, memory.cuda(), max_steps=steps, gate_threshold=0.5, seed=seed, prec=3)
    d = O.Dims.from_hparams(hp)
    d.spk = d.emo = 0
    sd = {"decoder." + k: v.detach().float().cpu() for k, v in dec.state_dict().items()}
    fired = 0
    for b in range(B):
        def keep_fn(t, b=b):      # row b of the batch masks (element numbering b * P + unit)
            return tuple(torch.from_numpy(PX.keep_mask(seed, s, t, (B, hp.prenet_dim), 0.5))[b:b + 1]
                         for s in (PX.SITE_PRENET1, PX.SITE_PRENET2))
        with torch.no_grad():
            om, og, oa, ol, _ = O.decoder_inference(sd, memory[b:b + 1], d, keep_fn, max_steps=steps)
        Lb = int(ol[0])
        assert Lb == plan[b], f"row {b}: oracle stop frame {Lb}, planned {plan[b]}"
        assert int(lengths[b]) == Lb, f"row {b}: stop frame {int(lengths[b])} vs oracle {Lb}"
        fired += Lb < steps
        assert_close(mel[b:b + 1, :, :Lb], om, f"mel row {b}", atol=2e-4, rtol=1e-3)
        assert_close(gate[b:b + 1, :Lb], og, f"gate row {b}", atol=2e-4, rtol=1e-3)
        assert_close(align[b:b + 1, :Lb], oa, f"align row {b}", atol=2e-5, rtol=1e-3)
    assert 0 < fired < B and n == steps


def test_infer_batch8_loop_ends_when_all_rows_fired():
    """All rows fire (at different frames): the launch ends at the LAST row's stop frame, well before max_decoder_steps."""
    from gantron_b200 import decoder_ops
    B, T_in, steps, seed = 8, 29, 40, 33
    hp, dec, memory, plan = _gate_setup(B, T_in, steps, seed, 0)
    om, og, oa, ol, _ = TI._oracle(hp, dec, memory, steps, seed, force=False)
    assert ol.tolist() == plan
    mel, gate, align, lengths, n = decoder_ops.decoder_inference(
        dec, memory.cuda(), max_steps=steps, gate_threshold=0.5, seed=seed, prec=3)
    assert n == om.shape[2] == int(ol.max()) and n < steps
    assert lengths.cpu().tolist() == ol.tolist() and len(set(ol.tolist())) > 1
    assert_close(mel, om, "mel", atol=2e-4, rtol=1e-3)
    assert_close(align, oa, "alignments", atol=2e-5, rtol=1e-3)


def test_decoder_inference_module_api_b256_shapes():
    """Decoder.inference (the module API bench.py times) at B = 256 in bf16 for a few frames against the oracle."""
    hp, dec, memory = TI._setup(256, 150, 4, "bf16", seed=9)
    dec._forced_seed = 77
    enc, style = memory[:, :, :512].cuda(), memory[:, :, 512:].cuda()
    with torch.no_grad():
        mel, gate, align = dec.inference(enc, style)
    om, og, oa, _, _ = TI._oracle(hp, dec, memory, 4, 77, force=False)
    assert mel.shape == (256, 80, 4) and gate.shape == (256, 4, 1) and align.shape == (256, 4, 150)
    assert_close(align, oa, "alignments bf16", atol=2e-2, rtol=0)
    assert_close(mel, om, "mel bf16", atol=5e-2, rtol=5e-2)
