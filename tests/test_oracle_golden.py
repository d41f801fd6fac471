"""CPU: the oracle (oracle/gantron_oracle.py) against the committed golden fixtures, which are outputs of the
UNMODIFIED reference run in the build container (oracle/make_golden.py).  This is what pins the oracle: the
reference has no tests or golden vectors of its own."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import gantron_oracle as O
from oracle.make_golden import GOLDEN_DIR, unpack_masks
from gantron_b200.hparams import HParams

TOL = dict(atol=2e-5, rtol=2e-4)       # fp32 CPU arithmetic on both sides; only summation order differs


def _load(name):
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    if not os.path.exists(path):
        pytest.fail(f"golden fixture {path} is missing (regenerate with python -m oracle.make_golden)")
    return np.load(path, allow_pickle=False)


def _close(got, ref, name, atol=TOL["atol"], rtol=TOL["rtol"]):
    got = torch.as_tensor(got).detach().double()
    ref = torch.as_tensor(ref).double()
    assert got.shape == ref.shape, f"{name}: shape {tuple(got.shape)} vs {tuple(ref.shape)}"
    err = (got - ref).abs()
    tol = atol + rtol * ref.abs()
    assert bool((err <= tol).all()), f"{name}: max err {float(err.max()):.3g} (ref absmax {float(ref.abs().max()):.3g})"


def _check_grads(z, grads: dict, prefix_filter=None, rel=1e-3):
    checked = 0
    for key in z.files:
        if not key.startswith("gnorm/"):
            continue
        name = key[len("gnorm/"):]
        if prefix_filter and not name.startswith(prefix_filter):
            continue
        g = grads[name].detach().float().reshape(-1)
        ref_norm = float(z[key])
        scale = max(ref_norm / max(g.numel(), 1) ** 0.5, 1e-12)       # rms of the reference gradient
        if f"grad/{name}" in z.files:
            ref = torch.from_numpy(z[f"grad/{name}"])
            got = g
        else:
            stride = int(z[f"gstride/{name}"])
            ref = torch.from_numpy(z[f"gsample/{name}"])
            got = g[::stride]
        err = float((got - ref).abs().max())
        # atol: gradients that are analytically zero (conv bias in front of BatchNorm) are pure rounding noise ~1e-7
        assert err <= rel * max(float(ref.abs().max()), scale) + 1e-6, f"grad {name}: max err {err:.3g} vs absmax {float(ref.abs().max()):.3g}"
        assert abs(float(g.double().norm()) - ref_norm) <= rel * ref_norm + 1e-5, f"grad norm {name}"
        checked += 1
    return checked


@pytest.mark.parametrize("name", ["train_vanilla", "train_combined"])
def test_oracle_train_step_matches_reference_golden(name):
    """Tacotron2.forward from the encoder outputs on + losses + frozen-D adversarial loss + backward."""
    z = _load(name)
    hp = HParams(str(z["hp_string"]))
    d = O.Dims.from_hparams(hp)
    shapes_g = {k: tuple(v) for k, v in json.loads(str(z["sd/G"])).items()}
    shapes_d = {k: tuple(v) for k, v in json.loads(str(z["sd/D"])).items()}
    wseed = int(z["wseed"])
    sd = {k: (v.clone().requires_grad_(True) if v.dtype.is_floating_point and "running" not in k else v)
          for k, v in O.seeded_state_dict(shapes_g, seed=wseed).items()}
    sdd = O.seeded_state_dict(shapes_d, seed=wseed + 1)
    masks = unpack_masks(z)
    n_g = int(z["n_g_masks"])
    T = z["in/mel_tgt"].shape[2]
    enc_masks, rest = masks[:3], masks[3:n_g]
    assert len(rest) == 2 + 2 * T + 5
    mk = dict(pre1=rest[0], pre2=rest[1], att=torch.stack(rest[2:2 + 2 * T:2]), dec=torch.stack(rest[3:3 + 2 * T:2]),
              post=rest[2 + 2 * T:])
    dk = masks[n_g:]
    assert len(dk) == 4
    memory = torch.from_numpy(z["in/memory"]).requires_grad_(True)
    mel_tgt, gate_tgt = torch.from_numpy(z["in/mel_tgt"]), torch.from_numpy(z["in/gate_tgt"])
    in_len, out_len = torch.from_numpy(z["in/in_len"]), torch.from_numpy(z["in/out_len"])
    mel, mel_post, gate, align = O.generator_from_memory(sd, memory, mel_tgt, in_len, out_len, d, mk, training=True)
    for n, t in (("mel", mel), ("mel_post", mel_post), ("gate", gate), ("align", align)):
        _close(t, z[f"out/{n}"], f"{name}:{n}")
    mel_loss, gate_loss = O.tacotron2_losses(mel, mel_post, gate, mel_tgt, gate_tgt)
    attn = O.attention_loss(align, in_len, out_len)
    adv = O.adversarial_loss(sdd, mel_post, out_len, d, dk)
    for n, t in (("mel_loss", mel_loss), ("gate_loss", gate_loss), ("attn_loss", attn), ("adv", adv)):
        _close(t, z[f"out/{n}"], f"{name}:{n}", atol=1e-6, rtol=1e-4)
    (mel_loss + gate_loss + adv + 10 * attn).backward()
    grads = {k: v.grad for k, v in sd.items() if getattr(v, "grad", None) is not None}
    grads["enc_out"] = memory.grad[:, :, :d.enc_dim + d.spk + d.emo]
    n = _check_grads(z, grads)
    assert n >= 40, f"only {n} gradients were compared"


def test_oracle_postnet_running_stats_match_reference_golden():
    z = _load("train_vanilla")
    hp = HParams(str(z["hp_string"]))
    d = O.Dims.from_hparams(hp)
    shapes_g = {k: tuple(v) for k, v in json.loads(str(z["sd/G"])).items()}
    sd = O.seeded_state_dict(shapes_g, seed=int(z["wseed"]))
    masks = unpack_masks(z)
    n_g = int(z["n_g_masks"])
    stats = {}
    # mel input of the postnet = the reference's UNMASKED decoder output; rebuild it with the oracle
    T = z["in/mel_tgt"].shape[2]
    rest = masks[3:n_g]
    mel, _, _ = O.decoder_teacher_forced(sd, torch.from_numpy(z["in/memory"]), torch.from_numpy(z["in/mel_tgt"]),
                                         torch.from_numpy(z["in/in_len"]), d, rest[0], rest[1],
                                         torch.stack(rest[2:2 + 2 * T:2]), torch.stack(rest[3:3 + 2 * T:2]))
    O.postnet(sd, mel, d, rest[2 + 2 * T:], True, stats_out=stats)
    for i in range(d.postnet_n):
        m, v = stats[i]
        rm = 0.9 * sd[f"postnet.convolutions.{i}.1.running_mean"] + 0.1 * m
        rv = 0.9 * sd[f"postnet.convolutions.{i}.1.running_var"] + 0.1 * v
        _close(rm, z[f"bn/postnet.convolutions.{i}.1.running_mean"], f"running_mean{i}")
        _close(rv, z[f"bn/postnet.convolutions.{i}.1.running_var"], f"running_var{i}")


def test_oracle_inference_matches_reference_golden():
    """Decoder.inference at B = 1 for the first N frames with identical prenet dropout masks, then eval Postnet."""
    z = _load("infer_vanilla_b1")
    hp = HParams(str(z["hp_string"]))
    d = O.Dims.from_hparams(hp)
    import gantron_b200.model as M
    shapes = {k: tuple(v.shape) for k, v in M.Tacotron2(HParams(str(z["hp_string"]))).state_dict().items()}
    sd = O.seeded_state_dict(shapes, seed=int(z["wseed"]))
    masks = unpack_masks(z)
    steps = int(z["steps"])
    assert len(masks) == 2 * steps
    memory = torch.cat((torch.from_numpy(z["in/memory"]), torch.from_numpy(z["in/style"])), -1)
    mel, gate, align, lengths, _ = O.decoder_inference(sd, memory, d, lambda s: (masks[2 * s], masks[2 * s + 1]), max_steps=steps)
    assert mel.shape[2] == steps and int(lengths[0]) == steps
    _close(mel, z["out/mel"], "mel")
    _close(gate, z["out/gate"], "gate")
    _close(align, z["out/align"], "align")
    post = mel + O.postnet(sd, mel, d, None, training=False)
    _close(post, z["out/mel_post"], "mel_post")


def test_oracle_discriminator_matches_reference_golden():
    z = _load("disc_vanilla")
    d = O.Dims.from_hparams(HParams("use_noise=True,use_labels=False"))
    import gantron_b200.model as M
    shapes = {k: tuple(v.shape) for k, v in M.Discriminator(HParams("use_noise=True,use_labels=False")).state_dict().items()}
    sd = {k: v.clone().requires_grad_(True) for k, v in O.seeded_state_dict(shapes, seed=int(z["wseed"])).items()}
    masks = unpack_masks(z)
    assert len(masks) == 8           # adversarial_loss pass + the extra forward pass
    mel = torch.from_numpy(z["in/mel"]).requires_grad_(True)
    lengths = torch.from_numpy(z["in/lengths"])
    loss = O.adversarial_loss(sd, mel, lengths, d, masks[:4])
    scores = O.discriminator_forward(sd, mel.transpose(1, 2), d, masks[4:])
    _close(loss, z["out/loss"], "loss", atol=1e-7, rtol=1e-4)
    _close(scores, z["out/scores"], "scores")
    loss.backward()
    grads = {k: v.grad for k, v in sd.items()}
    grads["mel"] = mel.grad
    assert _check_grads(z, grads) == 11


def test_oracle_linear_discriminator_matches_reference_golden():
    """LinearDiscriminator.adversarial_loss (model.py:567-583), windows drawn with the seeded module-level `random`."""
    import random
    z = _load("lin_disc_vanilla")
    hp = HParams("use_noise=True,use_labels=False")
    d = O.Dims.from_hparams(hp)
    import gantron_b200.model as M
    shapes = {k: tuple(v.shape) for k, v in M.LinearDiscriminator(hp).state_dict().items()}
    sd = {k: v.clone().requires_grad_(True) for k, v in O.seeded_state_dict(shapes, seed=int(z["wseed"])).items()}
    keeps = unpack_masks(z)
    assert len(keeps) == 3 and keeps[0].shape[0] == int(z["out/rows"])
    mel = torch.from_numpy(z["in/mel"]).requires_grad_(True)
    random.seed(int(z["pyseed"]))
    loss = O.linear_adversarial_loss(sd, mel, torch.from_numpy(z["in/lengths"]), d, keeps)
    _close(loss, z["out/loss"], "loss", atol=1e-7, rtol=1e-4)
    loss.backward()
    grads = {k: v.grad for k, v in sd.items()}
    grads["mel"] = mel.grad
    assert _check_grads(z, grads) == 9


def test_state_dict_contract_matches_reference_golden():
    """The drop-in boundary: gantron_b200.model has exactly the reference's state_dict keys and shapes
    (strict load_state_dict of reference checkpoints, train.py:139,147; hparams.ignore_layers names)."""
    import gantron_b200.model as M
    for name in ("train_vanilla", "train_combined"):
        z = _load(name)
        hp = HParams(str(z["hp_string"]))
        ours_g = {k: list(v.shape) for k, v in M.Tacotron2(hp).state_dict().items()}
        ours_d = {k: list(v.shape) for k, v in M.Discriminator(hp).state_dict().items()}
        assert ours_g == json.loads(str(z["sd/G"])), name
        assert ours_d == json.loads(str(z["sd/D"])), name
        assert list(ours_g) == list(json.loads(str(z["sd/G"]))), "key order differs"


def test_oracle_train_step_long_matches_reference_golden():
    """The benchmark's sequence shape (T_in = 150 ragged, 208 frames): inputs, encoder stand-in and Philox masks are
    regenerated from the recorded seeds; the reference's noise draw order (model.py:275) is checked on the way."""
    from oracle.make_golden import philox_train_masks, seeded_encoder_outputs
    import gantron_b200.model as M
    z = _load("train_vanilla_long")
    hp = HParams(str(z["hp_string"]))
    d = O.Dims.from_hparams(hp)
    B, T_in, T_out, wseed, mseed = (int(z[k]) for k in ("B", "T_in", "T_out", "wseed", "mseed"))
    shapes_g = {k: tuple(v.shape) for k, v in M.Tacotron2(HParams(str(z["hp_string"]))).state_dict().items()}
    shapes_d = {k: tuple(v.shape) for k, v in M.Discriminator(hp).state_dict().items()}
    sd = {k: (v.clone().requires_grad_(True) if v.dtype.is_floating_point and "running" not in k else v)
          for k, v in O.seeded_state_dict(shapes_g, seed=wseed).items()}
    sdd = O.seeded_state_dict(shapes_d, seed=wseed + 1)
    text, in_len, mel_tgt, gate_tgt, spk, emo, out_len = O.synthetic_batch(B, T_in, T_out, seed=wseed + 2, ragged=True,
                                                                           n_symbols=hp.n_symbols)
    enc = seeded_encoder_outputs(B, T_in, d.enc_dim, wseed + 3).requires_grad_(True)
    torch.manual_seed(mseed + 1)
    noise = torch.rand(B, 1, d.noise)                        # Decoder.append_noise: first CPU draw of the pass
    _close(noise[:, 0], z["in/noise"], "noise draw", atol=0, rtol=0)
    memory = torch.cat((enc, noise.repeat_interleave(T_in, dim=1)), -1)
    mk, dk = philox_train_masks(mseed, B, T_out, d, (T_out + d.disc_window - 1) // d.disc_window)
    mel, mel_post, gate, align = O.generator_from_memory(sd, memory, mel_tgt, in_len, out_len, d, mk, training=True)
    s = int(z["align_stride"])
    for n, t in (("mel", mel), ("mel_post", mel_post), ("gate", gate), ("align_strided", align[:, ::s]),
                 ("align_colsum", align.sum(1))):
        _close(t, z[f"out/{n}"], f"long:{n}", atol=5e-5, rtol=5e-4)
    mel_loss, gate_loss = O.tacotron2_losses(mel, mel_post, gate, mel_tgt, gate_tgt)
    attn = O.attention_loss(align, in_len, out_len)
    adv = O.adversarial_loss(sdd, mel_post, out_len, d, dk)
    for n, t in (("mel_loss", mel_loss), ("gate_loss", gate_loss), ("attn_loss", attn), ("adv", adv)):
        _close(t, z[f"out/{n}"], f"long:{n}", atol=1e-6, rtol=1e-4)
    (mel_loss + gate_loss + adv + 10 * attn).backward()
    grads = {k: v.grad for k, v in sd.items() if getattr(v, "grad", None) is not None}
    grads["enc_out"] = enc.grad
    assert _check_grads(z, grads) >= 40


def test_oracle_wrapper_inference_matches_reference_golden():
    """Tacotron2.inference, combined labels+noise config: the wrapper's random emotion draw (model.py:691-692), speaker
    and emotion columns (model.py:704-712) and the decoder's own style draw (model.py:468-469), in that RNG order."""
    import gantron_b200.model as M
    from oracle import philox as PX
    z = _load("infer_combined_wrapper")
    hp = HParams(str(z["hp_string"]))
    shapes = {k: tuple(v.shape) for k, v in M.Tacotron2(HParams(str(z["hp_string"]))).state_dict().items()}
    d = O.Dims.from_hparams(hp)
    assert (d.spk, d.emo, d.noise, d.d_mem) == (64, 5, 512, 1093)
    sd = O.seeded_state_dict(shapes, seed=int(z["wseed"]))
    mseed, steps = int(z["mseed"]), int(z["steps"])
    enc = torch.from_numpy(z["in/enc"])
    T_in = enc.shape[1]
    torch.manual_seed(mseed + 1)
    emotions = torch.rand(1, 5)
    spk = sd["speaker_embedding.weight"][torch.from_numpy(z["in/speaker"])]
    mem_in = torch.cat((enc, spk[:, None].repeat(1, T_in, 1), emotions[:, None].repeat(1, T_in, 1)), 2)
    _close(mem_in, z["in/decoder_memory"], "decoder memory (speaker / emotion columns)", atol=0, rtol=0)
    style = torch.rand(1, 1, d.noise).repeat_interleave(T_in, dim=1)
    memory = torch.cat((mem_in, style), -1)
    keep = lambda s: (torch.from_numpy(PX.keep_mask(mseed, PX.SITE_PRENET1, s, (1, d.prenet), 0.5)),
                      torch.from_numpy(PX.keep_mask(mseed, PX.SITE_PRENET2, s, (1, d.prenet), 0.5)))
    mel, gate, align, lengths, _ = O.decoder_inference(sd, memory, d, keep, max_steps=steps)
    _close(mel, z["out/mel"], "mel")
    _close(gate, z["out/gate"], "gate")
    _close(align, z["out/align"], "align")
    _close(mel + O.postnet(sd, mel, d, None, training=False), z["out/mel_post"], "mel_post")


@pytest.mark.parametrize("name", ["train_vanilla", "train_combined"])
def test_oracle_encoder_matches_reference_golden(name):
    """Text encoder: oracle.encoder_forward on the fixture's text batch reproduces the encoder outputs the
    unmodified reference fed to its decoder (the first 512 columns of the captured `memory`)."""
    z = _load(name)
    hp = HParams(str(z["hp_string"]))
    shapes_g = {k: tuple(v) for k, v in json.loads(str(z["sd/G"])).items()}
    sd = O.seeded_state_dict(shapes_g, seed=int(z["wseed"]))
    masks = unpack_masks(z)[:3]
    text, in_len = torch.from_numpy(z["in/text"]), torch.from_numpy(z["in/in_len"])
    emb = torch.nn.functional.embedding(text, sd["embedding.weight"]).transpose(1, 2)
    with torch.no_grad():
        enc = O.encoder_forward(sd, emb, in_len, masks, training=True)
    _close(enc, z["in/memory"][:, :, :hp.encoder_embedding_dim], f"{name}: encoder outputs")
