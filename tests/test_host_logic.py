"""CPU: host-side logic of the drop-in surface -- operand packing, hparams parsing, the synthetic batch
contract, module construction / attribute paths the reference's callers rely on."""
import torch

from gantron_b200 import packing
from gantron_b200.hparams import HParams
from gantron_b200.synth import synthetic_batch


def test_three_plane_split_is_fp32_accurate():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(64, 100, generator=g) * 3
    hi, lo, lo2 = packing.split3(x)
    rec = hi.float() + lo.float() + lo2.float()
    assert float((rec - x).abs().max()) <= 2 ** -22 * float(x.abs().max())


def test_expand_layout_reproduces_fp32_product():
    g = torch.Generator().manual_seed(1)
    w, x = torch.randn(16, 40, generator=g), torch.randn(8, 40, generator=g)
    for prec, tol in ((1, 3e-2), (3, 2e-6)):
        we, xe = packing.expand(w, 64, prec, True), packing.expand(x, 64, prec, False)
        assert we.shape == (16, 64 * packing.mult_of(prec)) and we.dtype == torch.bfloat16
        got = we.double() @ xe.double().t()
        ref = w.double() @ x.double().t()
        assert float((got - ref).abs().max()) <= tol * float(ref.abs().max()), prec


def test_pack_weight_segments_and_row_padding():
    w = torch.arange(6 * 10, dtype=torch.float32).reshape(6, 10)
    p = packing.pack_weight([(w[:, :3], 32), (w[:, 3:], 32)], prec=1, row_pad=16)
    assert p.shape == (16, 64)
    assert torch.equal(p[:6, :3].float(), w[:, :3]) and torch.equal(p[:6, 32:39].float(), w[:, 3:])
    assert float(p[6:].abs().sum()) == 0 and float(p[:, 3:32].abs().sum()) == 0


def test_hparams_string_overrides_and_ctor_mutation():
    hp = HParams("use_noise=True,use_labels=False,batch_size=64,vesus_path=data/vesus,b200_precision=fp32")
    assert hp.use_noise is True and hp.batch_size == 64 and hp.vesus_path == "data/vesus" and hp.b200_precision == "fp32"
    from gantron_b200 import model as M
    hp2 = HParams("use_noise=False,use_labels=True")
    M.Tacotron2(hp2)
    # the reference ctor mutates hparams (model.py:589-591)
    assert hp2.noise_size == 0 and hp2.use_labels is False


def test_decoder_dims_per_configuration():
    from gantron_b200 import model as M
    assert M.Decoder(HParams("use_noise=True,use_labels=False")).decoder_in_dim == 1024
    hp = HParams("use_noise=True,use_labels=True,vesus_path=x")
    g = M.Tacotron2(hp)
    assert g.decoder.decoder_in_dim == 1093 and g.decoder.attention_rnn.weight_ih.shape == (4096, 1349)
    assert hasattr(g, "speaker_embedding") and g.decoder.attention_layer.score_mask_value == -float("inf")
    hp = HParams("use_noise=False,use_labels=True,vesus_path=x")
    assert M.Tacotron2(hp).decoder.decoder_in_dim == 581


def test_synthetic_batch_follows_the_collate_contract():
    text, in_len, mel, gate, spk, emo, out_len = synthetic_batch(5, 30, 60, seed=3, ragged=True)
    assert text.shape == (5, 30) and mel.shape == (5, 80, 60) and gate.shape == (5, 60) and emo.shape == (5, 5)
    assert bool((in_len[:-1] >= in_len[1:]).all()) and int(in_len.max()) == 30      # sorted by text length
    for b in range(5):
        assert float(mel[b, :, out_len[b]:].abs().sum()) == 0
        assert float(gate[b, :out_len[b] - 1].sum()) == 0 and float(gate[b, out_len[b] - 1:].min()) == 1
        assert int(text[b, in_len[b]:].abs().sum()) == 0
    assert float(mel.max()) <= 2.0 and float(mel.min()) >= -11.5


def test_parse_output_masks_through_data():
    from gantron_b200 import model as M
    g = M.Tacotron2(HParams("use_noise=True,use_labels=False"))
    mel = torch.ones(2, 80, 6, requires_grad=True) * 1.0
    post, gate, al = mel * 2, torch.zeros(2, 6), torch.zeros(2, 6, 4)
    out = g.parse_output([mel, post, gate, al], torch.tensor([6, 4]))
    assert float(out[0][1, :, 4:].abs().sum()) == 0 and float(out[1][1, :, 4:].abs().sum()) == 0
    assert float(out[2][1, 4:].min()) == 1e3 and float(out[2][0].abs().sum()) == 0
    out[1].sum().backward()            # `.data` masking: the graph does not see it (model.py:638-640)


def test_collate_pads_sorts_and_marks_the_stop_frames():
    """gantron_b200.data_utils.TextMelCollate (reference data_utils.py:82-131): decreasing text length, zero padding, frames
    rounded up to n_frames_per_step, gate = 1 from the last real frame on."""
    import torch
    from gantron_b200.data_utils import TextMelCollate
    g = torch.Generator().manual_seed(0)
    items = [[torch.randint(1, 50, (L,), generator=g), torch.randn(80, F_, generator=g), float(i), torch.rand(5, generator=g)]
             for i, (L, F_) in enumerate([(5, 30), (9, 44), (2, 21), (9, 25)])]
    text, in_len, mel, gate, spk, emo, out_len = TextMelCollate(4, pin_memory=False)(items)
    assert in_len.tolist() == sorted([5, 9, 2, 9], reverse=True)
    assert mel.shape == (4, 80, 44) and text.shape == (4, 9) and gate.shape == (4, 44)
    order = [int(s) for s in spk.tolist()]
    assert sorted(order) == [0, 1, 2, 3] and [items[o][0].numel() for o in order] == in_len.tolist()
    for row, o in enumerate(order):
        L, F_ = items[o][0].numel(), items[o][1].size(1)
        assert torch.equal(text[row, :L], items[o][0]) and int(text[row, L:].abs().sum()) == 0
        assert torch.equal(mel[row, :, :F_], items[o][1]) and float(mel[row, :, F_:].abs().sum()) == 0.0
        assert gate[row, :F_ - 1].sum() == 0 and bool((gate[row, F_ - 1:] == 1).all())
        assert int(out_len[row]) == F_ and torch.equal(emo[row], items[o][3])
    assert TextMelCollate(5, pin_memory=False)(items)[2].shape[-1] == 45
