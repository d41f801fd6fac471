"""CPU, build container only (skipped where /root/reference is absent, e.g. on the GPU box): re-run the
UNMODIFIED reference live and check that (a) the committed golden fixtures are reproducible and (b) the
drop-in module surface matches the reference's (state_dict keys, ctor attribute paths)."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_loader as RL
from oracle import make_golden as MG

pytestmark = pytest.mark.skipif(not RL.available(), reason="reference tree not present")


def test_golden_fixture_is_reproducible_from_the_live_reference():
    torch.set_num_threads(max(1, (os.cpu_count() or 2) // 2))
    with RL.reference_modules() as ref, RL.cpu_shims():
        live = MG.make_disc_case(ref, B=3, T=47, wseed=41, mseed=401)
        live_inf = MG.make_infer_case(ref, B=1, T_in=13, steps=6, wseed=31, mseed=301)
    for name, data in (("disc_vanilla", live), ("infer_vanilla_b1", live_inf)):
        z = np.load(os.path.join(MG.GOLDEN_DIR, name + ".npz"))
        for k in z.files:
            if k.startswith(("out/", "grad/", "gsample/")):
                a, b = np.asarray(data[k], dtype=np.float64), np.asarray(z[k], dtype=np.float64)
                assert np.allclose(a, b, atol=1e-6, rtol=1e-5), f"{name}:{k}"


def test_drop_in_surface_matches_live_reference():
    from gantron_b200 import model as M
    from gantron_b200.hparams import HParams
    with RL.reference_modules() as ref, RL.cpu_shims():
        for s in ("use_noise=True,use_labels=False", "use_noise=True,use_labels=True,vesus_path=x",
                  "use_noise=False,use_labels=True,vesus_path=x", "use_noise=True,use_labels=True,encoder_inputs=True,vesus_path=x"):
            rg, rd = ref.model.Tacotron2(ref.hparams.HParams(s)), ref.model.Discriminator(ref.hparams.HParams(s))
            og, od = M.Tacotron2(HParams(s)), M.Discriminator(HParams(s))
            assert {k: tuple(v.shape) for k, v in rg.state_dict().items()} == {k: tuple(v.shape) for k, v in og.state_dict().items()}, s
            assert {k: tuple(v.shape) for k, v in rd.state_dict().items()} == {k: tuple(v.shape) for k, v in od.state_dict().items()}, s
            og.load_state_dict(rg.state_dict(), strict=True)
            for path in ("decoder.attention_layer.score_mask_value", "decoder.max_decoder_steps", "decoder.gate_threshold",
                         "encoder", "postnet", "mask_padding", "n_mel_channels"):
                obj_r, obj_o = rg, og
                for part in path.split("."):
                    obj_r, obj_o = getattr(obj_r, part), getattr(obj_o, part)
        # reference HParams objects are accepted as-is (duck typing)
        M.Tacotron2(ref.hparams.HParams("use_noise=True,use_labels=False"))


def _items(n, seed, labelled):
    g = torch.Generator().manual_seed(seed)
    items = []
    for i in range(n):
        L = int(torch.randint(3, 40, (1,), generator=g))
        F_ = int(torch.randint(21, 90, (1,), generator=g))
        it = [torch.randint(1, 100, (L,), generator=g), torch.randn(80, F_, generator=g)]
        if labelled:
            it += [float(i % 3), torch.rand(5, generator=g)]
        items.append(it)
    return items


@pytest.mark.parametrize("labelled", [False, True])
@pytest.mark.parametrize("step", [1, 3])
def test_collate_matches_live_reference(labelled, step):
    """gantron_b200.data_utils.TextMelCollate against data_utils.py:82-131 on the same items."""
    from gantron_b200.data_utils import TextMelCollate
    items = _items(7, 5 + step, labelled)
    with RL.reference_modules(), RL.cpu_shims():
        import data_utils as ref_du
        want = ref_du.TextMelCollate(step)(items)
    got = TextMelCollate(step, pin_memory=False)(items)
    names = ("text", "input_lengths", "mel", "gate", "speaker", "emotions", "output_lengths")
    for name, a, b in zip(names, got, want):
        if not labelled and name in ("speaker", "emotions"):
            continue                                   # uninitialised in the reference without labels
        assert a.dtype == b.dtype and a.shape == b.shape, name
        assert torch.equal(a, b), name
