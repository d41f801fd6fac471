"""GPU: the batch producer in front of the path -- pinned collate + one-batch-ahead upload -- hands the training
step exactly what Tacotron2.parse_batch would, and a discriminator_type='linear' iteration trains through it."""
import random

import pytest
import torch

pytestmark = pytest.mark.gpu


def _items(n, seed):
    g = torch.Generator().manual_seed(seed)
    return [[torch.randint(1, 100, (int(torch.randint(5, 30, (1,), generator=g)),), generator=g),
             (torch.randn(80, int(torch.randint(40, 90, (1,), generator=g)), generator=g) * 2 - 5).clamp(-11.5, 2.0)]
            for _ in range(n)]


def test_prefetcher_yields_parse_batch_results():
    from gantron_b200.data_utils import TextMelCollate, DevicePrefetcher
    from gantron_b200.hparams import HParams
    from gantron_b200 import model as M
    G = M.Tacotron2(HParams("use_noise=True,use_labels=False")).cuda()
    collate = TextMelCollate(1)
    batches = [collate(_items(4, s)) for s in range(5)]
    assert all(t.is_pinned() for t in batches[0])
    got = list(DevicePrefetcher(batches, G))
    assert len(got) == 5
    for b, (x, y) in zip(batches, got):
        xr, yr = G.parse_batch(b)
        for a, r in zip(list(x) + list(y), list(xr) + list(yr)):
            if torch.is_tensor(a):
                assert a.is_cuda and a.dtype == r.dtype and torch.equal(a, r)
            else:
                assert a == r


def test_linear_discriminator_iteration_trains():
    """One G and one D iteration of train.py:301-406 with discriminator_type='linear' (LinearDiscriminator on the kernels),
    batches through the collate + prefetcher: finite losses, every trainable parameter receives a gradient."""
    from gantron_b200.data_utils import TextMelCollate, DevicePrefetcher
    from gantron_b200.hparams import HParams
    from gantron_b200 import model as M
    from gantron_b200.loss_function import Tacotron2Loss
    torch.manual_seed(0)
    random.seed(0)
    hp = HParams("use_noise=True,use_labels=False,discriminator_type=linear")
    G, D = M.Tacotron2(hp).cuda().train(), M.LinearDiscriminator(hp).cuda().train()
    crit = Tacotron2Loss()
    collate = TextMelCollate(1)
    for x, y in DevicePrefetcher([collate(_items(4, 9))], G):
        for p in D.parameters():
            p.requires_grad = False
        y_pred = G(x)
        mel_loss, gate_loss, _ = crit(y_pred, y, x[1], x[-1])
        total = mel_loss + gate_loss + D.adversarial_loss(y_pred[1], x[-1])
        total.backward()
        assert torch.isfinite(total) and all(p.grad is not None for p in G.decoder.parameters())
        for p in D.parameters():
            p.requires_grad = True
        d_loss = D.adversarial_loss(y[0], x[-1]) - D.adversarial_loss(y_pred[1].detach(), x[-1])
        d_loss.backward()
        assert torch.isfinite(d_loss) and all(p.grad is not None and torch.isfinite(p.grad).all() for p in D.parameters())
