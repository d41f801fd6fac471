"""Load the UNMODIFIED reference (/root/reference) in this container (TEST INFRASTRUCTURE).

Used only by ``oracle/make_golden.py`` and ``tests/test_oracle_vs_reference.py``; both skip
when /root/reference is absent (it does not exist on the GPU box).  Nothing here is copied
from the reference: we import its modules in place.

What is needed to import and run it on CPU:
  * empty stub modules for librosa / unidecode / inflect (imported at module scope by
    layers.py:2, stft.py, text/cleaners.py:16, text/numbers.py:3, utils.py:6; never touched on
    the hot path),
  * `.cuda()` -> identity and `torch.cuda.LongTensor` -> `torch.LongTensor` shims because the
    reference hard-codes CUDA (utils.py:22, model.py:187,275,469),
  * a dropout hook so masks can be injected in call order (model.py:104,148,149,393,408;
    layers.py:55).
"""
from __future__ import annotations

import contextlib
import os
import sys
import types
from typing import Callable, List, Optional

import torch

REFERENCE_DIR = os.environ.get("GANTRON_REFERENCE_DIR", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_DIR, "model.py"))


def _install_stubs():
    def mod(name, **attrs):
        if name in sys.modules:
            return sys.modules[name]
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    nope = lambda *a, **k: (_ for _ in ()).throw(RuntimeError("stubbed module called"))
    lib = mod("librosa", load=nope)
    lib.filters = mod("librosa.filters", mel=nope)
    lib.util = mod("librosa.util", pad_center=nope, tiny=nope, normalize=nope)
    mod("unidecode", unidecode=lambda s: s)
    mod("inflect", engine=lambda: types.SimpleNamespace(number_to_words=lambda *a, **k: ""))


_REF_MODULE_NAMES = ("model", "layers", "utils", "hparams", "loss_function", "distributed", "stft", "data_utils",
                     "audio_processing", "text", "text.cleaners", "text.numbers", "text.symbols",
                     "text.cmudict")


@contextlib.contextmanager
def reference_modules():
    """Context manager yielding a namespace with the reference's `model`, `hparams`, `layers`,
    `loss_function`, `distributed` modules imported from REFERENCE_DIR.  Module names such as
    `model` are removed from sys.modules on exit so they cannot shadow this repo's own modules."""
    if not available():
        raise RuntimeError(f"reference not found at {REFERENCE_DIR}")
    _install_stubs()
    saved = {n: sys.modules.pop(n) for n in list(sys.modules) if n in _REF_MODULE_NAMES}
    sys.path.insert(0, REFERENCE_DIR)
    try:
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            import model as ref_model            # noqa
            import hparams as ref_hparams        # noqa
            import layers as ref_layers          # noqa
            import loss_function as ref_loss     # noqa
            import distributed as ref_dist       # noqa
        yield types.SimpleNamespace(model=ref_model, hparams=ref_hparams, layers=ref_layers,
                                    loss_function=ref_loss, distributed=ref_dist)
    finally:
        sys.path.remove(REFERENCE_DIR)
        for n in list(sys.modules):
            if n in _REF_MODULE_NAMES:
                del sys.modules[n]
        sys.modules.update(saved)


@contextlib.contextmanager
def cpu_shims():
    """Make the reference's hard-coded `.cuda()` calls no-ops so it runs on CPU."""
    t_cuda, m_cuda = torch.Tensor.cuda, torch.nn.Module.cuda
    saved_types = {n: getattr(torch.cuda, n, None) for n in ("LongTensor", "IntTensor", "FloatTensor")}
    t_zeros, t_linspace = torch.zeros, torch.linspace

    def no_cuda_device(fn):
        # loss_function.py:28,47,49 hard-code device='cuda'
        def wrapped(*a, **k):
            if str(k.get("device", "")) .startswith("cuda"):
                k.pop("device")
            return fn(*a, **k)
        return wrapped

    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.nn.Module.cuda = lambda self, *a, **k: self
    torch.cuda.LongTensor, torch.cuda.IntTensor, torch.cuda.FloatTensor = torch.LongTensor, torch.IntTensor, torch.FloatTensor
    torch.zeros, torch.linspace = no_cuda_device(t_zeros), no_cuda_device(t_linspace)
    try:
        yield
    finally:
        torch.Tensor.cuda, torch.nn.Module.cuda = t_cuda, m_cuda
        torch.zeros, torch.linspace = t_zeros, t_linspace
        for n, v in saved_types.items():
            if v is not None:
                setattr(torch.cuda, n, v)


@contextlib.contextmanager
def injected_dropout(provider: Callable[[torch.Size, float], torch.Tensor], log: Optional[List] = None):
    """Replace torch.nn.functional.dropout so that every *active* call (training=True, p>0) takes its
    keep mask from provider(shape, p), in call order.  Inactive calls pass through."""
    import torch.nn.functional as F
    orig = F.dropout

    def patched(input, p=0.5, training=True, inplace=False):
        if not training or p == 0.0:
            return input
        keep = provider(input.shape, p).to(input.dtype)
        if log is not None:
            log.append((tuple(input.shape), p))
        return input * keep * (1.0 / (1.0 - p))

    F.dropout = patched
    try:
        yield
    finally:
        F.dropout = orig


def queue_provider(masks: List[torch.Tensor]):
    """Provider that hands out a fixed list of masks in order, checking shapes."""
    it = iter(masks)

    def provider(shape, p):
        m = next(it)
        assert tuple(m.shape) == tuple(shape), f"mask shape {tuple(m.shape)} != dropout input {tuple(shape)}"
        return m
    return provider
