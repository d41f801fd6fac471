import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def describe_mismatch(got, ref, name="tensor", atol=0.0, rtol=0.0):
    """Human-readable summary of how two tensors differ (goes into assertion messages so one GPU run
    is enough to see the *structure* of an error)."""
    import torch
    got = got.detach().float().cpu()
    ref = ref.detach().float().cpu()
    if got.shape != ref.shape:
        return f"{name}: shape {tuple(got.shape)} vs {tuple(ref.shape)}"
    err = (got - ref).abs()
    tol = atol + rtol * ref.abs()
    bad = err > tol
    nbad = int(bad.sum())
    msg = (f"{name}: shape {tuple(got.shape)} max_abs_err {float(err.max()):.4g} ref_absmax {float(ref.abs().max()):.4g} "
           f"bad {nbad}/{bad.numel()} nan_got {int(torch.isnan(got).sum())}")
    if nbad:
        idx = bad.nonzero()[:6].tolist()
        msg += " first_bad " + "; ".join(f"{tuple(i)}: got {float(got[tuple(i)]):.5g} ref {float(ref[tuple(i)]):.5g}" for i in idx)
        if bad.dim() >= 2:
            rows = bad.reshape(-1, bad.shape[-1]).any(1).sum().item()
            cols = bad.reshape(-1, bad.shape[-1]).any(0).sum().item()
            msg += f" bad_rows {rows} bad_cols {cols}"
    return msg


def assert_close(got, ref, name="tensor", atol=1e-5, rtol=1e-5):
    import torch
    got_f = got.detach().float().cpu()
    ref_f = ref.detach().float().cpu()
    ok = got_f.shape == ref_f.shape and bool(((got_f - ref_f).abs() <= atol + rtol * ref_f.abs()).all()) \
        and not bool(torch.isnan(got_f).any())
    assert ok, describe_mismatch(got, ref, name, atol, rtol)
