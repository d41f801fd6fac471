"""CPU: the C-ABI shared library loads and exports every symbol include/gantron_b200.h declares; the ctypes
binding lists exactly those symbols; ABI struct sizes agree.  No compute calls (there is no GPU here)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gantron_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"typedef struct.*?\}\s*\w+;", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gtn_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from gantron_b200 import build, _lib
    build.build()
    return _lib.load()


def test_header_declares_the_expected_families():
    names = declared_functions()
    for must in ("gtn_decoder_forward2", "gtn_decoder_backward2", "gtn_gemm_bf16", "gtn_conv1d_bf16", "gtn_last_error"):
        assert must in names
    assert len(names) >= 25


def test_library_exports_every_declared_symbol(lib):
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, f"declared in include/gantron_b200.h but not exported: {missing}"


def test_binding_signatures_cover_the_header():
    from gantron_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_functions()


def test_struct_sizes_match_the_binding(lib):
    from gantron_b200 import _lib
    assert lib.gtn_sizeof(0) == C.sizeof(_lib.DecoderFwdParams)
    assert lib.gtn_sizeof(2) == C.sizeof(_lib.DecoderBwd2Params)
    assert lib.gtn_sizeof(3) == C.sizeof(_lib.DecoderInfer1Params)
    assert lib.gtn_sizeof(4) == C.sizeof(_lib.BiLstmParams)
    assert lib.gtn_version() >= 100


def test_errors_are_reported_not_swallowed(lib):
    from gantron_b200 import _lib
    rc = lib.gtn_decoder_forward2(None, None)
    assert rc != 0 and "null" in _lib.last_error().lower()


def test_no_cpu_fallback():
    """The product path must fail loudly without CUDA tensors (no silent eager fallback)."""
    import torch
    from gantron_b200.hparams import HParams
    from gantron_b200 import model as M
    dec = M.Decoder(HParams("use_noise=True,use_labels=False"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        dec(torch.zeros(1, 4, 512), torch.zeros(1, 80, 3), torch.tensor([4]))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        M.Discriminator(HParams()).adversarial_loss(torch.zeros(1, 80, 40), torch.tensor([40]))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "gantron_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{fn} imports oracle/"
