"""Tensor-level wrappers over the C ABI (thin: shape checks, pointer extraction, launches).

Every function here launches hand-written sm_100a kernels from libgantron_b200.so on torch's
current stream; none has a torch/CPU fallback.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib

ACT_NONE, ACT_RELU, ACT_TANH = 0, 1, 2
_BF16 = torch.bfloat16


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return 0
    if t.dtype == _BF16:
        return 1
    raise TypeError(f"output dtype must be fp32 or bf16, got {t.dtype}")


def _rows_k(t: torch.Tensor, name: str) -> Tuple[int, int, int, int, int]:
    """(batch, rows, K, ld, batch_stride) of a K-contiguous bf16 matrix (rows,K) or (Z,rows,K)."""
    assert t.dtype == _BF16 and t.is_cuda, f"{name}: bf16 CUDA tensor expected"
    assert t.stride(-1) == 1, f"{name}: K must be contiguous"
    if t.dim() == 2:
        return 1, t.shape[0], t.shape[1], t.stride(0), 0
    assert t.dim() == 3
    return t.shape[0], t.shape[1], t.shape[2], t.stride(1), t.stride(0)


def gemm(A: torch.Tensor, B: torch.Tensor, out: torch.Tensor, *, trans_out: bool = False,
         out2: Optional[torch.Tensor] = None, trans_out2: bool = False,
         bias_m: Optional[torch.Tensor] = None, bias_n: Optional[torch.Tensor] = None, act: int = ACT_NONE,
         accumulate: bool = False, dropout: Optional[Tuple[float, int, int, int]] = None,
         stats: Optional[torch.Tensor] = None, block_n: int = 128) -> torch.Tensor:
    """out[z,m,n] = sum_k A[z,m,k] B[z,n,k] with the fused epilogue of gtn_gemm_bf16.

    out is (M,N) / (Z,M,N) (or (N,M) / (Z,N,M) when trans_out); dropout = (p, seed, site, step)."""
    za, M, K, lda, sa = _rows_k(A, "A")
    zb, N, Kb, ldb, sb = _rows_k(B, "B")
    assert K == Kb and za == zb, f"GEMM shape mismatch A{tuple(A.shape)} B{tuple(B.shape)}"
    Z = za

    def out_geom(o, trans):
        assert o.is_cuda and o.stride(-1) == 1
        exp = (N, M) if trans else (M, N)
        assert tuple(o.shape[-2:]) == exp, f"output shape {tuple(o.shape)} != {exp}"
        if o.dim() == 2:
            assert Z == 1
            return o.stride(0), 0
        assert o.shape[0] == Z
        return o.stride(1), o.stride(0)

    ldc, sc = out_geom(out, trans_out)
    if out2 is not None:
        ldc2, sc2 = out_geom(out2, trans_out2)
    else:
        ldc2, sc2 = 0, 0
    p, seed, site, step = dropout if dropout is not None else (0.0, 0, 0, 0)
    for b in (bias_m, bias_n, stats):
        assert b is None or (b.dtype == torch.float32 and b.is_cuda and b.is_contiguous())
    lib = _lib.load()
    rc = lib.gtn_gemm_bf16(
        _lib.ptr(A), _lib.ptr(B), M, N, K, lda, ldb, Z, sa, sb,
        _lib.ptr(out), _dt(out), ldc, sc, int(trans_out),
        _lib.ptr(out2), _dt(out2) if out2 is not None else 0, ldc2, sc2, int(trans_out2),
        _lib.ptr(bias_m), _lib.ptr(bias_n), act, int(accumulate),
        float(p), int(seed), int(site), int(step), _lib.ptr(stats), block_n, _lib.stream())
    _lib.check(rc, "gtn_gemm_bf16")
    return out


def pack_conv_weight(w: torch.Tensor, transpose_for_dgrad: bool = False) -> Tuple[torch.Tensor, int]:
    """(Cout,Cin,k) fp32 conv weight -> bf16 [Cout][k*cin_pad] K-major pack for gtn_conv1d_bf16.

    transpose_for_dgrad: build the data-gradient operator instead, W'[ci][tap][co] = W[co][ci][k-1-tap]
    (so that dgrad is the same implicit-GEMM kernel run on dY)."""
    if transpose_for_dgrad:
        w = w.flip(-1).permute(1, 0, 2)
    co, ci, k = w.shape
    cin_pad = (ci + 63) // 64 * 64
    wp = torch.zeros(co, k, cin_pad, device=w.device, dtype=_BF16)
    wp[:, :, :ci] = w.permute(0, 2, 1).to(_BF16)
    return wp.reshape(co, k * cin_pad), cin_pad


def conv1d(wp: torch.Tensor, cin: int, cin_pad: int, taps: int, dil: int, pad: int, x_cl: torch.Tensor,
           out_cl: Optional[torch.Tensor], out_cf: Optional[torch.Tensor], bias: Optional[torch.Tensor] = None,
           act: int = ACT_NONE, stats: Optional[torch.Tensor] = None, block_n: int = 128,
           dropout: Optional[Tuple[float, int, int]] = None, drop_first: bool = False) -> None:
    """Implicit-GEMM conv1d.  x_cl channels-last bf16 (B,T,>=cin) -> out_cl (B,T,Cout) and/or out_cf (B,Cout,T).
    dropout = (p, seed, site), applied after the activation, or before it when drop_first."""
    assert x_cl.dtype == _BF16 and x_cl.dim() == 3 and x_cl.stride(2) == 1
    Bn, T, _ = x_cl.shape
    cout = wp.shape[0]
    assert wp.is_contiguous() and wp.shape[1] == taps * cin_pad
    if out_cl is not None:
        assert tuple(out_cl.shape[:2]) == (Bn, T) and out_cl.shape[2] >= cout and out_cl.stride(2) == 1
    if out_cf is not None:
        assert tuple(out_cf.shape[:2]) == (Bn, cout) and out_cf.shape[2] >= T and out_cf.stride(2) == 1
    dp, dseed, dsite = dropout if dropout is not None else (0.0, 0, 0)
    lib = _lib.load()
    rc = lib.gtn_conv1d_bf16(
        _lib.ptr(wp), cout, cin, cin_pad, taps, dil, pad,
        _lib.ptr(x_cl), Bn, T, x_cl.stride(1), x_cl.stride(0),
        _lib.ptr(out_cl), _dt(out_cl) if out_cl is not None else 0,
        out_cl.stride(1) if out_cl is not None else 0, out_cl.stride(0) if out_cl is not None else 0,
        _lib.ptr(out_cf), _dt(out_cf) if out_cf is not None else 0,
        out_cf.stride(1) if out_cf is not None else 0, out_cf.stride(0) if out_cf is not None else 0,
        _lib.ptr(bias), act, float(dp), int(dseed), int(dsite), int(drop_first),
        _lib.ptr(stats), block_n, _lib.stream())
    _lib.check(rc, "gtn_conv1d_bf16")


def conv1d_wgrad(dy_cl: torch.Tensor, x_cl: torch.Tensor, cin: int, taps: int, dil: int, pad: int,
                 dw: torch.Tensor, accumulate: bool = False, block_n: int = 128) -> torch.Tensor:
    """dw[tap][co][ci] (+)= sum_{b,t} dy_cl[b,t,co] x_cl[b,t+tap*dil-pad,ci]; both channels-last bf16."""
    assert dy_cl.dtype == _BF16 and x_cl.dtype == _BF16 and dy_cl.stride(2) == 1 and x_cl.stride(2) == 1
    Bn, T, cout = dy_cl.shape
    assert x_cl.shape[0] == Bn and x_cl.shape[1] == T and x_cl.shape[2] >= cin
    assert dw.dtype == torch.float32 and tuple(dw.shape) == (taps, cout, cin) and dw.is_contiguous()
    lib = _lib.load()
    rc = lib.gtn_conv1d_wgrad_bf16(
        _lib.ptr(dy_cl), _lib.ptr(x_cl), Bn, cout, cin, T,
        dy_cl.stride(1), dy_cl.stride(0), x_cl.stride(1), x_cl.stride(0),
        taps, dil, pad, _lib.ptr(dw), dw.stride(1), int(accumulate), block_n, _lib.stream())
    _lib.check(rc, "gtn_conv1d_wgrad_bf16")
    return dw


def gemm_tn(A: torch.Tensor, B: torch.Tensor, out: torch.Tensor, *, trans_out: bool = False,
            bias_m: Optional[torch.Tensor] = None, bias_n: Optional[torch.Tensor] = None,
            accumulate: bool = False, block_n: int = 128) -> torch.Tensor:
    """out[z,m,n] (+)= sum_k A[z,k,m] B[z,k,n]; A (K,M)/(Z,K,M), B (K,N)/(Z,K,N) bf16 with the last dim contiguous."""
    for t in (A, B):
        assert t.dtype == _BF16 and t.is_cuda and t.stride(-1) == 1
    if A.dim() == 2:
        Z, (K, M), lda, sa = 1, A.shape, A.stride(0), 0
        (Kb, N), ldb, sb = B.shape, B.stride(0), 0
    else:
        Z, K, M = A.shape
        lda, sa = A.stride(1), A.stride(0)
        Zb, Kb, N = B.shape
        assert Zb == Z
        ldb, sb = B.stride(1), B.stride(0)
    assert K == Kb
    exp = (N, M) if trans_out else (M, N)
    assert tuple(out.shape[-2:]) == exp and out.stride(-1) == 1
    ldc, sc = (out.stride(0), 0) if out.dim() == 2 else (out.stride(1), out.stride(0))
    lib = _lib.load()
    rc = lib.gtn_gemm_tn_bf16(_lib.ptr(A), _lib.ptr(B), M, N, K, lda, ldb, Z, sa, sb,
                              _lib.ptr(out), _dt(out), ldc, sc, int(trans_out),
                              _lib.ptr(bias_m), _lib.ptr(bias_n), int(accumulate), block_n, _lib.stream())
    _lib.check(rc, "gtn_gemm_tn_bf16")
    return out


# ==========================================================================================
# elementwise / reduction kernels (gtn_eltwise.cu); activations channels-last (B,T,C) or (rows,C)
# ==========================================================================================
def _rows_c(t: torch.Tensor, name: str) -> Tuple[int, int, int]:
    """(rows, C, ld) of a channels-last activation whose leading dims collapse to uniform rows."""
    assert t.is_cuda and t.stride(-1) == 1 and t.dtype in (torch.float32, _BF16), f"{name}: bad tensor"
    C = t.shape[-1]
    ld = t.stride(-2)
    rows = 1
    for i in range(t.dim() - 1):
        rows *= t.shape[i]
    # leading dims must be collapsible: stride(i) == stride(i+1) * shape(i+1)
    for i in range(t.dim() - 2):
        assert t.stride(i) == t.stride(i + 1) * t.shape[i + 1], f"{name}: leading dims are not collapsible"
    return rows, C, ld


def bn_finalize(stats: torch.Tensor, count: int, eps: float, momentum: float,
                running_mean: Optional[torch.Tensor], running_var: Optional[torch.Tensor]):
    C = stats.shape[1]
    mean = torch.empty(C, device=stats.device)
    rstd = torch.empty(C, device=stats.device)
    rc = _lib.load().gtn_bn_finalize(_lib.ptr(stats), C, int(count), float(eps), float(momentum), _lib.ptr(mean),
                                     _lib.ptr(rstd), _lib.ptr(running_mean), _lib.ptr(running_var), _lib.stream())
    _lib.check(rc, "gtn_bn_finalize")
    return mean, rstd


def bn_act_fwd(y: torch.Tensor, mean, rstd, gamma, beta, act: int, dropout: Tuple[float, int, int],
               out: torch.Tensor) -> torch.Tensor:
    rows, C, ldy = _rows_c(y, "y")
    rows_o, C_o, ldo = _rows_c(out, "out")
    assert rows == rows_o and C == C_o
    p, seed, site = dropout
    rc = _lib.load().gtn_bn_act_fwd(_lib.ptr(y), _dt(y), ldy, rows, C, _lib.ptr(mean), _lib.ptr(rstd), _lib.ptr(gamma),
                                    _lib.ptr(beta), act, float(p), int(seed), int(site), _lib.ptr(out), _dt(out), ldo,
                                    _lib.stream())
    _lib.check(rc, "gtn_bn_act_fwd")
    return out


def bn_act_bwd(dx: torch.Tensor, y: torch.Tensor, mean, rstd, gamma, beta, act: int, dropout: Tuple[float, int, int],
               dy: torch.Tensor, want_dbias: bool = True, batch_stats: bool = True):
    rows, C, lddx = _rows_c(dx, "dx")
    r2, C2, ldy = _rows_c(y, "y")
    r3, C3, lddy = _rows_c(dy, "dy")
    assert rows == r2 == r3 and C == C2 == C3
    dev = y.device
    dgamma, dbeta = torch.empty(C, device=dev), torch.empty(C, device=dev)
    dbias = torch.empty(C, device=dev) if want_dbias else None
    p, seed, site = dropout
    fn = _lib.load().gtn_bn_act_bwd if batch_stats else _lib.load().gtn_bn_act_bwd_eval
    rc = fn(_lib.ptr(dx), _dt(dx), lddx, _lib.ptr(y), _dt(y), ldy, rows, C, _lib.ptr(mean),
                                    _lib.ptr(rstd), _lib.ptr(gamma), _lib.ptr(beta), act, float(p), int(seed), int(site),
                                    _lib.ptr(dgamma), _lib.ptr(dbeta), _lib.ptr(dy), _dt(dy), lddy, _lib.ptr(dbias),
                                    _lib.stream())
    _lib.check(rc, "gtn_bn_act_bwd")
    return dgamma, dbeta, dbias


def drop_tanh_bwd(dout: torch.Tensor, a: torch.Tensor, dropout: Tuple[float, int, int], dpre: torch.Tensor,
                  want_dbias: bool = True):
    rows, C, lddo = _rows_c(dout, "dout")
    r2, C2, lda = _rows_c(a, "a")
    r3, C3, lddp = _rows_c(dpre, "dpre")
    assert rows == r2 == r3 and C == C2 == C3
    dbias = torch.empty(C, device=a.device) if want_dbias else None
    p, seed, site = dropout
    rc = _lib.load().gtn_drop_tanh_bwd(_lib.ptr(dout), _dt(dout), lddo, _lib.ptr(a), _dt(a), lda, rows, C, float(p),
                                       int(seed), int(site), _lib.ptr(dpre), _dt(dpre), lddp, _lib.ptr(dbias),
                                       _lib.stream())
    _lib.check(rc, "gtn_drop_tanh_bwd")
    return dbias


def disc_head_fwd(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    rows, C, ldx = _rows_c(x, "x")
    out = torch.empty(x.shape[:-1], device=x.device)
    rc = _lib.load().gtn_disc_head_fwd(_lib.ptr(x), _dt(x), ldx, rows, C, _lib.ptr(w), _lib.ptr(bias), _lib.ptr(out),
                                       _lib.stream())
    _lib.check(rc, "gtn_disc_head_fwd")
    return out


def disc_head_bwd(dout: torch.Tensor, x: torch.Tensor, w: torch.Tensor, dx: torch.Tensor, want_wgrad: bool):
    rows, C, ldx = _rows_c(x, "x")
    r2, C2, lddx = _rows_c(dx, "dx")
    assert rows == r2 and C == C2 and dout.is_contiguous() and dout.numel() == rows
    dw = torch.empty(C, device=x.device) if want_wgrad else None
    db = torch.empty(1, device=x.device) if want_wgrad else None
    rc = _lib.load().gtn_disc_head_bwd(_lib.ptr(dout), _lib.ptr(x), _dt(x), ldx, rows, C, _lib.ptr(w), _lib.ptr(dx),
                                       _dt(dx), lddx, _lib.ptr(dw), _lib.ptr(db), _lib.stream())
    _lib.check(rc, "gtn_disc_head_bwd")
    return dw, db
