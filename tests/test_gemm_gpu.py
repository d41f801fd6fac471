"""GPU parity of the tcgen05 GEMM / implicit-conv family against torch fp32 math on the same
bf16-rounded operands (the only rounding difference is fp32 accumulation order)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import assert_close

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).cuda()


@pytest.mark.parametrize("M,N,K,bn", [(128, 128, 64, 128), (256, 384, 512, 128), (256, 384, 512, 64),
                                      (256, 512, 512, 256), (200, 100, 136, 128), (81, 2048, 1096, 128),
                                      (4096, 256, 1024, 128)])
def test_gemm_plain(M, N, K, bn):
    from gantron_b200 import ops
    A = _rand(M, K, seed=1).to(BF)
    B = _rand(N, K, seed=2).to(BF)
    ref = A.float() @ B.float().t()
    out = torch.empty(M, N, device="cuda")
    ops.gemm(A, B, out, block_n=bn)
    torch.cuda.synchronize()
    assert_close(out, ref, f"gemm {M}x{N}x{K} bn{bn}", atol=2e-3 * K ** 0.5, rtol=1e-3)


def test_gemm_batched_strided_and_epilogue():
    from gantron_b200 import ops
    Z, M, N, K = 3, 150, 200, 264
    A = _rand(Z, M, K + 8, seed=3).to(BF)[:, :, :K]          # ld > K
    B = _rand(Z, N, K, seed=4).to(BF)
    bm, bn_ = _rand(M, seed=5), _rand(N, seed=6)
    ref = torch.einsum("zmk,znk->zmn", A.float(), B.float()) + bm[None, :, None] + bn_[None, None, :]
    out = torch.empty(Z, M, N, device="cuda")
    ops.gemm(A, B, out, bias_m=bm, bias_n=bn_)
    assert_close(out, ref, "batched+bias", atol=0.05, rtol=1e-3)
    # relu, transposed bf16 primary, fp32 secondary, stats
    out_t = torch.empty(Z, N, M, device="cuda", dtype=BF)
    out2 = torch.empty(Z, M, N, device="cuda")
    stats = torch.zeros(2, M, device="cuda")
    ops.gemm(A, B, out_t, trans_out=True, out2=out2, bias_m=bm, act=ops.ACT_RELU, stats=stats)
    r = torch.relu(torch.einsum("zmk,znk->zmn", A.float(), B.float()) + bm[None, :, None])
    assert_close(out2, r, "relu fp32 secondary", atol=0.05, rtol=1e-3)
    assert_close(out_t.float(), r.transpose(1, 2), "relu bf16 transposed", atol=0.3, rtol=1e-2)
    assert_close(stats[0], r.sum((0, 2)), "stats sum", atol=2.0, rtol=1e-3)
    assert_close(stats[1], (r * r).sum((0, 2)), "stats sumsq", atol=50.0, rtol=2e-3)
    # tanh + accumulate
    base = _rand(Z, M, N, seed=7)
    acc = base.clone()
    ops.gemm((A * 0.05).to(BF), B, acc, act=ops.ACT_TANH, accumulate=True)
    r2 = torch.tanh(torch.einsum("zmk,znk->zmn", (A * 0.05).to(BF).float(), B.float()) + base)
    assert_close(acc, r2, "tanh accumulate", atol=2e-3, rtol=1e-3)


def test_gemm_dropout_matches_oracle_philox():
    from gantron_b200 import ops
    from oracle import philox
    M, N, K = 96, 256, 80
    A = _rand(M, K, seed=8).to(BF)
    B = _rand(N, K, seed=9).to(BF)
    out = torch.empty(M, N, device="cuda")
    seed, site, step, p = 0x1234ABCD5678, philox.SITE_PRENET1, 0, 0.5
    ops.gemm(A, B, out, act=ops.ACT_RELU, dropout=(p, seed, site, step))
    keep = torch.from_numpy(philox.keep_mask(seed, site, step, (M, N), p)).cuda()
    ref = torch.relu(A.float() @ B.float().t()) * keep * 2.0
    assert_close(out, ref, "gemm+relu+dropout", atol=0.05, rtol=1e-3)


@pytest.mark.parametrize("cin,cout,k,dil,T,bn", [(80, 512, 5, 1, 100, 128), (512, 512, 5, 1, 300, 128),
                                                 (512, 80, 5, 1, 64, 64), (1600, 1024, 5, 1, 40, 64),
                                                 (1024, 512, 5, 2, 40, 64), (80, 1, 1, 1, 40, 64)])
def test_conv1d_forward(cin, cout, k, dil, T, bn):
    from gantron_b200 import ops
    Bn = 3
    x = _rand(Bn, cin, T, seed=10)
    w = _rand(cout, cin, k, seed=11, scale=(cin * k) ** -0.5)
    b = _rand(cout, seed=12)
    pad = dil * (k - 1) // 2
    xb, wb = x.to(BF), w.to(BF)
    ref = F.conv1d(xb.float(), wb.float(), b, padding=pad, dilation=dil)      # (B,cout,T)
    wp, cin_pad = ops.pack_conv_weight(w)
    x_cl = xb.transpose(1, 2).contiguous()
    if cin % 8:
        x_cl = F.pad(x_cl, (0, 8 - cin % 8))
    Tp = (T + 7) // 8 * 8
    out_cl = torch.empty(Bn, T, cout, device="cuda")
    out_cf = torch.zeros(Bn, cout, Tp, device="cuda", dtype=BF)
    stats = torch.zeros(2, cout, device="cuda")
    ops.conv1d(wp, cin, cin_pad, k, dil, pad, x_cl, out_cl, out_cf, bias=b, stats=stats, block_n=bn)
    assert_close(out_cl.transpose(1, 2), ref, "conv channels-last", atol=0.02, rtol=1e-3)
    assert_close(out_cf[:, :, :T].float(), ref, "conv channels-first bf16", atol=0.05, rtol=1e-2)
    assert_close(stats[0], ref.sum((0, 2)), "conv stats sum", atol=0.5, rtol=1e-3)
    assert_close(stats[1], (ref * ref).sum((0, 2)), "conv stats sumsq", atol=0.5, rtol=2e-3)


@pytest.mark.parametrize("cin,cout,k,dil,T", [(80, 512, 5, 1, 100), (512, 512, 5, 1, 130), (1024, 512, 5, 2, 40)])
def test_conv1d_dgrad_and_wgrad(cin, cout, k, dil, T):
    from gantron_b200 import ops
    Bn = 2
    pad = dil * (k - 1) // 2
    x = _rand(Bn, cin, T, seed=13).to(BF).float().requires_grad_(True)
    w = _rand(cout, cin, k, seed=14, scale=(cin * k) ** -0.5).to(BF).float().requires_grad_(True)
    dy = _rand(Bn, cout, T, seed=15).to(BF).float()
    y = F.conv1d(x, w, None, padding=pad, dilation=dil)
    y.backward(dy)
    # dgrad = same kernel on dY with the flipped/transposed pack
    wpt, cpad = ops.pack_conv_weight(w.detach(), transpose_for_dgrad=True)
    dy_cl = dy.to(BF).transpose(1, 2).contiguous()
    dx_cl = torch.empty(Bn, T, cin, device="cuda")
    ops.conv1d(wpt, cout, cpad, k, dil, pad, dy_cl, dx_cl, None)
    assert_close(dx_cl.transpose(1, 2), x.grad, "conv dgrad", atol=0.02, rtol=1e-3)
    # wgrad straight from the channels-last buffers (MN-major operands)
    x_cl = x.detach().to(BF).transpose(1, 2).contiguous()
    dw = torch.empty(k, cout, cin, device="cuda")
    ops.conv1d_wgrad(dy_cl, x_cl, cin, k, dil, pad, dw)
    assert_close(dw.permute(1, 2, 0), w.grad, "conv wgrad", atol=0.05, rtol=2e-3)
    ops.conv1d_wgrad(dy_cl, x_cl, cin, k, dil, pad, dw, accumulate=True, block_n=64)
    assert_close(dw.permute(1, 2, 0), 2 * w.grad, "conv wgrad accumulate bn64", atol=0.1, rtol=2e-3)


@pytest.mark.parametrize("M,N,K,bn", [(128, 128, 64, 128), (4096, 2304, 640, 128), (4096, 3072, 520, 256),
                                      (81, 2048, 200, 64), (150, 1024, 72, 128)])
def test_gemm_tn(M, N, K, bn):
    from gantron_b200 import ops
    Mp, Np = (M + 7) // 8 * 8, (N + 7) // 8 * 8
    A = _rand(K, Mp, seed=21).to(BF)[:, :M]
    B = _rand(K, Np, seed=22).to(BF)[:, :N]
    ref = A.float().t() @ B.float()
    out = torch.empty(M, N, device="cuda")
    ops.gemm_tn(A, B, out, block_n=bn)
    assert_close(out, ref, f"gemm_tn {M}x{N}x{K}", atol=2e-3 * K ** 0.5, rtol=1e-3)
    out_t = torch.zeros(N, M, device="cuda")
    ops.gemm_tn(A, B, out_t, trans_out=True, block_n=bn)
    ops.gemm_tn(A, B, out_t, trans_out=True, accumulate=True, block_n=bn)
    assert_close(out_t, 2 * ref.t(), "gemm_tn transposed+accumulate", atol=4e-3 * K ** 0.5, rtol=1e-3)


def test_gemm_tn_batched():
    from gantron_b200 import ops
    Z, M, N, K = 4, 152, 1024, 96     # d(memory)[b] = align[b]^T . dctx[b]
    A = _rand(Z, K, M, seed=23).to(BF)
    B = _rand(Z, K, N, seed=24).to(BF)
    out = torch.empty(Z, M, N, device="cuda")
    ops.gemm_tn(A, B, out)
    assert_close(out, torch.einsum("zkm,zkn->zmn", A.float(), B.float()), "gemm_tn batched", atol=0.03, rtol=1e-3)
