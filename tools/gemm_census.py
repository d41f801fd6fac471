"""Every tensor-core GEMM / conv launch of one training step with its shape, duration and TFLOP/s (dev tool;
feeds profiles/r2_gemm_census.md).  Wraps the four host entry points of gantron_b200.ops with CUDA events."""
import os
import sys
import traceback

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from gantron_b200 import ops

LOG = []
ON = [False]


def caller():
    for fs in reversed(traceback.extract_stack()[:-2]):
        if "gantron_b200" in fs.filename and os.path.basename(fs.filename) != "ops.py":
            return f"{os.path.basename(fs.filename)}:{fs.lineno}"
    return "?"


def wrap(name, fn, shape_of):
    def inner(*a, **k):
        if not ON[0]:
            return fn(*a, **k)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = fn(*a, **k)
        e1.record()
        LOG.append((name, shape_of(*a, **k), e0, e1, caller()))
        return r
    return inner


def gemm_shape(A, B, out, **k):
    Z = A.shape[0] if A.dim() == 3 else 1
    return Z, A.shape[-2], B.shape[-2], A.shape[-1]


def tn_shape(A, B, out, **k):
    Z = A.shape[0] if A.dim() == 3 else 1
    return Z, A.shape[-1], B.shape[-1], A.shape[-2]


def conv_shape(wp, cin, cin_pad, taps, dil, pad, x_cl, *a, **k):
    return 1, wp.shape[0], x_cl.shape[0] * x_cl.shape[1], taps * cin_pad


def wgrad_shape(dy, x, cin, taps, *a, **k):
    return taps, dy.shape[2], cin, dy.shape[0] * dy.shape[1]


ops.gemm = wrap("gemm", ops.gemm, gemm_shape)
ops.gemm_tn = wrap("gemm_tn", ops.gemm_tn, tn_shape)
ops.conv1d = wrap("conv1d", ops.conv1d, conv_shape)
ops.conv1d_wgrad = wrap("conv1d_wgrad", ops.conv1d_wgrad, wgrad_shape)

tr = bench.Trainer(False)
batch = bench._pinned_batch(bench.B_PER_GPU, seed=1234)
x, y = tr.G.parse_batch(batch)
for _ in range(3):
    tr.step(x, y)
torch.cuda.synchronize()
ON[0] = True
tr.step(x, y)
torch.cuda.synchronize()
tot_ms = tot_fl = 0.0
rows = []
for name, (Z, M, N, K), e0, e1, who in LOG:
    ms = e0.elapsed_time(e1)
    fl = 2.0 * Z * M * N * K
    tot_ms += ms
    tot_fl += fl
    rows.append((ms, name, Z, M, N, K, fl / ms / 1e9, who))
print(f"{len(rows)} launches, {tot_ms:.2f} ms (event-to-event, includes launch gaps), {tot_fl / 1e12:.2f} TFLOP, {tot_fl / tot_ms / 1e9:.0f} TFLOP/s")
print("| ms | entry | Z | M | N | K | TFLOP/s | caller |\n|---:|---|---:|---:|---:|---:|---:|---|")
for r in sorted(rows, reverse=True)[:60]:
    print("| %.3f | %s | %d | %d | %d | %d | %.0f | %s |" % r)
