"""Every aten::copy_ / elementwise launch of one training step with its shapes and the enclosing autograd node (dev tool)."""
import os
import sys
import collections

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench

tr = bench.Trainer(False)
batch = bench._pinned_batch(bench.B_PER_GPU, seed=1234)
x, y = tr.G.parse_batch(batch)
for _ in range(3):
    tr.step(x, y)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
    tr.step(x, y)
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0.0, 0])
for e in prof.events():
    dt = getattr(e, "self_device_time_total", 0)
    if dt <= 0 or not e.name.startswith("aten::"):
        continue
    par, names = e.cpu_parent, []
    while par is not None:
        if not par.name.startswith("aten::"):
            names.append(par.name)
        par = par.cpu_parent
    key = (e.name, str(e.input_shapes)[:90], (names[0] if names else "-")[:40])
    agg[key][0] += dt
    agg[key][1] += 1
rows = sorted(agg.items(), key=lambda kv: -kv[1][0])
print(f"aten device time {sum(v[0] for v in agg.values()) / 1e3:.2f} ms")
for (name, shp, par), (dt, n) in rows[:70]:
    print(f"{dt / 1e3:7.3f} ms {n:3d} {name:28s} {shp:90s} {par}")
