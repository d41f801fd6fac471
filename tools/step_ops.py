"""Which torch ops (and which lines of gantron_b200) the non-decoder device time of one training step belongs to:
torch.profiler over one step, CUDA time per op and per kernel (dev tool; feeds profiles/r2_step_launches_summary.md)."""
import os
import sys

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
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
    tr.step(x, y)
    torch.cuda.synchronize()
ka = prof.key_averages(group_by_stack_n=30)
rows = []
for e in ka:
    dt = getattr(e, "self_device_time_total", 0)
    if dt > 8:
        stack = [s for s in e.stack if ("gantron_b200/" in s or "bench.py" in s) and "tools/" not in s][:2]
        if not e.key.startswith("aten::") and not e.key.startswith("Mem"):
            continue
        rows.append((dt, e.count, e.key, " <- ".join(s.split("/")[-1] for s in stack)))
rows.sort(reverse=True)
tot = sum(r[0] for r in rows)
print(f"total device self time of listed ops {tot / 1e3:.2f} ms")
for dt, n, k, st in rows[:90]:
    print(f"{dt / 1e3:8.3f} ms {n:4d}  {k[:48]:48s} {st}")
