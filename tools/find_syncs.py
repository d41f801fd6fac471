"""List the host<->device synchronisations of one training step (torch sync debug mode) and the host time the step's
launch loop takes against its device time: the host-bound share is what grows under N ranks sharing the box's cores."""
import sys, os, time, warnings
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

tr = bench.Trainer(False)
batch = bench._pinned_batch(bench.B_PER_GPU, seed=1234)
x, y = tr.G.parse_batch(batch)
for _ in range(3):
    tr.step(x, y)
torch.cuda.synchronize()
torch.cuda.set_sync_debug_mode("warn")
with warnings.catch_warnings(record=True) as w:
    warnings.simplefilter("always")
    import traceback
    orig = warnings.showwarning
    tr.step(x, y)
torch.cuda.set_sync_debug_mode("default")
print("syncs in one step:", len(w))
for x_ in w:
    print("  ", x_.filename.replace(os.getcwd() + "/", ""), x_.lineno, str(x_.message)[:80])
# host launch-loop time with the GPU idle at the start vs device time
for rep in range(3):
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    tr.step(x, y)
    e1.record(); t1 = time.perf_counter()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"host returns after {1e3*(t1-t0):.2f} ms; device {e0.elapsed_time(e1):.2f} ms; wall {1e3*(t2-t0):.2f} ms")
