"""One profiled training step (dev tool for ncu): 2 warm-up steps, then exactly one step between
cudaProfilerStart/Stop (run ncu with --profile-from-start off).  Usage: profile_step.py [B] [T]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 800
    from gantron_b200.synth import synthetic_batch
    torch.cuda.set_device(0)
    tr = bench.Trainer(False)
    batch = synthetic_batch(B, 150, T, seed=1234)
    x, y = tr.G.parse_batch(batch)
    for _ in range(2):
        tr.step(x, y)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    tr.step(x, y)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    print("profiled one step", B, T)


main()
