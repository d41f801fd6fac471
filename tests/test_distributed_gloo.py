"""CPU, world_size 2 over gloo: the drop-in apply_gradient_allreduce (replaces distributed.py:126-173)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _Net(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(7, 5)
        self.bn = torch.nn.BatchNorm1d(5)
        self.b = torch.nn.Linear(5, 3)
        self.unused = torch.nn.Linear(2, 2)       # never receives a gradient

    def forward(self, x):
        return self.b(torch.tanh(self.bn(self.a(x))))


def _worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from gantron_b200.distributed import apply_gradient_allreduce, reduce_tensor
        import gantron_b200.distributed as GD
        torch.manual_seed(100 + rank)               # different initial weights per rank
        net = _Net()
        GD.BUCKET_BYTES = 128                        # force several buckets
        net = apply_gradient_allreduce(net)
        assert apply_gradient_allreduce(net) is net  # idempotent (train.py applies it twice)
        # 1. parameters AND buffers were broadcast from rank 0
        flat = torch.cat([p.detach().reshape(-1) for p in net.parameters()] + [net.bn.running_var.reshape(-1)])
        ref = flat.clone()
        dist.broadcast(ref, 0)
        assert torch.equal(flat, ref), "state not broadcast"
        assert len(net._gtn_reducer.buckets) > 1
        # 2. gradients are world-averaged after backward, for two consecutive steps
        for step in range(2):
            net.zero_grad()
            g = torch.Generator().manual_seed(1000 + 10 * step + rank)
            x = torch.randn(6, 7, generator=g)
            net(x).square().sum().backward()
            local = _Net()
            local.load_state_dict(net.state_dict())
            local.train()
            local(x).square().sum().backward()
            for (n, p), (_, q) in zip(net.named_parameters(), local.named_parameters()):
                if q.grad is None:
                    assert p.grad is None or float(p.grad.abs().sum()) == 0, n
                    continue
                want = q.grad.clone()
                dist.all_reduce(want)
                want /= world
                assert torch.allclose(p.grad, want, atol=1e-6, rtol=1e-5), f"step {step} {n}"
        # 2b. a backward pass that raises half way (after some gradients were bucketed) must not poison later passes
        class _Boom(torch.autograd.Function):
            @staticmethod
            def forward(ctx, x):
                return x.clone()

            @staticmethod
            def backward(ctx, g):
                raise RuntimeError("boom")
        net.zero_grad()
        x = torch.randn(6, 7, generator=torch.Generator().manual_seed(5 + rank))
        h = torch.tanh(net.bn(_Boom.apply(net.a(x))))
        with pytest.raises(RuntimeError, match="boom"):
            net.b(h).square().sum().backward()          # net.b got its gradients, net.a never did
        net.zero_grad()
        net(x).square().sum().backward()                 # a clean step afterwards is averaged again
        local = _Net()
        local.load_state_dict(net.state_dict())
        local.train()
        local(x).square().sum().backward()
        for (n, p), (_, q) in zip(net.named_parameters(), local.named_parameters()):
            if q.grad is None:
                continue
            want = q.grad.clone()
            dist.all_reduce(want)
            want /= world
            assert torch.allclose(p.grad, want, atol=1e-6, rtol=1e-5), f"after aborted backward: {n}"
        # 3. a module whose forward hook never fires (the reference's Discriminator bug) still reduces
        assert net._gtn_reducer.launched_early >= 1
        t = reduce_tensor(torch.tensor([float(rank + 1)]), world)
        assert abs(float(t) - (1 + world) / 2) < 1e-6
        open(os.path.join(tmp, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_apply_gradient_allreduce_world2_gloo(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(tmp_path / f"ok{r}") for r in range(world))
