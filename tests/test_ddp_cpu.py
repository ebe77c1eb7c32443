"""world_size-2 checks of b200net.parallel.DistributedDataParallel on CPU tensors over gloo: construction-time
broadcast, bucket layout (reverse registration order, 1 MiB first bucket), gradient averaging equal to stock
torch DDP and to the hand-computed mean, unused parameters, no_sync, buffer broadcast.  The CUDA/NCCL path uses
the same bucket logic; it is exercised by tests/test_ddp_gpu.py on a multi-GPU box."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.a = nn.Linear(300, 600)          # 180k + 600 params
        self.bn = nn.BatchNorm1d(600)
        self.b = nn.Linear(600, 500)          # 300k params -> its own bucket after the 1 MiB first bucket
        self.unused = nn.Linear(4, 4)         # never touched in forward
        self.c = nn.Linear(500, 10)

    def forward(self, x):
        return self.c(torch.relu(self.b(torch.relu(self.bn(self.a(x))))))


def _data(rank, step):
    g = torch.Generator().manual_seed(100 * step + rank)
    return torch.randn(8, 300, generator=g), torch.randint(0, 10, (8,), generator=g)


def _worker(rank, world, port, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import b200net
        from b200net.parallel import DistributedDataParallel as DDP
        torch.manual_seed(1234 + rank)                 # different initial weights per rank on purpose
        ours = Net()
        ref = Net()
        torch.manual_seed(1234)                        # what rank 0 holds
        rank0 = Net()
        ddp = DDP(ours, device_ids=None, output_device=None, bucket_cap_mb=1, find_unused_parameters=True)
        # (1) construction broadcast: every rank now holds rank 0's parameters and buffers
        for (k, v), (_, v0) in zip(ours.state_dict().items(), rank0.state_dict().items()):
            assert torch.equal(v, v0), k
        ref.load_state_dict(rank0.state_dict())
        stock = torch.nn.parallel.DistributedDataParallel(ref, find_unused_parameters=True)
        # (2) bucket layout: reverse registration order, first bucket <= 1 MiB
        layout = ddp.bucket_layout()
        flat_sizes = [n for _, sizes in layout for n in sizes]
        assert flat_sizes == [p.numel() for p in reversed(list(ours.parameters()))]
        assert sum(layout[0][1]) * 4 <= 1024 * 1024 and len(layout) >= 2
        crit = nn.CrossEntropyLoss()
        for step in range(3):
            x, y = _data(rank, step)
            for m in (ddp, stock):
                m.zero_grad()
                crit(m(x), y).backward()
            # (3) same averaged gradients as stock DDP; grads are views of the flat buckets
            for (n1, p1), (n2, p2) in zip(ours.named_parameters(), ref.named_parameters()):
                if n1.startswith("unused"):
                    assert p1.grad is None or float(p1.grad.abs().sum()) == 0.0
                    continue
                assert torch.allclose(p1.grad, p2.grad, rtol=1e-5, atol=1e-6), (step, n1)
                assert p1.grad.data_ptr() == ddp.grad_view(p1).data_ptr()
            # (4) and the same as the hand-computed mean over both ranks' batches
            if step == 0:
                local = Net()
                local.load_state_dict(rank0.state_dict())
                acc = None
                for r in range(world):
                    local.zero_grad()
                    local.train()
                    local.load_state_dict(rank0.state_dict())
                    xr, yr = _data(r, 0)
                    crit(local(xr), yr).backward()
                    g = [p.grad.clone() if p.grad is not None else None for p in local.parameters()]
                    acc = g if acc is None else [a + b if a is not None else None for a, b in zip(acc, g)]
                for (n1, p1), a in zip(ours.named_parameters(), acc):
                    if a is not None:
                        assert torch.allclose(p1.grad, a / world, rtol=1e-5, atol=1e-6), n1
            with torch.no_grad():
                for p1, p2 in zip(ours.parameters(), ref.parameters()):
                    if p1.grad is not None and p2.grad is not None:
                        p1 -= 0.1 * p1.grad
                        p2 -= 0.1 * p2.grad
        # (5) BN running statistics follow rank 0 (broadcast_buffers) on every rank
        gathered = [torch.zeros_like(ours.bn.running_mean) for _ in range(world)]
        ddp(_data(rank, 9)[0])   # triggers the buffer broadcast
        dist.all_gather(gathered, ours.bn.running_mean.clone())
        # (6) no_sync: gradients stay local
        with ddp.no_sync():
            ddp.zero_grad()
            x, y = _data(rank, 7)
            crit(ddp(x), y).backward()
        g_local = ours.c.weight.grad.clone()
        others = [torch.zeros_like(g_local) for _ in range(world)]
        dist.all_gather(others, g_local)
        assert not torch.allclose(others[0], others[1])
        # (7) gradient accumulation under no_sync when a parameter is used in the first micro-step only: the second
        # backward must not wipe what the first one left in that parameter's bucket slice
        class Two(nn.Module):
            def __init__(self):
                super().__init__()
                self.a = nn.Linear(6, 6)
                self.b = nn.Linear(6, 6)

            def forward(self, x, use_b):
                return self.b(self.a(x)) if use_b else self.a(x)

        torch.manual_seed(7)
        two = Two()
        d2 = DDP(two, device_ids=None, output_device=None, find_unused_parameters=True)
        xs = torch.randn(4, 6, generator=torch.Generator().manual_seed(50 + rank))
        d2.zero_grad()
        with d2.no_sync():
            d2(xs, True).sum().backward()
            gb = two.b.weight.grad.clone()
            assert float(gb.abs().sum()) > 0
            d2(xs, False).sum().backward()
            assert torch.equal(two.b.weight.grad, gb), "no_sync micro-step wiped an accumulated gradient"
        # (8) without find_unused_parameters an unused parameter is an error, as in torch's DDP
        torch.manual_seed(7)
        three = Two()
        d3 = DDP(three, device_ids=None, output_device=None)
        raised = False
        try:
            d3(xs, False).sum().backward()
        except RuntimeError as e:
            raised = "find_unused_parameters" in str(e)
        assert raised
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_ddp_world2_gloo():
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(_worker, args=(world, port, results), nprocs=world, join=True)
    assert dict(results) == {0: "ok", 1: "ok"}
