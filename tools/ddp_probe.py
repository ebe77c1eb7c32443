"""Where does the DDP step go?  torchrun --nproc-per-node 2 tools/ddp_probe.py : ResNet-50 step time for
(a) no DDP wrapper, (b) DDP with no_sync (no collectives), (c) DDP without buffer broadcast, (d) full DDP."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import b200net
from oracle import torch_oracle as O

rank = int(os.environ.get("RANK", 0)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)


def build():
    torch.manual_seed(42)
    m = b200net.models.resnet50()
    m.fc = b200net.nn.Sequential(b200net.nn.Linear(2048, 512), b200net.nn.ReLU(), b200net.nn.Dropout(0.2),
                                 b200net.nn.Linear(512, 10), b200net.nn.LogSoftmax(dim=1))
    return m.to(dev).train()


x, y = O.make_batch(256, 160, 1000 + rank, dev)
crit = b200net.nn.NLLLoss()


def run(tag, net, model, ctx=None):
    opt = b200net.optim.SGD(model.parameters(), lr=0.001)

    def step():
        opt.zero_grad()
        if ctx is not None:
            with ctx():
                crit(net(x), y).backward()
        else:
            crit(net(x), y).backward()
        opt.step()
    for _ in range(8):
        step()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        step()
    e1.record(); torch.cuda.synchronize()
    if rank == 0:
        print("%-40s %.2f ms/step" % (tag, e0.elapsed_time(e1) / 20), flush=True)


m = build(); run("no wrapper", m, m)
m = build(); d = b200net.parallel.DistributedDataParallel(m, device_ids=[local]); run("DDP no_sync", d, m, d.no_sync)
m = build(); d = b200net.parallel.DistributedDataParallel(m, device_ids=[local], broadcast_buffers=False); run("DDP, no buffer broadcast", d, m)
m = build(); d = b200net.parallel.DistributedDataParallel(m, device_ids=[local]); run("DDP", d, m)
dist.destroy_process_group()
