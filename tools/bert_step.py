"""Times the C5 per-GPU step (bert-base, N=32, S=256, bf16, dropout 0.1; AdamW 2e-5 + clip 1.0 + linear schedule:
pytorch_on_language_distr.py:257-275) on one GPU.  usage: python tools/bert_step.py [--steps K] [--n N] [--s S]"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import b200net  # noqa: E402
from b200net import bert  # noqa: E402
from oracle import bert_oracle as B  # noqa: E402  (data generator only)

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--n", type=int, default=32)
ap.add_argument("--s", type=int, default=256)
ap.add_argument("--fwd-only", action="store_true")
args = ap.parse_args()
dev = torch.device("cuda:0")
torch.manual_seed(42)
model = bert.BertForSequenceClassification(bert.BertConfig(num_labels=2)).to(dev).train()
opt = bert.AdamW(model.parameters(), lr=2e-5, eps=1e-8)
sched = bert.get_linear_schedule_with_warmup(opt, 0, 1000)
ids, mask, labels = B.make_batch(args.n, args.s, 0, dev)


def step():
    model.zero_grad()
    out = model(ids, token_type_ids=None, attention_mask=mask, labels=labels)
    loss = out[0]
    if not args.fwd_only:
        loss.backward()
        b200net.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        sched.step()
    return loss


for _ in range(args.warmup):
    step()
torch.cuda.synchronize()
lib = b200net._C.lib
l0 = lib.b200_launch_count()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for _ in range(args.steps):
    loss = step()
e1.record()
t_enq = time.perf_counter() - t0
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / args.steps
print("loss %.4f  %.2f ms/step  %.1f seq/s  %.0f tok/s  enqueue %.2f ms/step  launches/step %d  peak mem %.1f GB"
      % (loss.item(), ms, args.n / ms * 1e3, args.n * args.s / ms * 1e3, t_enq / args.steps * 1e3,
         (lib.b200_launch_count() - l0) // args.steps, torch.cuda.max_memory_allocated() / 2**30))
