"""Generates tests/golden/bert_small.pt (run in the build container): outputs of oracle/bert_oracle.py -- the
transformers BertForSequenceClassification on CPU executing the reference's statements
(pytorch_on_language_distr.py:257-275) -- for a small seeded case: bert-base width, 2 layers, N=4, S=32, dropout 0,
3 training steps (AdamW 2e-5 / eps 1e-8, clip 1.0, linear schedule) + one eval pass; fp32 (the reference's
arithmetic) and fp64 (the arbiter, key suffix _f64).  Stored: losses, logits per step, eval logits, the clipped
gradients of step 0 and the after-step values of a watch-list of parameters (strided digests).

usage: python oracle/make_golden_bert.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import bert_oracle as B  # noqa: E402
from oracle.torch_oracle import digest, cpu_threads  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
N, S, LAYERS, STEPS = 4, 32, 2, 3


def main():
    import transformers
    torch.set_num_threads(cpu_threads())
    out = {"torch": torch.__version__, "transformers": transformers.__version__, "n": N, "s": S, "layers": LAYERS,
           "steps": STEPS}
    for suffix, dtype in (("", torch.float32), ("_f64", torch.float64)):
        model = B.build_reference_model(num_layers=LAYERS, dtype=dtype)
        g = B.run_case(model, N, S, 1, keep_grads=True)
        grads = {k: digest(g["grads"][k]) for k in B.WATCH}
        model = B.build_reference_model(num_layers=LAYERS, dtype=dtype)
        res = B.run_case(model, N, S, STEPS)
        sd = model.state_dict()
        res["weights"] = {k: digest(sd[k]) for k in B.WATCH}
        res["grads_step0"] = grads
        out["case" + suffix] = res
    torch.save(out, os.path.join(GOLD, "bert_small.pt"))
    print("losses", out["case"]["losses"], out["case_f64"]["losses"])
    print("size", os.path.getsize(os.path.join(GOLD, "bert_small.pt")))


if __name__ == "__main__":
    main()
