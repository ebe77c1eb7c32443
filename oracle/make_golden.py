"""Generates tests/golden/* (run in the build container, where /root/reference is mounted):

  resnet50_head_repr.txt   the `repr(model)` the reference stored in its notebook output
                           (pytorch_training_inference_on_image.ipynb, the cell whose source ends with
                           `model.to(device)`, raw lines 454-635) -- a golden vector held by the reference itself
  oracle_small.pt          numeric outputs of the oracle (stock torch CPU) for small seeded cases:
                           Mode A (all layers, SGD 0.001) and Mode F (frozen backbone, Adam 0.003), 2 steps each,
                           N=8, 96x96 -- logits, losses, eval logits / argmax, digests of after-step weights & buffers,
                           in fp32 (the reference's arithmetic) and in fp64 (the arbiter, key suffix _f64)

usage: python oracle/make_golden.py
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import torch_oracle as O  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
N, HW = 8, 96
NB = "/root/reference/pytorch_training_inference_on_image.ipynb"


def stored_repr():
    nb = json.load(open(NB))
    for c in nb["cells"]:
        src = "".join(c.get("source", []))
        if "model.fc = nn.Sequential" in src and "resnet" not in src.lower() and src.rstrip().endswith("model.to(device)"):
            for o in c.get("outputs", []):
                txt = "".join(o.get("data", {}).get("text/plain", [])) or "".join(o.get("text", []))
                if txt.startswith("ResNet("):
                    return txt
    raise SystemExit("stored repr not found")


def main():
    os.makedirs(GOLD, exist_ok=True)
    if os.path.exists(NB):
        open(os.path.join(GOLD, "resnet50_head_repr.txt"), "w").write(stored_repr())
    torch.set_num_threads(O.cpu_threads())
    out = {"torch": torch.__version__, "n": N, "hw": HW}
    for mode in ("A", "F"):
        for suffix, dtype in (("", torch.float32), ("_f64", torch.float64)):
            model = O.build_reference_model(dropout=False).to(dtype)
            res = O.run_case(model, mode, n=N, hw=HW, steps=2)
            res["weights"] = {k: O.digest(v) for k, v in model.state_dict().items() if v.dtype.is_floating_point}
            out["mode_" + mode + suffix] = res
    torch.save(out, os.path.join(GOLD, "oracle_small.pt"))
    print("wrote", os.listdir(GOLD))


if __name__ == "__main__":
    main()
