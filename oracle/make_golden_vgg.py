"""Generates tests/golden/vgg_small.pt (run in the build container): the oracle (torchvision vgg16 on CPU executing
the reference's statements, another_neural_net.py:244-300) for a small seeded case -- N=4, 64x64 (features end at
2x2, AdaptiveAvgPool2d(7) upsamples), dropout 0, 2 training steps + one eval pass -- in Mode F (the reference's recipe:
frozen pretrained part, Adam on classifier[6]) and Mode A (all layers, SGD 0.001); fp32 and fp64 (key suffix _f64).

usage: python oracle/make_golden_vgg.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import torch_oracle as O  # noqa: E402

N, HW, STEPS = 4, 64, 2


def main():
    torch.set_num_threads(O.cpu_threads())
    out = {"torch": torch.__version__, "n": N, "hw": HW, "steps": STEPS}
    for mode in ("A", "F"):
        for suffix, dtype in (("", torch.float32), ("_f64", torch.float64)):
            model = O.build_vgg_reference(dropout=False).to(dtype)
            res = O.run_vgg_case(model, mode, N, HW, STEPS)
            sd = model.state_dict()
            res["weights"] = {k: O.digest(sd[k]) for k in O.VGG_WATCH}
            out["mode_" + mode + suffix] = res
            print(mode, suffix, res["losses"])
    path = os.path.join(ROOT, "tests", "golden", "vgg_small.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path))


if __name__ == "__main__":
    main()
