"""Per-layer-class timing of the bf16 conv kernels (fprop / dgrad / wgrad) against their rooflines.
Classes = SURVEY.md 8(a') at N images of 160x160 (ResNet-50), or with a third argument `vgg` the twelve 64+-channel
3x3 layers of VGG16 at 224x224.  CUDA events, 3 warm-up + 10 timed, inputs > L2 rotate is not
needed here: every class at N=256 moves more than the 126 MB L2 or is tensor-bound.  usage: conv_table.py [N] [filter] [vgg]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200net
from b200net import ops

PK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))) if os.path.exists(
    os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0, "hbm_gbs": 6650.0}
CLASSES = [  # name, H, C, K, R, stride, pad, count
    ("a", 40, 64, 64, 1, 1, 0, 1), ("b", 40, 64, 64, 3, 1, 1, 3), ("c", 40, 64, 256, 1, 1, 0, 4),
    ("d", 40, 256, 64, 1, 1, 0, 2), ("e", 40, 256, 128, 1, 1, 0, 1), ("f", 40, 128, 128, 3, 2, 1, 1),
    ("g", 20, 128, 512, 1, 1, 0, 4), ("h", 40, 256, 512, 1, 2, 0, 1), ("i", 20, 512, 128, 1, 1, 0, 3),
    ("j", 20, 128, 128, 3, 1, 1, 3), ("k", 20, 512, 256, 1, 1, 0, 1), ("l", 20, 256, 256, 3, 2, 1, 1),
    ("m", 10, 256, 1024, 1, 1, 0, 6), ("n", 20, 512, 1024, 1, 2, 0, 1), ("o", 10, 1024, 256, 1, 1, 0, 5),
    ("p", 10, 256, 256, 3, 1, 1, 5), ("q", 10, 1024, 512, 1, 1, 0, 1), ("r", 10, 512, 512, 3, 2, 1, 1),
    ("s", 5, 512, 2048, 1, 1, 0, 3), ("t", 10, 1024, 2048, 1, 2, 0, 1), ("u", 5, 2048, 512, 1, 1, 0, 2),
    ("v", 5, 512, 512, 3, 1, 1, 2),
]
VGG_CLASSES = [
    ("A", 224, 64, 64, 3, 1, 1, 1), ("B", 112, 64, 128, 3, 1, 1, 1), ("C", 112, 128, 128, 3, 1, 1, 1),
    ("D", 56, 128, 256, 3, 1, 1, 1), ("E", 56, 256, 256, 3, 1, 1, 2), ("F", 28, 256, 512, 3, 1, 1, 1),
    ("G", 28, 512, 512, 3, 1, 1, 2), ("H", 14, 512, 512, 3, 1, 1, 3),
]


GRAPH = os.environ.get("CONV_TABLE_EAGER") is None


def timeit(fn, iters=10):
    """us per call.  By default the `iters` calls are captured into one CUDA graph and the graph is replayed (that is
    how the training step runs them, and it keeps host launch overhead -- tensor-map encodes, ctypes -- out of the
    number); CONV_TABLE_EAGER=1 times eager launches instead."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if GRAPH:
        ops.NO_WORKSPACE_GROWTH[0] = True
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(iters):
                    fn()
        finally:
            ops.NO_WORKSPACE_GROWTH[0] = False
        g.replay()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(3):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / (3 * iters) * 1e3
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    filt = sys.argv[2] if len(sys.argv) > 2 else ""
    classes = VGG_CLASSES if (len(sys.argv) > 3 and sys.argv[3] == "vgg") else CLASSES
    if filt == "-":
        filt = ""
    dev = torch.device("cuda:0")
    peak, hbm = PK["bf16_tflops"] * 1e12, PK["hbm_gbs"] * 1e9
    tot = {"fprop": [0, 0], "dgrad": [0, 0], "wgrad": [0, 0]}
    print("%-3s %-26s %9s | %8s %8s %5s | %8s %8s %5s | %8s %8s %5s" % ("cls", "shape", "MFLOP/img", "fprop us", "roof us", "frac", "dgrad us", "roof us", "frac", "wgrad us", "roof us", "frac"))
    for name, H, C, K, R, st, pad, cnt in classes:
        if filt and name not in filt:
            continue
        Ho = (H + 2 * pad - R) // st + 1
        x = torch.randn(N, H, H, C, device=dev).to(torch.bfloat16)
        w = torch.randn(K, C, R, R, device=dev) / (C * R * R) ** 0.5
        wp = ops.pack_weight(w, torch.bfloat16)
        dy = torch.randn(N, Ho, Ho, K, device=dev).to(torch.bfloat16)
        dx = torch.empty_like(x)
        flops = 2.0 * N * Ho * Ho * K * C * R * R
        by_f = 2.0 * (x.numel() + dy.numel() + wp.numel())
        by_w = 2.0 * (x.numel() + dy.numel()) + 4.0 * w.numel()
        roof_f = max(flops / peak, by_f / hbm) * 1e6
        roof_w = max(flops / peak, by_w / hbm) * 1e6
        tf = timeit(lambda: ops.conv_fprop(x, wp, K, R, R, st, pad))
        td = timeit(lambda: ops.conv_dgrad(dy, wp, x.shape, K, R, R, st, pad, out=dx))
        tw = timeit(lambda: ops.conv_wgrad(x, dy, K, R, R, st, pad))
        print("%-3s %-26s %9.1f | %8.1f %8.1f %5.2f | %8.1f %8.1f %5.2f | %8.1f %8.1f %5.2f" % (
            name, "%dx%d %d->%d k%d s%d x%d" % (H, H, C, K, R, st, cnt), flops / N / 1e6, tf, roof_f, roof_f / tf, td, roof_f, roof_f / td, tw, roof_w, roof_w / tw))
        for k, t, r in (("fprop", tf, roof_f), ("dgrad", td, roof_f), ("wgrad", tw, roof_w)):
            tot[k][0] += t * cnt
            tot[k][1] += r * cnt
    for k, (t, r) in tot.items():
        print("total %s: measured %.2f ms, per-layer roofline %.2f ms (%.2f)" % (k, t / 1e3, r / 1e3, r / max(t, 1e-9)))


if __name__ == "__main__":
    main()
