"""Host-side drop-in fidelity that needs no GPU: module tree, repr, state_dict, pickling, optimizer protocol."""
import io
import os

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _head():
    return torch.nn.Sequential(torch.nn.Linear(2048, 512), torch.nn.ReLU(), torch.nn.Dropout(0.2),
                               torch.nn.Linear(512, 10), torch.nn.LogSoftmax(dim=1))


def test_repr_matches_the_reference_notebook_output():
    """pytorch_training_inference_on_image.ipynb:454-635 stores repr(model) after the head swap."""
    import b200net
    stored = open(os.path.join(ROOT, "tests", "golden", "resnet50_head_repr.txt")).read()
    m = b200net.models.resnet50(pretrained=False)
    m.fc = _head()
    assert repr(m) == stored


def test_oracle_module_tree_matches_the_reference_notebook_output():
    from oracle import torch_oracle as O
    stored = open(os.path.join(ROOT, "tests", "golden", "resnet50_head_repr.txt")).read()
    assert repr(O.build_reference_model()) == stored


def test_state_dict_is_torchvision_compatible():
    import b200net
    import torchvision
    ours = b200net.models.resnet50()
    ref = torchvision.models.resnet50(weights=None)
    so, sr = ours.state_dict(), ref.state_dict()
    assert list(so.keys()) == list(sr.keys())
    assert all(so[k].shape == sr[k].shape and so[k].dtype == sr[k].dtype for k in so)
    ours.load_state_dict(sr)  # OIHW fp32 in, no conversion at the boundary
    assert sum(p.numel() for p in ours.parameters()) == 25557032


def test_freeze_and_head_swap_protocol():
    """...on_image.ipynb:436-446"""
    import b200net
    model = b200net.models.resnet50(pretrained=False)
    for param in model.parameters():
        param.requires_grad = False
    model.fc = _head()
    opt = b200net.optim.Adam(model.fc.parameters(), lr=0.003)
    assert sum(p.numel() for p in model.parameters() if p.requires_grad) == 1054218
    assert len(opt.param_groups[0]["params"]) == 4 and opt.param_groups[0]["lr"] == 0.003


def test_torch_save_roundtrip():
    """torch.save(model, path) / torch.load(path): ...on_image.ipynb:700,852"""
    import b200net
    m = b200net.models.resnet50()
    m.fc = _head()
    buf = io.BytesIO()
    torch.save(m, buf)
    buf.seek(0)
    m2 = torch.load(buf, weights_only=False)
    assert repr(m2) == repr(m)
    assert torch.equal(m2.layer3[2].conv2.weight, m.layer3[2].conv2.weight)


def test_cpu_tensors_are_rejected_not_silently_computed():
    import b200net
    m = b200net.models.resnet50()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.rand(1, 3, 32, 32))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        b200net.nn.Linear(4, 4)(torch.rand(2, 4))


def test_optimizers_follow_the_torch_protocol():
    """LambdaLR / get_linear_schedule_with_warmup drive param_groups[i]['lr'] (pytorch_on_language_distr.py:181-183)."""
    import b200net
    w = torch.nn.Parameter(torch.zeros(4))
    for cls, kw in ((b200net.optim.SGD, dict(lr=0.001)), (b200net.optim.Adam, dict(lr=0.003)),
                    (b200net.optim.AdamW, dict(lr=2e-5, eps=1e-8))):
        opt = cls([w], **kw)
        assert isinstance(opt, torch.optim.Optimizer)
        sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: max(0.0, (10 - s) / 10))
        opt.step()      # no grads yet: nothing to launch, must not touch the GPU
        sched.step()
        assert abs(opt.param_groups[0]["lr"] - kw["lr"] * 0.9) < 1e-12
        opt.zero_grad()
        sd = opt.state_dict()
        assert "param_groups" in sd
    assert b200net.optim.AdamW([w]).defaults["weight_decay"] == 1e-2


def test_precision_switch():
    import b200net
    prev = b200net.get_precision()
    b200net.set_precision("fp32")
    assert b200net.functional.act_dtype() == torch.float32
    b200net.set_precision("bf16")
    assert b200net.functional.act_dtype() == torch.bfloat16
    with pytest.raises(ValueError):
        b200net.set_precision("fp8")
    b200net.set_precision(prev)


def test_torch_save_roundtrip_bert_and_vgg():
    """torch.save(model, path) / torch.load(path) (...on_image.ipynb:700,852) for the other two model families."""
    import b200net
    from b200net import bert
    for m in (bert.BertForSequenceClassification(bert.BertConfig(num_hidden_layers=1)),
              b200net.models.vgg16(pretrained=False)):
        buf = io.BytesIO()
        torch.save(m, buf)
        buf.seek(0)
        m2 = torch.load(buf, weights_only=False)
        assert repr(m2) == repr(m)
        a, b = next(iter(m.parameters())), next(iter(m2.parameters()))
        assert torch.equal(a, b)


def test_bench_per_layer_roofline_constant():
    """bench.py's per-layer convolution roofline (sum over the 53 convolutions x {fprop, dgrad, wgrad} of
    max(tensor time, HBM time)) counts 53 layers and 12.4 GFLOP per image, scales linearly with the batch, and equals
    the pure tensor time when HBM is infinitely fast / the pure HBM time when the tensor pipe is."""
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_for_test", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert sum(c[-1] for c in bench.RESNET50_CONV_CLASSES) == 52          # + the stem = 53 convolutions
    fwd = sum(2.0 * ((H + 2 * p - R) // s + 1) ** 2 * K * C * R * R * n for H, C, K, R, s, p, n in bench.RESNET50_CONV_CLASSES)
    fwd += 2.0 * 80 * 80 * 64 * 147
    assert abs(fwd / 1e9 - 4.1727 + 2 * (2048 * 512 + 512 * 10) / 1e9) < 0.02      # 4.173 GFLOP/img incl. head
    a = bench.conv_per_layer_roofline_ms(256, 1402.4, 6550.1)
    assert abs(bench.conv_per_layer_roofline_ms(512, 1402.4, 6550.1) / a - 2.0) < 0.02
    tensor_only = bench.conv_per_layer_roofline_ms(256, 1402.4, 1e12)
    stem = 2.0 * 256 * 80 * 80 * 64 * 147
    flops = 3.0 * 256 * (fwd - 2.0 * 80 * 80 * 64 * 147) + 2.0 * stem             # no dgrad for the stem
    assert abs(tensor_only - flops / 1402.4e12 * 1e3) / tensor_only < 1e-6
    assert a >= tensor_only and a >= bench.conv_per_layer_roofline_ms(256, 1e9, 6550.1)
