"""VGG16 path without a GPU: the oracle reproduces the committed golden vectors and the b200net module tree is a
drop-in for torchvision's (repr, state_dict keys, the reference's classifier[6] swap)."""
import os

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_vgg_oracle_reproduces_golden():
    from oracle import torch_oracle as O
    gold = torch.load(os.path.join(ROOT, "tests", "golden", "vgg_small.pt"), weights_only=False)
    g = gold["mode_F"]
    model = O.build_vgg_reference(dropout=False)
    res = O.run_vgg_case(model, "F", gold["n"], gold["hw"], gold["steps"])
    assert abs(res["losses"][0] - g["losses"][0]) <= 1e-5 * abs(g["losses"][0])
    assert torch.allclose(res["logps"][0], g["logps"][0], rtol=1e-4, atol=1e-5)
    assert torch.equal(res["eval_top"], g["eval_top"])
    sd = model.state_dict()
    for k, d in g["weights"].items():
        assert torch.allclose(O.digest(sd[k])["samples"], d["samples"], rtol=1e-3, atol=1e-6), k
    # frozen part really frozen
    ref = O.build_vgg_reference(dropout=False).state_dict()
    assert torch.equal(sd["features.0.weight"], ref["features.0.weight"])
    assert not torch.equal(sd["classifier.6.0.weight"], ref["classifier.6.0.weight"])


def test_vgg_module_tree_is_torchvision_compatible():
    import torchvision
    import b200net
    ours, ref = b200net.models.vgg16(pretrained=False), torchvision.models.vgg16(weights=None)
    assert repr(ours) == repr(ref)
    so, sr = ours.state_dict(), ref.state_dict()
    assert list(so.keys()) == list(sr.keys()) and all(so[k].shape == sr[k].shape for k in so)
    ours.load_state_dict(sr)
    assert sum(p.numel() for p in ours.parameters()) == 138357544
    # the reference's recipe (another_neural_net.py:247-258)
    for param in ours.parameters():
        param.requires_grad = False
    ours.classifier[6] = torch.nn.Sequential(torch.nn.Linear(4096, 256), torch.nn.ReLU(), torch.nn.Dropout(0.4),
                                             torch.nn.Linear(256, 10), torch.nn.LogSoftmax(dim=1))
    opt = b200net.optim.Adam(ours.parameters())
    assert sum(p.numel() for p in ours.parameters() if p.requires_grad) == 4096 * 256 + 256 + 2560 + 10
    assert len(opt.param_groups[0]["params"]) == 34
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ours(torch.rand(1, 3, 32, 32))
