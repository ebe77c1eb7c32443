"""GPU half of the checkpoint tests: after loading the files written by tests/test_checkpoint_cpu.py through the
reference's own calls, eval-mode outputs match the oracle module that wrote them (fp32 mode: 1e-4 of the output scale,
argmax identical; bf16 mode: 2e-2)."""
import pytest
import torch

from tests.test_checkpoint_cpu import write_bert_checkpoint, write_resnet_checkpoint

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def test_resnet50_checkpoint_file_eval_parity(tmp_path, monkeypatch):
    import b200net
    ref = write_resnet_checkpoint(str(tmp_path)).eval()
    monkeypatch.setenv("TORCH_HOME", str(tmp_path))
    x = torch.rand(8, 3, 160, 160, generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        want = ref(x)
    for prec, tol in (("fp32", 1e-4), ("bf16", 2e-2)):
        b200net.set_precision(prec)
        try:
            model = b200net.models.resnet50(pretrained=True).to(DEV).eval()
            with torch.no_grad():
                got = model(x.to(DEV))
            assert got.shape == (8, 1000)
            assert _rel(got, want) < tol, prec
            if prec == "fp32":
                assert torch.equal(got.argmax(1).cpu(), want.argmax(1))
        finally:
            b200net.set_precision("bf16")


@pytest.mark.parametrize("safetensors", [True, False])
def test_bert_checkpoint_directory_eval_parity(tmp_path, safetensors):
    import b200net
    from b200net import bert
    from oracle import bert_oracle as B
    d = str(tmp_path / "ckpt")
    ref = write_bert_checkpoint(d, layers=2, safetensors=safetensors).eval()
    ids, mask, _ = B.make_batch(4, 128, 5)
    with torch.no_grad():
        want = ref(ids, token_type_ids=None, attention_mask=mask)[0]
    b200net.set_precision("fp32")
    try:
        model = bert.BertForSequenceClassification.from_pretrained(d, num_labels=2, output_attentions=False,
                                                                   output_hidden_states=False).to(DEV).eval()
        with torch.no_grad():
            got = model(ids.to(DEV), token_type_ids=None, attention_mask=mask.to(DEV))[0]
        assert _rel(got, want) < 1e-4
    finally:
        b200net.set_precision("bf16")
