"""Checkpoint compatibility, proven on files: a torchvision-format `resnet50-*.pth` and Hugging Face-format BERT
directories (`model.safetensors` / `pytorch_model.bin` + `config.json`), written here from the ORACLE's randomly
initialised modules (no download needed), are loaded through the calls the reference makes --
`models.resnet50(pretrained=True)` (pytorch_training_inference_on_image.ipynb:389) and
`BertForSequenceClassification.from_pretrained(dir, num_labels=2, ...)` (pytorch_on_language_distr.py:155-161) --
and every tensor arrives bit-exact.  The GPU half (same files, eval-logit parity) is tests/test_checkpoint_gpu.py."""
import os

import pytest
import torch


def write_resnet_checkpoint(root, seed=7):
    import torchvision
    torch.manual_seed(seed)
    ref = torchvision.models.resnet50(weights=None)
    # non-trivial BatchNorm buffers, as a trained checkpoint has
    for m in ref.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.running_mean.normal_(0, 0.1)
            m.running_var.uniform_(0.5, 1.5)
            m.num_batches_tracked.fill_(123)
    d = os.path.join(root, "hub", "checkpoints")
    os.makedirs(d, exist_ok=True)
    torch.save(ref.state_dict(), os.path.join(d, "resnet50-0676ba61.pth"))     # torchvision's IMAGENET1K_V1 file name
    return ref


def write_bert_checkpoint(path, layers=2, safetensors=True, seed=9):
    import transformers
    torch.manual_seed(seed)
    cfg = transformers.BertConfig(num_labels=2, num_hidden_layers=layers)
    ref = transformers.BertForSequenceClassification(cfg)
    if safetensors:
        ref.save_pretrained(path)                                  # config.json + model.safetensors
    else:                                                          # the older layout: config.json + pytorch_model.bin
        os.makedirs(path, exist_ok=True)
        cfg.save_pretrained(path)
        torch.save(ref.state_dict(), os.path.join(path, "pytorch_model.bin"))
    return ref


def test_resnet50_pretrained_true_loads_a_torchvision_checkpoint_file(tmp_path, monkeypatch):
    import b200net
    ref = write_resnet_checkpoint(str(tmp_path))
    monkeypatch.setenv("TORCH_HOME", str(tmp_path))
    model = b200net.models.resnet50(pretrained=True)            # the reference's call
    sd, sr = model.state_dict(), ref.state_dict()
    assert list(sd.keys()) == list(sr.keys())
    for k in sr:
        assert sd[k].dtype == sr[k].dtype and torch.equal(sd[k], sr[k]), k
    # and back: a checkpoint written by our module loads into torchvision
    import torchvision
    t2 = torchvision.models.resnet50(weights=None)
    t2.load_state_dict(model.state_dict())


def test_resnet50_pretrained_true_without_a_file_is_an_error(tmp_path, monkeypatch):
    import b200net
    monkeypatch.setenv("TORCH_HOME", str(tmp_path))
    with pytest.raises(RuntimeError):
        b200net.models.resnet50(pretrained=True)


@pytest.mark.parametrize("safetensors", [True, False])
def test_bert_from_pretrained_reads_huggingface_directories(tmp_path, safetensors):
    from b200net import bert
    d = str(tmp_path / "ckpt")
    ref = write_bert_checkpoint(d, layers=2, safetensors=safetensors)
    assert os.path.isfile(os.path.join(d, "model.safetensors" if safetensors else "pytorch_model.bin"))
    model = bert.BertForSequenceClassification.from_pretrained(d, num_labels=2, output_attentions=False,
                                                               output_hidden_states=False)
    assert model.config.num_hidden_layers == 2                   # read from the checkpoint's config.json
    sd, sr = model.state_dict(), ref.state_dict()
    for k, v in sr.items():
        if k.endswith("position_ids"):
            continue
        assert torch.equal(sd[k], v), k


def test_bert_from_pretrained_rejects_an_empty_directory(tmp_path):
    from b200net import bert
    d = tmp_path / "empty"
    d.mkdir()
    with pytest.raises(OSError):
        bert.BertForSequenceClassification.from_pretrained(str(d), num_labels=2)
