"""The scripts under scripts/ -- the reference's loops verbatim with the model import swapped -- run end to end on one
GPU, launched the way the reference launches its own (python -m torch.distributed.launch --nproc_per_node=N script,
another_neural_net.py:392-393), on small synthetic data."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(cmd, timeout=600):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", PYTHONPATH=ROOT)
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


def test_image_script_under_torch_distributed_launch():
    out = _run([sys.executable, "-m", "torch.distributed.launch", "--nproc_per_node=1", "--master_addr=127.0.0.1",
                "--master_port=29611", "scripts/image_train_infer.py", "--synthetic", "--size", "160", "--batch", "32",
                "--epochs", "1", "--infer-images", "16", "--ddp", "--all-layers"])
    assert "image_train_infer: OK" in out and "Test accuracy" in out


def test_image_script_plain_frozen_backbone():
    out = _run([sys.executable, "scripts/image_train_infer.py", "--synthetic", "--size", "160", "--batch", "32",
                "--epochs", "1", "--infer-images", "16"])
    assert "image_train_infer: OK" in out


def test_bert_script_under_torch_distributed_launch():
    out = _run([sys.executable, "-m", "torch.distributed.launch", "--nproc_per_node=1", "--master_addr=127.0.0.1",
                "--master_port=29612", "scripts/bert_imdb_distr.py", "--layers", "2", "--reviews", "160"])
    assert "bert_imdb_distr: OK" in out and "Accuracy" in out
