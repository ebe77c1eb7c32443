"""The preprocessing oracle (oracle/preprocess_oracle.py, a numpy restatement of Pillow's 8-bit bilinear resampling +
ToTensor) is pinned against the reference's own transform executed here: PIL / torchvision on CPU, byte for byte."""
import numpy as np
import pytest
import torch

from oracle import preprocess_oracle as P


@pytest.mark.parametrize("h,w", [(375, 500), (120, 90), (333, 777), (160, 160), (64, 48)])
@pytest.mark.parametrize("size", [(160, 160), (224, 224)])
def test_oracle_matches_pil_and_the_reference_compose(h, w, size):
    from PIL import Image
    from torchvision import transforms
    img = np.random.RandomState(h * 1000 + w).randint(0, 256, (h, w, 3)).astype(np.uint8)
    want_u8 = np.asarray(Image.fromarray(img).resize((size[1], size[0]), Image.BILINEAR))
    assert np.array_equal(P.resize_u8(img, *size), want_u8)
    ref_tf = transforms.Compose([transforms.Resize(size), transforms.ToTensor()])     # ...on_image.ipynb:325-326
    want = ref_tf(Image.fromarray(img))
    assert torch.equal(torch.from_numpy(P.resize_to_tensor(img, *size)), want)
