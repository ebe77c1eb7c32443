"""GPU input pipeline against the reference's own transform run on the CPU:
transforms.Compose([transforms.Resize(size), transforms.ToTensor()]) on PIL images
(pytorch_training_inference_on_image.ipynb:325-326) and against the numpy oracle.  Integer work: BIT-EXACT (the kernel
restates Pillow's 22-bit fixed-point resampling)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _images(sizes, seed=0):
    rng = np.random.RandomState(seed)
    out = []
    for h, w in sizes:
        # smooth structure + noise, so that antialiasing matters and values span the whole byte range
        yy, xx = np.mgrid[0:h, 0:w]
        base = 127 + 100 * np.sin(yy / 7.0)[..., None] * np.cos(xx / 11.0)[..., None] * np.array([1.0, 0.7, -0.8])
        img = np.clip(base + rng.randint(-40, 40, size=(h, w, 3)), 0, 255).astype(np.uint8)
        out.append(img)
    return out


@pytest.mark.parametrize("size", [(160, 160), (224, 224)])
def test_resize_totensor_matches_the_reference_transform(size):
    import b200net
    from PIL import Image
    from torchvision import transforms
    sizes = [(375, 500), (160, 160), (120, 90), (333, 777), (500, 375), (224, 224)]
    imgs = _images(sizes)
    ref_tf = transforms.Compose([transforms.Resize(size), transforms.ToTensor()])        # the reference's lines
    want = torch.stack([ref_tf(Image.fromarray(im)) for im in imgs])
    prep = b200net.utils.GpuPreprocess(size, DEV)
    got = prep(imgs, layout="nchw")
    assert got.shape == want.shape and got.dtype == torch.float32
    from oracle import preprocess_oracle as P
    d = (got.cpu() - want).abs()
    print("resize to %s: max |diff| %.3e (%.2f levels), mismatching values %d of %d" % (size, d.max().item(), d.max().item() * 255, int((d > 0).sum()), d.numel()))
    assert torch.equal(got.cpu(), want)                                           # the reference transform, bit for bit
    assert torch.equal(got.cpu()[2], torch.from_numpy(P.resize_to_tensor(imgs[2], *size)))   # and the numpy oracle
    packed = prep(imgs, layout="packed")
    assert packed.shape == (len(imgs), size[0], size[1], 4) and packed.dtype == torch.bfloat16
    assert torch.equal(packed[..., :3], got.permute(0, 2, 3, 1).to(torch.bfloat16))
    assert float(packed[..., 3].abs().max()) == 0.0


def test_model_accepts_the_packed_batch():
    import b200net
    b200net.set_precision("bf16")
    torch.manual_seed(0)
    model = b200net.models.resnet50(pretrained=False).to(DEV).eval()
    imgs = _images([(200, 260)] * 4, seed=3)
    prep = b200net.utils.GpuPreprocess((160, 160), DEV)
    with torch.no_grad():
        a = model(prep(imgs, layout="nchw"))
        b = model(prep(imgs))                    # the stem's own layout: no layout kernel in front of the model
    assert torch.equal(a, b)
