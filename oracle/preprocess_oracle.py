"""ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by the product package).

CPU restatement (numpy, integer arithmetic) of the image transform the reference applies to every picture,

    transforms.Compose([transforms.Resize((224, 224)), transforms.ToTensor()])
    (pytorch_training_inference_on_image.ipynb:325-326, :847-848; another_neural_net.py:38-46),

whose arithmetic lives in Pillow (Image.resize, BILINEAR with antialiasing; src/libImaging/Resample.c of the Pillow
release torchvision calls -- not under /root/reference, not pinned by it; Pillow 12.2 in this image): per axis a
triangle filter widened by the down-scale factor, weights normalised in double precision and quantised to 22-bit
fixed point, products accumulated in an integer that starts at one half, shifted down and clipped to a byte --
horizontal pass, byte image, vertical pass -- and then ToTensor()'s byte / 255.

Pinned: tests/test_preprocess_cpu.py checks this restatement BYTE FOR BYTE against PIL itself (and against the
torchvision Compose above) on seeded images of several sizes, up- and down-scaling."""
import numpy as np

PRECISION_BITS = 22


def axis_coefficients(in_size, out_size):
    """[(first source index, [integer weights])] for every output index (Resample.c: precompute_coeffs +
    normalize_coeffs_8bpc, bilinear filter, support 1)."""
    scale = in_size / out_size
    fscale = max(scale, 1.0)
    support = 1.0 * fscale
    out = []
    for o in range(out_size):
        center = (o + 0.5) * scale
        ss = 1.0 / fscale
        lo = max(int(center - support + 0.5), 0)
        hi = min(int(center + support + 0.5), in_size)
        w = []
        for j in range(lo, hi):
            x = abs((j - center + 0.5) * ss)
            w.append(1.0 - x if x < 1.0 else 0.0)
        ww = sum(w)
        out.append((lo, [int(0.5 + (v / ww) * (1 << PRECISION_BITS)) for v in w]))
    return out


def resize_u8(img, out_h, out_w):
    """uint8 [H,W,3] -> uint8 [out_h,out_w,3], Pillow's two-pass 8-bit resampling."""
    img = np.asarray(img)
    H, W, _ = img.shape
    kx, ky = axis_coefficients(W, out_w), axis_coefficients(H, out_h)
    half = 1 << (PRECISION_BITS - 1)
    tmp = np.zeros((H, out_w, 3), np.int64)
    for ox, (lo, k) in enumerate(kx):
        acc = np.full((H, 3), half, np.int64)
        for j, kk in enumerate(k):
            acc += img[:, lo + j, :].astype(np.int64) * kk
        tmp[:, ox, :] = np.clip(acc >> PRECISION_BITS, 0, 255)
    out = np.zeros((out_h, out_w, 3), np.int64)
    for oy, (lo, k) in enumerate(ky):
        acc = np.full((out_w, 3), half, np.int64)
        for j, kk in enumerate(k):
            acc += tmp[lo + j, :, :] * kk
        out[oy] = np.clip(acc >> PRECISION_BITS, 0, 255)
    return out.astype(np.uint8)


def resize_to_tensor(img, out_h, out_w):
    """The whole transform: float32 [3,out_h,out_w] in [0,1]."""
    return resize_u8(img, out_h, out_w).transpose(2, 0, 1).astype(np.float32) / np.float32(255.0)
