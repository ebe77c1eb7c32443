"""Gradient utilities with torch.nn.utils signatures."""
import torch

from ._C import lib, check, ptr_array, i64_array


def clip_grad_norm_(parameters, max_norm, norm_type=2.0):
    """`torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)` (pytorch_on_language_distr.py:273): global L2 norm
    over all gradients (deterministic two-stage reduction), then an in-place scale by min(1, max_norm/(norm+1e-6)).
    The scale is read on the device: no host synchronisation.  Returns the total norm as a 0-d tensor."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    if float(norm_type) != 2.0:
        raise NotImplementedError("b200net clip_grad_norm_: L2 norm only")
    grads = [p.grad for p in parameters if p.grad is not None]
    if not grads:
        return torch.tensor(0.0)
    for g in grads:
        if not g.is_cuda or g.dtype != torch.float32 or not g.is_contiguous():
            raise RuntimeError("b200net clip_grad_norm_: gradients must be contiguous float32 CUDA tensors")
    dev = grads[0].device
    st = torch.cuda.current_stream().cuda_stream
    nbytes = int(lib.b200_l2norm_workspace_bytes())
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    norm_sq = torch.empty(1, dtype=torch.float32, device=dev)
    ptrs = ptr_array([g.data_ptr() for g in grads])
    sizes = i64_array([g.numel() for g in grads])
    check(lib.b200_multi_tensor_l2norm_sq(len(grads), ptrs, sizes, norm_sq.data_ptr(), ws.data_ptr(), nbytes, st),
          "l2norm")
    check(lib.b200_multi_tensor_clip_scale(len(grads), ptrs, sizes, norm_sq.data_ptr(), float(max_norm), st),
          "clip_scale")
    return norm_sq.sqrt().reshape(())


class DevicePrefetcher(object):
    """Wraps any iterable of (inputs, labels) host batches -- the reference's `trainloader`
    (pytorch_training_inference_on_image.ipynb:664, another_neural_net.py:123) -- and hands out the same batches
    already on the device: batch i+1 is copied host->device on a side stream while step i computes, so the loop's
    own `inputs.to(device), labels.to(device)` become no-ops and the copy no longer sits in front of every step
    (78 MB per 256-image batch: ~1.5 ms of a 14 ms step).  Pinned host tensors make the copies truly asynchronous;
    pageable ones still work.

        trainloader = b200net.utils.DevicePrefetcher(trainloader, device)
        for inputs, labels in trainloader:          # unchanged loop
            inputs, labels = inputs.to(device), labels.to(device)
    """

    def __init__(self, loader, device):
        self.loader = loader
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)

    def __len__(self):
        return len(self.loader)

    def _to_device(self, item):
        if isinstance(item, torch.Tensor):
            return item.to(self.device, non_blocking=True)
        if isinstance(item, (list, tuple)):
            return type(item)(self._to_device(t) for t in item)
        return item

    def __iter__(self):
        it = iter(self.loader)
        cur = torch.cuda.current_stream(self.device)

        def fetch():
            try:
                batch = next(it)
            except StopIteration:
                return None
            with torch.cuda.stream(self.stream):
                return self._to_device(batch)

        nxt = fetch()
        while nxt is not None:
            cur.wait_stream(self.stream)
            batch = nxt
            for t in (batch if isinstance(batch, (list, tuple)) else [batch]):
                if isinstance(t, torch.Tensor):
                    t.record_stream(cur)
            nxt = fetch()
            yield batch


class GpuPreprocess(object):
    """The reference's image transform on the GPU -- `transforms.Compose([transforms.Resize(size),
    transforms.ToTensor()])` (pytorch_training_inference_on_image.ipynb:325-326, :847-848) -- for a batch of decoded
    uint8 HWC images (numpy arrays, PIL images or uint8 tensors; sizes may differ within a batch).

        prep = b200net.utils.GpuPreprocess((160, 160), device)
        inputs = prep(images)                    # -> model-ready batch; model(inputs) as usual
        inputs = prep(images, layout="nchw")     # -> [N,3,H,W] fp32, exactly what ToTensor() would have produced

    The default output is the stem's own input layout ([N,H,W,4] bf16, 8-byte pixels), which `model.forward` accepts
    directly in bf16 mode: the host -> device copy moves the 3-byte source pixels instead of 12-byte resized floats
    and the model's first layout kernel disappears.  JPEG decoding stays with PIL on the host."""

    def __init__(self, size, device="cuda", antialias=True):
        self.size = (size, size) if isinstance(size, int) else tuple(size)
        self.device = torch.device(device)
        self.antialias = bool(antialias)

    def __call__(self, images, layout="packed"):
        import numpy as np
        from ._C import lib, check
        ts = []
        for im in images:
            if not isinstance(im, torch.Tensor):
                im = torch.from_numpy(np.ascontiguousarray(np.asarray(im)))
            if im.dtype != torch.uint8 or im.dim() != 3 or im.shape[2] != 3:
                raise ValueError("GpuPreprocess: images must be uint8 [H, W, 3] (RGB)")
            ts.append(im.contiguous().to(self.device, non_blocking=True))
        n = len(ts)
        ptrs = torch.tensor([t.data_ptr() for t in ts], dtype=torch.int64).to(self.device, non_blocking=True)
        hs = torch.tensor([t.shape[0] for t in ts], dtype=torch.int32).to(self.device, non_blocking=True)
        ws = torch.tensor([t.shape[1] for t in ts], dtype=torch.int32).to(self.device, non_blocking=True)
        rs = torch.tensor([t.shape[1] * 3 for t in ts], dtype=torch.int64).to(self.device, non_blocking=True)
        ho, wo = self.size
        if layout == "packed":
            out = torch.empty((n, ho, wo, 4), dtype=torch.bfloat16, device=self.device)
            code = 0
        elif layout == "nchw":
            out = torch.empty((n, 3, ho, wo), dtype=torch.float32, device=self.device)
            code = 1
        else:
            raise ValueError("layout must be 'packed' or 'nchw'")
        with torch.cuda.device(self.device):
            check(lib.b200_preprocess_u8(ptrs.data_ptr(), hs.data_ptr(), ws.data_ptr(), rs.data_ptr(), n, ho, wo,
                                         out.data_ptr(), code, 1 if self.antialias else 0,
                                         torch.cuda.current_stream().cuda_stream), "preprocess_u8")
        for t in ts:                       # the kernel reads them on the current stream
            t.record_stream(torch.cuda.current_stream())
        return out
