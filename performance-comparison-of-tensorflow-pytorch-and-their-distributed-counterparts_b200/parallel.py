"""DistributedDataParallel with the stock constructor signature -- the call the reference leaves commented out at
pytorch_on_language_distr.py:220-221:

    model = DistributedDataParallel(model, device_ids=[args.local_rank], output_device=args.local_rank)

Semantics follow torch.nn.parallel.DistributedDataParallel: rank-0 parameters and buffers are broadcast at
construction, buffers (BatchNorm running statistics) are re-broadcast before every training forward
(`broadcast_buffers=True`), gradients of parameters that require grad are averaged across ranks in buckets laid out
in REVERSE registration order (1 MiB first bucket, then `bucket_cap_mb`), BatchNorm batch statistics stay per rank.

CUDA graphs: the fused backbone / encoder nodes still replay their captured graphs under this wrapper.  The copies
into the bucket slices are part of the backward graph, which is captured in a few segments (ResNet: layer4 | layer3 |
rest; BERT: three encoder layers each); after each segment's replay the buckets it completed start their all-reduce
on the communication stream while the next segment runs, in the same fixed bucket order.

B200 design: gradients LIVE in the flat fp32 buckets (`param.grad` is a view of its bucket slice, so there is no
unpack and the optimizer reads reduced values in place); a bucket is all-reduced (ncclAvg over NVLink/NVSwitch) on a
dedicated communication stream as soon as its last gradient is written, overlapping the rest of backward.  The fused
backbone node (functional._ResNetBackbone) reports each weight gradient the moment its wgrad kernel is enqueued, so
the overlap is per layer even though the backbone is one autograd node.  The order in which buckets are launched is
fixed by construction (bucket i is only launched after buckets 0..i-1), so every rank issues the same collective
sequence.  On CPU tensors (gloo; used by the tests) the same logic runs with synchronous collectives.
"""
import torch
import torch.distributed as dist
import torch.nn as _nn

from . import functional as F
from ._C import lib, check, ptr_array, i64_array


class _Bucket(object):
    __slots__ = ("index", "params", "offsets", "flat", "pending", "launched", "event", "work")

    def __init__(self, index):
        self.index = index
        self.params = []
        self.offsets = []
        self.flat = None
        self.pending = 0
        self.launched = False
        self.event = None
        self.work = None


class DistributedDataParallel(_nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, dim=0, broadcast_buffers=True, process_group=None,
                 bucket_cap_mb=25, find_unused_parameters=False, gradient_as_bucket_view=True):
        super().__init__()
        if not dist.is_available() or not dist.is_initialized():
            raise RuntimeError("DistributedDataParallel: call torch.distributed.init_process_group first "
                               "(pytorch_on_language_distr.py:26)")
        self.module = module
        self.device_ids = device_ids
        self.output_device = output_device
        self.broadcast_buffers = broadcast_buffers
        self.find_unused_parameters = bool(find_unused_parameters)
        self.process_group = process_group if process_group is not None else dist.group.WORLD
        self.world_size = dist.get_world_size(self.process_group)
        self.require_backward_grad_sync = True
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any parameter that "
                               "requires a gradient.")
        self.device = params[0].device
        self.on_cuda = self.device.type == "cuda"
        self._comm_stream = torch.cuda.Stream(device=self.device, priority=-1) if self.on_cuda else None
        self._verify_shapes(list(module.parameters()))
        self._sync_module_states()
        self._flatten_buffers()
        self._build_buckets(params, int(bucket_cap_mb * 1024 * 1024))
        self._callback_queued = False
        self._next_to_launch = 0
        for p in params:
            p.register_post_accumulate_grad_hook(self._post_acc_hook)

    # ---------------------------------------------------------------- construction helpers
    def _verify_shapes(self, params):
        sizes = torch.tensor([p.numel() for p in params], dtype=torch.int64, device=self.device)
        ref = sizes.clone()
        dist.broadcast(ref, 0, group=self.process_group)
        if not torch.equal(ref, sizes):
            raise RuntimeError("DistributedDataParallel: parameter shapes differ across ranks")

    def _coalesced_broadcast(self, tensors):
        """rank-0 values -> every rank, one collective per dtype."""
        by_dtype = {}
        for t in tensors:
            by_dtype.setdefault(t.dtype, []).append(t)
        for dtype, ts in by_dtype.items():
            if dtype == torch.float32 and self.on_cuda and all(t.is_contiguous() for t in ts):
                sizes = i64_array([t.numel() for t in ts])
                ptrs = ptr_array([t.data_ptr() for t in ts])
                flat = torch.empty(sum(t.numel() for t in ts), dtype=dtype, device=self.device)
                st = torch.cuda.current_stream().cuda_stream
                check(lib.b200_bucket_pack(len(ts), ptrs, sizes, flat.data_ptr(), 1.0, st), "bucket_pack")
                dist.broadcast(flat, 0, group=self.process_group)
                check(lib.b200_bucket_unpack(len(ts), ptrs, sizes, flat.data_ptr(), 1.0, st), "bucket_unpack")
            else:
                flat = torch.cat([t.detach().reshape(-1) for t in ts])
                dist.broadcast(flat, 0, group=self.process_group)
                off = 0
                with torch.no_grad():
                    for t in ts:
                        t.copy_(flat[off:off + t.numel()].view_as(t))
                        off += t.numel()

    def _flatten_buffers(self):
        """Re-homes every fp32 buffer (BatchNorm running statistics) into ONE flat tensor, so that the buffer broadcast
        before every training forward (torch DDP's `broadcast_buffers=True`) is a single collective on that tensor --
        no pack / unpack kernels, no allocation, no Python walk over 106 tensors.  The module's buffers stay the same
        tensor objects (state_dict, load_state_dict and in-place updates are unaffected); if the user re-allocates
        them (`module.to(...)`), the aliasing check in forward() falls back to the coalesced broadcast.
        (Issuing the broadcast right AFTER the forward, on the communication stream, would hide its ~40 us entirely,
        but then a buffer modified between two forwards -- load_state_dict on a resumed run -- would not be replaced
        by rank 0's value, which is what torch does; tests/test_ddp_gpu.py checks exactly that.)"""
        self._buf_flat = None
        if not (self.on_cuda and self.broadcast_buffers and self.world_size > 1):
            return
        bufs = [b for b in self.module.buffers() if b.dtype == torch.float32 and b.is_contiguous() and b.numel() > 0]
        if not bufs:
            return
        offs, total = [], 0
        for b in bufs:
            offs.append(total)
            total += (b.numel() + 3) // 4 * 4            # 16-byte aligned slices
        flat = torch.zeros(total, dtype=torch.float32, device=self.device)
        with torch.no_grad():
            for b, off in zip(bufs, offs):
                view = flat[off:off + b.numel()].view(b.shape)
                view.copy_(b)
                b.data = view
        self._buf_flat, self._bufs = flat, bufs
        self._buf_ptrs = [b.data_ptr() for b in bufs]

    def _buffers_still_flat(self):
        return all(b.data_ptr() == p for b, p in zip(self._bufs, self._buf_ptrs))

    def _sync_module_states(self):
        with torch.no_grad():
            self._coalesced_broadcast([p.data for p in self.module.parameters()] + list(self.module.buffers()))
        for p in self.module.parameters():
            torch.autograd.graph.increment_version(p)

    def _build_buckets(self, params, cap_bytes):
        """Reverse registration order ~ order in which backward produces gradients; small first bucket so
        communication starts early (same policy as torch's Reducer: 1 MiB, then bucket_cap_mb)."""
        self.buckets = []
        self._where = {}
        cur, cur_bytes, limit = _Bucket(0), 0, 1024 * 1024
        for p in reversed(params):
            nbytes = p.numel() * 4
            if cur.params and cur_bytes + nbytes > limit:
                self.buckets.append(cur)
                cur, cur_bytes, limit = _Bucket(len(self.buckets)), 0, cap_bytes
            cur.offsets.append(cur_bytes // 4)
            cur.params.append(p)
            cur_bytes += nbytes
        self.buckets.append(cur)
        for b in self.buckets:
            total = b.offsets[-1] + b.params[-1].numel()
            b.flat = torch.zeros(total, dtype=torch.float32, device=self.device)
            for p, off in zip(b.params, b.offsets):
                self._where[id(p)] = (b, off)
            if self.on_cuda:
                b.event = torch.cuda.Event()
        self._reset()

    def bucket_layout(self):
        """[(bucket index, [parameter numel, ...])] -- used by tests and by the scaling report."""
        return [(b.index, [p.numel() for p in b.params]) for b in self.buckets]

    def _reset(self):
        for b in self.buckets:
            b.pending = len(b.params)
            b.launched = False
            b.work = None
        self._next_to_launch = 0
        self._callback_queued = False
        self._ready = set()

    # ---------------------------------------------------------------- gradient path
    def grad_view(self, p):
        """The bucket slice that holds p's gradient (shape of p)."""
        b, off = self._where[id(p)]
        return b.flat[off:off + p.numel()].view_as(p)

    def _mark_ready(self, p):
        if id(p) in self._ready:
            return
        self._ready.add(id(p))
        if not self._callback_queued:
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize)
        b, _ = self._where[id(p)]
        b.pending -= 1
        # launch in bucket order only: identical collective sequence on every rank
        while self._next_to_launch < len(self.buckets) and self.buckets[self._next_to_launch].pending == 0:
            self._launch(self.buckets[self._next_to_launch])
            self._next_to_launch += 1

    def _launch(self, b):
        b.launched = True
        if not self.require_backward_grad_sync:
            return
        if self.on_cuda:
            b.event.record(torch.cuda.current_stream())          # gradients of this bucket are enqueued up to here
            self._comm_stream.wait_event(b.event)
            with torch.cuda.stream(self._comm_stream):
                try:
                    dist.all_reduce(b.flat, op=dist.ReduceOp.AVG, group=self.process_group)
                except Exception:
                    dist.all_reduce(b.flat, op=dist.ReduceOp.SUM, group=self.process_group)
                    b.flat.div_(self.world_size)
                b.event.record(self._comm_stream)
        else:
            dist.all_reduce(b.flat, op=dist.ReduceOp.SUM, group=self.process_group)
            b.flat.div_(self.world_size)

    def _sink(self, p, grad):
        """Called by the fused backbone as soon as a gradient tensor exists: move it into its bucket slice and hand
        back the slice, which becomes `p.grad`."""
        if id(p) not in self._where or p.grad is not None:
            return grad          # accumulation into an existing .grad goes through the post-accumulate hook
        view = self.grad_view(p)
        if grad.data_ptr() != view.data_ptr():
            if self.on_cuda and grad.is_contiguous() and grad.dtype == torch.float32:
                check(lib.b200_bucket_pack(1, ptr_array([grad.data_ptr()]), i64_array([grad.numel()]),
                                           view.data_ptr(), 1.0, torch.cuda.current_stream().cuda_stream), "bucket_pack")
            else:
                view.copy_(grad)
        self._mark_ready(p)
        return view

    # ---- CUDA-graph mode of the fused nodes (functional.resnet_backbone, bert.bert_encoder): the pack into the
    # bucket slice is recorded inside the backward graph; after every replay the node reports the parameters in the
    # recorded order, which launches the buckets' all-reduces in the usual fixed order on the communication stream.
    def _sink_capture(self, p, grad, order, pending):
        """Graph capture: remember (bucket slice, gradient); flush_pending() records one multi-tensor copy for all
        of them (a copy kernel per parameter costs ~2.5 us each inside a graph: 0.4 ms for ResNet-50's 161)."""
        if id(p) not in self._where:
            return grad
        view = self.grad_view(p)
        if grad.data_ptr() != view.data_ptr():
            pending.append((view, grad if grad.is_contiguous() else grad.contiguous()))
        order.append(p)
        return view

    def flush_pending(self, pending):
        if pending:
            check(lib.b200_multi_tensor_copy(len(pending), ptr_array([v.data_ptr() for v, _ in pending]),
                                             ptr_array([g.data_ptr() for _, g in pending]),
                                             i64_array([g.numel() for _, g in pending]), 1.0,
                                             torch.cuda.current_stream().cuda_stream), "multi_tensor_copy")
            del pending[:]

    def _replayed(self, order):
        for p in order:
            self._mark_ready(p)

    def graph_ok(self):
        """Fused nodes may replay graphs under this wrapper: CUDA, gradient sync on, no stale .grad tensors (an
        accumulating step has to take the eager path, whose sink leaves accumulation to autograd)."""
        return (self.on_cuda and self.require_backward_grad_sync
                and all(q.grad is None for b in self.buckets for q in b.params))

    def _post_acc_hook(self, p):
        view = self.grad_view(p)
        if id(p) in self._ready:
            # the sink already moved this gradient into the bucket (and the bucket may be in flight): autograd may
            # have cloned the tensor it was handed, so re-alias .grad to the slice holding the reduced values
            if p.grad.data_ptr() != view.data_ptr():
                p.grad = view
            return
        g = p.grad
        if g.data_ptr() != view.data_ptr():
            with torch.no_grad():
                if self.on_cuda and g.is_contiguous() and g.dtype == torch.float32:
                    check(lib.b200_bucket_pack(1, ptr_array([g.data_ptr()]), i64_array([g.numel()]), view.data_ptr(),
                                               1.0, torch.cuda.current_stream().cuda_stream), "bucket_pack")
                else:
                    view.copy_(g)
            p.grad = view
        self._mark_ready(p)

    def _finalize(self):
        # Parameters that produced no gradient in this backward.  Under no_sync() nothing is reduced and their
        # bucket slices may hold gradients accumulated by earlier micro-steps (p.grad is a view of the slice): leave
        # them alone.  With synchronisation on, torch's DDP raises unless find_unused_parameters=True; with it, the
        # unused parameters take part in the collective as zeros -- unless they still carry an accumulated gradient.
        unused = [(b, p, off) for b in self.buckets if not b.launched
                  for p, off in zip(b.params, b.offsets) if id(p) not in self._ready]
        if unused and self.require_backward_grad_sync:
            if not self.find_unused_parameters:
                self._reset()
                raise RuntimeError(
                    "DistributedDataParallel: %d parameter(s) that require grad received no gradient in this backward "
                    "pass. Pass find_unused_parameters=True to the constructor if parts of the module are "
                    "legitimately unused (same rule as torch.nn.parallel.DistributedDataParallel)." % len(unused))
            for b, p, off in unused:
                if p.grad is None:
                    b.flat[off:off + p.numel()].zero_()
        while self._next_to_launch < len(self.buckets):
            self._launch(self.buckets[self._next_to_launch])
            self._next_to_launch += 1
        if self.on_cuda and self.require_backward_grad_sync:
            cur = torch.cuda.current_stream()
            for b in self.buckets:
                cur.wait_event(b.event)
        self._reset()

    # ---------------------------------------------------------------- module protocol
    def forward(self, *inputs, **kwargs):
        if self.broadcast_buffers and self.module.training and self.world_size > 1:
            if self._buf_flat is not None and self._buffers_still_flat():
                dist.broadcast(self._buf_flat, 0, group=self.process_group)      # rank 0's statistics -> every rank
            else:
                bufs = [b for b in self.module.buffers() if b.dtype.is_floating_point]
                if bufs:
                    with torch.no_grad():
                        self._coalesced_broadcast(bufs)
        if torch.is_grad_enabled() and self.require_backward_grad_sync:
            F.set_grad_sink(self._sink)
        else:
            F.set_grad_sink(None)
        return self.module(*inputs, **kwargs)

    class _NoSync(object):
        def __init__(self, ddp):
            self.ddp = ddp

        def __enter__(self):
            self.prev = self.ddp.require_backward_grad_sync
            self.ddp.require_backward_grad_sync = False

        def __exit__(self, *a):
            self.ddp.require_backward_grad_sync = self.prev

    def no_sync(self):
        return DistributedDataParallel._NoSync(self)
