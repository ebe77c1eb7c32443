"""BERT-base sequence classifier with the HuggingFace object protocol the reference drives:

    model = BertForSequenceClassification.from_pretrained("bert-base-uncased", num_labels=2,
                output_attentions=False, output_hidden_states=False)            pytorch_on_language_distr.py:155-161
    outputs = model(b_input_ids, token_type_ids=None, attention_mask=b_input_mask, labels=b_labels)   :258-261
    loss = outputs[0]  /  logits = outputs[0] without labels                                         :265, :318, :364

Module and parameter names equal transformers' (`bert.encoder.layer.3.attention.self.query.weight`, ... 201 tensors,
109 483 778 parameters for num_labels=2), so a HuggingFace state_dict loads with load_state_dict and
`model.parameters()` enumerates in the same order (optimizer state, clip_grad_norm_).

The modules are parameter containers: the forward of the whole encoder (embeddings + 12 layers) is ONE autograd node
(`_BertEncoderFn`) whose forward/backward are hand-scheduled launches of b200net kernels -- tcgen05 implicit-GEMM for
the Linears (QKV fused into one [2304 x 768] operand, bias in the epilogue), fused (dropout+)residual+LayerNorm,
erf-GELU, attention, embedding gather/scatter.  The pooler / classifier / cross-entropy run on the fp32 head kernels.
Nothing goes through ATen's matmul, layer_norm, softmax or embedding.
"""
import math
import os
import weakref
import warnings

import torch
from torch import nn
from torch.autograd.function import once_differentiable

from . import functional as F
from . import ops
from ._C import lib, check


class BertConfig(object):
    """Field names and defaults of transformers.BertConfig (bert-base-uncased)."""

    def __init__(self, vocab_size=30522, hidden_size=768, num_hidden_layers=12, num_attention_heads=12,
                 intermediate_size=3072, hidden_act="gelu", hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1,
                 max_position_embeddings=512, type_vocab_size=2, initializer_range=0.02, layer_norm_eps=1e-12,
                 pad_token_id=0, num_labels=2, classifier_dropout=None, **unused):
        self.vocab_size = vocab_size
        self.hidden_size = hidden_size
        self.num_hidden_layers = num_hidden_layers
        self.num_attention_heads = num_attention_heads
        self.intermediate_size = intermediate_size
        self.hidden_act = hidden_act
        self.hidden_dropout_prob = hidden_dropout_prob
        self.attention_probs_dropout_prob = attention_probs_dropout_prob
        self.max_position_embeddings = max_position_embeddings
        self.type_vocab_size = type_vocab_size
        self.initializer_range = initializer_range
        self.layer_norm_eps = layer_norm_eps
        self.pad_token_id = pad_token_id
        self.num_labels = num_labels
        self.classifier_dropout = classifier_dropout
        self.output_attentions = bool(unused.get("output_attentions", False))
        self.output_hidden_states = bool(unused.get("output_hidden_states", False))

    @classmethod
    def from_any(cls, cfg):
        """Accepts our BertConfig or a transformers.BertConfig (duck-typed)."""
        if isinstance(cls, type) and isinstance(cfg, cls):
            return cfg
        keys = ("vocab_size hidden_size num_hidden_layers num_attention_heads intermediate_size hidden_act "
                "hidden_dropout_prob attention_probs_dropout_prob max_position_embeddings type_vocab_size "
                "initializer_range layer_norm_eps pad_token_id num_labels classifier_dropout").split()
        return cls(**{k: getattr(cfg, k) for k in keys if hasattr(cfg, k)})


# ------------------------------------------------------------------ module tree (names = transformers')
class BertEmbeddings(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.word_embeddings = nn.Embedding(c.vocab_size, c.hidden_size, padding_idx=c.pad_token_id)
        self.position_embeddings = nn.Embedding(c.max_position_embeddings, c.hidden_size)
        self.token_type_embeddings = nn.Embedding(c.type_vocab_size, c.hidden_size)
        self.LayerNorm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)
        self.dropout = nn.Dropout(c.hidden_dropout_prob)


class BertSelfAttention(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.num_attention_heads = c.num_attention_heads
        self.attention_head_size = c.hidden_size // c.num_attention_heads
        self.query = nn.Linear(c.hidden_size, c.hidden_size)
        self.key = nn.Linear(c.hidden_size, c.hidden_size)
        self.value = nn.Linear(c.hidden_size, c.hidden_size)
        self.dropout = nn.Dropout(c.attention_probs_dropout_prob)


class BertSelfOutput(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.dense = nn.Linear(c.hidden_size, c.hidden_size)
        self.LayerNorm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)
        self.dropout = nn.Dropout(c.hidden_dropout_prob)


class BertAttention(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.self = BertSelfAttention(c)
        self.output = BertSelfOutput(c)


class BertIntermediate(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.dense = nn.Linear(c.hidden_size, c.intermediate_size)


class BertOutput(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.dense = nn.Linear(c.intermediate_size, c.hidden_size)
        self.LayerNorm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)
        self.dropout = nn.Dropout(c.hidden_dropout_prob)


class BertLayer(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.attention = BertAttention(c)
        self.intermediate = BertIntermediate(c)
        self.output = BertOutput(c)


class BertEncoder(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.layer = nn.ModuleList([BertLayer(c) for _ in range(c.num_hidden_layers)])


class BertPooler(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.dense = nn.Linear(c.hidden_size, c.hidden_size)
        self.activation = nn.Tanh()


class BertModel(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.config = c
        self.embeddings = BertEmbeddings(c)
        self.encoder = BertEncoder(c)
        self.pooler = BertPooler(c)


class SequenceClassifierOutput(object):
    """Indexable like transformers' ModelOutput: fields that are None are skipped, so outputs[0] is the loss when
    labels were given and the logits otherwise (pytorch_on_language_distr.py:265, :318)."""

    def __init__(self, loss=None, logits=None):
        self.loss = loss
        self.logits = logits

    def to_tuple(self):
        return tuple(v for v in (self.loss, self.logits) if v is not None)

    def __getitem__(self, i):
        if isinstance(i, str):
            return getattr(self, i)
        return self.to_tuple()[i]

    def __iter__(self):
        return iter(self.to_tuple())

    def __len__(self):
        return len(self.to_tuple())


# ------------------------------------------------------------------ operand cache (dtype casts / QKV concatenation)
_PK = {}
_CALLS = [0]
_STATUS = {}
_CHECK_IDS = os.environ.get("B200_CHECK_IDS") is not None
_M64 = 0xFFFFFFFFFFFFFFFF
_FLASH = os.environ.get("B200_NO_FLASH_ATTENTION") is None
_CAPTURING = [False]
_CAPTURE_PK = {}


def _cast_into(src, dst):
    check(lib.b200_cast(src.data_ptr(), dst.data_ptr(), src.numel(), ops._dt(src), ops._dt(dst),
                        ops._stream()), "cast")


def _operand(ws, dtype):
    """One contiguous [sum(out_i)][in] (or [sum(out_i)]) tensor in `dtype` from fp32 parameters; rebuilt when a
    parameter's storage or version counter changes (the optimizers bump the version after their in-place update)."""
    if len(ws) == 1 and dtype == torch.float32:
        return ws[0].detach()
    key = tuple(id(w) for w in ws) + (dtype,)
    if _CAPTURING[0]:
        # inside a graph capture the cast is recorded (the weights change between replays) into a buffer owned by
        # the graph, once per operand: the backward graph of the step reads the forward graph's copies
        out = _CAPTURE_PK.get(key)
        if out is None:
            rows = sum(w.shape[0] for w in ws)
            out = torch.empty((rows,) + tuple(ws[0].shape[1:]), dtype=dtype, device=ws[0].device)
            r = 0
            for w in ws:
                _cast_into(w.detach(), out[r:r + w.shape[0]])
                r += w.shape[0]
            _CAPTURE_PK[key] = out
        return out
    sig = tuple((w.data_ptr(), w._version) for w in ws)
    e = _PK.get(key)
    # id() values are recycled across models: the entry also remembers which parameter objects it was built from
    if e is not None and e[0] == sig and all(r() is w for r, w in zip(e[2], ws)):
        return e[1]
    rows = sum(w.shape[0] for w in ws)
    out = torch.empty((rows,) + tuple(ws[0].shape[1:]), dtype=dtype, device=ws[0].device)
    r = 0
    for w in ws:
        _cast_into(w.detach(), out[r:r + w.shape[0]])
        r += w.shape[0]
    _PK[key] = (sig, out, tuple(weakref.ref(w, lambda _r, k=key: _PK.pop(k, None)) for w in ws))
    return out


def _prepack_operands(model, dtype):
    """Every GEMM operand of the encoder that is stale (first call, or the optimizer stepped) is rebuilt by ONE
    multi-tensor launch per 48 source tensors (b200_pack_weight_multi: a Linear weight [out][in] is a 1x1 convolution
    weight) instead of one cast kernel each -- 72 + 36 launches per training step before."""
    groups = []
    for L in model.bert.encoder.layer:
        a = L.attention
        groups.append(((a.self.query.weight, a.self.key.weight, a.self.value.weight), dtype))
        groups.append(((a.self.query.bias, a.self.key.bias, a.self.value.bias), torch.float32))
        if dtype != torch.float32:
            groups += [((a.output.dense.weight,), dtype), ((L.intermediate.dense.weight,), dtype),
                       ((L.output.dense.weight,), dtype)]
    todo = {}            # output dtype -> [(source, destination rows)]
    for ws, dt in groups:
        key = tuple(id(w) for w in ws) + (dt,)
        if _CAPTURING[0]:
            if key in _CAPTURE_PK:
                continue
        else:
            e = _PK.get(key)
            sig = tuple((w.data_ptr(), w._version) for w in ws)
            if e is not None and e[0] == sig and all(r() is w for r, w in zip(e[2], ws)):
                continue
        if any(w.dtype != torch.float32 or not w.is_contiguous() or w.shape[-1] % 8 != 0 for w in ws):
            continue                                   # _operand's own path handles it
        rows = sum(w.shape[0] for w in ws)
        out = torch.empty((rows,) + tuple(ws[0].shape[1:]), dtype=dt, device=ws[0].device)
        r = 0
        for w in ws:
            todo.setdefault(dt, []).append((w.detach(), out[r:r + w.shape[0]]))
            r += w.shape[0]
        if _CAPTURING[0]:
            _CAPTURE_PK[key] = out
        else:
            _PK[key] = (tuple((w.data_ptr(), w._version) for w in ws), out,
                        tuple(weakref.ref(w, lambda _r, k=key: _PK.pop(k, None)) for w in ws))
    for dt, pairs in todo.items():
        ops.cast_rows_multi(pairs, dt)


def _status(device):
    t = _STATUS.get(device)
    if t is None:
        t = torch.zeros(1, dtype=torch.int32, device=device)
        _STATUS[device] = t
    return t


def _encoder_params(model):
    e = model.bert.embeddings
    ps = [e.word_embeddings.weight, e.position_embeddings.weight, e.token_type_embeddings.weight, e.LayerNorm.weight,
          e.LayerNorm.bias]
    for L in model.bert.encoder.layer:
        a, o = L.attention, L.output
        ps += [a.self.query.weight, a.self.query.bias, a.self.key.weight, a.self.key.bias, a.self.value.weight,
               a.self.value.bias, a.output.dense.weight, a.output.dense.bias, a.output.LayerNorm.weight,
               a.output.LayerNorm.bias, L.intermediate.dense.weight, L.intermediate.dense.bias, o.dense.weight,
               o.dense.bias, o.LayerNorm.weight, o.LayerNorm.bias]
    return ps


class _EncoderState(object):
    pass


def _site(call, site):
    # disjoint Philox counter ranges: 2^34 groups of four per dropout site, 64 sites per forward call
    return ((call * 64 + site) << 34) & _M64


def _encoder_fwd(model, ids, type_ids, mask, save, training, call=None):
    """call=None: eager (the call index is folded into the Philox offsets passed by value); call=0 under graph
    capture (a device-resident counter registered with b200_dropout_counter supplies the call index at replay)."""
    c = model.config
    dt = F.act_dtype()
    N, S = ids.shape
    H, heads = c.hidden_size, c.num_attention_heads
    p_h = float(c.hidden_dropout_prob) if training else 0.0
    p_a = float(c.attention_probs_dropout_prob) if training else 0.0
    seed = torch.cuda.initial_seed() & _M64
    if call is None:
        call = _CALLS[0]
        _CALLS[0] += 1
    e = model.bert.embeddings
    status = _status(ids.device)
    s0, h, m0, r0 = ops.embed_ln_fwd(ids, type_ids, e.word_embeddings.weight.detach(),
                                     e.position_embeddings.weight.detach(), e.token_type_embeddings.weight.detach(),
                                     e.LayerNorm.weight.detach(), e.LayerNorm.bias.detach(), e.LayerNorm.eps, p_h, seed,
                                     _site(call, 0), dt, status)
    if _CHECK_IDS and not _CAPTURING[0] and int(status.item()) != 0:
        status.zero_()
        raise IndexError("b200net BertEmbeddings: input id out of range")
    st = _EncoderState()
    st.layers = []
    # bf16 at S in {128, 256}: fused tcgen05 attention (saves the log-sum-exp); otherwise the fp32-arithmetic kernel
    # (saves the probabilities)
    flash = _FLASH and ops.attention_flash_ok(dt, S)
    _prepack_operands(model, dt)
    for li, L in enumerate(model.bert.encoder.layer):
        a, o = L.attention, L.output
        wqkv = _operand((a.self.query.weight, a.self.key.weight, a.self.value.weight), dt)
        bqkv = _operand((a.self.query.bias, a.self.key.bias, a.self.value.bias), torch.float32)
        qkv = ops.dense_fwd(h, wqkv, bqkv)
        if flash:
            probs, ctx = ops.attention_flash_fwd(qkv, mask, N, S, heads, p_a, seed, _site(call, 1 + 3 * li))
        else:
            probs, ctx = ops.attention_fwd(qkv, mask, N, S, heads, p_a, seed, _site(call, 1 + 3 * li))
        ao = ops.dense_fwd(ctx, _operand((a.output.dense.weight,), dt), a.output.dense.bias.detach())
        ln1 = a.output.LayerNorm
        s1, h1, m1, r1 = ops.ln_fwd(ao, h, ln1.weight.detach(), ln1.bias.detach(), ln1.eps, p_h, seed,
                                    _site(call, 2 + 3 * li))
        pre = ops.dense_fwd(h1, _operand((L.intermediate.dense.weight,), dt), L.intermediate.dense.bias.detach())
        act = ops.gelu_fwd(pre)
        f2 = ops.dense_fwd(act, _operand((o.dense.weight,), dt), o.dense.bias.detach())
        ln2 = o.LayerNorm
        s2, h2, m2, r2 = ops.ln_fwd(f2, h1, ln2.weight.detach(), ln2.bias.detach(), ln2.eps, p_h, seed,
                                    _site(call, 3 + 3 * li))
        if save:
            st.layers.append((h, qkv, probs, ctx, s1, m1, r1, h1, pre, act, s2, m2, r2))
        h = h2
    cls = ops.gather_rows(h, N, S)
    if save:
        st.emb = (s0, m0, r0)
        st.ids, st.type_ids = ids, type_ids
        st.N, st.S, st.dtype = N, S, dt
        st.p_h, st.p_a, st.seed, st.call = p_h, p_a, seed, call
        st.flash, st.mask = flash, mask
    return cls, st


def _encoder_bwd(model, st, dcls, emit, release=True, cut=None, cut_every=0):
    """Reverse walk over the saved activations; emit(param, grad) is called as soon as a gradient is enqueued.
    release=False keeps the saved activations referenced (a captured graph replays over them)."""
    c = model.config
    N, S, dt = st.N, st.S, st.dtype
    H, heads, FF = c.hidden_size, c.num_attention_heads, c.intermediate_size
    p_h, p_a, seed, call = st.p_h, st.p_a, st.seed, st.call
    dh_a = ops.scatter_rows(dcls.contiguous().float(), N * S, S, dt)
    dh_b = None
    layers = model.bert.encoder.layer
    for li in range(len(layers) - 1, -1, -1):
        if cut is not None and cut_every and li != len(layers) - 1 and (len(layers) - 1 - li) % cut_every == 0:
            cut()         # graph capture under DDP: a new segment starts here
        L = layers[li]
        a, o = L.attention, L.output
        h_in, qkv, probs, ctx, s1, m1, r1, h1, pre, act, s2, m2, r2 = st.layers[li]
        if release:
            st.layers[li] = None
        # BertOutput: LayerNorm(dropout(dense(act)) + h1)
        dres2, df2, dg, db, dbias = ops.ln_bwd(dh_a, dh_b, s2, m2, r2, o.LayerNorm.weight.detach(), p_h, seed,
                                               _site(call, 3 + 3 * li))
        emit(o.LayerNorm.weight, dg)
        emit(o.LayerNorm.bias, db)
        emit(o.dense.bias, dbias)
        emit(o.dense.weight, ops.dense_bwd_weight(df2, act))
        dact = ops.dense_bwd_input(df2, _operand((o.dense.weight,), dt), FF)
        dpre, db1 = ops.gelu_bwd_bias(dact, pre, out=dact)
        emit(L.intermediate.dense.bias, db1)
        emit(L.intermediate.dense.weight, ops.dense_bwd_weight(dpre, h1))
        dh1_b = ops.dense_bwd_input(dpre, _operand((L.intermediate.dense.weight,), dt), H)
        # BertSelfOutput: LayerNorm(dropout(dense(ctx)) + h_in)
        dres1, dao, dg, db, dbias = ops.ln_bwd(dres2, dh1_b, s1, m1, r1, a.output.LayerNorm.weight.detach(), p_h,
                                               seed, _site(call, 2 + 3 * li))
        emit(a.output.LayerNorm.weight, dg)
        emit(a.output.LayerNorm.bias, db)
        emit(a.output.dense.bias, dbias)
        emit(a.output.dense.weight, ops.dense_bwd_weight(dao, ctx))
        dctx = ops.dense_bwd_input(dao, _operand((a.output.dense.weight,), dt), H)
        if st.flash:
            dqkv = ops.attention_flash_bwd(qkv, st.mask, probs, ctx, dctx, N, S, heads, p_a)
        else:
            dqkv = ops.attention_bwd(qkv, probs, dctx, N, S, heads, p_a, seed, _site(call, 1 + 3 * li))
        dbqkv = ops.colsum_any(dqkv)
        dwqkv = ops.dense_bwd_weight(dqkv, h_in)
        for j, lin in enumerate((a.self.query, a.self.key, a.self.value)):
            emit(lin.weight, dwqkv[j * H:(j + 1) * H])
            emit(lin.bias, dbqkv[j * H:(j + 1) * H])
        wqkv = _operand((a.self.query.weight, a.self.key.weight, a.self.value.weight), dt)
        dh_b = ops.dense_bwd_input(dqkv, wqkv, H)
        dh_a = dres1
    e = model.bert.embeddings
    s0, m0, r0 = st.emb
    ds0, _, dg, db, _ = ops.ln_bwd(dh_a, dh_b, s0, m0, r0, e.LayerNorm.weight.detach(), p_h, seed, _site(call, 0),
                                   drop_after=True, want_dbias=False)
    emit(e.LayerNorm.weight, dg)
    emit(e.LayerNorm.bias, db)
    dword, dpos, dtyp = ops.embed_bwd(st.ids, st.type_ids, ds0, c.vocab_size, c.max_position_embeddings)
    emit(e.word_embeddings.weight, dword)
    emit(e.position_embeddings.weight, dpos)
    emit(e.token_type_embeddings.weight, dtyp)


# ------------------------------------------------------------------ CUDA-graph replay of the encoder
# One training step is ~520 kernel launches (~12 ms of Python / ctypes enqueue work against ~12 ms of GPU work).
# After GRAPH_WARMUP eager calls with the same (model, batch shape, mode, precision, trainable set) the encoder's
# forward and backward are captured once and replayed; the reference's loop body
# (pytorch_on_language_distr.py:257-275) is unchanged.  Dropout stays fresh per step through the device-resident
# call counter (b200_dropout_counter).  B200_CUDA_GRAPHS=0 disables it.  Under DistributedDataParallel the copies
# into the gradient buckets are part of the backward graph and the all-reduces are launched after the replay.
GRAPH_WARMUP = 3
GRAPH_CACHE = 4
_BG = {}
_BG_SEEN = {}
F._GRAPH_CACHES.extend([_BG, _BG_SEEN])


class _BertGraph(object):
    __slots__ = ("fwd", "bwd", "ids", "tt", "mask", "cls", "dcls", "grads", "ctr", "state", "ptrs", "keep", "uses",
                 "ddp", "order", "segments", "gen", "live")


def _graph_key(model, ids, tt, mask, params):
    needs = tuple(p.requires_grad for p in params) if torch.is_grad_enabled() else ()
    # forward-only graphs read operands cast outside the graph: valid for the captured weight versions only
    wver = -1 if any(needs) else sum(p._version for p in params)
    return (id(model), tuple(ids.shape), ids.device.index, F.get_precision(), model.training, tt is not None,
            mask is not None, needs, wver, id(F.grad_sink_owner()))


def _param_ptrs(params):
    return tuple(p.data_ptr() for p in params)


def _capture(model, ids, tt, mask, params, key):
    needs = key[7]
    save = any(needs)
    rec = _BertGraph()
    rec.ids = ids.clone()
    rec.tt = tt.clone() if tt is not None else None
    rec.mask = mask.clone() if mask is not None else None
    rec.ctr = torch.zeros(1, dtype=torch.int64, device=ids.device)
    rec.ptrs = _param_ptrs(params)
    rec.uses = 0
    rec.gen, rec.live = 0, None
    pool = torch.cuda.graph_pool_handle()
    _CAPTURE_PK.clear()
    _CAPTURING[0] = save          # forward-only graphs use the eager operand cache (validity is part of the key)
    ops.NO_WORKSPACE_GROWTH[0] = True
    check(lib.b200_dropout_counter(rec.ctr.data_ptr()), "dropout_counter")
    try:
        torch.cuda.synchronize()
        rec.fwd = torch.cuda.CUDAGraph()
        with torch.no_grad():
            with torch.cuda.graph(rec.fwd, pool=pool):
                rec.cls, rec.state = _encoder_fwd(model, rec.ids, rec.tt, rec.mask, save, model.training, call=0)
        rec.bwd, rec.grads, rec.dcls, rec.ddp, rec.order, rec.segments = None, None, None, None, None, None
        if save:
            rec.dcls = torch.zeros_like(rec.cls)
            grads = {}
            rec.ddp, rec.order = F.grad_sink_owner(), []
            pending = []

            def emit(p, g):
                if p.requires_grad:
                    grads[id(p)] = g if rec.ddp is None else rec.ddp._sink_capture(p, g, rec.order, pending)

            if rec.ddp is None:
                rec.bwd = torch.cuda.CUDAGraph()
                with torch.no_grad():
                    with torch.cuda.graph(rec.bwd, pool=pool):
                        _encoder_bwd(model, rec.state, rec.dcls, emit, release=False)
            else:
                # DDP: one graph per three encoder layers; the buckets a segment completed are all-reduced while the
                # next segment runs
                def cut():
                    rec.ddp.flush_pending(pending)      # this segment's gradients -> bucket slices, one launch per 48
                    sc.cut(lambda: len(rec.order))

                with torch.no_grad():
                    with F._SegmentedCapture(pool) as sc:
                        _encoder_bwd(model, rec.state, rec.dcls, emit, release=False, cut=cut, cut_every=3)
                        rec.ddp.flush_pending(pending)
                bounds = [0] + list(sc.marks) + [len(rec.order)]
                rec.segments = [(g, rec.order[bounds[i]:bounds[i + 1]]) for i, g in enumerate(sc.graphs)]
                rec.bwd = sc.graphs[0]
            rec.grads = [grads.get(id(p)) for p in params]
        rec.keep = (dict(ops._WS), list(_CAPTURE_PK.values()), [e[1] for e in _PK.values()])
    finally:
        check(lib.b200_dropout_counter(None), "dropout_counter")
        _CAPTURING[0] = False
        ops.NO_WORKSPACE_GROWTH[0] = False
        _CAPTURE_PK.clear()
    return rec


class _GraphedEncoder(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rec, ids, tt, mask, *params):
        rec.ids.copy_(ids)
        if tt is not None:
            rec.tt.copy_(tt)
        if mask is not None:
            rec.mask.copy_(mask)
        rec.ctr.fill_(_CALLS[0])
        _CALLS[0] += 1
        rec.fwd.replay()
        F.graph_replay_begin(rec, ctx)
        ctx.rec = rec
        ctx.saved_params = params
        return rec.cls.clone()

    @staticmethod
    @once_differentiable
    def backward(ctx, dcls):
        rec = ctx.rec
        if rec.bwd is None:
            raise RuntimeError("b200net: this encoder graph was captured without a backward pass")
        F.graph_replay_check(rec, ctx)
        rec.dcls.copy_(dcls)
        if rec.ddp is not None:
            for g, ready in rec.segments:
                g.replay()
                rec.ddp._replayed(ready)      # completed buckets start their all-reduce while the next segment runs
            return (None, None, None, None) + tuple(rec.grads)
        rec.bwd.replay()
        # gradients live in graph-owned memory that the next replay overwrites; with zero_grad -> backward -> step
        # they are consumed before that.  A parameter that still holds a .grad gets a private copy.
        out = [None if g is None else (g if p.grad is None else g.clone())
               for p, g in zip(ctx.saved_params, rec.grads)]
        return (None, None, None, None) + tuple(out)


def bert_encoder(model, ids, tt, mask):
    """[N,S] ids -> final hidden state of the first token [N][hidden] fp32 (embeddings + every encoder layer)."""
    params = _encoder_params(model)
    owner = F.grad_sink_owner()
    if not (F._GRAPHS_ON and not torch.cuda.is_current_stream_capturing()
            and (F._GRAD_SINK is None or (owner is not None and owner.graph_ok()))):
        return _BertEncoderFn.apply(model, torch.is_grad_enabled(), ids, tt, mask, *params)
    key = _graph_key(model, ids, tt, mask, params)
    rec = _BG.get(key)
    if rec is not None and rec != "failed":
        if rec.ptrs == _param_ptrs(params):
            if F.graph_record_busy(rec):   # an earlier replayed forward still waits for its backward
                return _BertEncoderFn.apply(model, torch.is_grad_enabled(), ids, tt, mask, *params)
            rec.uses += 1
            return _GraphedEncoder.apply(rec, ids, tt, mask, *params)
        del _BG[key]
        _BG_SEEN[key] = 0
        rec = None
    if rec is None:
        seen = _BG_SEEN.get(key, 0)
        if seen >= GRAPH_WARMUP:
            try:
                if len(_BG) >= GRAPH_CACHE:
                    victim = min((k for k in _BG if _BG[k] != "failed"), key=lambda k: _BG[k].uses, default=None)
                    if victim is not None:
                        del _BG[victim]
                rec = _capture(model, ids, tt, mask, params, key)
                _BG[key] = rec
                rec.uses = 1
                return _GraphedEncoder.apply(rec, ids, tt, mask, *params)
            except Exception as e:                    # capture is an optimisation: never fatal
                if "workspace would have to grow" in str(e):
                    _BG_SEEN[key] = 0                 # the eager calls so far never ran the backward: try again later
                else:
                    warnings.warn("b200net: CUDA-graph capture of the BERT encoder failed (%s); staying on the "
                                  "eager path" % e)
                    _BG[key] = "failed"
        else:
            if len(_BG_SEEN) > 64:
                _BG_SEEN.clear()
            _BG_SEEN[key] = seen + 1
    return _BertEncoderFn.apply(model, torch.is_grad_enabled(), ids, tt, mask, *params)


class _BertEncoderFn(torch.autograd.Function):
    """Embeddings + all encoder layers as one autograd node; returns the final hidden state of the first token of
    every sequence ([N][hidden] fp32), which is all the pooler reads."""

    @staticmethod
    def forward(ctx, model, grad_on, ids, type_ids, mask, *params):
        save = grad_on and any(ctx.needs_input_grad[5:])
        cls, st = _encoder_fwd(model, ids, type_ids, mask, save, model.training)
        if save:
            ctx.model = model
            ctx.state = st
            ctx.params = params
        return cls

    @staticmethod
    @once_differentiable
    def backward(ctx, dcls):
        grads = {}
        sink = F._GRAD_SINK

        def emit(p, g):
            if p.requires_grad:
                grads[id(p)] = sink(p, g) if sink is not None else g

        _encoder_bwd(ctx.model, ctx.state, dcls, emit)
        out = tuple(grads.get(id(p)) for p in ctx.params)
        ctx.state = None
        ctx.model = None
        ctx.params = None
        return (None, None, None, None, None) + out


class BertForSequenceClassification(nn.Module):
    """transformers.BertForSequenceClassification as the reference uses it (pytorch_on_language_distr.py:155-161)."""

    def __init__(self, config=None, **kw):
        super().__init__()
        c = BertConfig(**kw) if config is None else BertConfig.from_any(config)
        if c.hidden_act != "gelu":
            raise NotImplementedError("b200net BERT: hidden_act must be 'gelu' (erf), got %r" % (c.hidden_act,))
        if c.hidden_size // c.num_attention_heads != 64:
            raise NotImplementedError("b200net BERT: attention head size must be 64")
        if c.type_vocab_size != 2:
            raise NotImplementedError("b200net BERT: type_vocab_size must be 2")
        self.config = c
        self.num_labels = c.num_labels
        self.bert = BertModel(c)
        p = c.classifier_dropout if c.classifier_dropout is not None else c.hidden_dropout_prob
        self.dropout = nn.Dropout(p)
        self.classifier = nn.Linear(c.hidden_size, c.num_labels)
        self.apply(self._init_weights)

    def _init_weights(self, m):
        # transformers' BertPreTrainedModel._init_weights
        std = self.config.initializer_range
        if isinstance(m, nn.Linear):
            m.weight.data.normal_(mean=0.0, std=std)
            if m.bias is not None:
                m.bias.data.zero_()
        elif isinstance(m, nn.Embedding):
            m.weight.data.normal_(mean=0.0, std=std)
            if m.padding_idx is not None:
                m.weight.data[m.padding_idx].zero_()
        elif isinstance(m, nn.LayerNorm):
            m.bias.data.zero_()
            m.weight.data.fill_(1.0)

    @classmethod
    def from_pretrained(cls, name_or_path, num_labels=2, **kw):
        """A local directory holding `pytorch_model.bin` or `model.safetensors` (the two files Hugging Face
        checkpoints ship) is loaded; a directory that exists but holds neither is an error.  A hub name cannot be
        fetched on an offline box, so it yields the architecture with transformers' random initialisation (and says
        so)."""
        path = str(name_or_path)
        cfg_file = os.path.join(path, "config.json")
        cfg_kw = {}
        if os.path.isfile(cfg_file):              # the checkpoint's own architecture (layers, widths, vocabulary)
            import json
            known = ("vocab_size hidden_size num_hidden_layers num_attention_heads intermediate_size hidden_act "
                     "hidden_dropout_prob attention_probs_dropout_prob max_position_embeddings type_vocab_size "
                     "initializer_range layer_norm_eps pad_token_id classifier_dropout").split()
            with open(cfg_file) as fh:
                raw = json.load(fh)
            cfg_kw = {k: raw[k] for k in known if k in raw}
        cfg_kw.update(kw)
        model = cls(BertConfig(num_labels=num_labels, **cfg_kw))
        ckpt, st_file = os.path.join(path, "pytorch_model.bin"), os.path.join(path, "model.safetensors")
        sd = None
        if os.path.isfile(ckpt):
            sd = torch.load(ckpt, map_location="cpu")
        elif os.path.isfile(st_file):
            from safetensors.torch import load_file
            sd = load_file(st_file, device="cpu")
        elif os.path.isdir(path):
            raise OSError("b200net from_pretrained: %r holds neither pytorch_model.bin nor model.safetensors" % path)
        if sd is not None:
            # checkpoints of the bare encoder (BertModel / BertForPreTraining) store keys without / with other prefixes
            if not any(k.startswith("bert.") for k in sd):
                sd = {"bert." + k: v for k, v in sd.items()}
            sd = {k.replace("LayerNorm.gamma", "LayerNorm.weight").replace("LayerNorm.beta", "LayerNorm.bias"): v
                  for k, v in sd.items()}
            missing, unexpected = model.load_state_dict(sd, strict=False)
            head = [k for k in missing if not k.startswith("classifier.")]
            if head:
                raise RuntimeError("b200net from_pretrained: checkpoint lacks %s" % head[:5])
        else:
            warnings.warn("b200net: %r is not a local checkpoint directory and there is no network access; "
                          "using randomly initialised bert-base weights" % (name_or_path,))
        return model

    def forward(self, input_ids=None, attention_mask=None, token_type_ids=None, position_ids=None, head_mask=None,
                inputs_embeds=None, labels=None, output_attentions=None, output_hidden_states=None,
                return_dict=None):
        if input_ids is None or inputs_embeds is not None or position_ids is not None or head_mask is not None:
            raise NotImplementedError("b200net BERT: call as model(input_ids, token_type_ids=..., "
                                      "attention_mask=..., labels=...)")
        if output_attentions or output_hidden_states:
            raise NotImplementedError("b200net BERT: output_attentions / output_hidden_states are not produced")
        if not input_ids.is_cuda:
            raise RuntimeError("b200net kernels run on CUDA tensors only (no CPU fallback)")
        ids = input_ids.long().contiguous()
        if ids.dim() != 2:
            raise ValueError("input_ids must be [batch, seq_len]")
        if ids.shape[1] % 4 != 0:
            raise NotImplementedError("b200net BERT: sequence length must be a multiple of 4 (pad the batch)")
        tt = token_type_ids.long().contiguous() if token_type_ids is not None else None
        mask = attention_mask.long().contiguous() if attention_mask is not None else None
        cls_rows = bert_encoder(self, ids, tt, mask)
        pool = self.bert.pooler.dense
        pooled = F.tanh(F.linear(cls_rows, pool.weight, pool.bias))
        pooled = F.dropout(pooled, self.dropout.p, self.training)
        logits = F.linear(pooled, self.classifier.weight, self.classifier.bias)
        loss = None
        if labels is not None:
            loss = F.nll_loss(F.log_softmax(logits.view(-1, self.num_labels)), labels.view(-1))
        if return_dict is False:
            return SequenceClassifierOutput(loss, logits).to_tuple()
        return SequenceClassifierOutput(loss, logits)


# ------------------------------------------------------------------ the optimizer objects of the reference's script
from . import optim as _optim  # noqa: E402


class AdamW(_optim.AdamW):
    """`transformers.AdamW(model.parameters(), lr=2e-5, eps=1e-8)` (pytorch_on_language_distr.py:167-170): the
    transformers signature and defaults (eps=1e-6, weight_decay=0.0, correct_bias=True) over the multi-tensor
    Adam kernel, with the PUBLISHED transformers update rule: p -= lr*sqrt(1-b2^t)/(1-b1^t) * m / (sqrt(v) + eps)
    (eps is added before the bias correction of v; torch.optim.AdamW adds it after, an effective eps that is
    sqrt(1-b2^t) ~ 0.03x smaller at step 1).  One deviation remains and only matters when weight_decay != 0 (the
    reference uses the default 0.0): transformers applies p -= lr*wd*p AFTER the Adam update, this kernel -- like
    torch -- before it; the two differ by O(lr^2 * wd)."""
    _eps_before_bias_correction = True

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0, correct_bias=True):
        if not correct_bias:
            raise NotImplementedError("b200net AdamW: correct_bias=False is not implemented")
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)


def get_linear_schedule_with_warmup(optimizer, num_warmup_steps, num_training_steps, last_epoch=-1):
    """transformers.get_linear_schedule_with_warmup (:181-183): host-side LambdaLR driving param_group['lr']."""
    def lr_lambda(step):
        if step < num_warmup_steps:
            return float(step) / float(max(1, num_warmup_steps))
        return max(0.0, float(num_training_steps - step) / float(max(1, num_training_steps - num_warmup_steps)))
    return torch.optim.lr_scheduler.LambdaLR(optimizer, lr_lambda, last_epoch)
