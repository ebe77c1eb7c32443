"""BERT / IMDB path on the GPU, through the C-ABI.

Per-kernel parity against float64 torch on the same seeded inputs; whole-model parity against the oracle
(transformers' BertForSequenceClassification on CPU executing the reference's statements,
pytorch_on_language_distr.py:257-275) and the committed golden vectors tests/golden/bert_small.pt.
Tolerances: fp32 path logits/loss <= 1e-4 relative, after-step weights <= 1e-3 (north star); bf16 path 2e-2 of the
output scale per kernel."""
import math
import os

import pytest
import torch
import torch.nn.functional as TF

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DT = {"fp32": torch.float32, "bf16": torch.bfloat16}
TOL = {"fp32": 2e-5, "bf16": 2e-2}


def _dev():
    return torch.device("cuda:0")


def _rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def _rand(shape, dtype, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype).to(_dev())


# ---------------------------------------------------------------------------------------------- LayerNorm
@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("rows,H", [(37, 768), (1024, 768), (5, 256), (130, 1024)])
def test_layernorm_fwd_bwd(prec, rows, H):
    from b200net import ops
    dt = DT[prec]
    x = _rand((rows, H), dt, 1)
    res = _rand((rows, H), dt, 2)
    gamma = _rand((H,), torch.float32, 3) * 0.5 + 1.0
    beta = _rand((H,), torch.float32, 4)
    dy_a = _rand((rows, H), dt, 5)
    dy_b = _rand((rows, H), dt, 6)
    s, y, mean, rstd = ops.ln_fwd(x.clone(), res, gamma, beta, 1e-12)
    sd = (x.double() + res.double()).to(dt).double().requires_grad_(True)     # the kernel rounds s to the storage dtype
    gd, bd = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
    yd = TF.layer_norm(sd, (H,), gd, bd, 1e-12)
    assert _rel(s, sd) < 1e-6
    assert _rel(y, yd) < TOL[prec]
    assert _rel(mean, sd.mean(1)) < 1e-4 + TOL[prec] * 0.01
    dyd = dy_a.double() + dy_b.double()
    yd.backward(dyd)
    dres, dx, dg, db, dbias = ops.ln_bwd(dy_a, dy_b, s, mean, rstd, gamma)
    assert dx is dres
    assert _rel(dres, sd.grad) < TOL[prec]
    assert _rel(dg, gd.grad) < TOL[prec]
    assert _rel(db, bd.grad) < TOL[prec]
    assert _rel(dbias, dres.double().sum(0)) < 1e-5


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_layernorm_dropout_mask_replay(prec):
    """y = LN(dropout(x) + res): the kept fraction is 1-p, kept elements are scaled by 1/(1-p), and the backward
    regenerates the very same mask (dx is zero exactly where the forward dropped)."""
    from b200net import ops
    dt = DT[prec]
    rows, H, p = 512, 768, 0.1
    x = _rand((rows, H), dt, 1).abs() + 1.0     # keep x away from zero so dropped elements are identifiable
    zero = torch.zeros_like(x)
    gamma = torch.ones(H, device=_dev())
    beta = torch.zeros(H, device=_dev())
    s, y, mean, rstd = ops.ln_fwd(x.clone(), zero, gamma, beta, 1e-12, p=p, seed=1234, offset=77 << 34,
                                  inplace_sum=False)
    kept = s != 0
    frac = kept.float().mean().item()
    assert abs(frac - (1 - p)) < 0.01
    assert _rel(s[kept], (x.double() / (1 - p))[kept]) < TOL[prec]
    s2, _, _, _ = ops.ln_fwd(x.clone(), zero, gamma, beta, 1e-12, p=p, seed=1234, offset=78 << 34, inplace_sum=False)
    assert not torch.equal(s2 != 0, kept)       # another site, another mask
    dy = _rand((rows, H), dt, 5)
    dres, dx, dg, db, dbias = ops.ln_bwd(dy, None, s, mean, rstd, gamma, p=p, seed=1234, offset=77 << 34)
    assert torch.equal(dx != 0, kept & (dres != 0))
    assert _rel(dx[kept], (dres.double() / (1 - p))[kept]) < TOL[prec]
    assert _rel(dbias, dx.double().sum(0)) < 1e-5
    # embedding form: dropout AFTER the LayerNorm
    d2, _, _, _, _ = ops.ln_bwd(dy, None, s, mean, rstd, gamma, p=p, seed=1234, offset=77 << 34, drop_after=True,
                                want_dbias=False)
    sd = s.double().requires_grad_(True)
    yd = TF.layer_norm(sd, (H,), gamma.double(), beta.double(), 1e-12)
    yd.backward(dy.double() * kept.double() / (1 - p))
    assert _rel(d2, sd.grad) < TOL[prec]


# ---------------------------------------------------------------------------------------------- embeddings
@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("with_types", [False, True])
def test_embeddings_fwd_bwd(prec, with_types):
    from b200net import ops
    dt = DT[prec]
    N, S, H, V, P = 3, 20, 768, 1000, 64
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(0, V, (N, S), generator=g).to(_dev())
    ids[0, :5] = 7                                            # repeated ids: the scatter must accumulate
    tt = torch.randint(0, 2, (N, S), generator=g).to(_dev()) if with_types else None
    word, pos, typ = _rand((V, H), torch.float32, 1), _rand((P, H), torch.float32, 2), _rand((2, H), torch.float32, 3)
    gamma, beta = _rand((H,), torch.float32, 4) * 0.3 + 1, _rand((H,), torch.float32, 5)
    status = torch.zeros(1, dtype=torch.int32, device=_dev())
    s, y, mean, rstd = ops.embed_ln_fwd(ids, tt, word, pos, typ, gamma, beta, 1e-12, 0.0, 0, 0, dt, status)
    tref = typ[tt] if with_types else typ[0].expand(N, S, H)
    sref = ((word[ids] + tref) + pos[:S].unsqueeze(0)).reshape(N * S, H)
    assert _rel(s, sref) < (1e-6 if prec == "fp32" else 4e-3)
    yref = TF.layer_norm(s.double(), (H,), gamma.double(), beta.double(), 1e-12)
    assert _rel(y, yref) < TOL[prec]
    assert int(status.item()) == 0
    ds = _rand((N * S, H), dt, 6)
    dword, dpos, dtyp = ops.embed_bwd(ids, tt, ds, V, P)
    dsd = ds.double()
    rw = torch.zeros(V, H, dtype=torch.float64, device=_dev()).index_add_(0, ids.reshape(-1), dsd)
    rp = torch.zeros(P, H, dtype=torch.float64, device=_dev())
    rp[:S] = dsd.reshape(N, S, H).sum(0)
    tflat = tt.reshape(-1) if with_types else torch.zeros(N * S, dtype=torch.long, device=_dev())
    rt = torch.zeros(2, H, dtype=torch.float64, device=_dev()).index_add_(0, tflat, dsd)
    assert _rel(dword, rw) < 1e-5
    assert _rel(dpos, rp) < 1e-5
    assert _rel(dtyp, rt) < 1e-5
    bad = ids.clone()
    bad[1, 3] = V + 5
    ops.embed_ln_fwd(bad, tt, word, pos, typ, gamma, beta, 1e-12, 0.0, 0, 0, dt, status)
    assert int(status.item()) == 1


# ---------------------------------------------------------------------------------------------- GELU / colsum / tanh
@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_gelu_colsum_tanh(prec):
    from b200net import ops
    dt = DT[prec]
    x = _rand((200, 3072), dt, 1, 2.0)
    dy = _rand((200, 3072), dt, 2)
    xd = x.double().requires_grad_(True)
    yd = TF.gelu(xd)
    yd.backward(dy.double())
    assert _rel(ops.gelu_fwd(x), yd) < TOL[prec]
    assert _rel(ops.gelu_bwd(dy, x), xd.grad) < TOL[prec]
    dx, db = ops.gelu_bwd_bias(dy, x)
    assert torch.equal(dx, ops.gelu_bwd(dy, x))
    assert _rel(db, dx.double().sum(0)) < 1e-5
    for rows, cols in ((200, 3072), (8192, 768), (3, 2304), (1000, 8)):
        m = _rand((rows, cols), dt, 3)
        assert _rel(ops.colsum_any(m), m.double().sum(0)) < 1e-5
    t = _rand((32, 768), torch.float32, 4)
    td = t.double().requires_grad_(True)
    yd = torch.tanh(td)
    yd.backward(torch.ones_like(yd) * 0.5)
    y = ops.tanh_fwd(t)
    assert _rel(y, yd) < 1e-6
    assert _rel(ops.tanh_bwd(torch.full_like(y, 0.5), y), td.grad) < 1e-5
    h = _rand((4 * 16, 768), dt, 5)
    cls = ops.gather_rows(h, 4, 16)
    assert torch.equal(cls, h[::16].float())
    back = ops.scatter_rows(cls, 64, 16, dt)
    assert torch.equal(back[::16], h[::16]) and back.float().abs().sum().item() == pytest.approx(
        h[::16].float().abs().sum().item())


# ---------------------------------------------------------------------------------------------- attention
def _attention_ref(qkv, mask, N, S, heads, keep=None, p=0.0):
    D = 64
    q, k, v = qkv.double().reshape(N, S, 3, heads, D).permute(2, 0, 3, 1, 4)
    sc = q @ k.transpose(-1, -2) / 8.0
    if mask is not None:
        sc = sc + (1.0 - mask.double())[:, None, None, :] * -1e30
    pr = torch.softmax(sc, -1)
    pd = pr if keep is None else pr * keep.double() / (1 - p)
    ctx = (pd @ v).permute(0, 2, 1, 3).reshape(N * S, heads * D)
    return pr, ctx


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("N,S,heads", [(2, 128, 12), (3, 36, 2), (1, 256, 4), (1, 512, 1)])
def test_attention_fwd_bwd(prec, N, S, heads):
    from b200net import ops
    dt = DT[prec]
    qkv = _rand((N * S, 3 * heads * 64), dt, 1)
    g = torch.Generator().manual_seed(3)
    lens = torch.randint(S // 2, S + 1, (N,), generator=g)
    mask = (torch.arange(S)[None, :] < lens[:, None]).long().to(_dev())
    probs, ctx = ops.attention_fwd(qkv, mask, N, S, heads)
    qd = qkv.double().requires_grad_(True)
    pr, cref = _attention_ref(qd, mask, N, S, heads)
    assert _rel(probs, pr) < TOL[prec]
    assert _rel(ctx, cref) < TOL[prec]
    assert probs[0, 0, 0, lens[0]:].float().abs().sum().item() == 0.0     # padded keys get exactly zero weight
    dctx = _rand((N * S, heads * 64), dt, 2)
    cref.backward(dctx.double())
    dqkv = ops.attention_bwd(qkv, probs, dctx, N, S, heads)
    assert _rel(dqkv, qd.grad) < (1e-4 if prec == "fp32" else 3e-2)
    # no mask
    probs2, ctx2 = ops.attention_fwd(qkv, None, N, S, heads)
    _, cref2 = _attention_ref(qkv, None, N, S, heads)
    assert _rel(ctx2, cref2) < TOL[prec]


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_attention_dropout_replay(prec):
    """Forward with dropout equals the reference with the mask inferred from a p=0 run; the backward uses the
    same mask (checked through autograd on the reference with that mask)."""
    from b200net import ops
    dt = DT[prec]
    N, S, heads, p = 2, 64, 3, 0.1
    qkv = _rand((N * S, 3 * heads * 64), dt, 1)
    # make V the identity-like probe: ctx = P_dropped @ V with V = one-hot rows reveals the dropped pattern
    probe = qkv.clone().reshape(N, S, 3, heads, 64)
    probe[:, :, 2] = 0
    for j in range(64):
        probe[:, j, 2, :, j] = 1.0
    probe = probe.reshape(N * S, -1).contiguous()
    probs, ctx = ops.attention_fwd(probe, None, N, S, heads, p=p, seed=99, offset=5 << 34)
    pd = ctx.reshape(N, S, heads, 64).permute(0, 2, 1, 3)          # = dropped probabilities [N,h,q,k]
    keep = pd != 0
    frac = keep.float().mean().item()
    assert abs(frac - (1 - p)) < 0.02
    assert _rel(pd[keep], (probs.double() / (1 - p))[keep]) < TOL[prec]
    # same site on the real qkv: mask must be the same (it depends on seed/offset/index only)
    probs, ctx = ops.attention_fwd(qkv, None, N, S, heads, p=p, seed=99, offset=5 << 34)
    qd = qkv.double().requires_grad_(True)
    _, cref = _attention_ref(qd, None, N, S, heads, keep=keep, p=p)
    assert _rel(ctx, cref) < TOL[prec]
    dctx = _rand((N * S, heads * 64), dt, 2)
    cref.backward(dctx.double())
    dqkv = ops.attention_bwd(qkv, probs, dctx, N, S, heads, p=p, seed=99, offset=5 << 34)
    assert _rel(dqkv, qd.grad) < (1e-4 if prec == "fp32" else 3e-2)


@pytest.mark.parametrize("N,S,heads", [(2, 128, 12), (3, 256, 4), (1, 256, 1)])
@pytest.mark.parametrize("p", [0.0, 0.1])
def test_attention_flash_matches_reference_path(N, S, heads, p):
    """tcgen05 fused attention (no S x S tensor in HBM) against float64 torch (p = 0) and against the
    fp32-arithmetic kernels on the same (seed, offset): both derive the dropout mask from the same Philox indices."""
    from b200net import ops
    dt = torch.bfloat16
    assert ops.attention_flash_ok(dt, S)
    qkv = _rand((N * S, 3 * heads * 64), dt, 1)
    g = torch.Generator().manual_seed(3)
    lens = torch.randint(S // 2, S + 1, (N,), generator=g)
    mask = (torch.arange(S)[None, :] < lens[:, None]).long().to(_dev())
    dctx = _rand((N * S, heads * 64), dt, 2)
    seed, off = 4242, 9 << 34
    lse, ctx = ops.attention_flash_fwd(qkv, mask, N, S, heads, p, seed, off)
    probs, ctx_ref = ops.attention_fwd(qkv, mask, N, S, heads, p, seed, off)
    assert _rel(ctx, ctx_ref) < 2e-2
    dq = ops.attention_flash_bwd(qkv, mask, lse, ctx, dctx, N, S, heads, p)
    dq_ref = ops.attention_bwd(qkv, probs, dctx, N, S, heads, p, seed, off)
    for j, name in enumerate("qkv"):
        a = dq.reshape(N * S, 3, heads * 64)[:, j]
        b = dq_ref.reshape(N * S, 3, heads * 64)[:, j]
        assert _rel(a, b) < 3e-2, name
    if p == 0.0:
        qd = qkv.double().requires_grad_(True)
        pr, cref = _attention_ref(qd, mask, N, S, heads)
        assert _rel(ctx, cref) < 2e-2
        lse_ref = torch.logsumexp((torch.log(pr.clamp_min(1e-300))), -1)   # == 0: probabilities are normalised
        assert lse_ref.abs().max().item() < 1e-9
        cref.backward(dctx.double())
        assert _rel(dq, qd.grad) < 3e-2
    # padded keys: exactly zero gradient for their K and V rows
    pad = (mask == 0).reshape(-1)
    assert dq.reshape(N * S, 3, -1)[pad][:, 1:].float().abs().sum().item() == 0.0


# ---------------------------------------------------------------------------------------------- dense layers
@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("M,K,N", [(512, 768, 2304), (300, 768, 768), (256, 3072, 768), (128, 768, 3072)])
def test_dense_fwd_bwd(prec, M, K, N):
    from b200net import ops
    dt = DT[prec]
    x = _rand((M, K), dt, 1)
    w = _rand((N, K), torch.float32, 2, 0.05)
    b = _rand((N,), torch.float32, 3)
    wp = w.to(dt)
    y = ops.dense_fwd(x, wp, b)
    ref = x.double() @ wp.double().t() + b.double()
    assert _rel(y, ref) < TOL[prec] * (1 if prec == "bf16" else math.sqrt(K))
    dy = _rand((M, N), dt, 4)
    dx = ops.dense_bwd_input(dy, wp, K)
    assert _rel(dx, dy.double() @ wp.double()) < TOL[prec] * (1 if prec == "bf16" else math.sqrt(N))
    dw = ops.dense_bwd_weight(dy, x)
    assert dw.shape == (N, K) and dw.dtype == torch.float32
    assert _rel(dw, dy.double().t() @ x.double()) < (1e-3 if prec == "bf16" else 2e-5 * math.sqrt(M))


# ---------------------------------------------------------------------------------------------- whole model
@pytest.fixture(scope="module")
def bert_golden():
    return torch.load(os.path.join(ROOT, "tests", "golden", "bert_small.pt"), weights_only=False)


def _our_model(ref_model, dropout=0.0):
    """b200net model holding the oracle model's weights (state_dict keys are identical)."""
    from b200net import bert
    c = ref_model.config
    m = bert.BertForSequenceClassification(bert.BertConfig(
        num_labels=c.num_labels, num_hidden_layers=c.num_hidden_layers, hidden_dropout_prob=dropout,
        attention_probs_dropout_prob=dropout))
    m.load_state_dict({k: v.float() for k, v in ref_model.state_dict().items()})
    return m.to(_dev())


def _our_optimizer(model, total_steps, lr=2e-5, eps=1e-8):
    import b200net
    opt = b200net.optim.AdamW(model.parameters(), lr=lr, eps=eps, weight_decay=0.0)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda t: max(0.0, float(total_steps - t) / float(max(1, total_steps))))
    return opt, sched


def _digest_rel(t, d):
    from oracle.torch_oracle import digest
    mine = digest(t)
    scale = d["samples"].abs().max().clamp_min(1e-4)
    return ((mine["samples"].double() - d["samples"].double()).abs().max() / scale).item()


def test_bert_fp32_matches_golden(bert_golden):
    """fp32 mode against the committed oracle output (2 layers, N=4, S=32, 3 steps of the reference's loop body)."""
    import b200net
    from oracle import bert_oracle as B
    g, g64 = bert_golden["case"], bert_golden["case_f64"]
    b200net.set_precision("fp32")
    try:
        ref = B.build_reference_model(num_layers=bert_golden["layers"])
        model = _our_model(ref)
        # step-0 gradients (after clip_grad_norm_, as the golden run stores them)
        res = B.run_case(model, bert_golden["n"], bert_golden["s"], 1, make_opt=_our_optimizer,
                         clip=b200net.utils.clip_grad_norm_, device=_dev(), keep_grads=True)
        for k, d in g["grads_step0"].items():
            assert _digest_rel(res["grads"][k], d) < 1e-3, k
        model = _our_model(ref)
        res = B.run_case(model, bert_golden["n"], bert_golden["s"], bert_golden["steps"], make_opt=_our_optimizer,
                         clip=b200net.utils.clip_grad_norm_, device=_dev())
        for i in range(bert_golden["steps"]):
            assert abs(res["losses"][i] - g["losses"][i]) <= 1e-4 * abs(g["losses"][i]), (i, res["losses"], g["losses"])
            assert _rel(res["logits"][i], g["logits"][i]) <= 1e-4 + 3 * _rel(g["logits"][i], g64["logits"][i])
        assert _rel(res["eval_logits"], g["eval_logits"]) <= 1e-4 + 3 * _rel(g["eval_logits"], g64["eval_logits"])
        assert torch.equal(res["eval_logits"].argmax(1), g["eval_logits"].argmax(1))
        sd = model.state_dict()
        for k, d in g["weights"].items():
            assert _digest_rel(sd[k], d) < 1e-3, k
    finally:
        b200net.set_precision("bf16")


@pytest.mark.parametrize("S", [128, 256])
def test_bert_fp32_matches_oracle_at_reference_lengths(S):
    """Same statements on the oracle (CPU) and on b200net (GPU) at the reference's MAX_LEN=128
    (pytorch_on_language_distr.py:69) and at C5's S=256; 2 layers keep the CPU oracle to seconds."""
    import b200net
    from oracle import bert_oracle as B
    b200net.set_precision("fp32")
    try:
        ref = B.build_reference_model(num_layers=2)
        model = _our_model(ref)
        r0 = B.run_case(ref, 2, S, 2)
        r1 = B.run_case(model, 2, S, 2, make_opt=_our_optimizer, clip=b200net.utils.clip_grad_norm_, device=_dev())
        for i in range(2):
            assert abs(r0["losses"][i] - r1["losses"][i]) <= 1e-4 * abs(r0["losses"][i])
            assert _rel(r1["logits"][i], r0["logits"][i]) <= 2e-4
        assert _rel(r1["eval_logits"], r0["eval_logits"]) <= 2e-4
        sd0, sd1 = ref.state_dict(), model.state_dict()
        for k in B.WATCH:
            a, b = sd1[k].double().cpu(), sd0[k].double()
            assert ((a - b).abs().max() / b.abs().max().clamp_min(1e-4)).item() < 1e-3, k
    finally:
        b200net.set_precision("bf16")


def test_bert_bf16_tracks_fp32_curve():
    """bf16 tensor-core mode against our fp32 mode: 12 steps of the loop body on one batch, loss within 1e-2."""
    import b200net
    from oracle import bert_oracle as B
    ref = B.build_reference_model(num_layers=2)
    losses = {}
    for prec in ("fp32", "bf16"):
        b200net.set_precision(prec)
        try:
            model = _our_model(ref)
            r = B.run_case(model, 4, 128, 12, make_opt=_our_optimizer, clip=b200net.utils.clip_grad_norm_,
                           device=_dev())
            losses[prec] = r["losses"]
        finally:
            b200net.set_precision("bf16")
    for a, b in zip(losses["bf16"], losses["fp32"]):
        assert abs(a - b) <= 1e-2 * max(1.0, abs(b)), (losses["bf16"], losses["fp32"])


def test_bert_bf16_12_layers_matches_transformers_oracle():
    """bert-base depth (12 layers), N=2, S=128 (the reference's MAX_LEN), bf16 tensor-core mode, compared directly with
    the transformers-on-CPU oracle (NOT with our own fp32 mode): two steps of the reference's loop body.  Stated
    tolerances: loss 1e-2 relative (the north-star's bf16 figure); logits 5e-2 of the logit scale.  The logits of the
    untrained model are tiny (~0.3) and this implementation keeps the residual stream, LayerNorm outputs and attention
    probabilities in bf16 STORAGE (half the HBM traffic of every bandwidth kernel), where stock bf16 autocast keeps
    them in fp32 and only rounds GEMM operands: measured 3.0e-2 here against 6e-3 for the stock autocast witness,
    which is printed next to it."""
    import b200net
    from oracle import bert_oracle as B
    b200net.set_precision("bf16")
    ref = B.build_reference_model(num_layers=12)
    model = _our_model(ref)
    r0 = B.run_case(ref, 2, 128, 2)
    r1 = B.run_case(model, 2, 128, 2, make_opt=_our_optimizer, clip=b200net.utils.clip_grad_norm_, device=_dev())
    wit = B.build_reference_model(num_layers=12).to(_dev()).train()     # stock bf16 witness (step-0 weights)
    ids, mask, labels = B.make_batch(2, 128, B.SEED_DATA, _dev())
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        lw = wit(ids, token_type_ids=None, attention_mask=mask, labels=labels)[1].double().cpu()
    werr = _rel(lw, r0["logits"][0])
    tol = 5e-2
    for i in range(2):
        print("bert-12 bf16 step %d: loss ours %.6f oracle %.6f ; logits rel %.3e (stock bf16 autocast at step 0: "
              "%.3e)" % (i, r1["losses"][i], r0["losses"][i], _rel(r1["logits"][i], r0["logits"][i]), werr))
        assert abs(r0["losses"][i] - r1["losses"][i]) <= 1e-2 * abs(r0["losses"][i]), (r0["losses"], r1["losses"])
        assert _rel(r1["logits"][i], r0["logits"][i]) <= tol
    assert _rel(r1["eval_logits"], r0["eval_logits"]) <= tol


def test_bert_full_size_properties():
    """bert-base (12 layers), N=32, S=256, bf16, dropout 0.1 (config C5's per-GPU step): size-independent checks.
    (a) eval logits do not depend on the ids under the padding mask; (b) the training loss is finite, the step
    changes every parameter tensor, and re-seeding reproduces the same loss bit for bit (Philox replay);
    (c) outputs[0] is the loss with labels and the logits without (pytorch_on_language_distr.py:265, :318)."""
    import b200net
    from b200net import bert
    from oracle import bert_oracle as B
    torch.manual_seed(0)
    model = bert.BertForSequenceClassification.from_pretrained(os.path.join(ROOT, "no-such-dir"), num_labels=2,
                                                               output_attentions=False, output_hidden_states=False)
    model.to(_dev())
    ids, mask, labels = B.make_batch(32, 256, 0, _dev())
    model.eval()
    with torch.no_grad():
        out = model(ids, token_type_ids=None, attention_mask=mask)
        logits = out[0]
        assert logits.shape == (32, 2) and len(out) == 1
        ids2 = torch.where(mask.bool(), ids, torch.full_like(ids, 2047))
        logits2 = model(ids2, token_type_ids=None, attention_mask=mask)[0]
    assert torch.equal(logits, logits2)
    model.train()
    opt, sched = _our_optimizer(model, 30)
    before = [p.detach().clone() for p in model.parameters()]

    def step(seed):
        torch.manual_seed(seed)
        bert._CALLS[0] = 0
        b200net.functional._DROPOUT_CALLS[0] = 0      # the classifier dropout's Philox offset
        model.zero_grad()
        out = model(ids, token_type_ids=None, attention_mask=mask, labels=labels)
        assert len(out) == 2 and out[0].dim() == 0
        return out

    l1 = step(5)[0].item()
    l2 = step(5)[0].item()
    l3 = step(6)[0].item()
    assert math.isfinite(l1) and l1 == l2 and l1 != l3
    out = step(5)
    out[0].backward()
    gn = b200net.utils.clip_grad_norm_(model.parameters(), 1.0)
    assert math.isfinite(float(gn))
    opt.step()
    sched.step()
    changed = sum(int(not torch.equal(a, p.detach())) for a, p in zip(before, model.parameters()))
    # position rows >= S and unused vocabulary rows get a zero gradient but Adam still leaves them alone; every tensor moves
    assert changed == len(before)


def test_bert_graph_replay_matches_eager():
    """After GRAPH_WARMUP eager calls the encoder forward/backward replay as CUDA graphs inside the unchanged loop
    body; with dropout on, the device-resident call counter must give the replayed steps the same masks the eager
    path draws (same seed, same call index), so the loss curves coincide (embedding-gradient atomics aside)."""
    import b200net
    from b200net import bert
    from oracle import bert_oracle as B
    ref = B.build_reference_model(num_layers=2)
    curves = {}
    for graphs in (False, True):
        b200net.set_cuda_graphs(False)          # drop any cached graphs
        b200net.set_cuda_graphs(graphs)
        try:
            torch.manual_seed(11)
            bert._CALLS[0] = 0
            b200net.functional._DROPOUT_CALLS[0] = 0
            model = _our_model(ref, dropout=0.1)
            r = B.run_case(model, 4, 128, 8, make_opt=_our_optimizer, clip=b200net.utils.clip_grad_norm_,
                           device=_dev())
            curves[graphs] = r["losses"] + [float(r["eval_logits"].abs().sum())]
            if graphs:
                recs = [v for v in bert._BG.values() if v != "failed"]
                assert any(v.bwd is not None and v.uses >= 4 for v in recs), "training graph was not used"
        finally:
            b200net.set_cuda_graphs(True)
    for a, b in zip(curves[True], curves[False]):
        assert abs(a - b) <= 1e-4 * max(1.0, abs(b)), (curves[True], curves[False])
