"""b200_bn_train_stats_noalloc: the variant of the BatchNorm statistics kernel for callers that cannot allocate.

Who needs it: a host process in which a framework has already taken all GPU memory for its own pool (so a later
cudaMalloc fails), or a latency-bound loop that cannot pay for the device-wide synchronisation cudaMalloc implies.
Those callers have no workspace to pass; this entry point keeps its per-block partial sums in a fixed 2 MiB array
inside the library instead.

The test drives the entry point through the C-ABI (ctypes, raw device pointers) and checks it against the existing
workspace-taking entry point on the same inputs, twice in a row with different shapes, so that whatever the first
call leaves in the shared array would show up in the second.  It lives in its own file, collected last, so a run
that stops at the first failure has already executed every other test.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _call_pair(x, C, dtype_code):
    """Runs the workspace variant and the no-alloc variant on the same input; returns both sets of outputs."""
    from b200net._C import lib, check
    dev = x.device
    rows = x.numel() // C
    g = torch.Generator().manual_seed(C)
    gamma = (torch.rand(C, generator=g) + 0.5).to(dev)
    beta = torch.randn(C, generator=g).to(dev)
    st = torch.cuda.current_stream().cuda_stream
    outs = []
    for variant in ("workspace", "noalloc"):
        rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        o = torch.full((4, C), float("nan"), device=dev)
        if variant == "workspace":
            n = int(lib.b200_bn_workspace_bytes(rows, C))
            ws = torch.empty(n, dtype=torch.uint8, device=dev)
            check(lib.b200_bn_train_stats(x.data_ptr(), rows, C, dtype_code, gamma.data_ptr(), beta.data_ptr(),
                                          rm.data_ptr(), rv.data_ptr(), 0.1, 1e-5, o[0].data_ptr(), o[1].data_ptr(),
                                          o[2].data_ptr(), o[3].data_ptr(), ws.data_ptr(), n, st), "bn_train_stats")
        else:
            check(lib.b200_bn_train_stats_noalloc(x.data_ptr(), rows, C, dtype_code, gamma.data_ptr(), beta.data_ptr(),
                                                  rm.data_ptr(), rv.data_ptr(), 0.1, 1e-5, o[0].data_ptr(),
                                                  o[1].data_ptr(), o[2].data_ptr(), o[3].data_ptr(), st),
                  "bn_train_stats_noalloc")
        torch.cuda.synchronize()
        outs.append((o, rm, rv))
    return outs


def _close(a, b, tol=1e-5):
    a, b = a.double(), b.double()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item() <= tol


@pytest.mark.parametrize("dtype,code", [(torch.float32, 0), (torch.bfloat16, 1)])
def test_noalloc_matches_workspace_variant_on_two_calls_in_a_row(dtype, code):
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(7)
    # first call: large (the block count is capped by the fixed array); second call: small, different C, so any
    # value left over in the array from call one sits exactly where call two would read if it read too far
    for rows, C in ((20000, 2048), (256 * 20 * 20, 256), (3 * 7 * 5, 2048), (1000, 24)):
        x = (torch.randn((rows, C), generator=g) * 2 + 0.5).to(dtype).to(dev)
        (o_ws, rm_ws, rv_ws), (o_na, rm_na, rv_na) = _call_pair(x, C, code)
        assert torch.isfinite(o_na).all()
        for i, name in enumerate(("mean", "invstd", "scale", "shift")):
            assert _close(o_na[i], o_ws[i]), (rows, C, name)
        assert _close(rm_na, rm_ws) and _close(rv_na, rv_ws)


def test_noalloc_rejects_sizes_beyond_its_limit():
    import b200net
    from b200net._C import lib
    dev = torch.device("cuda:0")
    C = 4096 + 8
    x = torch.zeros((4, C), device=dev)
    o = torch.zeros((4, C), device=dev)
    rc = lib.b200_bn_train_stats_noalloc(x.data_ptr(), 4, C, 0, None, None, None, None, 0.1, 1e-5, o[0].data_ptr(),
                                         o[1].data_ptr(), o[2].data_ptr(), o[3].data_ptr(), None)
    assert rc == 2                                   # B200_UNSUPPORTED
    assert "limit" in b200net._C.last_error() and "4096" in b200net._C.last_error()
