"""Allocation-free C-ABI entry point (ssl4p_layernorm_bwd_static: scratch in a file-scope device array) against the
existing ssl4p_layernorm_bwd with a caller workspace.  Collected last on purpose (file name), so that a failure here
does not hide other tests under `pytest -x`.  Passes alone in a fresh process and after the other tests."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda:0")


def rel_err(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-20)).item()


def _inputs(M, D, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = (torch.randn(M, D, generator=g) * 2 + 0.5).to(dev())
    gamma = (1 + 0.1 * torch.randn(D, generator=g)).to(dev())
    dy = torch.randn(M, D, generator=g).bfloat16().to(dev())
    res = torch.randn(M, D, generator=g).to(dev())
    mean = x.mean(1)
    rstd = (x.var(1, unbiased=False) + 1e-6).rsqrt()
    return x, gamma, dy, res, mean.contiguous(), rstd.contiguous()


def _run(lib, static, M, D, seed, accumulate):
    """One backward through the raw C-ABI; every output buffer is allocated here, before the call."""
    from ssl4polyp_b200 import _lib
    x, gamma, dy, res, mean, rstd = _inputs(M, D, seed)
    dx = torch.empty(M, D, device=dev())
    dxb = torch.empty(M, D, dtype=torch.bfloat16, device=dev())
    dgamma = torch.full((D,), 0.25, device=dev())
    dbeta = torch.full((D,), -0.5, device=dev())
    colsum = torch.ones(D, device=dev())
    stream = torch.cuda.current_stream().cuda_stream
    args = [dy.data_ptr(), x.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(), res.data_ptr(), dx.data_ptr(), dxb.data_ptr(),
            dgamma.data_ptr(), dbeta.data_ptr(), colsum.data_ptr(), M, D, int(accumulate), int(accumulate)]
    if static:
        _lib.check(lib.ssl4p_layernorm_bwd_static(*args, stream), "layernorm_bwd_static")
    else:
        nb = lib.ssl4p_layernorm_bwd_workspace_bytes(M, D)
        ws = torch.empty(nb, dtype=torch.uint8, device=dev())
        _lib.check(lib.ssl4p_layernorm_bwd(*args, ws.data_ptr(), nb, stream), "layernorm_bwd")
    torch.cuda.synchronize()
    return dx, dxb, dgamma, dbeta, colsum


def test_layernorm_bwd_static_matches_workspace_entry_point_on_two_calls_in_a_row():
    from ssl4polyp_b200 import _lib
    lib = _lib.load()
    assert lib.ssl4p_layernorm_bwd_max_static_d() == 1024
    # second call: fewer rows and a narrower D than the first, so stale partial sums of call 1 sit in the array beyond
    # (and inside) what call 2 writes; third call: the maximum D, accumulating into dgamma/dbeta/colsum
    for k, (M, D, accumulate) in enumerate([(12800, 768, False), (300, 512, False), (5000, 1024, True), (1000, 200, True)]):
        want = _run(lib, False, M, D, 100 + k, accumulate)
        got = _run(lib, True, M, D, 100 + k, accumulate)
        for name, a, b in zip(("dx", "dxb", "dgamma", "dbeta", "dx_colsum"), got, want):
            assert torch.isfinite(a.float()).all(), (name, M, D)
            assert rel_err(a, b) < 1e-4, (name, M, D, rel_err(a, b))


def test_layernorm_bwd_static_rejects_sizes_beyond_the_static_limit():
    from ssl4polyp_b200 import _lib
    lib = _lib.load()
    M, D = 8, 2048
    z = torch.zeros(M * D, device=dev())
    zb = torch.zeros(M * D, dtype=torch.bfloat16, device=dev())
    rc = lib.ssl4p_layernorm_bwd_static(zb.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), None, z.data_ptr(), None,
                                        z.data_ptr(), z.data_ptr(), None, M, D, 0, 0, torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.ssl4p_last_error().decode()
    assert "static scratch limit" in msg and "1024" in msg, msg
