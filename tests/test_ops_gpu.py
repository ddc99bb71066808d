"""GPU parity tests of the individual C-ABI kernels against plain fp32 torch / the oracle (tests only)."""
import math
import os

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from oracle import mae_oracle as O  # noqa: E402  (checker only)


def dev():
    return torch.device("cuda:0")


def rel_err(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-20)).item()


# ---------------------------------------------------------------------------------------------- GEMM
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 512), (100, 768, 768), (394, 2304, 768), (1000, 192, 1024), (12800, 768, 3072)])
@pytest.mark.parametrize("bn", [0, 64, 128, 192, 256])
def test_gemm_nt_bias(M, N, K, bn):
    from ssl4polyp_b200 import ops
    g = torch.Generator(device="cpu").manual_seed(M + N + K)
    a = torch.randn(M, K, generator=g).to(dev()).bfloat16()
    w = (torch.randn(N, K, generator=g) / math.sqrt(K)).to(dev()).bfloat16()
    bias = torch.randn(N, generator=g).to(dev())
    ref = a.float() @ w.float().t() + bias
    out = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16, bias=bias, force_bn=bn)
    assert rel_err(out, ref) < 1e-2, rel_err(out, ref)
    out32 = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_F32, bias=bias, force_bn=bn)
    assert rel_err(out32, ref) < 2e-3, rel_err(out32, ref)


def test_gemm_nt_gelu_residual():
    from ssl4polyp_b200 import ops
    M, N, K = 394, 2048, 512
    a = torch.randn(M, K, device=dev()).bfloat16()
    w = (torch.randn(N, K, device=dev()) / math.sqrt(K)).bfloat16()
    bias = torch.randn(N, device=dev())
    h_ref = a.float() @ w.float().t() + bias
    dg, g = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_GELU_BF16, bias=bias)
    x = h_ref
    dgelu = 0.5 * (1 + torch.erf(x / math.sqrt(2))) + x * torch.exp(-0.5 * x * x) / math.sqrt(2 * math.pi)
    assert rel_err(g, F.gelu(h_ref)) < 4e-3                       # bf16 rounding only: the erf evaluation is 1.5e-7 accurate
    assert rel_err(dg, dgelu) < 4e-3
    assert (g.float() - F.gelu(h_ref)).abs().max().item() < 2e-2 and (dg.float() - dgelu).abs().max().item() < 1e-2
    res = torch.randn(M, N, device=dev())
    out = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_RESIDUAL_F32, bias=bias, aux=res)
    assert rel_err(out, h_ref + res) < 2e-3


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (394, 512, 2048), (100, 768, 2304), (12800, 768, 3072)])
@pytest.mark.parametrize("bn", [0, 64, 128, 192, 256])
def test_gemm_nn_dgrad(M, N, K, bn):
    from ssl4polyp_b200 import ops
    dy = torch.randn(M, K, device=dev()).bfloat16()
    w = (torch.randn(K, N, device=dev()) / math.sqrt(K)).bfloat16()          # [K=out_features, N=in_features]
    ref = dy.float() @ w.float()
    out = ops.gemm(dy, w, ops.GEMM_NN, ops.EPI_BIAS_BF16, force_bn=bn)
    assert rel_err(out, ref) < 1e-2, rel_err(out, ref)
    dact = torch.randn(M, N, device=dev()).bfloat16()            # the saved gelu'(h) of the forward epilogue
    out2 = ops.gemm(dy, w, ops.GEMM_NN, ops.EPI_DGELU_BF16, aux=dact, force_bn=bn)
    assert rel_err(out2, ref * dact.float()) < 1e-2
    # the same epilogue emits the column sums of its output (Mlp.fc1 bias gradient) with no second pass
    cs = torch.full((N,), 7.0, device=dev())
    out3 = ops.gemm(dy, w, ops.GEMM_NN, ops.EPI_DGELU_BF16, aux=dact, force_bn=bn, colsum_out=cs)
    assert torch.equal(out3, out2)
    want = (ref * dact.float()).sum(0)
    assert rel_err(cs, want) < 2e-3, rel_err(cs, want)
    ops.gemm(dy, w, ops.GEMM_NN, ops.EPI_DGELU_BF16, aux=dact, force_bn=bn, colsum_out=cs, colsum_accumulate=True)
    assert rel_err(cs, 2 * want) < 2e-3


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (768, 768, 100), (2304, 768, 394), (512, 2048, 12800), (768, 3072, 1000)])
@pytest.mark.parametrize("splits", [1, 3, 0])
@pytest.mark.parametrize("bn", [0, 64, 256])
def test_gemm_tn_wgrad(M, N, K, splits, bn):
    from ssl4polyp_b200 import ops
    dy = torch.randn(K, M, device=dev()).bfloat16()       # [tokens, out_features]
    x = torch.randn(K, N, device=dev()).bfloat16()        # [tokens, in_features]
    ref = dy.float().t() @ x.float()
    s = splits if splits > 0 else ops.pick_splits(M, N, K)
    out = ops.gemm(dy, x, ops.GEMM_TN, ops.EPI_PARTIAL_F32, splits=s, force_bn=bn)
    assert rel_err(out, ref) < 2e-3, (rel_err(out, ref), s)
    acc = torch.ones(M, N, device=dev())
    ops.gemm(dy, x, ops.GEMM_TN, ops.EPI_PARTIAL_F32, splits=s, accumulate_into=acc, force_bn=bn)
    assert rel_err(acc, ref + 1) < 2e-3


CG2 = 2 << 16          # force_bn bits 16..: CTA group 2 (tcgen05.mma.cta_group::2, a CTA pair per 256-row tile)


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (300, 512, 512), (394, 2304, 768), (1000, 192, 1024), (12800, 768, 3072), (25216, 768, 768)])
@pytest.mark.parametrize("bn", [CG2, CG2 | 128, CG2 | 192, CG2 | 256])
def test_gemm_cta_pair_forward_epilogues(M, N, K, bn):
    """Every NT epilogue through the CTA-pair kernel (ragged last row tile, N not a multiple of the tile, multi-wave)."""
    from ssl4polyp_b200 import ops
    g = torch.Generator(device="cpu").manual_seed(M + N + K + bn)
    a = torch.randn(M, K, generator=g).to(dev()).bfloat16()
    w = (torch.randn(N, K, generator=g) / math.sqrt(K)).to(dev()).bfloat16()
    bias = torch.randn(N, generator=g).to(dev())
    ref = a.float() @ w.float().t() + bias
    out = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16, bias=bias, force_bn=bn)
    assert rel_err(out, ref) < 1e-2, rel_err(out, ref)
    out32 = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_F32, bias=bias, force_bn=bn)
    assert rel_err(out32, ref) < 2e-3, rel_err(out32, ref)
    res = torch.randn(M, N, generator=g).to(dev())
    outr = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_RESIDUAL_F32, bias=bias, aux=res, force_bn=bn)
    assert rel_err(outr, ref + res) < 2e-3
    dg, ge = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_GELU_BF16, bias=bias, force_bn=bn)
    dgelu = 0.5 * (1 + torch.erf(ref / math.sqrt(2))) + ref * torch.exp(-0.5 * ref * ref) / math.sqrt(2 * math.pi)
    assert rel_err(ge, F.gelu(ref)) < 4e-3 and rel_err(dg, dgelu) < 4e-3
    # the pair and the single-CTA kernel run the same k order per output element: identical bits
    assert torch.equal(out, ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16, bias=bias, force_bn=(1 << 16) | (bn & 0xffff)))


@pytest.mark.parametrize("M,N,K", [(256, 128, 64), (394, 512, 2048), (300, 768, 2304), (12800, 768, 3072)])
@pytest.mark.parametrize("bn", [CG2, CG2 | 128, CG2 | 192, CG2 | 256])
def test_gemm_cta_pair_dgrad(M, N, K, bn):
    from ssl4polyp_b200 import ops
    dy = torch.randn(M, K, device=dev()).bfloat16()
    w = (torch.randn(K, N, device=dev()) / math.sqrt(K)).bfloat16()
    ref = dy.float() @ w.float()
    out = ops.gemm(dy, w, ops.GEMM_NN, ops.EPI_BIAS_BF16, force_bn=bn)
    assert rel_err(out, ref) < 1e-2, rel_err(out, ref)
    dact = torch.randn(M, N, device=dev()).bfloat16()
    cs = torch.full((N,), 7.0, device=dev())
    out2 = ops.gemm(dy, w, ops.GEMM_NN, ops.EPI_DGELU_BF16, aux=dact, force_bn=bn, colsum_out=cs)
    assert rel_err(out2, ref * dact.float()) < 1e-2
    assert rel_err(cs, (ref * dact.float()).sum(0)) < 2e-3


@pytest.mark.parametrize("M,N,K", [(256, 64, 64), (768, 768, 100), (2304, 768, 394), (512, 2048, 12800), (768, 3072, 1000), (512, 512, 50432)])
@pytest.mark.parametrize("splits", [1, 3, 0])
@pytest.mark.parametrize("bn", [CG2, CG2 | 128, CG2 | 256])
def test_gemm_cta_pair_wgrad(M, N, K, splits, bn):
    from ssl4polyp_b200 import ops
    dy = torch.randn(K, M, device=dev()).bfloat16()
    x = torch.randn(K, N, device=dev()).bfloat16()
    ref = dy.float().t() @ x.float()
    s = splits if splits > 0 else ops.pick_splits(M, N, K)
    out = ops.gemm(dy, x, ops.GEMM_TN, ops.EPI_PARTIAL_F32, splits=s, force_bn=bn)
    assert rel_err(out, ref) < 2e-3, (rel_err(out, ref), s)


def test_gemm_cta_group_policy_switch():
    """ssl4p_gemm_set_cta_group: both policies give the same bits on a model-shaped GEMM; the previous value comes back."""
    from ssl4polyp_b200 import ops, _lib
    lib = _lib.load()
    a = torch.randn(1024, 768, device=dev()).bfloat16()
    w = (torch.randn(2304, 768, device=dev()) / 28).bfloat16()
    prev = lib.ssl4p_gemm_set_cta_group(1)
    try:
        o1 = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16)
        assert lib.ssl4p_gemm_set_cta_group(2) == 1
        o2 = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16)
    finally:
        lib.ssl4p_gemm_set_cta_group(prev)
    assert torch.equal(o1, o2)


def test_gemm_dynamic_tile_scheduler_slots_reset_themselves():
    """ssl4p_gemm_set_scheduler_buffer: with a 4-slot buffer (2 for captured launches, 2 rotating for eager ones) every eager
    slot is reused after two launches, so the results only stay right if each kernel puts its {next tile, finished units}
    pair back to zero; static tile lists (NULL) give the same bits; a captured launch keeps its slot across replays."""
    from ssl4polyp_b200 import ops, _lib
    lib = _lib.load()
    a = torch.randn(12800, 768, device=dev()).bfloat16()
    w = (torch.randn(2304, 768, device=dev()) / 28).bfloat16()
    dy = torch.randn(12800, 768, device=dev()).bfloat16()
    ref = ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16)                  # whatever the process-wide setting is
    ref_w = ops.gemm(dy, a, ops.GEMM_TN, ops.EPI_PARTIAL_F32, splits=ops.pick_splits(768, 768, 12800))
    small = torch.zeros(4 * 32, dtype=torch.uint8, device=dev())
    try:
        _lib.check(lib.ssl4p_gemm_set_scheduler_buffer(small.data_ptr(), small.numel()), "set_scheduler_buffer")
        for _ in range(7):
            assert torch.equal(ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16), ref)
            assert torch.equal(ops.gemm(dy, a, ops.GEMM_TN, ops.EPI_PARTIAL_F32, splits=ops.pick_splits(768, 768, 12800)), ref_w)
        torch.cuda.synchronize()
        assert int(small.view(torch.int32).abs().sum()) == 0               # every slot is back to zero
        out = torch.empty_like(ref)
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16, out=out)       # warm-up outside the capture
            import gc
            gc.collect()                                                   # a graph of an earlier test freed mid-capture would
            gc.disable()                                                   # invalidate this one (see engine._replay)
            try:
                with torch.cuda.graph(g, stream=s):
                    ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16, out=out)
            finally:
                gc.enable()
        torch.cuda.current_stream().wait_stream(s)
        for _ in range(3):
            out.zero_()
            g.replay()
            torch.cuda.synchronize()
            assert torch.equal(out, ref)
        _lib.check(lib.ssl4p_gemm_set_scheduler_buffer(None, 0), "set_scheduler_buffer")
        assert torch.equal(ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16), ref)   # static tile lists
    finally:
        buf = ops._sched_buffers.get(torch.cuda.current_device())
        if buf is not None:
            lib.ssl4p_gemm_set_scheduler_buffer(buf.data_ptr(), buf.numel())
        else:
            lib.ssl4p_gemm_set_scheduler_buffer(None, 0)
        del g


def test_gemm_strided_views():
    """A operand as a row-strided view (qkv slices), C as a strided view."""
    from ssl4polyp_b200 import ops
    M, K, N = 300, 512, 256
    big = torch.randn(M, 3 * K, device=dev()).bfloat16()
    a = big[:, K:2 * K]
    w = (torch.randn(N, K, device=dev()) / math.sqrt(K)).bfloat16()
    outbig = torch.zeros(M, 2 * N, device=dev(), dtype=torch.bfloat16)
    ops.gemm(a, w, ops.GEMM_NT, ops.EPI_BIAS_BF16, out=outbig[:, N:])
    assert rel_err(outbig[:, N:], a.float() @ w.float().t()) < 1e-2
    assert outbig[:, :N].abs().max().item() == 0


# ---------------------------------------------------------------------------------------------- masking / tokens
@pytest.mark.parametrize("N,L,ratio", [(4, 196, 0.75), (256, 196, 0.75), (3, 16, 0.75), (5, 196, 0.6), (2, 1024, 0.5)])
def test_mask_from_noise_bit_exact(N, L, ratio):
    from ssl4polyp_b200 import ops
    noise = torch.rand(N, L, generator=torch.Generator().manual_seed(N * L))
    keep = O.len_keep_of(L, ratio)
    ids_keep_o, mask_o, ids_restore_o = O.masking_from_noise(noise, ratio)
    ids_restore, mask, ids_keep = ops.mask_from_noise(noise.to(dev()), keep)
    assert ids_restore.dtype == torch.int64 and mask.dtype == torch.float32
    assert torch.equal(ids_restore.cpu(), ids_restore_o)
    assert torch.equal(mask.cpu(), mask_o)
    assert torch.equal(ids_keep.cpu(), ids_keep_o)


def test_mask_from_noise_ties_golden():
    from ssl4polyp_b200 import ops
    import os
    gold = torch.load(os.path.join(os.path.dirname(__file__), "golden", "mae_golden.pt"))["ties"]
    ids_restore, mask, ids_keep = ops.mask_from_noise(gold["noise"].to(dev()), gold["ids_keep"].shape[1])
    assert torch.equal(ids_restore.cpu(), gold["ids_restore"])
    assert torch.equal(mask.cpu(), gold["mask"])
    assert torch.equal(ids_keep.cpu(), gold["ids_keep"])


@pytest.mark.parametrize("N,size", [(2, 224), (3, 64)])
def test_patchify_roundtrip(N, size):
    from ssl4polyp_b200 import ops
    imgs = torch.randn(N, 3, size, size)
    x = ops.patchify(imgs.to(dev()), 16)
    assert torch.equal(x.cpu(), O.patchify(imgs, 16))
    back = ops.unpatchify(x, 16)
    assert torch.equal(back.cpu(), imgs)


def test_im2col_assemble_gather():
    from ssl4polyp_b200 import ops
    N, L, D, keep = 3, 196, 768, 49
    imgs = torch.randn(N, 3, 224, 224)
    noise = torch.rand(N, L)
    ids_keep, _, ids_restore = O.masking_from_noise(noise, 0.75)
    cols = ops.im2col_patches(imgs.to(dev()), ids_keep.to(dev()), 16)
    ref = imgs.reshape(N, 3, 14, 16, 14, 16).permute(0, 2, 4, 1, 3, 5).reshape(N, L, 768)   # (c, pi, qi) column order
    ref = torch.gather(ref, 1, ids_keep.unsqueeze(-1).expand(-1, -1, 768)).reshape(N * keep, 768)
    assert torch.equal(cols.float().cpu(), ref.bfloat16().float())
    allcols = ops.im2col_patches(imgs.to(dev()), None, 16)
    assert torch.equal(allcols.float().cpu(), imgs.reshape(N, 3, 14, 16, 14, 16).permute(0, 2, 4, 1, 3, 5).reshape(N * L, 768).bfloat16().float())
    tok = torch.randn(N * keep, D); pos = torch.randn(1 + L, D); cls = torch.randn(D)
    x = ops.assemble_tokens(tok.to(dev()), ids_keep.to(dev()), pos.to(dev()), cls.to(dev()), N, keep, D)
    ref_x = torch.cat([(cls + pos[0]).expand(N, 1, D), tok.reshape(N, keep, D) + pos[1:][ids_keep]], dim=1)
    assert torch.allclose(x.cpu(), ref_x, atol=1e-6)
    xs = torch.randn(N, L, D)
    got = ops.gather_tokens(xs.to(dev()), ids_keep.to(dev()))
    assert torch.equal(got.cpu(), torch.gather(xs, 1, ids_keep.unsqueeze(-1).expand(-1, -1, D)))
    dx = torch.randn(N, keep + 1, D)
    dcls = torch.zeros(D, device=dev())
    dtok = ops.assemble_tokens_bwd(dx.to(dev()), dcls)
    assert torch.equal(dtok.float().cpu(), dx[:, 1:].reshape(N * keep, D).bfloat16().float())
    assert torch.allclose(dcls.cpu(), dx[:, 0].sum(0), atol=1e-5)


def test_decoder_unshuffle_fwd_bwd():
    from ssl4polyp_b200 import ops
    N, L, Dd, keep = 3, 196, 512, 49
    noise = torch.rand(N, L)
    ids_keep, mask, ids_restore = O.masking_from_noise(noise, 0.75)
    y = torch.randn(N, keep + 1, Dd).bfloat16()
    mt = torch.randn(Dd); dpos = torch.randn(1 + L, Dd)
    x = ops.decoder_unshuffle_fwd(y.to(dev()), mt.to(dev()), ids_restore.to(dev()), dpos.to(dev()))
    yf = y.float()
    body = torch.cat([yf[:, 1:], mt.expand(N, L - keep, Dd)], dim=1)
    body = torch.gather(body, 1, ids_restore.unsqueeze(-1).expand(-1, -1, Dd))
    ref = torch.cat([yf[:, :1], body], dim=1) + dpos
    assert torch.allclose(x.cpu(), ref, atol=1e-6)
    dx = torch.randn(N, 1 + L, Dd)
    dmt = torch.zeros(Dd, device=dev())
    dy = ops.decoder_unshuffle_bwd(dx.to(dev()), ids_keep.to(dev()), mask.to(dev()), dmt)
    ref_dy = torch.cat([dx[:, :1], torch.gather(dx[:, 1:], 1, ids_keep.unsqueeze(-1).expand(-1, -1, Dd))], dim=1)
    assert torch.equal(dy.float().cpu(), ref_dy.bfloat16().float())
    ref_dmt = (dx[:, 1:] * mask.unsqueeze(-1)).sum((0, 1))
    assert torch.allclose(dmt.cpu(), ref_dmt, atol=1e-3, rtol=1e-4)


# ---------------------------------------------------------------------------------------------- LayerNorm
@pytest.mark.parametrize("M,D", [(100, 768), (394, 512), (50, 1024), (37, 64), (20, 32), (5000, 512), (12803, 768), (6401, 1024)])
def test_layernorm_fwd_bwd(M, D):
    from ssl4polyp_b200 import ops
    x = (torch.randn(M, D) * 2 + 0.5).requires_grad_(True)
    gamma = (1 + 0.1 * torch.randn(D)).requires_grad_(True)
    beta = (0.1 * torch.randn(D)).requires_grad_(True)
    y_ref = F.layer_norm(x, (D,), gamma, beta, 1e-6)
    y, mean, rstd = ops.layernorm_fwd(x.detach().to(dev()), gamma.detach().to(dev()), beta.detach().to(dev()), 1e-6)
    assert rel_err(y.cpu(), y_ref.detach()) < 5e-3
    dy = torch.randn(M, D).bfloat16()
    y_ref.backward(dy.float())
    res = torch.randn(M, D)
    dgamma = torch.zeros(D, device=dev()); dbeta = torch.zeros(D, device=dev())
    colsum = torch.ones(D, device=dev())
    dx, dxb = ops.layernorm_bwd(dy.to(dev()), x.detach().to(dev()), mean, rstd, gamma.detach().to(dev()), res.to(dev()).clone(), dgamma, dbeta,
                                dx_colsum=colsum, accumulate_colsum=True)
    assert rel_err(dx.cpu(), x.grad + res) < 1e-4
    assert rel_err(colsum.cpu(), 1 + (x.grad + res).sum(0)) < 1e-4          # the upstream Linear's bias gradient, same pass
    assert rel_err(dxb.cpu(), x.grad + res) < 5e-3
    assert rel_err(dgamma.cpu(), gamma.grad) < 1e-4
    assert rel_err(dbeta.cpu(), beta.grad) < 1e-4
    dx2, none = ops.layernorm_bwd(dy.to(dev()), x.detach().to(dev()), mean, rstd, gamma.detach().to(dev()), None, dgamma, dbeta, want_bf16=False, accumulate=True)
    assert none is None and rel_err(dx2.cpu(), x.grad) < 1e-4
    assert rel_err(dgamma.cpu(), 2 * gamma.grad) < 1e-4


# ---------------------------------------------------------------------------------------------- attention
def _mma_check():
    """tests/libssl4p_check.so: the mma.sync attention kernels, an independent on-device implementation (test infrastructure,
    not part of the product library)."""
    import ctypes as C
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libssl4p_check.so")
    lib = C.CDLL(path)
    vp, i32, f32 = C.c_void_p, C.c_int, C.c_float
    lib.ssl4p_attention_fwd_mma.argtypes = [vp, vp, vp, i32, i32, i32, i32, f32, vp]
    lib.ssl4p_attention_bwd_mma.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, vp]
    lib.ssl4p_check_last_error.restype = C.c_char_p
    return lib


def _attention_mma(qkv, B, T, H, hd, scale, out=None, dout=None, lse=None):
    lib = _mma_check()
    st = torch.cuda.current_stream().cuda_stream
    if dout is None:
        out = torch.empty(B * T, H * hd, dtype=torch.bfloat16, device=qkv.device)
        lse = torch.empty(B, H, T, dtype=torch.float32, device=qkv.device)
        rc = lib.ssl4p_attention_fwd_mma(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), B, T, H, hd, scale, st)
        assert rc == 0, lib.ssl4p_check_last_error()
        return out, lse
    dqkv = torch.empty_like(qkv)
    rc = lib.ssl4p_attention_bwd_mma(qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), dqkv.data_ptr(), B, T, H, hd, scale, st)
    assert rc == 0, lib.ssl4p_check_last_error()
    return dqkv


@pytest.mark.parametrize("impl", ["tcgen05", "tcgen05-alias", "tcgen05-sep", "tcgen05-nopair", "mma"])
@pytest.mark.parametrize("B,T,H,hd", [(2, 50, 12, 64), (2, 197, 16, 32), (1, 197, 12, 64), (3, 5, 2, 32), (2, 17, 2, 32), (2, 64, 4, 64),
                                      (40, 50, 12, 64), (21, 197, 16, 32), (3, 128, 2, 64), (2, 129, 4, 32), (1, 256, 2, 32), (300, 208, 2, 32), (5, 209, 2, 64), (160, 16, 2, 32),
                                      (1, 50, 12, 64), (301, 50, 2, 64), (7, 63, 4, 32), (2, 65, 2, 64)])
def test_attention_fwd_bwd(B, T, H, hd, impl, monkeypatch):
    """Both attention kernels against fp32 torch on the same bf16 inputs (the larger batches put several work items
    on every persistent CTA; T = 128/129/256 sit on the tile boundaries)."""
    from ssl4polyp_b200 import ops
    if impl == "mma" and T > 208:
        pytest.skip("the mma.sync cross-check kernel stops at T = 208")
    if impl == "tcgen05-nopair":                             # T <= 64 normally packs two images per tile: also run one image per tile
        if T > 64:
            pytest.skip("pairing only exists for T <= 64")
        monkeypatch.setenv("SSL4P_ATTN_PAIR", "0")
        impl = "tcgen05"
    elif impl.startswith("tcgen05-"):                        # force one of the two TMEM plans of the forward for every T
        monkeypatch.setenv("SSL4P_ATTN_SEP", "0" if impl.endswith("alias") else "1")
        impl = "tcgen05"
    torch.manual_seed(B * 1000 + T)
    qkv = torch.randn(B, T, 3, H, hd).bfloat16()
    scale = hd ** -0.5
    q, k, v = [t.float().requires_grad_(True) for t in qkv.permute(2, 0, 3, 1, 4)]
    a = ((q @ k.transpose(-2, -1)) * scale).softmax(-1)
    o_ref = (a @ v).transpose(1, 2).reshape(B, T, H * hd)
    qkv_d = qkv.reshape(B * T, 3 * H * hd).to(dev())
    out, lse = _attention_mma(qkv_d, B, T, H, hd, scale) if impl == "mma" else ops.attention_fwd(qkv_d, B, T, H, hd, scale)
    assert rel_err(out.cpu().reshape(B, T, H * hd), o_ref.detach()) < 1e-2
    lse_ref = torch.logsumexp((q @ k.transpose(-2, -1)) * scale, dim=-1)
    assert torch.allclose(lse.cpu(), lse_ref.detach(), atol=2e-3, rtol=1e-3)
    do = torch.randn(B, T, H * hd).bfloat16()
    o_ref.backward(do.float())
    do_d = do.reshape(B * T, H * hd).to(dev())
    dqkv = _attention_mma(qkv_d, B, T, H, hd, scale, out=out, dout=do_d, lse=lse) if impl == "mma" else ops.attention_bwd(qkv_d, out, do_d, lse, B, T, H, hd, scale)
    dqkv = dqkv.cpu().float().reshape(B, T, 3, H, hd).permute(2, 0, 3, 1, 4)
    for i, (name, t) in enumerate((("dq", q), ("dk", k), ("dv", v))):
        assert rel_err(dqkv[i], t.grad) < 2e-2, (name, rel_err(dqkv[i], t.grad))


# ---------------------------------------------------------------------------------------------- loss
@pytest.mark.parametrize("norm_pix", [True, False])
@pytest.mark.parametrize("N,size,cls_row", [(2, 224, True), (3, 64, False)])
def test_masked_mse(norm_pix, N, size, cls_row):
    from ssl4polyp_b200 import ops
    L = (size // 16) ** 2
    imgs = torch.randn(N, 3, size, size)
    _, mask, _ = O.masking_from_noise(torch.rand(N, L), 0.75)
    pred_full = torch.randn(N, L + (1 if cls_row else 0), 768)
    pred = (pred_full[:, 1:] if cls_row else pred_full).clone().requires_grad_(True)
    loss_ref = O.masked_mse(imgs, pred, mask, norm_pix)
    loss_ref.backward()
    loss, ws = ops.masked_mse_fwd(imgs.to(dev()), pred_full.to(dev()), cls_row, mask.to(dev()), norm_pix, 16)
    assert abs(loss.item() - loss_ref.item()) / abs(loss_ref.item()) < 1e-5
    gout = torch.tensor(0.5, device=dev())
    dpred = ops.masked_mse_bwd(imgs.to(dev()), pred_full.to(dev()), cls_row, mask.to(dev()), norm_pix, 16, gout, ws)
    got = dpred.float().cpu()
    if cls_row:
        assert got[:, 0].abs().max().item() == 0
        got = got[:, 1:]
    assert rel_err(got, 0.5 * pred.grad) < 5e-3
    assert got[mask == 0].abs().max().item() == 0


# ---------------------------------------------------------------------------------------------- misc
def test_colsum_cast_adamw():
    from ssl4polyp_b200 import ops
    x = torch.randn(1000, 768).bfloat16()
    out = ops.colsum(x.to(dev()))
    assert rel_err(out.cpu(), x.float().sum(0)) < 1e-4
    # the 8-columns-per-lane kernel (N, ld multiples of 8) on ragged row counts, a strided view and an accumulate; the
    # 2-columns-per-lane kernel for the rest
    for M, N in ((50432, 1536), (12800, 2304), (37, 8), (1003, 264), (999, 770)):
        y = torch.randn(M, N, generator=torch.Generator().manual_seed(M)).bfloat16()
        assert rel_err(ops.colsum(y.to(dev())).cpu(), y.float().sum(0)) < 1e-4, (M, N)
    big = torch.randn(777, 3 * 512).bfloat16().to(dev())
    view = big[:, 512:1024]
    acc = torch.ones(512, device=dev())
    ops.colsum(view, out=acc, accumulate=True)
    assert rel_err(acc.cpu(), 1 + view.float().sum(0).cpu()) < 1e-4
    src = torch.randn(100003)
    assert torch.equal(ops.cast_bf16(src.to(dev())).float().cpu(), src.bfloat16().float())
    n = 256 * 40
    p = torch.randn(n); g = torch.randn(n)
    flags = (torch.arange(n // 256) % 2).to(torch.uint8)
    ref_p = torch.nn.Parameter(p.clone())
    opt = torch.optim.AdamW([{"params": [ref_p]}], lr=1e-2, betas=(0.9, 0.95), weight_decay=0.0)
    pd, md, vd = p.to(dev()), torch.zeros(n, device=dev()), torch.zeros(n, device=dev())
    pb = torch.empty(n, dtype=torch.bfloat16, device=dev())
    bad = torch.zeros(1, dtype=torch.int32, device=dev())
    wd = 0.05
    for step in (1, 2, 3):
        ref_p.grad = g.clone()
        # reference with per-chunk weight decay: decoupled decay applied manually
        with torch.no_grad():
            decay = flags.repeat_interleave(256).float()
            ref_p.mul_(1 - 1e-2 * wd * decay)
        opt.step()
        ops.adamw_flat(pd, g.to(dev()), md, vd, pb, flags.to(dev()), 1e-2, 0.9, 0.95, 1e-8, wd, step, nonfinite=bad)
    assert rel_err(pd.cpu(), ref_p.detach()) < 1e-5
    assert torch.equal(pb.float().cpu(), pd.cpu().bfloat16().float())
    assert bad.item() == 0
    gbad = g.clone(); gbad[5] = float("nan")
    ops.adamw_flat(pd, gbad.to(dev()), md, vd, pb, flags.to(dev()), 1e-2, 0.9, 0.95, 1e-8, wd, 4, nonfinite=bad)
    assert bad.item() >= 1


def test_pool_head():
    from ssl4polyp_b200 import ops
    N, T, D, ncls = 4, 197, 768, 2
    xn = torch.randn(N, T, D).bfloat16()
    W = torch.randn(ncls, D) * 0.05; b = torch.randn(ncls)
    for spatial in (False, True):
        xf = xn.float().requires_grad_(True); Wr = W.clone().requires_grad_(True); br = b.clone().requires_grad_(True)
        pooled = xf[:, 1:].mean(1) if spatial else xf[:, 0]
        ref = F.linear(pooled, Wr, br)
        logits = ops.pool_head_fwd(xn.to(dev()), W.to(dev()), b.to(dev()), N, T, D, spatial)
        assert torch.allclose(logits.cpu(), ref.detach(), atol=1e-3, rtol=1e-3)
        dl = torch.randn(N, ncls)
        ref.backward(dl)
        dW = torch.zeros(ncls, D, device=dev()); db = torch.zeros(ncls, device=dev())
        dxn = ops.pool_head_bwd(xn.to(dev()), W.to(dev()), dl.to(dev()), dW, db, N, T, D, spatial)
        assert rel_err(dxn.cpu().reshape(N, T, D), xf.grad) < 1e-2
        assert rel_err(dW.cpu(), Wr.grad) < 1e-4 and rel_err(db.cpu(), br.grad) < 1e-5


# ---------------------------------------------------------------------------------------------------------------------
# input transform (main_pretrain.py:156-160): bit-exact against Pillow/torchvision goldens and the oracle
def _input_golden():
    import numpy as np
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "input_golden.npz"))


def _run_crop(frames, params, size=224):
    import numpy as np
    from ssl4polyp_b200 import ops
    desc, chunks, off = [], [], 0
    for f, (i, j, h, w, flip) in zip(frames, params):
        desc.append([off, f.shape[1] * 3, i, j, h, w, int(flip), 0])
        chunks.append(np.ascontiguousarray(f).reshape(-1))
        off += f.size
    buf = torch.from_numpy(np.concatenate(chunks)).cuda()
    out = ops.crop_resize_normalize(buf, torch.tensor(desc, dtype=torch.int64).cuda(), size,
                                    max(max(p[2], p[3]) for p in params), (0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_crop_resize_normalize_bit_exact_against_pillow_golden():
    """All eight golden cases in ONE launch (mixed frame sizes, 37x41 up-scaling to 1072x1920 down-scaling, both
    skipped-pass cases, flips): fp32 output sha256 == what torchvision produced; uint8 resample recovered exactly."""
    import hashlib
    import numpy as np
    from oracle import input_oracle as IO
    g = _input_golden()
    frames, params = [], []
    for case in range(8):
        H, W, i, j, h, w, flip = (int(v) for v in g[f"params{case}"])
        frames.append(IO.synthetic_frame(case, H, W)); params.append((i, j, h, w, bool(flip)))
    out = _run_crop(frames, params)
    mean = np.asarray(IO.IMAGENET_MEAN, np.float32)[:, None, None]; std = np.asarray(IO.IMAGENET_STD, np.float32)[:, None, None]
    for case in range(8):
        t = out[case]
        want = IO.to_tensor_normalize(g[f"u8_{case}"], params[case][4])
        assert np.array_equal(t.view(np.uint32), want.view(np.uint32)), (case, float(np.abs(t - want).max()))
        assert hashlib.sha256(np.ascontiguousarray(t).tobytes()).digest() == g[f"tensor_sha256_{case}"].tobytes(), case


@pytest.mark.parametrize("H,W,box", [(300, 300, (0, 0, 300, 300)), (225, 223, (0, 0, 225, 223)), (64, 3000, (0, 0, 64, 3000)),
                                      (3000, 64, (10, 1, 2990, 63)), (1, 1, (0, 0, 1, 1)), (500, 700, (499, 0, 1, 700))])
def test_crop_resize_edge_geometries_match_the_oracle(H, W, box):
    """Extreme aspect ratios, one-pixel crops, crops touching the frame border, scale 13 (55 taps)."""
    import numpy as np
    from oracle import input_oracle as IO
    frame = IO.synthetic_frame(11, H, W)
    i, j, h, w = box
    for flip in (False, True):
        got = _run_crop([frame], [(i, j, h, w, flip)])[0]
        want = IO.to_tensor_normalize(IO.crop_resize_u8(frame, i, j, h, w), flip)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), float(np.abs(got - want).max())


def test_crop_resize_rejects_oversized_crops_and_handles_empty_batches():
    from ssl4polyp_b200 import ops
    buf = torch.zeros(16, dtype=torch.uint8, device="cuda")
    desc = torch.zeros(1, 8, dtype=torch.int64, device="cuda")
    with pytest.raises(RuntimeError, match="taps"):
        ops.crop_resize_normalize(buf, desc, 224, 224 * 16, (0, 0, 0), (1, 1, 1))
    out = ops.crop_resize_normalize(buf, torch.zeros(0, 8, dtype=torch.int64, device="cuda"), 224, 10, (0, 0, 0), (1, 1, 1))
    assert out.shape == (0, 3, 224, 224)


def test_train_transform_reproduces_the_compose_under_one_seed():
    """TrainTransform (host RNG draw + packed uint8 H2D + kernels) == the reference Compose restated by the oracle, image
    for image under one torch seed; only the crop rows travel to the device."""
    import numpy as np
    from oracle import input_oracle as IO
    from ssl4polyp_b200.input_pipeline import TrainTransform
    frames = [IO.synthetic_frame(20 + k, 240 + 53 * k, 320 + 71 * k) for k in range(6)]
    tf = TrainTransform(224)
    for rep in range(2):                                   # second call reuses the other staging buffer
        torch.manual_seed(77 + rep)
        got = tf(frames).cpu().numpy()
        torch.manual_seed(77 + rep)
        for k, f in enumerate(frames):
            want, p = IO.transform_train(f)
            assert tuple(tf.last_params[k]) == p
            assert np.array_equal(got[k].view(np.uint32), want.view(np.uint32)), (rep, k)
    assert tf.h2d_bytes < sum(f.size for f in frames) and tf.h2d_bytes * 1.0 > 0
    # the batch feeds the model as the DataLoader's batch would
    from ssl4polyp_b200 import models_mae
    torch.manual_seed(0)
    m = models_mae.mae_vit_base_patch16(norm_pix_loss=True).cuda()
    loss, pred, mask = m(tf(frames), mask_ratio=0.75)
    assert torch.isfinite(loss) and pred.shape == (6, 196, 768)


def test_eval_transform_bilinear_resize_bit_exact_against_pillow_golden():
    """classification/data/transforms.py:233-258 (val / test): T.Resize((224,224)) PIL bilinear -> ToTensor -> Normalize."""
    import hashlib
    import numpy as np
    from oracle import input_oracle as IO
    from ssl4polyp_b200.input_pipeline import EvalTransform
    g = _input_golden()
    seeds = (0, 1, 4, 6, 7)
    frames = []
    for sd in seeds:
        H, W = (int(v) for v in g[f"params{sd}"][:2])
        frames.append(IO.synthetic_frame(sd, H, W))
    got = EvalTransform(224)(frames).cpu().numpy()
    for k, sd in enumerate(seeds):
        assert hashlib.sha256(np.ascontiguousarray(got[k]).tobytes()).digest() == g[f"eval_sha256_{sd}"].tobytes(), sd
    small = IO.synthetic_frame(3, 224, 224)
    want = IO.transform_eval(small)
    assert np.array_equal(EvalTransform(224)([small]).cpu().numpy()[0].view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("size", [96, 225])
def test_crop_resize_other_output_sizes_match_the_oracle(size):
    """`--input_size` other than 224 takes the runtime-size instantiation of the kernel."""
    import numpy as np
    from oracle import input_oracle as IO
    frame = IO.synthetic_frame(31, 333, 411)
    got = _run_crop([frame], [(7, 9, 300, 380, True)], size=size)[0]
    want = IO.to_tensor_normalize(IO.crop_resize_u8(frame, 7, 9, 300, 380, size), True)
    assert got.shape == (3, size, size) and np.array_equal(got.view(np.uint32), want.view(np.uint32))


# ---------------------------------------------------------------------------------------------- fine-tune TRAIN augmentations
def _aug_params(n, order=(4, 4, 4, 4), hue=0, hflip=0, vflip=0, mode=1, coef=(0,) * 6, b=1.0, c=1.0, s=1.0):
    import numpy as np
    p = np.zeros((n, 32), np.int32)
    p[:, 0:4] = order
    p[:, 4], p[:, 5], p[:, 6], p[:, 7] = hue, hflip, vflip, mode
    p[:, 8:14] = coef
    p[:, 16:19] = np.array([b, c, s], np.float32).view(np.int32)
    return torch.from_numpy(p).cuda()


def test_hue_shift_all_16m_colours_bit_exact():
    """adjust_hue = Pillow's RGB -> HSV -> RGB round trip (Convert.c) with a uint8 wrap-around shift: every one of the 2^24
    colours, for a zero shift (still lossy) and for the shifts ColorJitter(hue=0.01) can draw (0, +-1, +-2)."""
    import numpy as np
    from oracle import input_oracle as IO
    from ssl4polyp_b200 import ops
    g = np.arange(256, dtype=np.uint8)
    allrgb = np.ascontiguousarray(np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(256, 256, 256, 3))   # 256 images of 256 x 256
    dev_img = torch.from_numpy(allrgb).cuda()
    hsv0 = IO.rgb2hsv_u8(allrgb.reshape(4096, 4096, 3))
    for shift in (0, 2, 254, 100):
        got = ops.color_jitter_u8(dev_img, _aug_params(256, order=(3, 4, 4, 4), hue=shift)).cpu().numpy()
        hsv = hsv0.copy()
        hsv[..., 0] = hsv[..., 0] + np.uint8(shift)
        want = IO.hsv2rgb_u8(hsv).reshape(256, 256, 256, 3)
        assert np.array_equal(got, want), (shift, int((got != want).any(-1).sum()))


@pytest.mark.parametrize("S", [224, 64])
def test_colour_operations_match_the_oracle(S):
    import numpy as np
    from oracle import input_oracle as IO
    from ssl4polyp_b200 import ops
    rng = np.random.default_rng(S)
    imgs = np.stack([IO.synthetic_frame(40 + k, S, S) if k % 2 else rng.integers(0, 256, (S, S, 3), dtype=np.uint8) for k in range(4)])
    imgs[3] = imgs[3] // 4 + 100                                         # low contrast: the blend extrapolates less
    d = torch.from_numpy(imgs).cuda()
    for f in (0.6, 0.9999, 1.0, 1.0001, 1.4, 0.5, 1.5, 0.0):
        for op, fn in ((0, IO.adjust_brightness), (1, IO.adjust_contrast), (2, IO.adjust_saturation)):
            got = ops.color_jitter_u8(d, _aug_params(4, order=(op, 4, 4, 4), b=f, c=f, s=f)).cpu().numpy()
            for k in range(4):
                assert np.array_equal(got[k], fn(imgs[k], float(np.float32(f)))), (op, f, k)
    # the four operations chained in every order ColorJitter.get_params can draw, per-image parameters, in place
    import itertools
    orders = list(itertools.permutations(range(4)))
    big = np.stack([imgs[k % 4] for k in range(24)])
    prm = _aug_params(24, b=0.7, c=1.3, s=0.8, hue=255)
    prm[:, 0:4] = torch.tensor(orders, dtype=torch.int32)
    dd = torch.from_numpy(big).cuda()
    got = ops.color_jitter_u8(dd, prm, out=dd).cpu().numpy()
    for k, o in enumerate(orders):
        want = IO.color_jitter(big[k], o, float(np.float32(0.7)), float(np.float32(1.3)), float(np.float32(0.8)), -0.005)   # int32(-0.005 * 255) = -1 -> 255
        assert np.array_equal(got[k], want), o


@pytest.mark.parametrize("S,sigmas", [(224, (0.001, 0.3, 1.0, 2.0)), (64, (0.7, 1.9)), (30, (1.3,))])
def test_gaussian_blur_matches_the_oracle_bit_for_bit(S, sigmas):
    """Same float64 operations in the same order as oracle.gaussian_blur_u8 (itself <= 1 grey level from torchvision)."""
    import numpy as np
    from oracle import input_oracle as IO
    from ssl4polyp_b200 import ops
    imgs = np.stack([IO.synthetic_frame(60 + k, S, S) for k in range(len(sigmas))])
    taps = torch.zeros(len(sigmas), 32)
    for k, sg in enumerate(sigmas):
        taps[k, :25] = torch.from_numpy(IO.gaussian_kernel1d(25, sg))
    got = ops.gaussian_blur_u8(torch.from_numpy(imgs).cuda(), taps.cuda(), 25).cpu().numpy()
    for k, sg in enumerate(sigmas):
        assert np.array_equal(got[k], IO.gaussian_blur_u8(imgs[k], IO.gaussian_kernel1d(25, sg))), sg


def test_flip_rotate_normalize_matches_the_oracle():
    import numpy as np
    from oracle import input_oracle as IO
    from ssl4polyp_b200 import ops
    from ssl4polyp_b200.input_pipeline import _rotate_coeffs
    S = 224
    angles = [0.0, 180.0, -180.0, 90.0, -90.0, 270.0, 45.0, 1e-9, 118.42686462402344, -0.3, 179.999, -91.5]
    rng = np.random.default_rng(3)
    angles += [float(a) for a in rng.uniform(-180, 180, 12)]
    n = len(angles)
    imgs = np.stack([IO.synthetic_frame(80 + (k % 3), S, S) for k in range(n)])
    prm = _aug_params(n)
    for k, a in enumerate(angles):
        mode, coef = _rotate_coeffs(a, S, S)
        assert (mode, tuple(coef)) == IO.rotate_coeffs(a, S, S)
        prm[k, 5], prm[k, 6], prm[k, 7] = k & 1, (k >> 1) & 1, mode
        prm[k, 8:14] = torch.tensor(coef, dtype=torch.int32)
    out, u8 = ops.flip_rotate_normalize_u8(torch.from_numpy(imgs).cuda(), prm, IO.IMAGENET_MEAN, IO.IMAGENET_STD, want_u8=True)
    out, u8 = out.cpu().numpy(), u8.cpu().numpy()
    for k, a in enumerate(angles):
        img = imgs[k]
        if k & 1:
            img = img[:, ::-1]
        if (k >> 1) & 1:
            img = img[::-1]
        want = IO.rotate_nearest(np.ascontiguousarray(img), a)
        assert np.array_equal(u8[k], want), a
        assert np.array_equal(out[k].view(np.uint32), IO.to_tensor_normalize(want, False).view(np.uint32)), a


def test_train_augment_reproduces_the_compose_under_one_seed():
    """input_pipeline.TrainAugment (host RNG draws + uint8 H2D + five kernels) against (a) the oracle's stage-'train'
    transform, bit for bit, and (b) what the installed torchvision Compose produced under the same seed
    (tests/golden/augment_golden.npz): the same draws, <= 1 grey level apart on exactly the recorded number of values
    (torchvision's float32 blur rounding)."""
    import numpy as np
    from oracle import input_oracle as IO
    from ssl4polyp_b200.input_pipeline import TrainAugment
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "augment_golden.npz"))
    seeds = (0, 1, 4, 6, 8, 9)
    frames = [IO.synthetic_frame(sd, *(int(v) for v in g[f"frame_{sd}"])) for sd in seeds]
    tf = TrainAugment(224)
    torch.manual_seed(int(g["seed"][0]))
    out, u8 = tf(frames, want_u8=True)
    out, u8 = out.cpu().numpy(), u8.cpu().numpy()
    torch.manual_seed(int(g["seed"][0]))
    bad = 0
    for k, sd in enumerate(seeds):
        want_t, want_u8, p = IO.transform_finetune_train(frames[k])
        assert tf.last_augment[k] == p
        assert np.array_equal(u8[k], want_u8), sd
        assert np.array_equal(out[k].view(np.uint32), want_t.view(np.uint32)), sd
        d = np.abs(u8[k].astype(int) - g[f"u8_{sd}"].astype(int))
        assert d.max() <= 1
        bad += int((d > 0).sum())
    assert bad == int(g["blur_mismatch_values"][0])
    # feeds the classifier like a DataLoader batch would
    assert out.shape == (6, 3, 224, 224) and np.isfinite(out).all()
