"""Tensor-level wrappers over the C-ABI (include/ssl4p.h).  PyTorch here is plumbing only: it owns the device
memory and the stream; every computation below is one of our sm_100a kernels.  No fallbacks: a CPU tensor or a
missing library raises.
"""
from __future__ import annotations

import os

import torch

from . import _lib

GEMM_NT, GEMM_NN, GEMM_TN = 0, 1, 2
EPI_BIAS_BF16, EPI_BIAS_GELU_BF16, EPI_BIAS_RESIDUAL_F32, EPI_BIAS_F32, EPI_DGELU_BF16, EPI_PARTIAL_F32 = range(6)
MAX_SPLITS = 16

bf16, f32, i64 = torch.bfloat16, torch.float32, torch.int64


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def _req(t: torch.Tensor, dtype, name: str, contiguous: bool = True):
    if not t.is_cuda:
        raise RuntimeError(f"{name}: expected a CUDA tensor (ssl4polyp_b200 has no CPU path)")
    if t.dtype != dtype:
        raise RuntimeError(f"{name}: expected dtype {dtype}, got {t.dtype}")
    if contiguous and not t.is_contiguous():
        raise RuntimeError(f"{name}: expected a contiguous tensor")
    return t


_ws_cache: dict = {}
_ws_retired: list = []      # superseded scratch buffers stay allocated: a captured CUDA graph may hold their addresses


def workspace(nbytes: int, device, tag: str = "ws") -> torch.Tensor:
    """Grow-only scratch arena per (device, tag); kernels on one stream are serialised, so reuse is safe.
    A buffer that has to grow is replaced, never freed: `engine._replay` bakes workspace pointers into its CUDA graph,
    and a later, larger eager call must not hand that memory back to the caching allocator under the graph.  Growth is
    geometric (at least 2x), so the retired buffers add up to less than the live one."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), tag)
    t = _ws_cache.get(key)
    if t is None or t.numel() < nbytes:
        if t is not None:
            _ws_retired.append(t)
            nbytes = max(nbytes, 2 * t.numel())
        t = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _ws_cache[key] = t
    return t


# ------------------------------------------------------------------------------------------------------------
# GEMM
# ------------------------------------------------------------------------------------------------------------
GEMM_PROFILE = None      # bench.py sets this to a list to time every GEMM launch with CUDA events (roofline leg)


_sched_buffers: dict = {}


def _ensure_gemm_scheduler(lib) -> None:
    """Once per process: hand the library the zeroed device buffer its dynamic tile scheduler draws launch slots from
    (include/ssl4p.h: ssl4p_gemm_set_scheduler_buffer).  SSL4P_GEMM_SCHED=0 keeps the static tile lists."""
    dev = torch.cuda.current_device()
    if dev in _sched_buffers:
        return
    if torch.cuda.is_current_stream_capturing():      # never allocate the buffer from a graph's private pool: static tile lists
        return                                        # for this capture, try again on the next eager call
    if os.environ.get("SSL4P_GEMM_SCHED", "1") == "0" or _sched_buffers:      # the library keeps one buffer (one device per process)
        _sched_buffers[dev] = None
        return
    buf = torch.zeros(2 << 20, dtype=torch.uint8, device=torch.device("cuda", dev))
    _lib.check(lib.ssl4p_gemm_set_scheduler_buffer(buf.data_ptr(), buf.numel()), "gemm_set_scheduler_buffer")
    _sched_buffers[dev] = buf


def _gemm_launch(lib, M, N, K, layout, epilogue, *args):
    _ensure_gemm_scheduler(lib)
    if GEMM_PROFILE is None:
        return lib.ssl4p_gemm_bf16(*args)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    rc = lib.ssl4p_gemm_bf16(*args)
    e1.record()
    GEMM_PROFILE.append((e0, e1, 2.0 * M * N * K, (M, N, K, layout, epilogue)))
    return rc


def gemm(a, b, layout, epilogue, *, bias=None, aux=None, out=None, out2=None, splits=1, force_bn=0,
         accumulate_into=None, colsum_out=None, colsum_accumulate=False):
    """C = op(A) . op(B) with a fused epilogue; see include/ssl4p.h.  a, b: bf16 2-D (last dim contiguous).
    layout NT: a [M,K], b [N,K];  NN: a [M,K], b [K,N];  TN: a [K,M], b [K,N].
    For TN (wgrad) the result is fp32; `splits`>1 goes through a deterministic split-K reduce, optionally
    accumulating into `accumulate_into` (gradient accumulation)."""
    lib = _lib.load()
    _req(a, bf16, "gemm A", contiguous=False)
    _req(b, bf16, "gemm B", contiguous=False)
    assert a.dim() == 2 and b.dim() == 2 and a.stride(1) == 1 and b.stride(1) == 1
    if layout == GEMM_NT:
        M, K = a.shape; N, K2 = b.shape
    elif layout == GEMM_NN:
        M, K = a.shape; K2, N = b.shape
    else:
        K, M = a.shape; K2, N = b.shape
    assert K == K2, f"gemm: inner dimensions differ ({K} vs {K2})"
    dev = a.device
    out_dtype = f32 if epilogue in (EPI_BIAS_RESIDUAL_F32, EPI_BIAS_F32, EPI_PARTIAL_F32) else bf16
    if epilogue == EPI_PARTIAL_F32:
        final = accumulate_into if accumulate_into is not None else (out if out is not None else torch.empty(M, N, dtype=f32, device=dev))
        _req(final, f32, "gemm out")
        assert final.shape == (M, N)
        if splits == 1 and accumulate_into is None:
            _lib.check(_gemm_launch(lib, M, N, K, layout, epilogue, a.data_ptr(), b.data_ptr(), final.data_ptr(), None, None, None, M, N, K,
                                           a.stride(0), b.stride(0), N, 0, 0, layout, epilogue, 1, force_bn, _stream()), "gemm")
            return final
        ws = workspace(splits * M * N * 4, dev, "splitk")
        _lib.check(_gemm_launch(lib, M, N, K, layout, epilogue, a.data_ptr(), b.data_ptr(), ws.data_ptr(), None, None, None, M, N, K,
                                       a.stride(0), b.stride(0), N, 0, 0, layout, epilogue, splits, force_bn, _stream()), "gemm")
        _lib.check(lib.ssl4p_splitk_reduce(ws.data_ptr(), splits, M * N, final.data_ptr(), M * N,
                                           1 if accumulate_into is not None else 0, _stream()), "splitk_reduce")
        return final
    if out is None:
        out = torch.empty(M, N, dtype=out_dtype, device=dev)
    _req(out, out_dtype, "gemm out", contiguous=False)
    assert out.shape == (M, N) and out.stride(1) == 1
    if epilogue == EPI_BIAS_GELU_BF16:
        if out2 is None:
            out2 = torch.empty(M, N, dtype=bf16, device=dev)
        _req(out2, bf16, "gemm out2", contiguous=False)
    if bias is not None:
        _req(bias, f32, "gemm bias")
        assert bias.numel() == N
    ldaux = 0
    if aux is not None:
        _req(aux, f32 if epilogue == EPI_BIAS_RESIDUAL_F32 else bf16, "gemm aux", contiguous=False)
        assert aux.shape == (M, N) and aux.stride(1) == 1
        ldaux = aux.stride(0)
    part = None
    if colsum_out is not None:
        # EPI_DGELU_BF16 only: the epilogue also emits the column sums of C per 32-row group (Mlp.fc1 bias gradient)
        if epilogue != EPI_DGELU_BF16:
            raise RuntimeError("gemm: colsum_out needs the DGELU epilogue")
        _req(colsum_out, f32, "gemm colsum_out")
        assert colsum_out.numel() == N
        G = (M + 31) // 32
        part = workspace(G * N * 4, dev, "gemm_colsum").view(f32)
    _lib.check(_gemm_launch(lib, M, N, K, layout, epilogue, a.data_ptr(), b.data_ptr(), out.data_ptr(),
                            part.data_ptr() if part is not None else _ptr(out2), _ptr(bias), _ptr(aux), M, N, K,
                                   a.stride(0), b.stride(0), out.stride(0), out2.stride(0) if out2 is not None else 0, ldaux,
                                   layout, epilogue, 1, force_bn, _stream()), "gemm")
    if part is not None:
        ws2 = workspace(32 * N * 4, dev, "reduce_rows")
        _lib.check(lib.ssl4p_reduce_rows_f32(part.data_ptr(), (M + 31) // 32, N, colsum_out.data_ptr(), int(colsum_accumulate),
                                             ws2.data_ptr(), ws2.numel(), _stream()), "reduce_rows")
    return (out, out2) if epilogue == EPI_BIAS_GELU_BF16 else out


def pick_splits(M, N, K) -> int:
    return int(_lib.load().ssl4p_gemm_pick_splits(M, N, K))


def colsum(x, out=None, accumulate=False):
    """out[j] (+)= sum_m x[m, j]; x bf16 [M, N]."""
    lib = _lib.load()
    _req(x, bf16, "colsum x", contiguous=False)
    M, N = x.shape
    if out is None:
        out = torch.empty(N, dtype=f32, device=x.device)
        accumulate = False
    nb = lib.ssl4p_colsum_workspace_bytes(M, N)
    ws = workspace(nb, x.device, "colsum")
    _lib.check(lib.ssl4p_colsum_bf16(x.data_ptr(), M, N, x.stride(0), out.data_ptr(), int(accumulate), ws.data_ptr(), ws.numel(), _stream()), "colsum")
    return out


def reduce_rows_f32(x, out=None, accumulate=False):
    """out[j] (+)= sum_g x[g, j]; x fp32 [G, N] contiguous (fixed-order two-stage reduce: deterministic)."""
    lib = _lib.load()
    _req(x, f32, "reduce_rows x")
    G, N = x.shape
    if out is None:
        out = torch.empty(N, dtype=f32, device=x.device)
        accumulate = False
    _req(out, f32, "reduce_rows out")
    assert out.numel() == N
    ws = workspace(32 * N * 4, x.device, "reduce_rows")
    _lib.check(lib.ssl4p_reduce_rows_f32(x.data_ptr(), G, N, out.data_ptr(), int(accumulate), ws.data_ptr(), ws.numel(), _stream()),
               "reduce_rows")
    return out


def crop_resize_normalize(frames, desc, size, max_crop, mean, std, out=None, filter="bicubic"):
    """uint8 HWC frames (one packed device buffer) -> crop / PIL-bicubic resize / flip / to_tensor / normalize ->
    fp32 [N,3,size,size] (main_pretrain.py:156-160).  desc int64 [N,8] on the device: byte offset, row pitch, top,
    left, height, width, flip, 0.  max_crop = largest crop extent in the batch (host int).  filter: "bicubic"
    (RandomResizedCrop(interpolation=3)) or "bilinear" (T.Resize of classification/data/transforms.py:233-235)."""
    import ctypes
    lib = _lib.load()
    _req(frames, torch.uint8, "crop_resize frames")
    _req(desc, torch.int64, "crop_resize desc")
    if desc.dim() != 2 or desc.shape[1] != 8:
        raise RuntimeError("crop_resize: desc must be int64 [N, 8]")
    N = desc.shape[0]
    if out is None:
        out = torch.empty(N, 3, size, size, dtype=f32, device=frames.device)
    _req(out, f32, "crop_resize out")
    assert out.shape == (N, 3, size, size)
    if N == 0:
        return out
    nb = lib.ssl4p_crop_resize_workspace_bytes(N, size)
    ws = workspace(nb, frames.device, "crop_resize")
    m3 = (ctypes.c_float * 3)(*[float(v) for v in mean])
    s3 = (ctypes.c_float * 3)(*[float(v) for v in std])
    _lib.check(lib.ssl4p_crop_resize_normalize_u8(frames.data_ptr(), desc.data_ptr(), N, size, int(max_crop),
                                                  {"bicubic": 0, "bilinear": 1}[filter], ctypes.addressof(m3),
                                                  ctypes.addressof(s3), out.data_ptr(), ws.data_ptr(), ws.numel(), _stream()), "crop_resize")
    return out


def crop_resize_u8(frames, desc, size, max_crop, filter="bilinear", out=None):
    """The resample of `crop_resize_normalize` with the uint8 HWC image as the result: [N,size,size,3] (T.Resize of
    classification/data/transforms.py:233-235 when the box is the whole frame)."""
    lib = _lib.load()
    _req(frames, torch.uint8, "crop_resize frames")
    _req(desc, torch.int64, "crop_resize desc")
    if desc.dim() != 2 or desc.shape[1] != 8:
        raise RuntimeError("crop_resize: desc must be int64 [N, 8]")
    N = desc.shape[0]
    if out is None:
        out = torch.empty(N, size, size, 3, dtype=torch.uint8, device=frames.device)
    _req(out, torch.uint8, "crop_resize_u8 out")
    assert out.shape == (N, size, size, 3)
    if N == 0:
        return out
    ws = workspace(lib.ssl4p_crop_resize_workspace_bytes(N, size), frames.device, "crop_resize")
    _lib.check(lib.ssl4p_crop_resize_u8(frames.data_ptr(), desc.data_ptr(), N, size, int(max_crop), {"bicubic": 0, "bilinear": 1}[filter],
                                        out.data_ptr(), ws.data_ptr(), ws.numel(), _stream()), "crop_resize_u8")
    return out


AUG_PARAM_INTS, AUG_MAX_TAPS = 32, 32


def color_jitter_u8(img, params, out=None):
    """ColorJitter on uint8 [N,S,S,3] with host-drawn parameters (include/ssl4p.h: int32 [N,32]); one CTA per image."""
    lib = _lib.load()
    _req(img, torch.uint8, "color_jitter img")
    _req(params, torch.int32, "color_jitter params")
    N, S = img.shape[0], img.shape[1]
    assert img.shape == (N, S, S, 3) and params.shape == (N, AUG_PARAM_INTS)
    if out is None:
        out = torch.empty_like(img)
    _req(out, torch.uint8, "color_jitter out")
    _lib.check(lib.ssl4p_color_jitter_u8(img.data_ptr(), out.data_ptr(), params.data_ptr(), N, S, _stream()), "color_jitter")
    return out


def gaussian_blur_u8(img, taps, ntaps, out=None):
    """GaussianBlur on uint8 [N,S,S,3]; taps float32 [N,32] (the first `ntaps` are the 1-D kernel of each image)."""
    lib = _lib.load()
    _req(img, torch.uint8, "gaussian_blur img")
    _req(taps, f32, "gaussian_blur taps")
    N, S = img.shape[0], img.shape[1]
    assert img.shape == (N, S, S, 3) and taps.shape == (N, AUG_MAX_TAPS)
    if out is None:
        out = torch.empty_like(img)
    _req(out, torch.uint8, "gaussian_blur out")
    if out.data_ptr() == img.data_ptr():
        raise RuntimeError("gaussian_blur: cannot run in place")
    ws = workspace(lib.ssl4p_gaussian_blur_workspace_bytes(N, S), img.device, "gaussian_blur")
    _lib.check(lib.ssl4p_gaussian_blur_u8(img.data_ptr(), out.data_ptr(), taps.data_ptr(), int(ntaps), N, S,
                                          ws.data_ptr(), ws.numel(), _stream()), "gaussian_blur")
    return out


def flip_rotate_normalize_u8(img, params, mean, std, out=None, want_u8=False):
    """flips + Image.rotate(NEAREST) as one gather, then ToTensor + Normalize: fp32 [N,3,S,S] (and the uint8 image if asked)."""
    import ctypes
    lib = _lib.load()
    _req(img, torch.uint8, "flip_rotate img")
    _req(params, torch.int32, "flip_rotate params")
    N, S = img.shape[0], img.shape[1]
    assert img.shape == (N, S, S, 3) and params.shape == (N, AUG_PARAM_INTS)
    if out is None:
        out = torch.empty(N, 3, S, S, dtype=f32, device=img.device)
    _req(out, f32, "flip_rotate out")
    u8 = torch.empty_like(img) if want_u8 else None
    m3 = (ctypes.c_float * 3)(*[float(v) for v in mean])
    s3 = (ctypes.c_float * 3)(*[float(v) for v in std])
    _lib.check(lib.ssl4p_flip_rotate_normalize_u8(img.data_ptr(), params.data_ptr(), N, S, ctypes.addressof(m3), ctypes.addressof(s3),
                                                  out.data_ptr(), _ptr(u8), _stream()), "flip_rotate_normalize")
    return (out, u8) if want_u8 else out


def cast_bf16(src, dst=None):
    lib = _lib.load()
    _req(src, f32, "cast src")
    if dst is None:
        dst = torch.empty(src.shape, dtype=bf16, device=src.device)
    _req(dst, bf16, "cast dst")
    _lib.check(lib.ssl4p_cast_f32_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _stream()), "cast")
    return dst


# ------------------------------------------------------------------------------------------------------------
# masking / patches / tokens
# ------------------------------------------------------------------------------------------------------------
def mask_from_noise(noise, len_keep):
    """-> (ids_restore int64 [N,L], mask fp32 [N,L], ids_keep int64 [N,len_keep])"""
    lib = _lib.load()
    _req(noise, f32, "noise")
    N, L = noise.shape
    ids_restore = torch.empty(N, L, dtype=i64, device=noise.device)
    mask = torch.empty(N, L, dtype=f32, device=noise.device)
    ids_keep = torch.empty(N, len_keep, dtype=i64, device=noise.device)
    _lib.check(lib.ssl4p_mask_from_noise(noise.data_ptr(), N, L, len_keep, ids_restore.data_ptr(), mask.data_ptr(),
                                         ids_keep.data_ptr(), _stream()), "mask_from_noise")
    return ids_restore, mask, ids_keep


def gather_tokens(x, ids_keep):
    lib = _lib.load()
    _req(x, f32, "x"); _req(ids_keep, i64, "ids_keep")
    N, L, D = x.shape
    keep = ids_keep.shape[1]
    out = torch.empty(N, keep, D, dtype=f32, device=x.device)
    _lib.check(lib.ssl4p_gather_tokens_f32(x.data_ptr(), ids_keep.data_ptr(), out.data_ptr(), N, L, keep, D, _stream()), "gather_tokens")
    return out


def patchify(imgs, p):
    lib = _lib.load()
    _req(imgs, f32, "imgs")
    N, C, H, W = imgs.shape
    out = torch.empty(N, (H // p) * (W // p), p * p * C, dtype=f32, device=imgs.device)
    _lib.check(lib.ssl4p_patchify(imgs.data_ptr(), out.data_ptr(), N, C, H, W, p, _stream()), "patchify")
    return out


def unpatchify(x, p, C=3):
    lib = _lib.load()
    _req(x, f32, "x")
    N, L, P = x.shape
    h = int(L ** .5)
    assert h * h == L and P == p * p * C
    out = torch.empty(N, C, h * p, h * p, dtype=f32, device=x.device)
    _lib.check(lib.ssl4p_unpatchify(x.data_ptr(), out.data_ptr(), N, C, h * p, h * p, p, _stream()), "unpatchify")
    return out


def im2col_patches(imgs, ids_keep, p):
    lib = _lib.load()
    _req(imgs, f32, "imgs")
    N, C, H, W = imgs.shape
    keep = ids_keep.shape[1] if ids_keep is not None else (H // p) * (W // p)
    cols = torch.empty(N * keep, C * p * p, dtype=bf16, device=imgs.device)
    _lib.check(lib.ssl4p_im2col_patches(imgs.data_ptr(), _ptr(ids_keep), cols.data_ptr(), N, C, H, W, p, keep, _stream()), "im2col")
    return cols


def assemble_tokens(tok, ids_keep, pos, cls, N, keep, D):
    lib = _lib.load()
    x = torch.empty(N, keep + 1, D, dtype=f32, device=tok.device)
    _lib.check(lib.ssl4p_assemble_tokens(tok.data_ptr(), _ptr(ids_keep), pos.data_ptr(), cls.data_ptr(), x.data_ptr(), N, keep, D, _stream()), "assemble_tokens")
    return x


def assemble_tokens_bwd(dx, dcls, accumulate=False):
    lib = _lib.load()
    N, T, D = dx.shape
    dtok = torch.empty(N * (T - 1), D, dtype=bf16, device=dx.device)
    _lib.check(lib.ssl4p_assemble_tokens_bwd(dx.data_ptr(), dtok.data_ptr(), dcls.data_ptr(), N, T - 1, D, int(accumulate), _stream()), "assemble_tokens_bwd")
    return dtok


def decoder_unshuffle_fwd(y, mask_token, ids_restore, dpos):
    lib = _lib.load()
    _req(y, bf16, "y")
    N, Te, Dd = y.shape
    L = ids_restore.shape[1]
    x = torch.empty(N, 1 + L, Dd, dtype=f32, device=y.device)
    _lib.check(lib.ssl4p_decoder_unshuffle_fwd(y.data_ptr(), mask_token.data_ptr(), ids_restore.data_ptr(), dpos.data_ptr(), x.data_ptr(),
                                               N, Te, L, Dd, _stream()), "decoder_unshuffle_fwd")
    return x


def decoder_unshuffle_bwd(dx, ids_keep, mask, dmask_token, accumulate=False):
    lib = _lib.load()
    N, T, Dd = dx.shape
    L = T - 1
    Te = ids_keep.shape[1] + 1
    dy = torch.empty(N, Te, Dd, dtype=bf16, device=dx.device)
    nb = lib.ssl4p_decoder_unshuffle_bwd_workspace_bytes(N, L, Dd)
    ws = workspace(nb, dx.device, "unshuffle")
    _lib.check(lib.ssl4p_decoder_unshuffle_bwd(dx.data_ptr(), ids_keep.data_ptr(), mask.data_ptr(), dy.data_ptr(), dmask_token.data_ptr(),
                                               N, Te, L, Dd, int(accumulate), ws.data_ptr(), ws.numel(), _stream()), "decoder_unshuffle_bwd")
    return dy


# ------------------------------------------------------------------------------------------------------------
# LayerNorm / attention / loss / optimiser
# ------------------------------------------------------------------------------------------------------------
def layernorm_fwd(x, gamma, beta, eps):
    """x fp32 [M, D] -> (y bf16 [M, D], mean fp32 [M], rstd fp32 [M])"""
    lib = _lib.load()
    _req(x, f32, "ln x")
    M, D = x.shape
    y = torch.empty(M, D, dtype=bf16, device=x.device)
    mean = torch.empty(M, dtype=f32, device=x.device)
    rstd = torch.empty(M, dtype=f32, device=x.device)
    _lib.check(lib.ssl4p_layernorm_fwd(x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), float(eps), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                       M, D, _stream()), "layernorm_fwd")
    return y, mean, rstd


def layernorm_bwd(dy, x, mean, rstd, gamma, residual_grad, dgamma, dbeta, *, want_bf16=True, accumulate=False, dx_out=None,
                  dx_colsum=None, accumulate_colsum=False):
    """-> (dx fp32 [M,D], dxb bf16 [M,D] | None).  dx = residual_grad + LN'(dy); may alias residual_grad (in place).
    dx_colsum (fp32 [D], optional) (+)= column sums of dx (the upstream Linear's bias gradient)."""
    lib = _lib.load()
    _req(dy, bf16, "ln dy")
    M, D = dy.shape
    if dx_out is None:
        dx_out = residual_grad if residual_grad is not None else torch.empty(M, D, dtype=f32, device=dy.device)
    dxb = torch.empty(M, D, dtype=bf16, device=dy.device) if want_bf16 else None
    nb = lib.ssl4p_layernorm_bwd_workspace_bytes(M, D)
    ws = workspace(nb, dy.device, "ln")
    _lib.check(lib.ssl4p_layernorm_bwd(dy.data_ptr(), x.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(), _ptr(residual_grad),
                                       dx_out.data_ptr(), _ptr(dxb), dgamma.data_ptr(), dbeta.data_ptr(), _ptr(dx_colsum), M, D, int(accumulate),
                                       int(accumulate_colsum), ws.data_ptr(), ws.numel(), _stream()), "layernorm_bwd")
    return dx_out, dxb


def attention_fwd(qkv, B, T, H, hd, scale):
    """qkv bf16 [B*T, 3*H*hd] -> (out bf16 [B*T, H*hd], lse fp32 [B,H,T])"""
    lib = _lib.load()
    _req(qkv, bf16, "qkv")
    out = torch.empty(B * T, H * hd, dtype=bf16, device=qkv.device)
    lse = torch.empty(B, H, T, dtype=f32, device=qkv.device)
    _lib.check(lib.ssl4p_attention_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), B, T, H, hd, float(scale), _stream()), "attention_fwd")
    return out, lse


def attention_bwd(qkv, out, dout, lse, B, T, H, hd, scale):
    lib = _lib.load()
    _req(dout, bf16, "dout")
    dqkv = torch.empty_like(qkv)
    _lib.check(lib.ssl4p_attention_bwd(qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), dqkv.data_ptr(), B, T, H, hd, float(scale),
                                       _stream()), "attention_bwd")
    return dqkv


def masked_mse_fwd(imgs, pred_buf, has_cls_row, mask, norm_pix, p):
    """pred_buf fp32 [N, L(+1), P].  -> (loss fp32 0-dim, workspace tensor to hand to masked_mse_bwd)"""
    lib = _lib.load()
    _req(imgs, f32, "imgs"); _req(pred_buf, f32, "pred"); _req(mask, f32, "mask")
    N, C, H, W = imgs.shape
    L = (H // p) * (W // p)
    P = p * p * C
    rows = pred_buf.shape[1]
    assert rows == L + (1 if has_cls_row else 0) and pred_buf.shape[2] == P
    ws = torch.empty(lib.ssl4p_masked_mse_workspace_bytes(N, L), dtype=torch.uint8, device=imgs.device)
    loss = torch.empty((), dtype=f32, device=imgs.device)
    pptr = pred_buf.data_ptr() + (P * 4 if has_cls_row else 0)
    _lib.check(lib.ssl4p_masked_mse_fwd(imgs.data_ptr(), pptr, rows * P, mask.data_ptr(), int(norm_pix), loss.data_ptr(), N, C, H, W, p,
                                        ws.data_ptr(), ws.numel(), _stream()), "masked_mse_fwd")
    return loss, ws


def masked_mse_bwd(imgs, pred_buf, has_cls_row, mask, norm_pix, p, grad_out, ws):
    """-> dpred bf16 with the same [N, L(+1), P] shape as pred_buf (cls rows and unmasked rows are zero)."""
    lib = _lib.load()
    N, C, H, W = imgs.shape
    P = p * p * C
    rows = pred_buf.shape[1]
    _req(grad_out, f32, "grad_out")
    dpred = torch.empty(pred_buf.shape, dtype=bf16, device=imgs.device)
    pptr = pred_buf.data_ptr() + (P * 4 if has_cls_row else 0)
    _lib.check(lib.ssl4p_masked_mse_bwd(imgs.data_ptr(), pptr, rows * P, mask.data_ptr(), int(norm_pix), grad_out.data_ptr(), dpred.data_ptr(),
                                        rows * P, N, C, H, W, p, ws.data_ptr(), ws.numel(), _stream()), "masked_mse_bwd")
    return dpred



def _pred_layout(pred, L, P):
    """pred: fp32 [N, L, P] whose rows are contiguous and whose batch stride may exceed L*P (a [:, 1:, :] view)."""
    _req(pred, f32, "pred", contiguous=False)
    if pred.dim() != 3 or pred.shape[1] != L or pred.shape[2] != P or pred.stride(2) != 1 or pred.stride(1) != P:
        raise RuntimeError("pred must be [N, L, P] with contiguous rows (stride(1) == P, stride(2) == 1)")
    return pred.data_ptr(), pred.stride(0)


def masked_mse_fwd_strided(imgs, pred, mask, norm_pix, p):
    """forward_loss on a pred that is either contiguous [N,L,P] or the [:,1:,:] view of a [N,1+L,P] buffer."""
    lib = _lib.load()
    _req(imgs, f32, "imgs"); _req(mask, f32, "mask")
    N, C, H, W = imgs.shape
    L = (H // p) * (W // p)
    P = p * p * C
    pptr, bstride = _pred_layout(pred, L, P)
    ws = torch.empty(lib.ssl4p_masked_mse_workspace_bytes(N, L), dtype=torch.uint8, device=imgs.device)
    loss = torch.empty((), dtype=f32, device=imgs.device)
    _lib.check(lib.ssl4p_masked_mse_fwd(imgs.data_ptr(), pptr, bstride, mask.data_ptr(), int(norm_pix), loss.data_ptr(), N, C, H, W, p,
                                        ws.data_ptr(), ws.numel(), _stream()), "masked_mse_fwd")
    return loss, ws


def masked_mse_bwd_strided(imgs, pred, mask, norm_pix, p, grad_out, ws):
    """-> dpred bf16 contiguous [N, L, P]."""
    lib = _lib.load()
    N, C, H, W = imgs.shape
    L = (H // p) * (W // p)
    P = p * p * C
    pptr, bstride = _pred_layout(pred, L, P)
    _req(grad_out, f32, "grad_out")
    dpred = torch.empty(N, L, P, dtype=bf16, device=imgs.device)
    _lib.check(lib.ssl4p_masked_mse_bwd(imgs.data_ptr(), pptr, bstride, mask.data_ptr(), int(norm_pix), grad_out.data_ptr(), dpred.data_ptr(),
                                        L * P, N, C, H, W, p, ws.data_ptr(), ws.numel(), _stream()), "masked_mse_bwd")
    return dpred


def adamw_flat(p, g, m, v, p_bf16, wd_flags, lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0, nonfinite=None):
    lib = _lib.load()
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    _lib.check(lib.ssl4p_adamw_flat(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), _ptr(p_bf16), _ptr(wd_flags), p.numel(),
                                    float(lr), float(beta1), float(beta2), float(eps), float(weight_decay), float(bc1), float(bc2),
                                    float(grad_scale), _ptr(nonfinite), _stream()), "adamw_flat")


def adamw_flat_dev(p, g, m, v, p_bf16, wd_flags, hyper, beta1, beta2, eps, weight_decay, grad_scale=1.0, nonfinite=None):
    """AdamW with lr = hyper[0] and the step counter hyper[1] on the device (CUDA-graph replayable)."""
    lib = _lib.load()
    _req(hyper, f32, "hyper")
    _lib.check(lib.ssl4p_adamw_flat_dev(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), _ptr(p_bf16), _ptr(wd_flags), p.numel(),
                                        hyper.data_ptr(), float(beta1), float(beta2), float(eps), float(weight_decay),
                                        float(grad_scale), _ptr(nonfinite), _stream()), "adamw_flat_dev")


def adamw_flat_dev_range(p, g, m, v, p_bf16, wd_flags, start, end, hyper, beta1, beta2, eps, weight_decay, grad_scale=1.0, nonfinite=None):
    """AdamW over arena elements [start, end) (multiples of 256) with lr / step read from `hyper`; the step counter is NOT
    advanced (adamw_advance_step does that once per optimiser step)."""
    lib = _lib.load()
    assert start % 256 == 0 and end % 256 == 0 and start <= end
    n = end - start
    _lib.check(lib.ssl4p_adamw_flat_dev_range(p.data_ptr() + 4 * start, g.data_ptr() + 4 * start, m.data_ptr() + 4 * start,
                                              v.data_ptr() + 4 * start, None if p_bf16 is None else p_bf16.data_ptr() + 2 * start,
                                              None if wd_flags is None else wd_flags.data_ptr() + start // 256, n, hyper.data_ptr(),
                                              float(beta1), float(beta2), float(eps), float(weight_decay), float(grad_scale),
                                              _ptr(nonfinite), _stream()), "adamw_flat_dev_range")


def adamw_advance_step(hyper):
    _lib.check(_lib.load().ssl4p_adamw_advance_step(hyper.data_ptr(), _stream()), "adamw_advance_step")


def pool_head_fwd(xn, W, b, N, T, D, spatial):
    lib = _lib.load()
    ncls = W.shape[0]
    logits = torch.empty(N, ncls, dtype=f32, device=xn.device)
    _lib.check(lib.ssl4p_pool_head_fwd(xn.data_ptr(), W.data_ptr(), _ptr(b), logits.data_ptr(), N, T, D, ncls, int(spatial), _stream()), "pool_head_fwd")
    return logits


def pool_head_bwd(xn, W, dlogits, dW, db, N, T, D, spatial, want_dx=True, accumulate=False):
    lib = _lib.load()
    ncls = W.shape[0]
    dxn = torch.empty(N * T, D, dtype=bf16, device=xn.device) if want_dx else None
    _lib.check(lib.ssl4p_pool_head_bwd(xn.data_ptr(), W.data_ptr(), dlogits.data_ptr(), _ptr(dxn), _ptr(dW), _ptr(db), N, T, D, ncls,
                                       int(spatial), int(accumulate), _stream()), "pool_head_bwd")
    return dxn
