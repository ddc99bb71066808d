/* ssl4p.h — C-ABI of libssl4p.so: the sm_100a (B200) kernels behind the MAE ViT hot path of irconde/SSL4POLYP.
 *
 * The reference has no native/FFI layer (pure Python): its boundary for this path is the
 * nn.Module API of src/ssl4polyp/models/mae/models_mae.py and src/ssl4polyp/models/models.py.  Each entry point
 * below replaces the stock ATen/cuBLAS/cuDNN call sequence behind the cited reference lines; the Python side
 * (ssl4polyp_b200/ops.py) binds them with ctypes and ssl4polyp_b200/models_mae.py mirrors the module API.
 *
 * Conventions
 *   - plain C types only; every pointer is a DEVICE pointer borrowed for the duration of the enqueue;
 *     the library never allocates or frees device memory and never synchronises the device;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*; NULL = legacy default stream);
 *   - return 0 on success, non-zero on failure; ssl4p_last_error() returns a thread-local message;
 *   - "bf16" = __nv_bfloat16, row-major, innermost dimension contiguous unless a leading dimension is given.
 */
#ifndef SSL4P_H_
#define SSL4P_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSL4P_ABI_VERSION 1
int ssl4p_abi_version(void);
/* Leave n SMs out of every persistent grid (GEMM, attention) so that a concurrent collective (the NCCL gradient
 * all-reduce overlapped with backward, main_pretrain.py:212-214 DDP) owns them; returns the previous value. */
int ssl4p_reserve_sms(int n);
const char* ssl4p_last_error(void);
/* number of kernels this library has launched in the calling process (for bench.py's gpu_launches) */
unsigned long long ssl4p_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * GEMM on tcgen05 tensor cores (bf16 in, fp32 accumulate in TMEM), TMA-fed, fused epilogues.
 * Replaces nn.Linear / Conv2d(k=s=16) forward+backward: timm 0.4.12 Attention.qkv/proj, Mlp.fc1/fc2 as
 * instantiated at models_mae.py:39-41,53-55; decoder_embed :47,174; decoder_pred :58,191; PatchEmbed.proj :33,152.
 *   layout NT: C[M,N] = A[M,K] . B[N,K]^T      (y = x W^T, forward)
 *   layout NN: C[M,N] = A[M,K] . B[K,N]        (dx = dy W, dgrad)
 *   layout TN: C[M,N] = A[K,M]^T . B[K,N]      (dW = dy^T x, wgrad; K = tokens; deterministic split-K)
 * ------------------------------------------------------------------------------------------------ */
enum { SSL4P_GEMM_NT = 0, SSL4P_GEMM_NN = 1, SSL4P_GEMM_TN = 2 };
enum {
    SSL4P_EPI_BIAS_BF16 = 0,         /* C(bf16)  = acc + bias (bias may be NULL)                          */
    SSL4P_EPI_BIAS_GELU_BF16 = 1,    /* h = acc + bias ; C2(bf16) = gelu_erf(h) ; C(bf16) = gelu_erf'(h) (Mlp.fc1: the derivative
                                        is what backward needs; both come from one erf evaluation)         */
    SSL4P_EPI_BIAS_RESIDUAL_F32 = 2, /* C(fp32)  = acc + bias + aux(fp32)     (attn.proj / Mlp.fc2 + skip) */
    SSL4P_EPI_BIAS_F32 = 3,          /* C(fp32)  = acc + bias                                             */
    SSL4P_EPI_DGELU_BF16 = 4,        /* C(bf16)  = acc * aux(bf16), aux = the saved gelu_erf'(h)   (dgrad through GELU);
                                        optional C2 = fp32 [ceil(M/32), N]: column sums of C per 32-row group (the Mlp.fc1 bias
                                        gradient without a second pass over C; finish with ssl4p_reduce_rows_f32)  */
    SSL4P_EPI_PARTIAL_F32 = 5        /* C(fp32)[split] = acc  (split-K partial planes, M*ldc apart)       */
};
#define SSL4P_MAX_SPLITS 16
int ssl4p_gemm_bf16(const void* A, const void* B, void* C, void* C2, const float* bias, const void* aux,
                    int M, int N, int K, int lda, int ldb, int ldc, int ldc2, int ldaux,
                    int layout, int epilogue, int splits, int force_bn, void* stream);
/* force_bn: low 16 bits = N tile (0 = choose; 64, 128, 192, 256), bits 16..23 = CTA group (0 = choose, 1, 2).
 * CTA group 2 = `tcgen05.mma.cta_group::2`: the two SMs of a TPC (a cluster of two CTAs) share one 256 x BN tile, each
 * staging its 128 rows of A and half of B.  ssl4p_gemm_set_cta_group sets the process-wide policy (0 = per shape,
 * 1 = never, 2 = wherever M > 128) and returns the previous one; the environment variable SSL4P_GEMM_CG seeds it. */
int ssl4p_gemm_set_cta_group(int mode);
/* CTA pairs the device can keep resident at once (cudaOccupancyMaxActiveClusters, known after the first pair launch; 0 before) */
int ssl4p_gemm_max_cta_pairs(void);
/* Dynamic tile scheduling for the persistent GEMM grids: hand the library a ZERO-INITIALISED device buffer (32-byte aligned,
 * 32 bytes per launch slot; 2 MB is plenty) that stays alive as long as GEMMs are launched on the current device; NULL
 * switches back to static tile lists (the default).  With it a CTA takes its tiles off a per-launch counter, so a CTA that
 * is placed late - the SMs it needs are held by the NCCL kernels of the overlapped gradient all-reduce - finds the queue
 * drained and exits instead of running its whole share as a second wave.  Results do not depend on the setting. */
int ssl4p_gemm_set_scheduler_buffer(void* zeroed_device_buffer, size_t bytes);
int ssl4p_gemm_pick_splits(int M, int N, int K);
size_t ssl4p_gemm_workspace_bytes(int M, int N, int K, int layout, int epilogue);
/* out[j] (+)= sum_g part[g*N + j], g < G  (two fixed-order stages: deterministic); workspace >= 32*N floats */
int ssl4p_reduce_rows_f32(const float* part, int G, int N, float* out, int accumulate, void* workspace, size_t workspace_bytes, void* stream);
/* out[i] (+)= sum_s parts[s*stride + i]   (fixed order: deterministic) */
int ssl4p_splitk_reduce(const float* parts, int splits, long long stride, float* out, long long n, int accumulate, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Pretraining input transform (main_pretrain.py:156-160: RandomResizedCrop(224, scale=(0.2,1), bicubic) ->
 * RandomHorizontalFlip -> ToTensor -> Normalize, a torchvision Compose run per image on DataLoader workers).
 * Replaces the Pillow resample (libImaging/Resample.c, 8-bit path) + torchvision to_tensor/normalize with two
 * kernels, bit-exact in the uint8 resample and in the fp32 result.  The crop boxes and flips are drawn by the
 * caller (host RNG, torchvision's get_params sequence).
 *   frames  device uint8: RGB frames, HWC, packed anywhere in one buffer
 *   desc    device int64 [N,8] per image: byte offset of the frame in `frames`, row pitch in bytes, crop top, left,
 *           height, width (pixels, inside the frame), flip (0/1), 0
 *   out     device fp32 [N,3,S,S]
 *   max_crop  the largest crop height or width in the batch (sizes the filter windows: at most
 *             SSL4P_RESAMPLE_MAX_TAPS taps per output, i.e. crops up to ~15 S)
 *   mean3 / std3  HOST pointers to 3 floats
 * ------------------------------------------------------------------------------------------------ */
#define SSL4P_RESAMPLE_MAX_TAPS 64
#define SSL4P_RESAMPLE_BICUBIC 0   /* PIL Image.BICUBIC, a = -0.5, support 2: RandomResizedCrop(interpolation=3)              */
#define SSL4P_RESAMPLE_BILINEAR 1  /* PIL Image.BILINEAR, support 1: T.Resize((S,S)) of the fine-tune / eval transforms
                                      (classification/data/transforms.py:233-235; box = the whole frame, flip = 0)       */
size_t ssl4p_crop_resize_workspace_bytes(int N, int S);
int ssl4p_crop_resize_normalize_u8(const void* frames, const long long* desc, int N, int S, int max_crop, int filter,
                                   const float* mean3, const float* std3, float* out,
                                   void* workspace, size_t workspace_bytes, void* stream);

/* The same resample with the uint8 HWC image as the result (no flip, no normalisation): out_u8 [N,S,S,3].  Feeds the
 * fine-tune TRAIN augmentations below, which the reference applies to the PIL image between Resize and ToTensor. */
int ssl4p_crop_resize_u8(const void* frames, const long long* desc, int N, int S, int max_crop, int filter,
                         void* out_u8, void* workspace, size_t workspace_bytes, void* stream);

/* Host side of the upload: copy N rectangles of uint8 rows (rectangle i: rows[i] rows of row_bytes[i] bytes, row r at
 * src[i] + r * src_pitch[i]) to dst + dst_offset[i], rows packed back to back, with memcpy on `threads` host threads (byte-
 * balanced, cut at row boundaries).  dst is the caller's pinned staging buffer.  The reference does this work per sample on
 * DataLoader workers (classification/data/transforms.py:233-258; models/mae/main_pretrain.py:156-175); here it is one call per
 * batch, so a ctypes caller releases its interpreter lock once and draws the batch's random parameters meanwhile.  No CUDA. */
int ssl4p_host_pack_rects(const void* const* src, const long long* src_pitch, const long long* rows, const long long* row_bytes,
                          void* dst, const long long* dst_offset, int n, int threads);

/* ------------------------------------------------------------------------------------------------
 * Fine-tune TRAIN augmentations (classification/data/transforms.py:233-258, stage 'train': ColorJitter(0.4, 0.5, 0.25,
 * 0.01) -> GaussianBlur((25,25), sigma=(0.001,2)) -> RandomHorizontalFlip -> RandomVerticalFlip -> RandomRotation(180) ->
 * ToTensor -> Normalize, a torchvision Compose on PIL images run per image on DataLoader workers).  Replaces Pillow's
 * ImageEnhance / Image.blend (Blend.c), convert('L'/'HSV'/'RGB') (Convert.c), Image.rotate NEAREST (Geometry.c affine_fixed)
 * and torchvision's adjust_hue / gaussian_blur / to_tensor / normalize; bit-identical to the CPU restatement in
 * oracle/input_oracle.py (pinned against the installed libraries; the blur sums in float64, see there).
 * All random draws are made by the caller with the reference's own torch RNG calls and passed per image:
 *   params  device int32 [N, SSL4P_AUG_PARAM_INTS]: [0..3] the order of the four colour operations as drawn by
 *           ColorJitter.get_params (0 brightness, 1 contrast, 2 saturation, 3 hue), [4] hue shift =
 *           uint8(int32(hue_factor * 255)), [5] hflip, [6] vflip, [7] rotation mode (0 = affine with the 16.16 fixed-point
 *           coefficients [8..13] = a0..a5 of Geometry.c affine_fixed, 1 = identity, 2 = 180, 3 = 90 ccw, 4 = 270),
 *           [16..18] the float32 bit patterns of the brightness / contrast / saturation factors
 *   taps    device float32 [N, SSL4P_AUG_MAX_TAPS]: the 1-D Gaussian of torchvision's _get_gaussian_kernel1d (host-computed)
 * img / out: uint8 [N,S,S,3] (HWC); color_jitter and gaussian_blur may run in place (img == out) only for color_jitter.
 * ------------------------------------------------------------------------------------------------ */
#define SSL4P_AUG_PARAM_INTS 32
#define SSL4P_AUG_MAX_TAPS 32
int ssl4p_color_jitter_u8(const void* img, void* out, const int* params, int N, int S, void* stream);
size_t ssl4p_gaussian_blur_workspace_bytes(int N, int S);
int ssl4p_gaussian_blur_u8(const void* img, void* out, const float* taps, int ntaps, int N, int S,
                           void* workspace, size_t workspace_bytes, void* stream);
/* flips + rotation as one gather, then ToTensor + Normalize: out fp32 [N,3,S,S] and/or out_u8 [N,S,S,3] (either may be
 * NULL); mean3 / std3 HOST pointers to 3 floats */
int ssl4p_flip_rotate_normalize_u8(const void* img, const int* params, int N, int S, const float* mean3, const float* std3,
                                   float* out, void* out_u8, void* stream);

/* ------------------------------------------------------------------------------------------------
 * random_masking (models_mae.py:123-148) as a function of the noise the caller drew with torch.rand.
 *   ids_restore[n,i] = rank of noise[n,i] in its row (stable ascending: ties -> lower index first)  int64
 *   mask[n,i]        = ids_restore[n,i] >= len_keep ? 1 : 0                                          fp32
 *   ids_keep[n,r]    = i such that ids_restore[n,i] == r, r < len_keep (= ids_shuffle[:, :len_keep]) int64
 * ------------------------------------------------------------------------------------------------ */
int ssl4p_mask_from_noise(const float* noise, int N, int L, int len_keep,
                          int64_t* ids_restore, float* mask, int64_t* ids_keep, void* stream);
/* x_masked[n,r,:] = x[n, ids_keep[n,r], :]   (models_mae.py:139-140), fp32 */
int ssl4p_gather_tokens_f32(const float* x, const int64_t* ids_keep, float* out, int N, int L, int keep, int D, void* stream);

/* patchify / unpatchify (models_mae.py:95-121): imgs[n,c,h*p+pi,w*p+qi] <-> x[n,h*W+w,(pi*p+qi)*C+c], fp32 */
int ssl4p_patchify(const float* imgs, float* x, int N, int C, int H, int W, int p, void* stream);
int ssl4p_unpatchify(const float* x, float* imgs, int N, int C, int H, int W, int p, void* stream);

/* PatchEmbed operand (timm PatchEmbed = Conv2d(k=s=p), models_mae.py:33,152) restricted to the kept patches:
 *   cols[n*keep+r, (c*p+pi)*p+qi] = bf16(imgs[n,c,h*p+pi,w*p+qi]),  (h,w) = divmod(ids_keep[n,r], W/p)
 * ids_keep == NULL -> all patches in order (keep must equal (H/p)*(W/p)). */
int ssl4p_im2col_patches(const float* imgs, const int64_t* ids_keep, void* cols_bf16,
                         int N, int C, int H, int W, int p, int keep, void* stream);
/* x[n,0,:] = cls + pos[0];  x[n,1+r,:] = tok[n*keep+r,:] + pos[1+ids_keep[n,r]]   (models_mae.py:155,161-163) */
int ssl4p_assemble_tokens(const float* tok, const int64_t* ids_keep, const float* pos, const float* cls,
                          float* x, int N, int keep, int D, void* stream);
/* backward of the above: dtok(bf16)[n*keep+r,:] = dx[n,1+r,:];  dcls[:] (+)= sum_n dx[n,0,:] */
int ssl4p_assemble_tokens_bwd(const float* dx, void* dtok_bf16, float* dcls, int N, int keep, int D, int accumulate, void* stream);

/* LayerNorm over the last dim (nn.LayerNorm, eps per models_mae.py:227), fp32 in, bf16 out, fp32 statistics saved */
int ssl4p_layernorm_fwd(const float* x, const float* gamma, const float* beta, float eps,
                        void* y_bf16, float* mean, float* rstd, int M, int D, void* stream);
/* dx_out(fp32) = (residual_grad ? residual_grad : 0) + LN'(dy);  dxb(bf16) = bf16(dx_out) if non-NULL;
 * dgamma/dbeta (+)= column sums (two-stage, deterministic; workspace >= ssl4p_layernorm_bwd_workspace_bytes);
 * dx_colsum (optional, [D]) (+)= column sums of dx_out = the bias gradient of the Linear whose output (plus skip) this
 * LayerNorm consumed (attn.proj / mlp.fc2 of the timm Block), produced in the same pass. */
int ssl4p_layernorm_bwd(const void* dy_bf16, const float* x, const float* mean, const float* rstd, const float* gamma,
                        const float* residual_grad, float* dx_out, void* dxb_bf16, float* dgamma, float* dbeta,
                        float* dx_colsum, int M, int D, int accumulate, int accumulate_colsum,
                        void* workspace, size_t workspace_bytes, void* stream);
size_t ssl4p_layernorm_bwd_workspace_bytes(int M, int D);
/* Allocation-free variant of ssl4p_layernorm_bwd.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops).  Same arguments and results
 * minus the workspace: the [grid][3*D] partial sums live in a file-scope device array (3.47 MiB per GPU that loads the
 * library, held until exit).  Limit: D <= ssl4p_layernorm_bwd_max_static_d() (1024); larger D returns SSL4P_ERR_ARG.
 * Calls no cudaMalloc/cudaFree.  All calls share the array: calls that overlap on the GPU (different streams) race.
 * The first call on a device resolves the array's device address (cudaGetSymbolAddress, checked, cached per device):
 * make one call per device during set-up, before any stream capture. */
int ssl4p_layernorm_bwd_static(const void* dy_bf16, const float* x, const float* mean, const float* rstd, const float* gamma,
                               const float* residual_grad, float* dx_out, void* dxb_bf16, float* dgamma, float* dbeta,
                               float* dx_colsum, int M, int D, int accumulate, int accumulate_colsum, void* stream);
int ssl4p_layernorm_bwd_max_static_d(void);

/* Multi-head self-attention of timm 0.4.12 Attention.forward (softmax(q k^T * scale) v), fused, per (batch, head).
 * qkv: bf16 [B, T, 3, H, hd] (the qkv Linear output as-is); out: bf16 [B, T, H*hd]; lse: fp32 [B, H, T]. */
int ssl4p_attention_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int hd, float scale, void* stream);
int ssl4p_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                        int B, int T, int H, int hd, float scale, void* stream);
/* Decoder input assembly (models_mae.py:177-183):
 *   x[n,0,:]   = y[n,0,:] + dpos[0]
 *   x[n,1+i,:] = (ids_restore[n,i] < Te-1 ? y[n,1+ids_restore[n,i],:] : mask_token) + dpos[1+i]
 * y: bf16 [N,Te,Dd] (decoder_embed output), x: fp32 [N,1+L,Dd]. */
int ssl4p_decoder_unshuffle_fwd(const void* y_bf16, const float* mask_token, const int64_t* ids_restore, const float* dpos,
                                float* x, int N, int Te, int L, int Dd, void* stream);
/* dy(bf16)[n,0,:] = dx[n,0,:]; dy[n,1+r,:] = dx[n,1+ids_keep[n,r],:]; dmask_token (+)= sum over masked positions */
int ssl4p_decoder_unshuffle_bwd(const float* dx, const int64_t* ids_keep, const float* mask, void* dy_bf16, float* dmask_token,
                                int N, int Te, int L, int Dd, int accumulate, void* workspace, size_t workspace_bytes, void* stream);
size_t ssl4p_decoder_unshuffle_bwd_workspace_bytes(int N, int L, int Dd);

/* forward_loss (models_mae.py:198-214) without materialising patchify(imgs):
 *   target = patch pixels (optionally (t-mean)/sqrt(var_unbiased+1e-6)); loss = sum_masked mean_j (pred-target)^2 / sum(mask)
 * pred: fp32, row (n,i) at pred + (n*pred_batch_stride + i*P) (lets pred live inside a [N,1+L,P] buffer). */
int ssl4p_masked_mse_fwd(const float* imgs, const float* pred, long long pred_batch_stride, const float* mask, int norm_pix,
                         float* loss, int N, int C, int H, int W, int p, void* workspace, size_t workspace_bytes, void* stream);
/* dpred(bf16)[n,i,:] = grad_out * 2 * mask * (pred - target) / (P * sum(mask)); rows with mask 0 are written as 0.
 * grad_out: device scalar (fp32).  dpred row (n,i) at dpred + n*dpred_batch_stride + i*P. */
int ssl4p_masked_mse_bwd(const float* imgs, const float* pred, long long pred_batch_stride, const float* mask, int norm_pix,
                         const float* grad_out, void* dpred_bf16, long long dpred_batch_stride,
                         int N, int C, int H, int W, int p, void* workspace, size_t workspace_bytes, void* stream);
size_t ssl4p_masked_mse_workspace_bytes(int N, int L);

/* out[j] (+)= sum_m x[m,j]  (bias gradients), two-stage deterministic */
int ssl4p_colsum_bf16(const void* x_bf16, int M, int N, int ld, float* out, int accumulate,
                      void* workspace, size_t workspace_bytes, void* stream);
size_t ssl4p_colsum_workspace_bytes(int M, int N);

/* dst(bf16)[i] = src(fp32)[i] */
int ssl4p_cast_f32_bf16(const float* src, void* dst_bf16, long long n, void* stream);

/* Fused flat AdamW (torch.optim.AdamW semantics, main_pretrain.py:217-218) over a flat parameter arena:
 *   g = grad*grad_scale; p *= 1 - lr*wd[chunk]; m,v update; p -= lr/bc1 * m / (sqrt(v)/sqrt(bc2) + eps)
 * wd_flags: one byte per 256-element chunk (1 = apply weight decay).  Also refreshes the bf16 shadow weights and
 * counts non-finite gradients into *nonfinite (models/mae/util/misc.py:387-400 detect_grad_anomalies). */
int ssl4p_adamw_flat(float* p, const float* g, float* m, float* v, void* p_bf16, const uint8_t* wd_flags, long long n,
                     float lr, float beta1, float beta2, float eps, float weight_decay, float bias_corr1, float bias_corr2,
                     float grad_scale, int* nonfinite, void* stream);
/* Same update with the per-step scalars resident on the device, so that the whole training step can be replayed
 * from a CUDA graph: hyper[0] = lr (util/lr_sched.py:9-21 writes it before each replay), hyper[1] = number of
 * optimiser steps completed so far (the kernel uses t = hyper[1]+1 for the bias corrections and a trailing
 * one-thread kernel stores t back). */
int ssl4p_adamw_flat_dev(float* p, const float* g, float* m, float* v, void* p_bf16, const uint8_t* wd_flags, long long n,
                         float* hyper, float beta1, float beta2, float eps, float weight_decay,
                         float grad_scale, int* nonfinite, void* stream);
/* The same update over ONE RANGE of the arena (a gradient bucket: the pointers are the arena bases advanced by the bucket
 * offset, wd_flags by offset/256) WITHOUT touching the step counter, so the optimiser can run bucket by bucket behind the
 * backward / the bucket all-reduces (reference: DDP bucketed averaging, main_pretrain.py:212-214, then optimizer.step,
 * engine_pretrain.py:67-70).  Call ssl4p_adamw_advance_step once after the last range of a step. */
int ssl4p_adamw_flat_dev_range(float* p, const float* g, float* m, float* v, void* p_bf16, const uint8_t* wd_flags, long long n,
                               const float* hyper, float beta1, float beta2, float eps, float weight_decay,
                               float grad_scale, int* nonfinite, void* stream);
int ssl4p_adamw_advance_step(float* hyper, void* stream);

/* ViT_from_MAE head (models.py:211-222): logits[n,:] = W . pooled[n,:] + b with pooled = x[n,0,:] ("cls") or
 * mean_i x[n,1+i,:] ("spatial"); x is the fp32 output of the final LayerNorm kept as bf16 [N,T,D]. */
int ssl4p_pool_head_fwd(const void* xn_bf16, const float* W, const float* b, float* logits, int N, int T, int D, int ncls,
                        int spatial, void* stream);
int ssl4p_pool_head_bwd(const void* xn_bf16, const float* W, const float* dlogits, void* dxn_bf16, float* dW, float* db,
                        int N, int T, int D, int ncls, int spatial, int accumulate, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SSL4P_H_ */
