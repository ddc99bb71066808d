// Token-level data movement of the MAE path: random masking from noise, patchify/unpatchify, the kept-patch
// im2col operand of PatchEmbed, token assembly (+pos, +cls), decoder un-shuffle, bias-gradient column sums.
// All HBM-bound: coalesced, 16-byte vectorised, shared-memory staged where a transpose is involved.
// Reference: /root/reference/src/ssl4polyp/models/mae/models_mae.py:95-148 (patchify, unpatchify, random_masking),
// :152-163 (patch embed + pos + cls), :174-183 (decoder input assembly).
#include "common.cuh"
#include "ssl4p.h"

namespace ssl4p {

// ---------------------------------------------------------------------------------------------------------
// random_masking as a function of the noise: rank-by-counting == stable ascending argsort twice.
// One CTA per row; the row (L <= 4096 floats) sits in shared memory; thread i counts the keys ordered before it.
// ---------------------------------------------------------------------------------------------------------
__global__ void mask_from_noise_kernel(const float* __restrict__ noise, int L, int len_keep,
                                       int64_t* __restrict__ ids_restore, float* __restrict__ mask, int64_t* __restrict__ ids_keep) {
    extern __shared__ float s_noise[];
    const int n = blockIdx.x;
    const float* row = noise + (size_t)n * L;
    for (int i = threadIdx.x; i < L; i += blockDim.x) s_noise[i] = row[i];
    __syncthreads();
    for (int i = threadIdx.x; i < L; i += blockDim.x) {
        const float v = s_noise[i];
        int rank = 0;
        for (int j = 0; j < L; ++j) {
            const float u = s_noise[j];                       // broadcast read
            rank += (u < v) || (u == v && j < i);
        }
        ids_restore[(size_t)n * L + i] = rank;
        mask[(size_t)n * L + i] = rank >= len_keep ? 1.0f : 0.0f;
        if (rank < len_keep) ids_keep[(size_t)n * len_keep + rank] = i;
    }
}

__global__ void gather_tokens_kernel(const float* __restrict__ x, const int64_t* __restrict__ ids_keep, float* __restrict__ out,
                                     int L, int keep, int D4) {
    const int nr = blockIdx.x;                                 // n*keep + r
    const int n = nr / keep;
    const int64_t src = ids_keep[nr];
    const float4* s = reinterpret_cast<const float4*>(x) + ((size_t)n * L + src) * D4;
    float4* d = reinterpret_cast<float4*>(out) + (size_t)nr * D4;
    for (int i = threadIdx.x; i < D4; i += blockDim.x) d[i] = s[i];
}

// ---------------------------------------------------------------------------------------------------------
// patchify / unpatchify: one CTA per (image, patch-row); the C x p x W strip is staged in shared memory so that
// both the NCHW side and the [patch, (pi,qi,c)] side are accessed with unit stride.
// ---------------------------------------------------------------------------------------------------------
template <bool kUnpatchify>
__global__ void patchify_kernel(const float* __restrict__ src, float* __restrict__ dst, int C, int H, int W, int p) {
    extern __shared__ float s_strip[];                         // [C][p][W] with channel stride p*W+1
    const int hp = H / p, wp = W / p;
    const int n = blockIdx.x / hp, h = blockIdx.x % hp;
    const int cs = p * W + 1;
    const int P = p * p * C;
    const int strip = C * p * W;
    const float* img = (kUnpatchify ? dst : src) + ((size_t)n * C * H + (size_t)h * p) * W;   // + c*H*W + pi*W + col
    const float* tok = (kUnpatchify ? src : dst) + ((size_t)n * hp * wp + (size_t)h * wp) * P;
    if (!kUnpatchify) {
        for (int i = threadIdx.x; i < strip; i += blockDim.x) {
            const int c = i / (p * W), r = i % (p * W);
            s_strip[c * cs + r] = img[(size_t)c * H * W + r];
        }
        __syncthreads();
        float* out = const_cast<float*>(tok);
        for (int o = threadIdx.x; o < wp * P; o += blockDim.x) {
            const int w = o / P, rem = o % P;
            const int c = rem % C, qi = (rem / C) % p, pi = rem / (C * p);
            out[o] = s_strip[c * cs + pi * W + w * p + qi];
        }
    } else {
        for (int o = threadIdx.x; o < wp * P; o += blockDim.x) {
            const int w = o / P, rem = o % P;
            const int c = rem % C, qi = (rem / C) % p, pi = rem / (C * p);
            s_strip[c * cs + pi * W + w * p + qi] = tok[o];
        }
        __syncthreads();
        float* out = const_cast<float*>(img);
        for (int i = threadIdx.x; i < strip; i += blockDim.x) {
            const int c = i / (p * W), r = i % (p * W);
            out[(size_t)c * H * W + r] = s_strip[c * cs + r];
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// PatchEmbed operand for the kept patches: one CTA per kept patch, float4 reads (4 pixels of one patch row),
// 8-byte bf16 writes.  Column order (c, pi, qi) == Conv2d weight viewed [D, C*p*p].
// ---------------------------------------------------------------------------------------------------------
__global__ void im2col_patches_kernel(const float* __restrict__ imgs, const int64_t* __restrict__ ids_keep,
                                      __nv_bfloat16* __restrict__ cols, int C, int H, int W, int p, int keep) {
    const int nr = blockIdx.x;
    const int n = nr / keep;
    const int wp = W / p;
    const int pid = ids_keep ? (int)ids_keep[nr] : (nr % keep);
    const int h = pid / wp, w = pid % wp;
    const int P4 = C * p * p / 4;
    const float* base = imgs + (size_t)n * C * H * W + (size_t)(h * p) * W + w * p;
    uint2* out = reinterpret_cast<uint2*>(cols + (size_t)nr * C * p * p);
    for (int i = threadIdx.x; i < P4; i += blockDim.x) {
        const int e = i * 4;
        const int c = e / (p * p), pi = (e / p) % p, qi = e % p;
        const float4 v = *reinterpret_cast<const float4*>(base + (size_t)c * H * W + (size_t)pi * W + qi);
        uint2 o; o.x = pack_bf16(v.x, v.y); o.y = pack_bf16(v.z, v.w);
        out[i] = o;
    }
}

__global__ void assemble_tokens_kernel(const float* __restrict__ tok, const int64_t* __restrict__ ids_keep, const float* __restrict__ pos,
                                       const float* __restrict__ cls, float* __restrict__ x, int keep, int D4) {
    const int row = blockIdx.x;                                // n*(keep+1) + t
    const int n = row / (keep + 1), t = row % (keep + 1);
    float4* out = reinterpret_cast<float4*>(x) + (size_t)row * D4;
    const float4* a; const float4* b;
    if (t == 0) { a = reinterpret_cast<const float4*>(cls); b = reinterpret_cast<const float4*>(pos); }
    else {
        const int r = n * keep + t - 1;
        const int64_t pid = ids_keep ? ids_keep[r] : (int64_t)(t - 1);
        a = reinterpret_cast<const float4*>(tok) + (size_t)r * D4;
        b = reinterpret_cast<const float4*>(pos) + (size_t)(1 + pid) * D4;
    }
    for (int i = threadIdx.x; i < D4; i += blockDim.x) {
        const float4 u = a[i], v = b[i];
        out[i] = make_float4(u.x + v.x, u.y + v.y, u.z + v.z, u.w + v.w);
    }
}

__global__ void assemble_tokens_bwd_kernel(const float* __restrict__ dx, __nv_bfloat16* __restrict__ dtok, int keep, int D4) {
    const int r = blockIdx.x;                                  // n*keep + j
    const int n = r / keep, j = r % keep;
    const float4* s = reinterpret_cast<const float4*>(dx) + ((size_t)n * (keep + 1) + 1 + j) * D4;
    uint2* d = reinterpret_cast<uint2*>(dtok) + (size_t)r * D4;
    for (int i = threadIdx.x; i < D4; i += blockDim.x) {
        const float4 v = s[i];
        uint2 o; o.x = pack_bf16(v.x, v.y); o.y = pack_bf16(v.z, v.w);
        d[i] = o;
    }
}
// dcls[d] (+)= sum_n dx[n,0,d]   (fixed order)
__global__ void cls_grad_kernel(const float* __restrict__ dx, float* __restrict__ dcls, int N, int T, int D, int accumulate) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= D) return;
    float acc = accumulate ? dcls[d] : 0.f;
    for (int n = 0; n < N; ++n) acc += dx[(size_t)n * T * D + d];
    dcls[d] = acc;
}

// ---------------------------------------------------------------------------------------------------------
// decoder input assembly and its backward
// ---------------------------------------------------------------------------------------------------------
__global__ void decoder_unshuffle_fwd_kernel(const __nv_bfloat16* __restrict__ y, const float* __restrict__ mask_token,
                                             const int64_t* __restrict__ ids_restore, const float* __restrict__ dpos,
                                             float* __restrict__ x, int Te, int L, int D4) {
    const int row = blockIdx.x;                                // n*(1+L) + t
    const int n = row / (1 + L), t = row % (1 + L);
    float4* out = reinterpret_cast<float4*>(x) + (size_t)row * D4;
    const float4* pe = reinterpret_cast<const float4*>(dpos) + (size_t)t * D4;
    int src = 0;                                               // row of y, or -1 for the mask token
    if (t > 0) {
        const int64_t r = ids_restore[(size_t)n * L + t - 1];
        src = r < Te - 1 ? (int)r + 1 : -1;
    }
    if (src >= 0) {
        const uint2* s = reinterpret_cast<const uint2*>(y) + ((size_t)n * Te + src) * D4;
        for (int i = threadIdx.x; i < D4; i += blockDim.x) {
            const uint2 u = s[i]; const float4 v = pe[i];
            const float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y);
            out[i] = make_float4(a.x + v.x, a.y + v.y, b.x + v.z, b.y + v.w);
        }
    } else {
        const float4* s = reinterpret_cast<const float4*>(mask_token);
        for (int i = threadIdx.x; i < D4; i += blockDim.x) {
            const float4 u = s[i], v = pe[i];
            out[i] = make_float4(u.x + v.x, u.y + v.y, u.z + v.z, u.w + v.w);
        }
    }
}

__global__ void decoder_unshuffle_bwd_kernel(const float* __restrict__ dx, const int64_t* __restrict__ ids_keep,
                                             __nv_bfloat16* __restrict__ dy, int Te, int L, int D4) {
    const int row = blockIdx.x;                                // n*Te + t
    const int n = row / Te, t = row % Te;
    const int src = t == 0 ? 0 : 1 + (int)ids_keep[(size_t)n * (Te - 1) + t - 1];
    const float4* s = reinterpret_cast<const float4*>(dx) + ((size_t)n * (1 + L) + src) * D4;
    uint2* d = reinterpret_cast<uint2*>(dy) + (size_t)row * D4;
    for (int i = threadIdx.x; i < D4; i += blockDim.x) {
        const float4 v = s[i];
        uint2 o; o.x = pack_bf16(v.x, v.y); o.y = pack_bf16(v.z, v.w);
        d[i] = o;
    }
}
// stage 1 of dmask_token: block g sums the masked rows g, g+G, ... ; thread = column
__global__ void mask_token_grad_partial_kernel(const float* __restrict__ dx, const float* __restrict__ mask, float* __restrict__ part,
                                               int NL, int L, int D) {
    const int d = blockIdx.y * blockDim.x + threadIdx.x;
    if (d >= D) return;
    // four independent accumulators (rows r, r+G, r+2G, r+3G of a round): the loads of a round are in flight together
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const int G = gridDim.x;
    for (int r0 = blockIdx.x; r0 < NL; r0 += 4 * G) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int r = r0 + k * G;
            if (r < NL && mask[r] != 0.f) {
                const int n = r / L, i = r % L;
                acc[k] += dx[((size_t)n * (1 + L) + 1 + i) * D + d];
            }
        }
    }
    part[(size_t)blockIdx.x * D + d] = (acc[0] + acc[1]) + (acc[2] + acc[3]);
}
__global__ void reduce_rows_kernel(const float* __restrict__ part, int G, int D, float* __restrict__ out, int accumulate) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= D) return;
    // four interleaved partial sums, eight loads in flight (the grid is a handful of CTAs: a serial chain of G dependent
    // L2 round trips took 10 us at G = 198); the order of the additions is fixed, so the result is deterministic
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    int g = 0;
    for (; g + 7 < G; g += 8) {
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = part[(size_t)(g + k) * D + d];
        a0 += v[0]; a1 += v[1]; a2 += v[2]; a3 += v[3];
        a0 += v[4]; a1 += v[5]; a2 += v[6]; a3 += v[7];
    }
    for (; g < G; ++g) a0 += part[(size_t)g * D + d];
    const float acc = (a0 + a1) + (a2 + a3);
    out[d] = accumulate ? out[d] + acc : acc;
}

// stage 1 of a tall reduce: block (128 columns) x slab of rows -> part2[slab][D]
__global__ void reduce_rows_slab_kernel(const float* __restrict__ part, int G, int D, int rows_per_slab, float* __restrict__ part2) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= D) return;
    const int g0 = blockIdx.y * rows_per_slab, g1 = min(G, g0 + rows_per_slab);
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    int g = g0;
    for (; g + 3 < g1; g += 4) {
        a0 += part[(size_t)g * D + d]; a1 += part[(size_t)(g + 1) * D + d];
        a2 += part[(size_t)(g + 2) * D + d]; a3 += part[(size_t)(g + 3) * D + d];
    }
    for (; g < g1; ++g) a0 += part[(size_t)g * D + d];
    part2[(size_t)blockIdx.y * D + d] = (a0 + a1) + (a2 + a3);
}

// ---------------------------------------------------------------------------------------------------------
// column sums of a bf16 matrix (bias gradients): stage 1 = slabs of rows, stage 2 = fixed-order reduce
// block (32, 8): lane -> 2 adjacent columns, ty -> row phase
// ---------------------------------------------------------------------------------------------------------
__global__ void colsum_partial_kernel(const __nv_bfloat16* __restrict__ x, int M, int N, int ld, float* __restrict__ part, int rows_per_slab) {
    __shared__ float2 s_acc[8][32];
    const int col = blockIdx.x * 64 + threadIdx.x * 2;
    const int r0 = blockIdx.y * rows_per_slab;
    const int r1 = min(M, r0 + rows_per_slab);
    float2 acc = make_float2(0.f, 0.f);
    if (col < N) {
        for (int r = r0 + threadIdx.y; r < r1; r += 8) {
            const uint32_t u = *reinterpret_cast<const uint32_t*>(x + (size_t)r * ld + col);
            const float2 f = unpack_bf16(u);
            acc.x += f.x; acc.y += f.y;
        }
    }
    s_acc[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && col < N) {
        float2 t = s_acc[0][threadIdx.x];
        for (int k = 1; k < 8; ++k) { t.x += s_acc[k][threadIdx.x].x; t.y += s_acc[k][threadIdx.x].y; }
        part[(size_t)blockIdx.y * N + col] = t.x;
        part[(size_t)blockIdx.y * N + col + 1] = t.y;
    }
}

// ---------------------------------------------------------------------------------------------------------
// ViT_from_MAE pooled head (models.py:211-222): tiny (N x D) . (D x ncls); one warp per (sample, class)
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float pooled_value(const __nv_bfloat16* xn, int n, int T, int D, int d, int spatial) {
    if (!spatial) return __bfloat162float(xn[(size_t)n * T * D + d]);
    float s = 0.f;
    for (int t = 1; t < T; ++t) s += __bfloat162float(xn[((size_t)n * T + t) * D + d]);
    return s / (float)(T - 1);
}
__global__ void pool_head_fwd_kernel(const __nv_bfloat16* __restrict__ xn, const float* __restrict__ W, const float* __restrict__ b,
                                     float* __restrict__ logits, int N, int T, int D, int ncls, int spatial) {
    const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (wid >= N * ncls) return;
    const int n = wid / ncls, c = wid % ncls;
    float acc = 0.f;
    for (int d = lane_id(); d < D; d += 32) acc += pooled_value(xn, n, T, D, d, spatial) * W[(size_t)c * D + d];
    acc = warp_sum(acc);
    if (lane_id() == 0) logits[wid] = acc + (b ? b[c] : 0.f);
}
// dxn[n,t,d]: cls -> row 0 gets sum_c dl[n,c] W[c,d], other rows 0; spatial -> rows 1.. get that / (T-1), row 0 gets 0
__global__ void pool_head_bwd_dx_kernel(const float* __restrict__ W, const float* __restrict__ dlogits, __nv_bfloat16* __restrict__ dxn,
                                        int T, int D, int ncls, int spatial) {
    const int row = blockIdx.x;                                // n*T + t
    const int n = row / T, t = row % T;
    const bool live = spatial ? (t > 0) : (t == 0);
    const float sc = spatial ? 1.f / (float)(T - 1) : 1.f;
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float g = 0.f;
        if (live) { for (int c = 0; c < ncls; ++c) g += dlogits[n * ncls + c] * W[(size_t)c * D + d]; g *= sc; }
        dxn[(size_t)row * D + d] = __float2bfloat16(g);
    }
}
__global__ void pool_head_bwd_dw_kernel(const __nv_bfloat16* __restrict__ xn, const float* __restrict__ dlogits, float* __restrict__ dW,
                                        float* __restrict__ db, int N, int T, int D, int ncls, int spatial, int accumulate) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;     // c*D + d
    if (idx < ncls * D) {
        const int c = idx / D, d = idx % D;
        float acc = accumulate ? dW[idx] : 0.f;
        for (int n = 0; n < N; ++n) acc += dlogits[n * ncls + c] * pooled_value(xn, n, T, D, d, spatial);
        dW[idx] = acc;
    }
    if (db && idx < ncls) {
        float acc = accumulate ? db[idx] : 0.f;
        for (int n = 0; n < N; ++n) acc += dlogits[n * ncls + idx];
        db[idx] = acc;
    }
}

// 8 columns (one 16-byte load) per lane, 4 rows in flight per thread: 64 bytes outstanding per thread instead of 4, which is what
// it takes to stream at HBM speed (the 2-column kernel above reached 3.5 TB/s, ncu r02d).  N and ld multiples of 8 only.
__global__ void __launch_bounds__(256)
colsum_partial_v8_kernel(const __nv_bfloat16* __restrict__ x, int M, int N, int ld, float* __restrict__ part, int rows_per_slab) {
    __shared__ float s_acc[8][32][9];
    const int col = blockIdx.x * 256 + threadIdx.x * 8;
    const int r0 = blockIdx.y * rows_per_slab;
    const int r1 = min(M, r0 + rows_per_slab);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (col < N) {
        const __nv_bfloat16* base = x + col;
        int r = r0 + threadIdx.y;
        for (; r + 24 < r1; r += 32) {
            uint4 u[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) u[k] = ldg_v4(base + (size_t)(r + 8 * k) * ld);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t w[4] = {u[k].x, u[k].y, u[k].z, u[k].w};
#pragma unroll
                for (int q = 0; q < 4; ++q) { const float2 f = unpack_bf16(w[q]); acc[2 * q] += f.x; acc[2 * q + 1] += f.y; }
            }
        }
        for (; r < r1; r += 8) {
            const uint4 u = ldg_v4(base + (size_t)r * ld);
            const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) { const float2 f = unpack_bf16(w[q]); acc[2 * q] += f.x; acc[2 * q + 1] += f.y; }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) s_acc[threadIdx.y][threadIdx.x][j] = acc[j];
    __syncthreads();
    // fixed-order sum over the 8 row phases: thread (ty, tx) finishes column tx * 8 + ty
    const int c = blockIdx.x * 256 + threadIdx.x * 8 + threadIdx.y;
    if (c < N) {
        float t = s_acc[0][threadIdx.x][threadIdx.y];
#pragma unroll
        for (int k = 1; k < 8; ++k) t += s_acc[k][threadIdx.x][threadIdx.y];
        part[(size_t)blockIdx.y * N + c] = t;
    }
}

static int colsum_slabs_v8(int M, int N) {
    const int colblocks = (N + 255) / 256;
    int s = (4 * num_sms() + colblocks - 1) / colblocks;
    const int max_s = (M + 63) / 64;
    if (s > max_s) s = max_s;
    if (s > 128) s = 128;
    if (s < 1) s = 1;
    return s;
}

static int colsum_slabs(int M, int N) {
    const int colblocks = (N + 63) / 64;
    int s = (4 * num_sms() + colblocks - 1) / colblocks;
    const int max_s = (M + 63) / 64;
    if (s > max_s) s = max_s;
    if (s > 256) s = 256;
    if (s < 1) s = 1;
    return s;
}

}  // namespace ssl4p

using namespace ssl4p;
#define STREAM ((cudaStream_t)stream)

extern "C" int ssl4p_mask_from_noise(const float* noise, int N, int L, int len_keep, int64_t* ids_restore, float* mask, int64_t* ids_keep, void* stream) {
    SSL4P_REQUIRE(noise && ids_restore && mask, "mask_from_noise: null pointer");
    SSL4P_REQUIRE(N >= 0 && L > 0 && L <= 8192, "mask_from_noise: unsupported L=%d", L);
    SSL4P_REQUIRE(len_keep >= 0 && len_keep <= L && (len_keep == 0 || ids_keep), "mask_from_noise: bad len_keep=%d", len_keep);
    if (N == 0) return SSL4P_OK;
    const int threads = L >= 256 ? 256 : ((L + 31) / 32) * 32;
    mask_from_noise_kernel<<<N, threads, L * sizeof(float), STREAM>>>(noise, L, len_keep, ids_restore, mask, ids_keep);
    return check_launch("mask_from_noise_kernel");
}

extern "C" int ssl4p_gather_tokens_f32(const float* x, const int64_t* ids_keep, float* out, int N, int L, int keep, int D, void* stream) {
    SSL4P_REQUIRE(x && ids_keep && out && D % 4 == 0, "gather_tokens: bad arguments (D %% 4 must be 0)");
    if (N * keep == 0) return SSL4P_OK;
    gather_tokens_kernel<<<N * keep, 128, 0, STREAM>>>(x, ids_keep, out, L, keep, D / 4);
    return check_launch("gather_tokens_kernel");
}

static int patchify_common(bool un, const float* src, float* dst, int N, int C, int H, int W, int p, cudaStream_t s) {
    SSL4P_REQUIRE(src && dst && p > 0 && H % p == 0 && W % p == 0 && H == W, "patchify: bad image geometry %dx%d p=%d", H, W, p);
    const size_t smem = (size_t)C * (p * W + 1) * sizeof(float);
    SSL4P_REQUIRE(smem <= 200 * 1024, "patchify: strip of %zu bytes exceeds shared memory", smem);
    if (N == 0) return SSL4P_OK;
    if (un) {
        if (smem > 48 * 1024) SSL4P_CUDA(cudaFuncSetAttribute(patchify_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        patchify_kernel<true><<<N * (H / p), 256, smem, s>>>(src, dst, C, H, W, p);
    } else {
        if (smem > 48 * 1024) SSL4P_CUDA(cudaFuncSetAttribute(patchify_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        patchify_kernel<false><<<N * (H / p), 256, smem, s>>>(src, dst, C, H, W, p);
    }
    return check_launch("patchify_kernel");
}
extern "C" int ssl4p_patchify(const float* imgs, float* x, int N, int C, int H, int W, int p, void* stream) {
    return patchify_common(false, imgs, x, N, C, H, W, p, STREAM);
}
extern "C" int ssl4p_unpatchify(const float* x, float* imgs, int N, int C, int H, int W, int p, void* stream) {
    return patchify_common(true, x, imgs, N, C, H, W, p, STREAM);
}

extern "C" int ssl4p_im2col_patches(const float* imgs, const int64_t* ids_keep, void* cols, int N, int C, int H, int W, int p, int keep, void* stream) {
    SSL4P_REQUIRE(imgs && cols && p % 4 == 0 && H % p == 0 && W % p == 0 && W % 4 == 0, "im2col: bad geometry");
    SSL4P_REQUIRE(ids_keep || keep == (H / p) * (W / p), "im2col: ids_keep == NULL needs keep == number of patches");
    if (N * keep == 0) return SSL4P_OK;
    im2col_patches_kernel<<<N * keep, 192, 0, STREAM>>>(imgs, ids_keep, (__nv_bfloat16*)cols, C, H, W, p, keep);
    return check_launch("im2col_patches_kernel");
}

extern "C" int ssl4p_assemble_tokens(const float* tok, const int64_t* ids_keep, const float* pos, const float* cls, float* x, int N, int keep, int D, void* stream) {
    SSL4P_REQUIRE(tok && pos && cls && x && D % 4 == 0, "assemble_tokens: bad arguments");
    if (N == 0) return SSL4P_OK;
    assemble_tokens_kernel<<<N * (keep + 1), 128, 0, STREAM>>>(tok, ids_keep, pos, cls, x, keep, D / 4);
    return check_launch("assemble_tokens_kernel");
}

extern "C" int ssl4p_assemble_tokens_bwd(const float* dx, void* dtok, float* dcls, int N, int keep, int D, int accumulate, void* stream) {
    SSL4P_REQUIRE(dx && dtok && dcls && D % 4 == 0, "assemble_tokens_bwd: bad arguments");
    if (N == 0) return SSL4P_OK;
    assemble_tokens_bwd_kernel<<<N * keep, 128, 0, STREAM>>>(dx, (__nv_bfloat16*)dtok, keep, D / 4);
    int rc = check_launch("assemble_tokens_bwd_kernel");
    if (rc) return rc;
    cls_grad_kernel<<<(D + 127) / 128, 128, 0, STREAM>>>(dx, dcls, N, keep + 1, D, accumulate);
    return check_launch("cls_grad_kernel");
}

extern "C" int ssl4p_decoder_unshuffle_fwd(const void* y, const float* mask_token, const int64_t* ids_restore, const float* dpos, float* x,
                                           int N, int Te, int L, int Dd, void* stream) {
    SSL4P_REQUIRE(y && mask_token && ids_restore && dpos && x && Dd % 4 == 0, "decoder_unshuffle_fwd: bad arguments");
    if (N == 0) return SSL4P_OK;
    decoder_unshuffle_fwd_kernel<<<N * (1 + L), 128, 0, STREAM>>>((const __nv_bfloat16*)y, mask_token, ids_restore, dpos, x, Te, L, Dd / 4);
    return check_launch("decoder_unshuffle_fwd_kernel");
}

constexpr int kMaskTokenGroups = 1184;         // stage-1 blocks of dmask_token (8 per SM: 42 rows each at B = 256)
extern "C" size_t ssl4p_decoder_unshuffle_bwd_workspace_bytes(int N, int L, int Dd) {
    (void)N; (void)L;
    return (size_t)(kMaskTokenGroups + 32) * Dd * sizeof(float);
}
extern "C" int ssl4p_decoder_unshuffle_bwd(const float* dx, const int64_t* ids_keep, const float* mask, void* dy, float* dmask_token,
                                           int N, int Te, int L, int Dd, int accumulate, void* workspace, size_t workspace_bytes, void* stream) {
    SSL4P_REQUIRE(dx && ids_keep && mask && dy && dmask_token && Dd % 4 == 0, "decoder_unshuffle_bwd: bad arguments");
    SSL4P_REQUIRE(workspace && workspace_bytes >= ssl4p_decoder_unshuffle_bwd_workspace_bytes(N, L, Dd), "decoder_unshuffle_bwd: workspace too small");
    if (N == 0) return SSL4P_OK;
    decoder_unshuffle_bwd_kernel<<<N * Te, 128, 0, STREAM>>>(dx, ids_keep, (__nv_bfloat16*)dy, Te, L, Dd / 4);
    int rc = check_launch("decoder_unshuffle_bwd_kernel");
    if (rc) return rc;
    const int G = kMaskTokenGroups;
    dim3 grid(G, (Dd + 127) / 128);
    float* part = (float*)workspace;
    mask_token_grad_partial_kernel<<<grid, 128, 0, STREAM>>>(dx, mask, part, N * L, L, Dd);
    rc = check_launch("mask_token_grad_partial_kernel");
    if (rc) return rc;
    const int rows_per_slab = (G + 31) / 32;                   // two fixed-order stages: 1184 -> 32 -> 1 rows
    const int S = (G + rows_per_slab - 1) / rows_per_slab;
    float* part2 = part + (size_t)G * Dd;
    reduce_rows_slab_kernel<<<dim3((Dd + 127) / 128, S), 128, 0, STREAM>>>(part, G, Dd, rows_per_slab, part2);
    rc = check_launch("reduce_rows_slab_kernel");
    if (rc) return rc;
    reduce_rows_kernel<<<(Dd + 127) / 128, 128, 0, STREAM>>>(part2, S, Dd, dmask_token, accumulate);
    return check_launch("reduce_rows_kernel");
}

extern "C" int ssl4p_reduce_rows_f32(const float* part, int G, int N, float* out, int accumulate, void* workspace, size_t workspace_bytes, void* stream) {
    SSL4P_REQUIRE(part && out && G > 0 && N > 0, "reduce_rows: bad arguments");
    if (G <= 32) {
        reduce_rows_kernel<<<(N + 127) / 128, 128, 0, STREAM>>>(part, G, N, out, accumulate);
        return check_launch("reduce_rows_kernel");
    }
    SSL4P_REQUIRE(workspace && workspace_bytes >= (size_t)32 * N * sizeof(float), "reduce_rows: workspace too small");
    const int rows_per_slab = (G + 31) / 32;
    const int S = (G + rows_per_slab - 1) / rows_per_slab;
    reduce_rows_slab_kernel<<<dim3((N + 127) / 128, S), 128, 0, STREAM>>>(part, G, N, rows_per_slab, (float*)workspace);
    int rc = check_launch("reduce_rows_slab_kernel");
    if (rc) return rc;
    reduce_rows_kernel<<<(N + 127) / 128, 128, 0, STREAM>>>((const float*)workspace, S, N, out, accumulate);
    return check_launch("reduce_rows_kernel");
}

extern "C" size_t ssl4p_colsum_workspace_bytes(int M, int N) { return (size_t)256 * (size_t)N * sizeof(float); (void)M; }
extern "C" int ssl4p_colsum_bf16(const void* x, int M, int N, int ld, float* out, int accumulate, void* workspace, size_t workspace_bytes, void* stream) {
    SSL4P_REQUIRE(x && out && M > 0 && N > 0 && N % 2 == 0 && ld % 2 == 0, "colsum: bad arguments");
    SSL4P_REQUIRE(workspace && workspace_bytes >= ssl4p_colsum_workspace_bytes(M, N), "colsum: workspace too small");
    const bool v8 = N % 8 == 0 && ld % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
    const int S = v8 ? colsum_slabs_v8(M, N) : colsum_slabs(M, N);
    const int rows_per_slab = (M + S - 1) / S;
    dim3 block(32, 8);
    if (v8) colsum_partial_v8_kernel<<<dim3((N + 255) / 256, S), block, 0, STREAM>>>((const __nv_bfloat16*)x, M, N, ld, (float*)workspace, rows_per_slab);
    else colsum_partial_kernel<<<dim3((N + 63) / 64, S), block, 0, STREAM>>>((const __nv_bfloat16*)x, M, N, ld, (float*)workspace, rows_per_slab);
    int rc = check_launch("colsum_partial_kernel");
    if (rc) return rc;
    reduce_rows_kernel<<<(N + 127) / 128, 128, 0, STREAM>>>((const float*)workspace, S, N, out, accumulate);
    return check_launch("reduce_rows_kernel");
}

extern "C" int ssl4p_pool_head_fwd(const void* xn, const float* W, const float* b, float* logits, int N, int T, int D, int ncls, int spatial, void* stream) {
    SSL4P_REQUIRE(xn && W && logits && N >= 0 && ncls > 0, "pool_head_fwd: bad arguments");
    if (N == 0) return SSL4P_OK;
    const int warps = N * ncls;
    pool_head_fwd_kernel<<<(warps * 32 + 127) / 128, 128, 0, STREAM>>>((const __nv_bfloat16*)xn, W, b, logits, N, T, D, ncls, spatial);
    return check_launch("pool_head_fwd_kernel");
}
extern "C" int ssl4p_pool_head_bwd(const void* xn, const float* W, const float* dlogits, void* dxn, float* dW, float* db,
                                   int N, int T, int D, int ncls, int spatial, int accumulate, void* stream) {
    SSL4P_REQUIRE(xn && W && dlogits && N >= 0 && ncls > 0, "pool_head_bwd: bad arguments");
    if (N == 0) return SSL4P_OK;
    int rc = SSL4P_OK;
    if (dxn) {
        pool_head_bwd_dx_kernel<<<N * T, 128, 0, STREAM>>>(W, dlogits, (__nv_bfloat16*)dxn, T, D, ncls, spatial);
        rc = check_launch("pool_head_bwd_dx_kernel");
        if (rc) return rc;
    }
    if (dW) {
        pool_head_bwd_dw_kernel<<<(ncls * D + 127) / 128, 128, 0, STREAM>>>((const __nv_bfloat16*)xn, dlogits, dW, db, N, T, D, ncls, spatial, accumulate);
        rc = check_launch("pool_head_bwd_dw_kernel");
    }
    return rc;
}
