// LayerNorm forward/backward over the last dimension (nn.LayerNorm(eps=1e-6) as passed at
// /root/reference/src/ssl4polyp/models/mae/models_mae.py:227,235; used as norm1/norm2 of every timm Block, `norm` :168
// and `decoder_norm` :188).  fp32 residual stream in, bf16 GEMM operand out.  HBM-bound: one warp per row, the row
// lives in registers (16-byte loads), statistics by warp shuffles; the backward fuses the residual-gradient add and
// produces the bf16 shadow of the gradient stream; dgamma/dbeta are reduced in two deterministic stages.
#include "common.cuh"
#include "ssl4p.h"

namespace ssl4p {

constexpr int kLnWarps = 8;

// NV = float4 per lane (D = 128*NV).  NV == 0: generic D (D % 4 == 0), re-reading the row from L1/L2.
template <int NV>
__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                     __nv_bfloat16* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out, int M, int D) {
    const int lane = lane_id();
    const int row = blockIdx.x * kLnWarps + (threadIdx.x >> 5);
    pdl_wait_then_trigger();
    if (row >= M) return;
    const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * D);
    const float4* g4 = reinterpret_cast<const float4*>(gamma);
    const float4* b4 = reinterpret_cast<const float4*>(beta);
    uint2* yr = reinterpret_cast<uint2*>(y + (size_t)row * D);
    const float inv_d = 1.0f / (float)D;
    if constexpr (NV > 0) {
        float4 v[NV];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) { v[i] = xr[lane + 32 * i]; s += (v[i].x + v[i].y) + (v[i].z + v[i].w); }
        const float mean = warp_sum(s) * inv_d;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
            q += (a * a + b * b) + (c * c + d * d);
        }
        const float rstd = rsqrtf(warp_sum(q) * inv_d + eps);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const float4 g = g4[lane + 32 * i], b = b4[lane + 32 * i];
            uint2 o;
            o.x = pack_bf16((v[i].x - mean) * rstd * g.x + b.x, (v[i].y - mean) * rstd * g.y + b.y);
            o.y = pack_bf16((v[i].z - mean) * rstd * g.z + b.z, (v[i].w - mean) * rstd * g.w + b.w);
            yr[lane + 32 * i] = o;
        }
        if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
    } else {
        const int D4 = D / 4;
        float s = 0.f;
        for (int i = lane; i < D4; i += 32) { const float4 v = xr[i]; s += (v.x + v.y) + (v.z + v.w); }
        const float mean = warp_sum(s) * inv_d;
        float q = 0.f;
        for (int i = lane; i < D4; i += 32) {
            const float4 v = xr[i];
            const float a = v.x - mean, b = v.y - mean, c = v.z - mean, d = v.w - mean;
            q += (a * a + b * b) + (c * c + d * d);
        }
        const float rstd = rsqrtf(warp_sum(q) * inv_d + eps);
        for (int i = lane; i < D4; i += 32) {
            const float4 v = xr[i], g = g4[i], b = b4[i];
            uint2 o;
            o.x = pack_bf16((v.x - mean) * rstd * g.x + b.x, (v.y - mean) * rstd * g.y + b.y);
            o.y = pack_bf16((v.z - mean) * rstd * g.z + b.z, (v.w - mean) * rstd * g.w + b.w);
            yr[i] = o;
        }
        if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
    }
}

// Backward.  Grid-stride over rows (persistent), each lane keeps its dgamma/dbeta partials in registers.
// part layout: [gridDim.x][2][D]  (dgamma then dbeta).
template <int NV>
__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ mean_in,
                     const float* __restrict__ rstd_in, const float* __restrict__ gamma, const float* __restrict__ res_grad,
                     float* __restrict__ dx_out, __nv_bfloat16* __restrict__ dxb, float* __restrict__ part, int M, int D) {
    extern __shared__ float s_red[];                           // [kLnWarps][3*D]
    const int lane = lane_id();
    const int warp = threadIdx.x >> 5;
    pdl_wait_then_trigger();
    const float inv_d = 1.0f / (float)D;
    const float4* g4 = reinterpret_cast<const float4*>(gamma);
    static_assert(NV > 0, "register-resident variant only");
    float4 dg[NV], db[NV], ds[NV], gam[NV];                    // ds = column sums of the dx written below (bias gradients upstream)
#pragma unroll
    for (int i = 0; i < NV; ++i) { dg[i] = make_float4(0, 0, 0, 0); db[i] = make_float4(0, 0, 0, 0); ds[i] = make_float4(0, 0, 0, 0); gam[i] = g4[lane + 32 * i]; }

    for (int row = blockIdx.x * kLnWarps + warp; row < M; row += gridDim.x * kLnWarps) {
        const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * D);
        const uint2* dyr = reinterpret_cast<const uint2*>(dy + (size_t)row * D);
        const float mean = mean_in[row], rstd = rstd_in[row];
        float4 xh[NV], gy[NV];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const float4 v = xr[lane + 32 * i];
            const uint2 u = dyr[lane + 32 * i];
            const float2 d01 = unpack_bf16(u.x), d23 = unpack_bf16(u.y);
            xh[i] = make_float4((v.x - mean) * rstd, (v.y - mean) * rstd, (v.z - mean) * rstd, (v.w - mean) * rstd);
            dg[i].x += d01.x * xh[i].x; dg[i].y += d01.y * xh[i].y; dg[i].z += d23.x * xh[i].z; dg[i].w += d23.y * xh[i].w;
            db[i].x += d01.x; db[i].y += d01.y; db[i].z += d23.x; db[i].w += d23.y;
            gy[i] = make_float4(d01.x * gam[i].x, d01.y * gam[i].y, d23.x * gam[i].z, d23.y * gam[i].w);
            c1 += (gy[i].x + gy[i].y) + (gy[i].z + gy[i].w);
            c2 += (gy[i].x * xh[i].x + gy[i].y * xh[i].y) + (gy[i].z * xh[i].z + gy[i].w * xh[i].w);
        }
        c1 = warp_sum(c1) * inv_d;
        c2 = warp_sum(c2) * inv_d;
        float4* dxr = reinterpret_cast<float4*>(dx_out + (size_t)row * D);
        const float4* rr = res_grad ? reinterpret_cast<const float4*>(res_grad + (size_t)row * D) : nullptr;
        uint2* dbr = dxb ? reinterpret_cast<uint2*>(dxb + (size_t)row * D) : nullptr;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float4 o;
            o.x = rstd * (gy[i].x - c1 - xh[i].x * c2); o.y = rstd * (gy[i].y - c1 - xh[i].y * c2);
            o.z = rstd * (gy[i].z - c1 - xh[i].z * c2); o.w = rstd * (gy[i].w - c1 - xh[i].w * c2);
            if (rr) { const float4 r = rr[lane + 32 * i]; o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w; }
            dxr[lane + 32 * i] = o;
            ds[i].x += o.x; ds[i].y += o.y; ds[i].z += o.z; ds[i].w += o.w;
            if (dbr) { uint2 b; b.x = pack_bf16(o.x, o.y); b.y = pack_bf16(o.z, o.w); dbr[lane + 32 * i] = b; }
        }
    }
    // block reduce of the per-warp partial sums (fixed order), then one partial row per block
    float4* sw = reinterpret_cast<float4*>(s_red + (size_t)warp * 3 * D);
#pragma unroll
    for (int i = 0; i < NV; ++i) { sw[lane + 32 * i] = dg[i]; sw[D / 4 + lane + 32 * i] = db[i]; sw[D / 2 + lane + 32 * i] = ds[i]; }
    __syncthreads();
    for (int j = threadIdx.x; j < 3 * D; j += blockDim.x) {
        float acc = 0.f;
#pragma unroll
        for (int w = 0; w < kLnWarps; ++w) acc += s_red[(size_t)w * 3 * D + j];
        part[(size_t)blockIdx.x * 3 * D + j] = acc;
    }
}

// Same arithmetic (bit-identical sums), but the dgamma/dbeta/column-sum partials live in the warp's shared-memory slice
// instead of 3*NV float4 registers per lane: <= 128 registers, two CTAs per SM, twice the loads in flight per SM.
template <int NV>
__global__ void __launch_bounds__(kLnWarps * 32, 2)
layernorm_bwd_smem_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ mean_in,
                          const float* __restrict__ rstd_in, const float* __restrict__ gamma, const float* __restrict__ res_grad,
                          float* __restrict__ dx_out, __nv_bfloat16* __restrict__ dxb, float* __restrict__ part, int M, int D) {
    extern __shared__ float s_red[];                           // [kLnWarps][3*D]
    const int lane = lane_id();
    const int warp = threadIdx.x >> 5;
    pdl_wait_then_trigger();
    const float inv_d = 1.0f / (float)D;
    const float4* g4 = reinterpret_cast<const float4*>(gamma);
    float4* sdg = reinterpret_cast<float4*>(s_red + (size_t)warp * 3 * D);
    float4* sdb = sdg + D / 4;
    float4* sds = sdg + D / 2;
    float4 gam[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        gam[i] = g4[lane + 32 * i];
        sdg[lane + 32 * i] = make_float4(0, 0, 0, 0); sdb[lane + 32 * i] = make_float4(0, 0, 0, 0); sds[lane + 32 * i] = make_float4(0, 0, 0, 0);
    }
    // every lane only ever touches its own shared-memory words: no synchronisation inside the row loop
    for (int row = blockIdx.x * kLnWarps + warp; row < M; row += gridDim.x * kLnWarps) {
        const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * D);
        const uint2* dyr = reinterpret_cast<const uint2*>(dy + (size_t)row * D);
        const float mean = mean_in[row], rstd = rstd_in[row];
        float4 xh[NV], gy[NV];
        uint2 u[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) { xh[i] = xr[lane + 32 * i]; u[i] = dyr[lane + 32 * i]; }
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const float2 d01 = unpack_bf16(u[i].x), d23 = unpack_bf16(u[i].y);
            xh[i] = make_float4((xh[i].x - mean) * rstd, (xh[i].y - mean) * rstd, (xh[i].z - mean) * rstd, (xh[i].w - mean) * rstd);
            float4 a = sdg[lane + 32 * i], b = sdb[lane + 32 * i];
            a.x += d01.x * xh[i].x; a.y += d01.y * xh[i].y; a.z += d23.x * xh[i].z; a.w += d23.y * xh[i].w;
            b.x += d01.x; b.y += d01.y; b.z += d23.x; b.w += d23.y;
            sdg[lane + 32 * i] = a; sdb[lane + 32 * i] = b;
            gy[i] = make_float4(d01.x * gam[i].x, d01.y * gam[i].y, d23.x * gam[i].z, d23.y * gam[i].w);
            c1 += (gy[i].x + gy[i].y) + (gy[i].z + gy[i].w);
            c2 += (gy[i].x * xh[i].x + gy[i].y * xh[i].y) + (gy[i].z * xh[i].z + gy[i].w * xh[i].w);
        }
        c1 = warp_sum(c1) * inv_d;
        c2 = warp_sum(c2) * inv_d;
        float4* dxr = reinterpret_cast<float4*>(dx_out + (size_t)row * D);
        const float4* rr = res_grad ? reinterpret_cast<const float4*>(res_grad + (size_t)row * D) : nullptr;
        uint2* dbr = dxb ? reinterpret_cast<uint2*>(dxb + (size_t)row * D) : nullptr;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float4 o;
            o.x = rstd * (gy[i].x - c1 - xh[i].x * c2); o.y = rstd * (gy[i].y - c1 - xh[i].y * c2);
            o.z = rstd * (gy[i].z - c1 - xh[i].z * c2); o.w = rstd * (gy[i].w - c1 - xh[i].w * c2);
            if (rr) { const float4 r = rr[lane + 32 * i]; o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w; }
            dxr[lane + 32 * i] = o;
            float4 s = sds[lane + 32 * i];
            s.x += o.x; s.y += o.y; s.z += o.z; s.w += o.w;
            sds[lane + 32 * i] = s;
            if (dbr) { uint2 b; b.x = pack_bf16(o.x, o.y); b.y = pack_bf16(o.z, o.w); dbr[lane + 32 * i] = b; }
        }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < 3 * D; j += blockDim.x) {
        float acc = 0.f;
#pragma unroll
        for (int w = 0; w < kLnWarps; ++w) acc += s_red[(size_t)w * 3 * D + j];
        part[(size_t)blockIdx.x * 3 * D + j] = acc;
    }
}

// generic-D backward (small test configurations): one warp per row, partial sums through shared-memory atomics are
// avoided by giving every block a private [2*D] accumulator updated by one warp at a time in fixed order.
__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_bwd_generic_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ mean_in,
                             const float* __restrict__ rstd_in, const float* __restrict__ gamma, const float* __restrict__ res_grad,
                             float* __restrict__ dx_out, __nv_bfloat16* __restrict__ dxb, float* __restrict__ part, int M, int D) {
    extern __shared__ float s_red[];                           // [kLnWarps][3*D]
    const int lane = lane_id();
    const int warp = threadIdx.x >> 5;
    pdl_wait_then_trigger();
    const float inv_d = 1.0f / (float)D;
    float* my = s_red + (size_t)warp * 3 * D;
    for (int j = lane; j < 3 * D; j += 32) my[j] = 0.f;
    __syncwarp();
    for (int row = blockIdx.x * kLnWarps + warp; row < M; row += gridDim.x * kLnWarps) {
        const float* xr = x + (size_t)row * D;
        const __nv_bfloat16* dyr = dy + (size_t)row * D;
        const float mean = mean_in[row], rstd = rstd_in[row];
        float c1 = 0.f, c2 = 0.f;
        for (int j = lane; j < D; j += 32) {
            const float xh = (xr[j] - mean) * rstd, d = __bfloat162float(dyr[j]);
            const float gy = d * gamma[j];
            c1 += gy; c2 += gy * xh;
            my[j] += d * xh; my[D + j] += d;
        }
        c1 = warp_sum(c1) * inv_d; c2 = warp_sum(c2) * inv_d;
        for (int j = lane; j < D; j += 32) {
            const float xh = (xr[j] - mean) * rstd, gy = __bfloat162float(dyr[j]) * gamma[j];
            float o = rstd * (gy - c1 - xh * c2);
            if (res_grad) o += res_grad[(size_t)row * D + j];
            dx_out[(size_t)row * D + j] = o;
            my[2 * D + j] += o;
            if (dxb) dxb[(size_t)row * D + j] = __float2bfloat16(o);
        }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < 3 * D; j += blockDim.x) {
        float acc = 0.f;
        for (int w = 0; w < kLnWarps; ++w) acc += s_red[(size_t)w * 3 * D + j];
        part[(size_t)blockIdx.x * 3 * D + j] = acc;
    }
}

// out[j] (+)= sum_g part[g][j]  over [G][3D] -> dgamma (j < D), dbeta (D <= j < 2D), column sums of dx (j >= 2D, optional).
// 32 columns per block, 8 row groups summed in parallel and combined in fixed order (deterministic).
__global__ void __launch_bounds__(256)
ln_param_grad_reduce_kernel(const float* __restrict__ part, int G, int D, float* __restrict__ dgamma,
                            float* __restrict__ dbeta, float* __restrict__ dxsum, int accumulate, int accumulate_dxsum) {
    __shared__ float sh[8][33];
    pdl_wait_then_trigger();
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int j = blockIdx.x * 32 + tx;
    const int W = 3 * D;
    float acc = 0.f;
    if (j < W)
        for (int g = ty; g < G; g += 8) acc += part[(size_t)g * W + j];
    sh[ty][tx] = acc;
    __syncthreads();
    if (ty == 0 && j < W) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += sh[k][tx];
        if (j < D) dgamma[j] = (accumulate ? dgamma[j] : 0.f) + t;
        else if (j < 2 * D) dbeta[j - D] = (accumulate ? dbeta[j - D] : 0.f) + t;
        else if (dxsum != nullptr) dxsum[j - 2 * D] = (accumulate_dxsum ? dxsum[j - 2 * D] : 0.f) + t;
    }
}

static int ln_bwd_grid(int M) {
    int g = 2 * num_sms();
    const int need = (M + kLnWarps - 1) / kLnWarps;
    return g < need ? g : need;
}

}  // namespace ssl4p

using namespace ssl4p;
#define STREAM ((cudaStream_t)stream)

extern "C" int ssl4p_layernorm_fwd(const float* x, const float* gamma, const float* beta, float eps, void* y, float* mean, float* rstd,
                                   int M, int D, void* stream) {
    SSL4P_REQUIRE(x && gamma && beta && y && mean && rstd, "layernorm_fwd: null pointer");
    SSL4P_REQUIRE(M >= 0 && D > 0 && D % 4 == 0, "layernorm_fwd: D must be a multiple of 4 (D=%d)", D);
    if (M == 0) return SSL4P_OK;
    const int grid = (M + kLnWarps - 1) / kLnWarps, threads = kLnWarps * 32;
    __nv_bfloat16* yb = (__nv_bfloat16*)y;
    auto kern = D == 512 ? layernorm_fwd_kernel<4> : D == 768 ? layernorm_fwd_kernel<6> : D == 1024 ? layernorm_fwd_kernel<8> : layernorm_fwd_kernel<0>;
    cudaError_t e = launch_ex(kern, dim3(grid), dim3(threads), (size_t)0, STREAM, 1, x, gamma, beta, eps, yb, mean, rstd, M, D);
    if (e != cudaSuccess) { set_error("layernorm_fwd_kernel launch failed: %s", cudaGetErrorString(e)); return SSL4P_ERR_CUDA; }
    return check_launch("layernorm_fwd_kernel");
}

extern "C" size_t ssl4p_layernorm_bwd_workspace_bytes(int M, int D) {
    (void)M;
    return (size_t)2 * 148 * 2 * 3 * (size_t)D * sizeof(float);     // up to 592 partial rows of [3*D]
}

extern "C" int ssl4p_layernorm_bwd(const void* dy, const float* x, const float* mean, const float* rstd, const float* gamma,
                                   const float* residual_grad, float* dx_out, void* dxb, float* dgamma, float* dbeta,
                                   float* dx_colsum, int M, int D, int accumulate, int accumulate_colsum,
                                   void* workspace, size_t workspace_bytes, void* stream) {
    SSL4P_REQUIRE(dy && x && mean && rstd && gamma && dx_out && dgamma && dbeta, "layernorm_bwd: null pointer");
    SSL4P_REQUIRE(M > 0 && D > 0 && D % 4 == 0, "layernorm_bwd: bad shape");
    const int grid = ln_bwd_grid(M);
    SSL4P_REQUIRE(workspace && workspace_bytes >= (size_t)grid * 3 * D * sizeof(float), "layernorm_bwd: workspace too small");
    const int threads = kLnWarps * 32;
    const size_t smem = (size_t)kLnWarps * 3 * D * sizeof(float);
    float* part = (float*)workspace;
    const __nv_bfloat16* dyb = (const __nv_bfloat16*)dy;
    __nv_bfloat16* dxbb = (__nv_bfloat16*)dxb;
    cudaError_t le = cudaSuccess;
#define LN_BWD(NV_)                                                                                                        \
    {                                                                                                                      \
        static bool attr = false;                                                                                          \
        if (!attr && smem > 48 * 1024) { SSL4P_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel<NV_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = true; } \
        le = launch_ex(layernorm_bwd_kernel<NV_>, dim3(grid), dim3(threads), smem, STREAM, 1, dyb, x, mean, rstd, gamma, residual_grad, dx_out, dxbb, part, M, D); \
    }
#define LN_BWD_SMEM(NV_)                                                                                                   \
    {                                                                                                                      \
        static bool attr = false;                                                                                          \
        if (!attr) { SSL4P_CUDA(cudaFuncSetAttribute(layernorm_bwd_smem_kernel<NV_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = true; } \
        le = launch_ex(layernorm_bwd_smem_kernel<NV_>, dim3(grid), dim3(threads), smem, STREAM, 1, dyb, x, mean, rstd, gamma, residual_grad, dx_out, dxbb, part, M, D); \
    }
    // SSL4P_LN_BWD_SMEM=0 keeps the register-resident sums (one CTA per SM) for D >= 768
    static const bool smem_acc = getenv("SSL4P_LN_BWD_SMEM") ? atoi(getenv("SSL4P_LN_BWD_SMEM")) != 0 : true;
    switch (D) {
        case 512: LN_BWD(4) break;
        case 768: if (smem_acc) LN_BWD_SMEM(6) else LN_BWD(6) break;
        case 1024: if (smem_acc) LN_BWD_SMEM(8) else LN_BWD(8) break;
        default: {
            SSL4P_REQUIRE(smem <= 48 * 1024, "layernorm_bwd: generic path supports D <= 512");
            le = launch_ex(layernorm_bwd_generic_kernel, dim3(grid), dim3(threads), smem, STREAM, 1, dyb, x, mean, rstd, gamma, residual_grad, dx_out, dxbb, part, M, D);
        }
    }
#undef LN_BWD
#undef LN_BWD_SMEM
    if (le != cudaSuccess) { set_error("layernorm_bwd_kernel launch failed: %s", cudaGetErrorString(le)); return SSL4P_ERR_CUDA; }
    int rc = check_launch("layernorm_bwd_kernel");
    if (rc) return rc;
    le = launch_ex(ln_param_grad_reduce_kernel, dim3((3 * D + 31) / 32), dim3(256), (size_t)0, STREAM, 1, (const float*)part, grid, D, dgamma, dbeta, dx_colsum, accumulate, accumulate_colsum);
    if (le != cudaSuccess) { set_error("ln_param_grad_reduce_kernel launch failed: %s", cudaGetErrorString(le)); return SSL4P_ERR_CUDA; }
    return check_launch("ln_param_grad_reduce_kernel");
}

// ---- allocation-free variant (no counterpart in the reference) ----------
// The [grid][3*D] partial-sum scratch of the backward lives in a file-scope device array, so callers that cannot call
// cudaMalloc (graph capture, tight serving loops) need no workspace.  Limit: D <= 1024 (and grid <= 296 partial rows), 3.47 MiB
// held by every GPU that loads the library.  The kernel overwrites every partial row it later reduces, so nothing has to
// be cleared between calls; all calls share the array, so calls that overlap on the GPU (different streams) race.
constexpr int kLnStaticMaxD = 1024;
constexpr int kLnStaticMaxRows = 2 * 148;
static __device__ __align__(256) float g_ln_bwd_scratch[(size_t)kLnStaticMaxRows * 3 * kLnStaticMaxD];

extern "C" int ssl4p_layernorm_bwd_max_static_d(void) { return kLnStaticMaxD; }

extern "C" int ssl4p_layernorm_bwd_static(const void* dy, const float* x, const float* mean, const float* rstd, const float* gamma,
                                          const float* residual_grad, float* dx_out, void* dxb, float* dgamma, float* dbeta,
                                          float* dx_colsum, int M, int D, int accumulate, int accumulate_colsum, void* stream) {
    SSL4P_REQUIRE(D > 0 && D <= kLnStaticMaxD, "layernorm_bwd_static: D=%d exceeds the static scratch limit (D <= %d); use ssl4p_layernorm_bwd with a caller workspace", D, kLnStaticMaxD);
    // the array's name is a host-side stand-in: resolve the device address once per device, with that device current
    constexpr int kMaxDevices = 64;
    static float* addr[kMaxDevices] = {};
    int dev = 0;
    SSL4P_CUDA(cudaGetDevice(&dev));
    SSL4P_REQUIRE(dev >= 0 && dev < kMaxDevices, "layernorm_bwd_static: device ordinal %d out of range", dev);
    if (addr[dev] == nullptr) {
        float* p = nullptr;
        SSL4P_CUDA(cudaGetSymbolAddress((void**)&p, g_ln_bwd_scratch));
        SSL4P_REQUIRE(p != nullptr, "layernorm_bwd_static: cudaGetSymbolAddress returned a null address");
        addr[dev] = p;
    }
    // same launches as ssl4p_layernorm_bwd; it rejects a grid whose partial rows would not fit in the array
    return ssl4p_layernorm_bwd(dy, x, mean, rstd, gamma, residual_grad, dx_out, dxb, dgamma, dbeta, dx_colsum, M, D, accumulate,
                               accumulate_colsum, addr[dev], sizeof(float) * (size_t)kLnStaticMaxRows * 3 * kLnStaticMaxD, stream);
}
