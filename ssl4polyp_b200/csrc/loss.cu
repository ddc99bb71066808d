// Normalised-pixel masked MSE loss and its backward without materialising patchify(imgs)
// (/root/reference/src/ssl4polyp/models/mae/models_mae.py:198-214; patch index map :95-107).
// One warp per patch.  The patch is read from the NCHW image with 16-byte loads in image order (16 contiguous pixels
// per (c, row)), re-ordered through shared memory into the (pi, qi, c) order of `pred`, normalised with the unbiased
// variance, and compared with the prediction row read with unit stride.  Unmasked patches are skipped (mask == 0
// contributes exactly 0 to the reference's sum).  The final sum is a fixed-order single-block reduction.
#include "common.cuh"
#include "ssl4p.h"

namespace ssl4p {

constexpr int kLossWarps = 8;

// loads the target patch (n, i) into s_t in pred order and returns (mean, rstd) for norm_pix (or (0,1))
__device__ __forceinline__ float2 load_target_patch(const float* __restrict__ imgs, int n, int i, int C, int H, int W, int p,
                                                    int norm_pix, float* s_t) {
    const int lane = lane_id();
    const int wp = W / p;
    const int h = i / wp, w = i % wp;
    const int P = p * p * C;
    const float* base = imgs + (size_t)n * C * H * W + (size_t)(h * p) * W + w * p;
    float s = 0.f;
    for (int e = lane * 4; e < P; e += 128) {
        const int c = e / (p * p), pi = (e / p) % p, qi = e % p;
        const float4 v = *reinterpret_cast<const float4*>(base + (size_t)c * H * W + (size_t)pi * W + qi);
        const int o = (pi * p + qi) * C + c;
        s_t[o] = v.x; s_t[o + C] = v.y; s_t[o + 2 * C] = v.z; s_t[o + 3 * C] = v.w;
        s += (v.x + v.y) + (v.z + v.w);
    }
    __syncwarp();
    if (!norm_pix) return make_float2(0.f, 1.f);
    const float mean = warp_sum(s) / (float)P;
    float q = 0.f;
    for (int e = lane; e < P; e += 32) { const float d = s_t[e] - mean; q += d * d; }
    const float var = warp_sum(q) / (float)(P - 1);            // torch.var default: unbiased
    return make_float2(mean, 1.0f / sqrtf(var + 1.0e-6f));
}

__global__ void __launch_bounds__(kLossWarps * 32)
masked_mse_fwd_kernel(const float* __restrict__ imgs, const float* __restrict__ pred, long long pred_bs, const float* __restrict__ mask,
                      int norm_pix, float* __restrict__ per_patch, int NL, int L, int C, int H, int W, int p) {
    extern __shared__ float s_patch[];                          // [kLossWarps][P]
    const int P = p * p * C;
    const int r = blockIdx.x * kLossWarps + (threadIdx.x >> 5);
    if (r >= NL) return;
    const int lane = lane_id();
    if (mask[r] == 0.f) { if (lane == 0) per_patch[r] = 0.f; return; }
    const int n = r / L, i = r % L;
    float* s_t = s_patch + (size_t)(threadIdx.x >> 5) * P;
    const float2 st = load_target_patch(imgs, n, i, C, H, W, p, norm_pix, s_t);
    const float4* pr = reinterpret_cast<const float4*>(pred + (size_t)n * pred_bs + (size_t)i * P);
    float acc = 0.f;
    for (int e4 = lane; e4 < P / 4; e4 += 32) {
        const float4 q = pr[e4];
        const float4 t = *reinterpret_cast<const float4*>(s_t + e4 * 4);
        const float a = q.x - (t.x - st.x) * st.y, b = q.y - (t.y - st.x) * st.y;
        const float c = q.z - (t.z - st.x) * st.y, d = q.w - (t.w - st.x) * st.y;
        acc += (a * a + b * b) + (c * c + d * d);
    }
    acc = warp_sum(acc);
    if (lane == 0) per_patch[r] = acc / (float)P * mask[r];
}

// loss = sum(per_patch) / sum(mask); also stores sum(mask) at per_patch[NL] for the backward.  One block, fixed order.
__global__ void masked_mse_finalize_kernel(float* __restrict__ per_patch, const float* __restrict__ mask, int NL, float* __restrict__ loss) {
    __shared__ float s_a[32], s_b[32];
    // four strided elements per round, all eight loads issued before the first add (one block: the loads' latency is the cost)
    float a = 0.f, b = 0.f;
    const int bd = blockDim.x;
    int i = threadIdx.x;
    for (; i + 3 * bd < NL; i += 4 * bd) {
        const float p0 = per_patch[i], p1 = per_patch[i + bd], p2 = per_patch[i + 2 * bd], p3 = per_patch[i + 3 * bd];
        const float m0 = mask[i], m1 = mask[i + bd], m2 = mask[i + 2 * bd], m3 = mask[i + 3 * bd];
        a += (p0 + p1) + (p2 + p3);
        b += (m0 + m1) + (m2 + m3);
    }
    for (; i < NL; i += bd) { a += per_patch[i]; b += mask[i]; }
    a = warp_sum(a); b = warp_sum(b);
    if (lane_id() == 0) { s_a[threadIdx.x >> 5] = a; s_b[threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float ta = 0.f, tb = 0.f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { ta += s_a[w]; tb += s_b[w]; }
        *loss = ta / tb;
        per_patch[NL] = tb;
    }
}

// one warp per row of dpred ([N, rows_per_img, P] bf16, rows_per_img = L or 1+L when pred sits behind a cls row)
__global__ void __launch_bounds__(kLossWarps * 32)
masked_mse_bwd_kernel(const float* __restrict__ imgs, const float* __restrict__ pred, long long pred_bs, const float* __restrict__ mask,
                      int norm_pix, const float* __restrict__ grad_out, const float* __restrict__ mask_sum,
                      __nv_bfloat16* __restrict__ dpred, long long dpred_bs, int NL, int L, int C, int H, int W, int p) {
    extern __shared__ float s_patch[];
    const int P = p * p * C;
    const int r = blockIdx.x * kLossWarps + (threadIdx.x >> 5);
    if (r >= NL) return;
    const int lane = lane_id();
    const int n = r / L, i = r % L;
    uint2* out = reinterpret_cast<uint2*>(dpred + (size_t)n * dpred_bs + (size_t)i * P);
    if (mask[r] == 0.f) {
        for (int e4 = lane; e4 < P / 4; e4 += 32) out[e4] = make_uint2(0u, 0u);
        return;
    }
    float* s_t = s_patch + (size_t)(threadIdx.x >> 5) * P;
    const float2 st = load_target_patch(imgs, n, i, C, H, W, p, norm_pix, s_t);
    const float scale = (*grad_out) * 2.0f * mask[r] / ((float)P * (*mask_sum));
    const float4* pr = reinterpret_cast<const float4*>(pred + (size_t)n * pred_bs + (size_t)i * P);
    for (int e4 = lane; e4 < P / 4; e4 += 32) {
        const float4 q = pr[e4];
        const float4 t = *reinterpret_cast<const float4*>(s_t + e4 * 4);
        uint2 o;
        o.x = pack_bf16(scale * (q.x - (t.x - st.x) * st.y), scale * (q.y - (t.y - st.x) * st.y));
        o.y = pack_bf16(scale * (q.z - (t.z - st.x) * st.y), scale * (q.w - (t.w - st.x) * st.y));
        out[e4] = o;
    }
}

__global__ void zero_rows_bf16_kernel(__nv_bfloat16* __restrict__ base, long long row_stride, int n_rows, int P4) {
    const int r = blockIdx.x;
    if (r >= n_rows) return;
    uint2* out = reinterpret_cast<uint2*>(base + (size_t)r * row_stride);
    for (int e = threadIdx.x; e < P4; e += blockDim.x) out[e] = make_uint2(0u, 0u);
}

}  // namespace ssl4p

using namespace ssl4p;
#define STREAM ((cudaStream_t)stream)

extern "C" size_t ssl4p_masked_mse_workspace_bytes(int N, int L) { return ((size_t)N * L + 4) * sizeof(float); }

static int loss_check(const float* imgs, const float* pred, const float* mask, int N, int C, int H, int W, int p, void* ws, size_t wsb) {
    SSL4P_REQUIRE(imgs && pred && mask, "masked_mse: null pointer");
    SSL4P_REQUIRE(p % 4 == 0 && H % p == 0 && W % p == 0 && W % 4 == 0, "masked_mse: bad geometry");
    SSL4P_REQUIRE((size_t)p * p * C * 4 * kLossWarps <= 160 * 1024, "masked_mse: patch too large for shared memory");
    SSL4P_REQUIRE(ws && wsb >= ssl4p_masked_mse_workspace_bytes(N, (H / p) * (W / p)), "masked_mse: workspace too small");
    return SSL4P_OK;
}

extern "C" int ssl4p_masked_mse_fwd(const float* imgs, const float* pred, long long pred_batch_stride, const float* mask, int norm_pix,
                                    float* loss, int N, int C, int H, int W, int p, void* workspace, size_t workspace_bytes, void* stream) {
    int rc = loss_check(imgs, pred, mask, N, C, H, W, p, workspace, workspace_bytes);
    if (rc) return rc;
    SSL4P_REQUIRE(loss && N > 0, "masked_mse_fwd: bad arguments");
    const int L = (H / p) * (W / p), NL = N * L, P = p * p * C;
    SSL4P_REQUIRE(pred_batch_stride % 4 == 0 && (reinterpret_cast<uintptr_t>(pred) & 15) == 0, "masked_mse: pred must be 16B aligned");
    const size_t smem = (size_t)kLossWarps * P * sizeof(float);
    static bool attr = false;
    if (!attr && smem > 48 * 1024) { SSL4P_CUDA(cudaFuncSetAttribute(masked_mse_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024)); attr = true; }
    float* per_patch = (float*)workspace;
    masked_mse_fwd_kernel<<<(NL + kLossWarps - 1) / kLossWarps, kLossWarps * 32, smem, STREAM>>>(imgs, pred, pred_batch_stride, mask, norm_pix,
                                                                                               per_patch, NL, L, C, H, W, p);
    rc = check_launch("masked_mse_fwd_kernel");
    if (rc) return rc;
    masked_mse_finalize_kernel<<<1, 1024, 0, STREAM>>>(per_patch, mask, NL, loss);
    return check_launch("masked_mse_finalize_kernel");
}

extern "C" int ssl4p_masked_mse_bwd(const float* imgs, const float* pred, long long pred_batch_stride, const float* mask, int norm_pix,
                                    const float* grad_out, void* dpred, long long dpred_batch_stride, int N, int C, int H, int W, int p,
                                    void* workspace, size_t workspace_bytes, void* stream) {
    int rc = loss_check(imgs, pred, mask, N, C, H, W, p, workspace, workspace_bytes);
    if (rc) return rc;
    SSL4P_REQUIRE(grad_out && dpred && N > 0, "masked_mse_bwd: bad arguments");
    const int L = (H / p) * (W / p), NL = N * L, P = p * p * C;
    SSL4P_REQUIRE(dpred_batch_stride == (long long)L * P || dpred_batch_stride == (long long)(L + 1) * P,
                  "masked_mse_bwd: dpred batch stride must be L*P or (1+L)*P");
    const size_t smem = (size_t)kLossWarps * P * sizeof(float);
    static bool attr = false;
    if (!attr && smem > 48 * 1024) { SSL4P_CUDA(cudaFuncSetAttribute(masked_mse_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024)); attr = true; }
    const float* per_patch = (const float*)workspace;
    __nv_bfloat16* dp = (__nv_bfloat16*)dpred;
    if (dpred_batch_stride == (long long)(L + 1) * P) {
        // dpred points at the cls row of a [N, 1+L, P] buffer: zero the cls rows, patches start one row later
        zero_rows_bf16_kernel<<<N, 128, 0, STREAM>>>(dp, dpred_batch_stride, N, P / 4);
        rc = check_launch("zero_rows_bf16_kernel");
        if (rc) return rc;
        dp += P;
    }
    masked_mse_bwd_kernel<<<(NL + kLossWarps - 1) / kLossWarps, kLossWarps * 32, smem, STREAM>>>(
        imgs, pred, pred_batch_stride, mask, norm_pix, grad_out, per_patch + NL, dp, dpred_batch_stride, NL, L, C, H, W, p);
    return check_launch("masked_mse_bwd_kernel");
}
