// Fused flat AdamW over the parameter arena (reference optimiser: torch.optim.AdamW over timm's
// add_weight_decay groups, /root/reference/src/ssl4polyp/models/mae/main_pretrain.py:217-218).
// One pass over p/g/m/v (fp32, 16-byte vectors), decoupled weight decay selected per 256-element chunk, bf16 shadow
// weights refreshed in the same pass, non-finite gradients counted (util/misc.py:387-400) without a host sync.
#include "common.cuh"
#include "ssl4p.h"

namespace ssl4p {

// hyper == nullptr: lr / inv_bc1 / inv_sqrt_bc2 are the host-computed scalars.  hyper != nullptr: lr = hyper[0] and the
// bias corrections follow from t = hyper[1] + 1 (device-resident step counter: CUDA-graph replay).
__global__ void __launch_bounds__(256)
adamw_flat_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                  __nv_bfloat16* __restrict__ pb, const uint8_t* __restrict__ wd_flags, long long n4,
                  float lr, float beta1, float beta2, float eps, float wd, float inv_bc1, float inv_sqrt_bc2, float grad_scale,
                  int* __restrict__ nonfinite, const float* __restrict__ hyper) {
    if (hyper != nullptr) {
        lr = hyper[0];
        const float t = hyper[1] + 1.0f;
        inv_bc1 = 1.0f / (1.0f - powf(beta1, t));
        inv_sqrt_bc2 = rsqrtf(1.0f - powf(beta2, t));
    }
    int bad = 0;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += step) {
        float4 pp = reinterpret_cast<float4*>(p)[i];
        const float4 gg = reinterpret_cast<const float4*>(g)[i];
        float4 mm = reinterpret_cast<float4*>(m)[i];
        float4 vv = reinterpret_cast<float4*>(v)[i];
        const float decay = (wd_flags == nullptr || wd_flags[i >> 6]) ? 1.0f - lr * wd : 1.0f;   // 64 float4 = 256 elements
        float* pa = reinterpret_cast<float*>(&pp);
        const float* ga = reinterpret_cast<const float*>(&gg);
        float* ma = reinterpret_cast<float*>(&mm);
        float* va = reinterpret_cast<float*>(&vv);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float gr = ga[k] * grad_scale;
            if (!isfinite(gr)) bad = 1;
            ma[k] = beta1 * ma[k] + (1.0f - beta1) * gr;
            va[k] = beta2 * va[k] + (1.0f - beta2) * gr * gr;
            const float denom = sqrtf(va[k]) * inv_sqrt_bc2 + eps;
            pa[k] = pa[k] * decay - lr * inv_bc1 * (ma[k] / denom);
        }
        reinterpret_cast<float4*>(p)[i] = pp;
        reinterpret_cast<float4*>(m)[i] = mm;
        reinterpret_cast<float4*>(v)[i] = vv;
        if (pb) {
            uint2 o; o.x = pack_bf16(pa[0], pa[1]); o.y = pack_bf16(pa[2], pa[3]);
            reinterpret_cast<uint2*>(pb)[i] = o;
        }
    }
    if (nonfinite && __syncthreads_or(bad) && threadIdx.x == 0) atomicAdd(nonfinite, 1);
}

__global__ void adamw_step_inc_kernel(float* hyper) { hyper[1] += 1.0f; }

}  // namespace ssl4p

using namespace ssl4p;

extern "C" int ssl4p_adamw_flat(float* p, const float* g, float* m, float* v, void* p_bf16, const uint8_t* wd_flags, long long n,
                                float lr, float beta1, float beta2, float eps, float weight_decay, float bias_corr1, float bias_corr2,
                                float grad_scale, int* nonfinite, void* stream) {
    SSL4P_REQUIRE(p && g && m && v, "adamw: null pointer");
    SSL4P_REQUIRE(n >= 0 && n % 4 == 0, "adamw: n must be a multiple of 4 (arena chunks are 256 elements)");
    SSL4P_REQUIRE(bias_corr1 > 0.f && bias_corr2 > 0.f, "adamw: bias corrections must be positive");
    if (n == 0) return SSL4P_OK;
    const long long n4 = n / 4;
    long long blocks = (n4 + 255) / 256;
    const long long cap = (long long)num_sms() * 8;
    if (blocks > cap) blocks = cap;
    adamw_flat_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, (__nv_bfloat16*)p_bf16, wd_flags, n4, lr, beta1, beta2, eps,
                                                                          weight_decay, 1.0f / bias_corr1, 1.0f / sqrtf(bias_corr2), grad_scale, nonfinite, nullptr);
    return check_launch("adamw_flat_kernel");
}

extern "C" int ssl4p_adamw_flat_dev(float* p, const float* g, float* m, float* v, void* p_bf16, const uint8_t* wd_flags, long long n,
                                    float* hyper, float beta1, float beta2, float eps, float weight_decay,
                                    float grad_scale, int* nonfinite, void* stream) {
    SSL4P_REQUIRE(p && g && m && v && hyper, "adamw_dev: null pointer");
    SSL4P_REQUIRE(n >= 0 && n % 4 == 0, "adamw_dev: n must be a multiple of 4 (arena chunks are 256 elements)");
    if (n > 0) {
        const long long n4 = n / 4;
        long long blocks = (n4 + 255) / 256;
        const long long cap = (long long)num_sms() * 8;
        if (blocks > cap) blocks = cap;
        adamw_flat_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, (__nv_bfloat16*)p_bf16, wd_flags, n4, 0.f, beta1, beta2, eps,
                                                                              weight_decay, 1.f, 1.f, grad_scale, nonfinite, hyper);
        int rc = check_launch("adamw_flat_kernel");
        if (rc) return rc;
    }
    adamw_step_inc_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(hyper);
    return check_launch("adamw_step_inc_kernel");
}

extern "C" int ssl4p_adamw_flat_dev_range(float* p, const float* g, float* m, float* v, void* p_bf16, const uint8_t* wd_flags, long long n,
                                          const float* hyper, float beta1, float beta2, float eps, float weight_decay,
                                          float grad_scale, int* nonfinite, void* stream) {
    SSL4P_REQUIRE(p && g && m && v && hyper, "adamw_dev_range: null pointer");
    SSL4P_REQUIRE(n >= 0 && n % 256 == 0, "adamw_dev_range: a range must be whole 256-element arena chunks");
    SSL4P_REQUIRE((reinterpret_cast<uintptr_t>(p) & 15) == 0 && (reinterpret_cast<uintptr_t>(g) & 15) == 0, "adamw_dev_range: unaligned range");
    if (n == 0) return SSL4P_OK;
    const long long n4 = n / 4;
    long long blocks = (n4 + 255) / 256;
    const long long cap = (long long)num_sms() * 8;
    if (blocks > cap) blocks = cap;
    adamw_flat_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, (__nv_bfloat16*)p_bf16, wd_flags, n4, 0.f, beta1, beta2, eps,
                                                                          weight_decay, 1.f, 1.f, grad_scale, nonfinite, hyper);
    return check_launch("adamw_flat_kernel");
}

extern "C" int ssl4p_adamw_advance_step(float* hyper, void* stream) {
    SSL4P_REQUIRE(hyper, "adamw_advance_step: null pointer");
    adamw_step_inc_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(hyper);
    return check_launch("adamw_step_inc_kernel");
}
