// Library plumbing: thread-local error message, launch accounting, device properties, split-K reduce, cast.
#include "common.cuh"
#include "ssl4p.h"
#include <atomic>
#include <stdarg.h>
#include <stdlib.h>

namespace ssl4p {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_launch(const char* what) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: launch failed: %s", what, cudaGetErrorString(e));
        return SSL4P_ERR_CUDA;
    }
    return SSL4P_OK;
}

// Programmatic dependent launch (GEMM, attention, LayerNorm kernels): measured on B200 with the step replayed from a CUDA
// graph it changes nothing (24.33 / 24.41 ms without, 24.49 / 24.47 ms with, profiles/r02g_pdl_ab.txt): the graph's
// kernel-to-kernel hand-off is already short against the 30-130 us kernels.  Off unless SSL4P_PDL=1.
bool pdl_enabled() {
    static const bool on = [] { const char* e = getenv("SSL4P_PDL"); return e ? atoi(e) != 0 : false; }();
    return on;
}

static std::atomic<int> g_reserved_sms{0};

// SMs our persistent kernels size their grids for: the device's count minus the ones set aside for a concurrent
// collective (ssl4p_reserve_sms).  A persistent grid that covers EVERY SM cannot finish while a long-running NCCL
// kernel holds a few of them (our CTAs need the whole shared memory of an SM), so under data parallelism the two
// are given disjoint SM sets instead of fighting for the same ones.
int num_sms() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
            sms = 148;
    }
    const int r = g_reserved_sms.load(std::memory_order_relaxed);
    return sms - r > 8 ? sms - r : 8;
}

// out[i] (+)= sum_s parts[s*stride + i], fixed summation order
__global__ void splitk_reduce_kernel(const float* __restrict__ parts, int splits, long long stride, float* __restrict__ out,
                                     long long n4, int accumulate) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    float4 acc = accumulate ? reinterpret_cast<const float4*>(out)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s = 0; s < splits; ++s) {
        const float4 v = *reinterpret_cast<const float4*>(parts + (long long)s * stride + i * 4);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    reinterpret_cast<float4*>(out)[i] = acc;
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n8) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (; i < n8; i += step) {
        const float4 a = reinterpret_cast<const float4*>(src)[2 * i];
        const float4 b = reinterpret_cast<const float4*>(src)[2 * i + 1];
        uint4 o;
        o.x = pack_bf16(a.x, a.y); o.y = pack_bf16(a.z, a.w); o.z = pack_bf16(b.x, b.y); o.w = pack_bf16(b.z, b.w);
        reinterpret_cast<uint4*>(dst)[i] = o;
    }
}
__global__ void cast_f32_bf16_tail_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long start, long long n) {
    const long long i = start + threadIdx.x;
    if (i < n) dst[i] = __float2bfloat16(src[i]);
}

}  // namespace ssl4p

using namespace ssl4p;

extern "C" int ssl4p_reserve_sms(int n) {
    if (n < 0) n = 0;
    if (n > 64) n = 64;
    return g_reserved_sms.exchange(n, std::memory_order_relaxed);
}

extern "C" int ssl4p_abi_version(void) { return SSL4P_ABI_VERSION; }
extern "C" const char* ssl4p_last_error(void) { return g_err; }
extern "C" unsigned long long ssl4p_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" int ssl4p_splitk_reduce(const float* parts, int splits, long long stride, float* out, long long n, int accumulate, void* stream) {
    SSL4P_REQUIRE(parts && out && splits >= 1 && n > 0, "splitk_reduce: bad arguments");
    SSL4P_REQUIRE(n % 4 == 0 && stride % 4 == 0, "splitk_reduce: n and stride must be multiples of 4");
    const long long n4 = n / 4;
    const int threads = 256;
    const long long blocks = (n4 + threads - 1) / threads;
    splitk_reduce_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(parts, splits, stride, out, n4, accumulate);
    return check_launch("splitk_reduce_kernel");
}

extern "C" int ssl4p_cast_f32_bf16(const float* src, void* dst, long long n, void* stream) {
    SSL4P_REQUIRE(src && dst && n >= 0, "cast: bad arguments");
    if (n == 0) return SSL4P_OK;
    SSL4P_REQUIRE((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0, "cast: pointers must be 16B aligned");
    const long long n8 = n / 8;
    if (n8 > 0) {
        long long blocks = (n8 + 255) / 256;
        const long long cap = (long long)num_sms() * 16;
        if (blocks > cap) blocks = cap;
        cast_f32_bf16_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, n8);
        int rc = check_launch("cast_f32_bf16_kernel");
        if (rc) return rc;
    }
    if (n8 * 8 < n) {
        cast_f32_bf16_tail_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, n8 * 8, n);
        return check_launch("cast_f32_bf16_tail_kernel");
    }
    return SSL4P_OK;
}
