// TEST INFRASTRUCTURE ONLY: built into tests/libssl4p_check.so (never into the product library libssl4p.so).  The first
// revision of the attention kernels, on warp-level mma.sync; kept as an independent ON-DEVICE cross-check of the tcgen05
// kernels (tests/test_ops_gpu.py::test_attention_fwd_bwd) and as the comparison arm of scripts/attn_bench.py.
//
// Fused multi-head self-attention forward/backward for the short sequences of the MAE path
// (timm 0.4.12 Attention.forward as used by /root/reference/src/ssl4polyp/models/mae/models_mae.py:166,186:
//  softmax(q k^T * hd^-0.5) v with T = 50 (encoder), 197 (decoder / classifier), hd = 64 / 32).
// One CTA per (batch, head): Q, K, V of the head live in shared memory, the whole score row lives in registers
// (single-pass softmax, no online rescale), nothing of size [B,H,T,T] ever touches HBM.
// Tensor-core path of this revision: warp-level mma.sync m16n8k16 bf16 (fp32 accumulate).  Attention is 3.8 % of the
// step's FLOPs; the tcgen05 version is tracked in DESIGN.md as the next optimisation.
// Backward recomputes P from the saved log-sum-exp in two register-resident passes (dQ by query tile, dK/dV by key
// tile), so there are no atomics and the result is deterministic.
#include "common.cuh"
#include "ssl4p_check.h"
#include <stdarg.h>

namespace ssl4p {

// the plumbing the product library keeps in api.cu, restated so that this library stands alone
static thread_local char g_check_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_check_err, sizeof(g_check_err), fmt, ap);
    va_end(ap);
}
int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("%s: launch failed: %s", what, cudaGetErrorString(e)); return SSL4P_ERR_CUDA; }
    return SSL4P_OK;
}

__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t lds32(const __nv_bfloat16* p) { return *reinterpret_cast<const uint32_t*>(p); }

// A fragment (16 rows x 16 k) of a row-major smem matrix with leading dimension ld
__device__ __forceinline__ void load_a_frag(uint32_t (&a)[4], const __nv_bfloat16* s, int ld, int row0, int k0) {
    const int g = lane_id() >> 2, t = lane_id() & 3;
    a[0] = lds32(s + (row0 + g) * ld + k0 + 2 * t);
    a[1] = lds32(s + (row0 + g + 8) * ld + k0 + 2 * t);
    a[2] = lds32(s + (row0 + g) * ld + k0 + 8 + 2 * t);
    a[3] = lds32(s + (row0 + g + 8) * ld + k0 + 8 + 2 * t);
}
// B fragment (16 k x 8 n) where B[k][n] = s[n0+n][k0+k]  (s row-major [n][k])
__device__ __forceinline__ void load_b_frag(uint32_t& b0, uint32_t& b1, const __nv_bfloat16* s, int ld, int n0, int k0) {
    const int g = lane_id() >> 2, t = lane_id() & 3;
    b0 = lds32(s + (n0 + g) * ld + k0 + 2 * t);
    b1 = lds32(s + (n0 + g) * ld + k0 + 8 + 2 * t);
}

constexpr float kLog2e = 1.4426950408889634f;

template <int HD>
struct AttnSmem {
    static constexpr int ldr = HD + 8;                       // row-major [T][HD] leading dimension (bank-conflict free)
    static __host__ __device__ int ldt(int Tpad) { return Tpad + 8; }   // transposed [HD][T]
};

// ---- cooperative load of one head into shared memory ----------------------------------------------------
// which: 0 = q, 1 = k, 2 = v slice of qkv [B,T,3,H,HD];  row-major copy into dst (zero padded to Tpad rows) and/or
// transposed copy into dst_t.
template <int HD>
__device__ __forceinline__ void load_head(const __nv_bfloat16* __restrict__ src, size_t row_stride, int T, int Tpad,
                                          __nv_bfloat16* dst, __nv_bfloat16* dst_t) {
    constexpr int CH = HD / 8;
    const int ldr = AttnSmem<HD>::ldr, ldt = AttnSmem<HD>::ldt(Tpad);
    for (int idx = threadIdx.x; idx < Tpad * CH; idx += blockDim.x) {
        const int t = idx / CH, ch = idx % CH;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (t < T) v = *reinterpret_cast<const uint4*>(src + (size_t)t * row_stride + ch * 8);
        if (dst) *reinterpret_cast<uint4*>(dst + t * ldr + ch * 8) = v;
        if (dst_t) {
            const __nv_bfloat16* e = reinterpret_cast<const __nv_bfloat16*>(&v);
#pragma unroll
            for (int j = 0; j < 8; ++j) dst_t[(ch * 8 + j) * ldt + t] = e[j];
        }
    }
}

// =========================================================================================================
// forward
// =========================================================================================================
template <int HD, int NT, int NWARPS>
__global__ void __launch_bounds__(NWARPS * 32)
attention_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out, float* __restrict__ lse,
                     int T, int H, float scale) {
    extern __shared__ __align__(16) uint8_t smem_attn[];
    constexpr int Tpad = NT * 8;
    constexpr int ldr = AttnSmem<HD>::ldr;
    const int ldt = AttnSmem<HD>::ldt(Tpad);
    __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smem_attn);
    __nv_bfloat16* sK = sQ + Tpad * ldr;
    __nv_bfloat16* sVt = sK + Tpad * ldr;

    const int b = blockIdx.x / H, h = blockIdx.x % H;
    const size_t row_stride = (size_t)3 * H * HD;
    const __nv_bfloat16* base = qkv + (size_t)b * T * row_stride + (size_t)h * HD;
    load_head<HD>(base, row_stride, T, Tpad, sQ, nullptr);
    load_head<HD>(base + (size_t)H * HD, row_stride, T, Tpad, sK, nullptr);
    load_head<HD>(base + (size_t)2 * H * HD, row_stride, T, Tpad, nullptr, sVt);
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = lane_id(), g = lane >> 2, t4 = lane & 3;
    const int q_tiles = (T + 15) / 16;
    const int nt_used = 2 * ((T + 15) / 16);                  // n-tiles of 8 keys actually needed (even)
    const float sl2 = scale * kLog2e;

    for (int qt = warp; qt < q_tiles; qt += NWARPS) {
        const int q0 = qt * 16;
        uint32_t aq[HD / 16][4];
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) load_a_frag(aq[ks], sQ, ldr, q0, ks * 16);

        float s[NT][4];
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
            if (j < nt_used) {
#pragma unroll
                for (int ks = 0; ks < HD / 16; ++ks) {
                    uint32_t b0, b1;
                    load_b_frag(b0, b1, sK, ldr, j * 8, ks * 16);
                    mma_bf16(s[j], aq[ks], b0, b1);
                }
            }
        }
        // mask padded keys, row max
        float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            const int c = j * 8 + 2 * t4;
            if (c >= T) { s[j][0] = -INFINITY; s[j][2] = -INFINITY; }
            if (c + 1 >= T) { s[j][1] = -INFINITY; s[j][3] = -INFINITY; }
            m0 = fmaxf(m0, fmaxf(s[j][0], s[j][1]));
            m1 = fmaxf(m1, fmaxf(s[j][2], s[j][3]));
        }
        m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
        m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
        float l0 = 0.f, l1 = 0.f;
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            s[j][0] = exp2f((s[j][0] - m0) * sl2); s[j][1] = exp2f((s[j][1] - m0) * sl2);
            s[j][2] = exp2f((s[j][2] - m1) * sl2); s[j][3] = exp2f((s[j][3] - m1) * sl2);
            l0 += s[j][0] + s[j][1]; l1 += s[j][2] + s[j][3];
        }
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);

        float o[HD / 8][4];
#pragma unroll
        for (int dn = 0; dn < HD / 8; ++dn) o[dn][0] = o[dn][1] = o[dn][2] = o[dn][3] = 0.f;
#pragma unroll
        for (int jj = 0; jj < NT / 2; ++jj) {
            if (2 * jj < nt_used) {
                uint32_t ap[4];
                ap[0] = pack_bf16(s[2 * jj][0], s[2 * jj][1]); ap[1] = pack_bf16(s[2 * jj][2], s[2 * jj][3]);
                ap[2] = pack_bf16(s[2 * jj + 1][0], s[2 * jj + 1][1]); ap[3] = pack_bf16(s[2 * jj + 1][2], s[2 * jj + 1][3]);
#pragma unroll
                for (int dn = 0; dn < HD / 8; ++dn) {
                    uint32_t b0, b1;
                    load_b_frag(b0, b1, sVt, ldt, dn * 8, jj * 16);
                    mma_bf16(o[dn], ap, b0, b1);
                }
            }
        }
        const float i0 = 1.0f / l0, i1 = 1.0f / l1;
        const int r0 = q0 + g, r1 = q0 + g + 8;
        __nv_bfloat16* o0 = out + ((size_t)(b * T + r0) * H + h) * HD;
        __nv_bfloat16* o1 = out + ((size_t)(b * T + r1) * H + h) * HD;
#pragma unroll
        for (int dn = 0; dn < HD / 8; ++dn) {
            if (r0 < T) *reinterpret_cast<uint32_t*>(o0 + dn * 8 + 2 * t4) = pack_bf16(o[dn][0] * i0, o[dn][1] * i0);
            if (r1 < T) *reinterpret_cast<uint32_t*>(o1 + dn * 8 + 2 * t4) = pack_bf16(o[dn][2] * i1, o[dn][3] * i1);
        }
        if (t4 == 0) {
            if (r0 < T) lse[((size_t)b * H + h) * T + r0] = m0 * scale + logf(l0);
            if (r1 < T) lse[((size_t)b * H + h) * T + r1] = m1 * scale + logf(l1);
        }
    }
}

// =========================================================================================================
// backward
// =========================================================================================================
template <int HD, int NT, int NWARPS>
__global__ void __launch_bounds__(NWARPS * 32)
attention_bwd_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ dout,
                     const float* __restrict__ lse, __nv_bfloat16* __restrict__ dqkv, int T, int H, float scale) {
    extern __shared__ __align__(16) uint8_t smem_attn[];
    constexpr int Tpad = NT * 8;
    constexpr int ldr = AttnSmem<HD>::ldr;
    const int ldt = AttnSmem<HD>::ldt(Tpad);
    __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smem_attn);
    __nv_bfloat16* sK = sQ + Tpad * ldr;
    __nv_bfloat16* sV = sK + Tpad * ldr;
    __nv_bfloat16* sdO = sV + Tpad * ldr;
    __nv_bfloat16* sQt = sdO + Tpad * ldr;
    __nv_bfloat16* sKt = sQt + HD * ldt;
    __nv_bfloat16* sdOt = sKt + HD * ldt;
    float* sLse = reinterpret_cast<float*>(sdOt + HD * ldt);
    float* sD = sLse + Tpad;

    const int b = blockIdx.x / H, h = blockIdx.x % H;
    const size_t row_stride = (size_t)3 * H * HD;
    const __nv_bfloat16* base = qkv + (size_t)b * T * row_stride + (size_t)h * HD;
    const size_t o_stride = (size_t)H * HD;
    const __nv_bfloat16* obase = out + (size_t)b * T * o_stride + (size_t)h * HD;
    const __nv_bfloat16* dobase = dout + (size_t)b * T * o_stride + (size_t)h * HD;
    load_head<HD>(base, row_stride, T, Tpad, sQ, sQt);
    load_head<HD>(base + (size_t)H * HD, row_stride, T, Tpad, sK, sKt);
    load_head<HD>(base + (size_t)2 * H * HD, row_stride, T, Tpad, sV, nullptr);
    load_head<HD>(dobase, o_stride, T, Tpad, sdO, sdOt);
    // D_i = sum_d dO[i][d] * O[i][d]; HD/8 consecutive threads share a row
    {
        constexpr int CH = HD / 8;
        for (int idx = threadIdx.x; idx < Tpad * CH; idx += blockDim.x) {
            const int t = idx / CH, ch = idx % CH;
            float acc = 0.f;
            if (t < T) {
                const uint4 a = *reinterpret_cast<const uint4*>(obase + (size_t)t * o_stride + ch * 8);
                const uint4 d = *reinterpret_cast<const uint4*>(dobase + (size_t)t * o_stride + ch * 8);
                const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, dw[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) { const float2 x = unpack_bf16(aw[j]), y = unpack_bf16(dw[j]); acc += x.x * y.x + x.y * y.y; }
            }
#pragma unroll
            for (int o = CH / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (ch == 0) { sD[t] = acc; sLse[t] = t < T ? lse[((size_t)b * H + h) * T + t] : 0.f; }
        }
    }
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = lane_id(), g = lane >> 2, t4 = lane & 3;
    const int tiles = (T + 15) / 16;
    const int chunks = (T + 63) / 64;                          // 64-wide chunks of the "other" dimension
    const float sl2 = scale * kLog2e;
    __nv_bfloat16* dq_base = dqkv + (size_t)b * T * row_stride + (size_t)h * HD;
    __nv_bfloat16* dk_base = dq_base + (size_t)H * HD;
    __nv_bfloat16* dv_base = dq_base + (size_t)2 * H * HD;

    // ---------------- pass A: dQ, one 16-query tile per warp iteration ----------------
    for (int qt = warp; qt < tiles; qt += NWARPS) {
        const int q0 = qt * 16;
        uint32_t aq[HD / 16][4], ado[HD / 16][4];
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) { load_a_frag(aq[ks], sQ, ldr, q0, ks * 16); load_a_frag(ado[ks], sdO, ldr, q0, ks * 16); }
        const float lse0 = sLse[q0 + g] * kLog2e, lse1 = sLse[q0 + g + 8] * kLog2e;
        const float D0 = sD[q0 + g], D1 = sD[q0 + g + 8];
        float dq[HD / 8][4];
#pragma unroll
        for (int dn = 0; dn < HD / 8; ++dn) dq[dn][0] = dq[dn][1] = dq[dn][2] = dq[dn][3] = 0.f;
        for (int kc = 0; kc < chunks; ++kc) {
            const int k0 = kc * 64;
            float s[8][4], dp[8][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
                dp[j][0] = dp[j][1] = dp[j][2] = dp[j][3] = 0.f;
                if (k0 + j * 8 < Tpad) {
#pragma unroll
                    for (int ks = 0; ks < HD / 16; ++ks) {
                        uint32_t b0, b1;
                        load_b_frag(b0, b1, sK, ldr, k0 + j * 8, ks * 16);
                        mma_bf16(s[j], aq[ks], b0, b1);
                        load_b_frag(b0, b1, sV, ldr, k0 + j * 8, ks * 16);
                        mma_bf16(dp[j], ado[ks], b0, b1);
                    }
                }
            }
            // dS = P * (dP - D) * scale  (P = exp(S*scale - lse)), packed straight into A fragments
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                if (k0 + jj * 16 < Tpad) {
                    uint32_t ads[4];
                    float e[2][4];
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const int j = 2 * jj + u;
                        const int c = k0 + j * 8 + 2 * t4;
                        const float p0 = c < T ? exp2f(s[j][0] * sl2 - lse0) : 0.f;
                        const float p1 = c + 1 < T ? exp2f(s[j][1] * sl2 - lse0) : 0.f;
                        const float p2 = c < T ? exp2f(s[j][2] * sl2 - lse1) : 0.f;
                        const float p3 = c + 1 < T ? exp2f(s[j][3] * sl2 - lse1) : 0.f;
                        e[u][0] = p0 * (dp[j][0] - D0) * scale; e[u][1] = p1 * (dp[j][1] - D0) * scale;
                        e[u][2] = p2 * (dp[j][2] - D1) * scale; e[u][3] = p3 * (dp[j][3] - D1) * scale;
                    }
                    ads[0] = pack_bf16(e[0][0], e[0][1]); ads[1] = pack_bf16(e[0][2], e[0][3]);
                    ads[2] = pack_bf16(e[1][0], e[1][1]); ads[3] = pack_bf16(e[1][2], e[1][3]);
#pragma unroll
                    for (int dn = 0; dn < HD / 8; ++dn) {
                        uint32_t b0, b1;
                        load_b_frag(b0, b1, sKt, ldt, dn * 8, k0 + jj * 16);
                        mma_bf16(dq[dn], ads, b0, b1);
                    }
                }
            }
        }
        const int r0 = q0 + g, r1 = q0 + g + 8;
#pragma unroll
        for (int dn = 0; dn < HD / 8; ++dn) {
            if (r0 < T) *reinterpret_cast<uint32_t*>(dq_base + (size_t)r0 * row_stride + dn * 8 + 2 * t4) = pack_bf16(dq[dn][0], dq[dn][1]);
            if (r1 < T) *reinterpret_cast<uint32_t*>(dq_base + (size_t)r1 * row_stride + dn * 8 + 2 * t4) = pack_bf16(dq[dn][2], dq[dn][3]);
        }
    }

    // ---------------- pass B: dK, dV, one 16-key tile per warp iteration (transposed problem) ----------------
    for (int kt = warp; kt < tiles; kt += NWARPS) {
        const int k0 = kt * 16;
        uint32_t ak[HD / 16][4], av[HD / 16][4];
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) { load_a_frag(ak[ks], sK, ldr, k0, ks * 16); load_a_frag(av[ks], sV, ldr, k0, ks * 16); }
        const bool key0_ok = k0 + g < T, key1_ok = k0 + g + 8 < T;
        float dk[HD / 8][4], dv[HD / 8][4];
#pragma unroll
        for (int dn = 0; dn < HD / 8; ++dn) { dk[dn][0] = dk[dn][1] = dk[dn][2] = dk[dn][3] = 0.f; dv[dn][0] = dv[dn][1] = dv[dn][2] = dv[dn][3] = 0.f; }
        for (int qc = 0; qc < chunks; ++qc) {
            const int q0 = qc * 64;
            float st[8][4], dpt[8][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                st[j][0] = st[j][1] = st[j][2] = st[j][3] = 0.f;
                dpt[j][0] = dpt[j][1] = dpt[j][2] = dpt[j][3] = 0.f;
                if (q0 + j * 8 < Tpad) {
#pragma unroll
                    for (int ks = 0; ks < HD / 16; ++ks) {
                        uint32_t b0, b1;
                        load_b_frag(b0, b1, sQ, ldr, q0 + j * 8, ks * 16);
                        mma_bf16(st[j], ak[ks], b0, b1);
                        load_b_frag(b0, b1, sdO, ldr, q0 + j * 8, ks * 16);
                        mma_bf16(dpt[j], av[ks], b0, b1);
                    }
                }
            }
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                if (q0 + jj * 16 < Tpad) {
                    uint32_t ap[4], ads[4];
                    float pv[2][4], ev[2][4];
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const int j = 2 * jj + u;
                        const int c = q0 + j * 8 + 2 * t4;                 // query index of columns c, c+1
                        const float la = sLse[c] * kLog2e, lb = sLse[c + 1] * kLog2e;
                        const float Da = sD[c], Db = sD[c + 1];
                        const bool qa = c < T, qb = c + 1 < T;
                        pv[u][0] = (qa && key0_ok) ? exp2f(st[j][0] * sl2 - la) : 0.f;
                        pv[u][1] = (qb && key0_ok) ? exp2f(st[j][1] * sl2 - lb) : 0.f;
                        pv[u][2] = (qa && key1_ok) ? exp2f(st[j][2] * sl2 - la) : 0.f;
                        pv[u][3] = (qb && key1_ok) ? exp2f(st[j][3] * sl2 - lb) : 0.f;
                        ev[u][0] = pv[u][0] * (dpt[j][0] - Da) * scale; ev[u][1] = pv[u][1] * (dpt[j][1] - Db) * scale;
                        ev[u][2] = pv[u][2] * (dpt[j][2] - Da) * scale; ev[u][3] = pv[u][3] * (dpt[j][3] - Db) * scale;
                    }
                    ap[0] = pack_bf16(pv[0][0], pv[0][1]); ap[1] = pack_bf16(pv[0][2], pv[0][3]);
                    ap[2] = pack_bf16(pv[1][0], pv[1][1]); ap[3] = pack_bf16(pv[1][2], pv[1][3]);
                    ads[0] = pack_bf16(ev[0][0], ev[0][1]); ads[1] = pack_bf16(ev[0][2], ev[0][3]);
                    ads[2] = pack_bf16(ev[1][0], ev[1][1]); ads[3] = pack_bf16(ev[1][2], ev[1][3]);
#pragma unroll
                    for (int dn = 0; dn < HD / 8; ++dn) {
                        uint32_t b0, b1;
                        load_b_frag(b0, b1, sdOt, ldt, dn * 8, q0 + jj * 16);
                        mma_bf16(dv[dn], ap, b0, b1);
                        load_b_frag(b0, b1, sQt, ldt, dn * 8, q0 + jj * 16);
                        mma_bf16(dk[dn], ads, b0, b1);
                    }
                }
            }
        }
        const int r0 = k0 + g, r1 = k0 + g + 8;
#pragma unroll
        for (int dn = 0; dn < HD / 8; ++dn) {
            if (r0 < T) {
                *reinterpret_cast<uint32_t*>(dk_base + (size_t)r0 * row_stride + dn * 8 + 2 * t4) = pack_bf16(dk[dn][0], dk[dn][1]);
                *reinterpret_cast<uint32_t*>(dv_base + (size_t)r0 * row_stride + dn * 8 + 2 * t4) = pack_bf16(dv[dn][0], dv[dn][1]);
            }
            if (r1 < T) {
                *reinterpret_cast<uint32_t*>(dk_base + (size_t)r1 * row_stride + dn * 8 + 2 * t4) = pack_bf16(dk[dn][2], dk[dn][3]);
                *reinterpret_cast<uint32_t*>(dv_base + (size_t)r1 * row_stride + dn * 8 + 2 * t4) = pack_bf16(dv[dn][2], dv[dn][3]);
            }
        }
    }
}

template <int HD, int NT>
static size_t fwd_smem() { return (size_t)(2 * NT * 8 * (HD + 8) + HD * (NT * 8 + 8)) * 2; }
template <int HD, int NT>
static size_t bwd_smem() { return (size_t)(4 * NT * 8 * (HD + 8) + 3 * HD * (NT * 8 + 8)) * 2 + 2 * NT * 8 * sizeof(float); }

template <int HD, int NT, int NWARPS>
static int launch_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, float scale, cudaStream_t s) {
    const size_t smem = fwd_smem<HD, NT>();
    auto kern = attention_fwd_kernel<HD, NT, NWARPS>;
    static bool attr = false;
    if (!attr) { SSL4P_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = true; }
    kern<<<B * H, NWARPS * 32, smem, s>>>((const __nv_bfloat16*)qkv, (__nv_bfloat16*)out, lse, T, H, scale);
    return check_launch("attention_fwd_kernel");
}
template <int HD, int NT, int NWARPS>
static int launch_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int B, int T, int H, float scale, cudaStream_t s) {
    const size_t smem = bwd_smem<HD, NT>();
    auto kern = attention_bwd_kernel<HD, NT, NWARPS>;
    static bool attr = false;
    if (!attr) { SSL4P_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = true; }
    kern<<<B * H, NWARPS * 32, smem, s>>>((const __nv_bfloat16*)qkv, (const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, lse,
                                          (__nv_bfloat16*)dqkv, T, H, scale);
    return check_launch("attention_bwd_kernel");
}

}  // namespace ssl4p

using namespace ssl4p;

extern "C" const char* ssl4p_check_last_error(void) { return g_check_err; }

static int attn_check(int B, int T, int H, int hd) {
    SSL4P_REQUIRE(B > 0 && H > 0 && T > 0, "attention: bad shape");
    SSL4P_REQUIRE(hd == 32 || hd == 64, "attention: head_dim %d not supported (32 or 64)", hd);
    SSL4P_REQUIRE(T <= 208, "attention: sequence length %d not supported (<= 208)", T);
    return SSL4P_OK;
}

extern "C" int ssl4p_attention_fwd_mma(const void* qkv, void* out, float* lse, int B, int T, int H, int hd, float scale, void* stream) {
    SSL4P_REQUIRE(qkv && out && lse, "attention_fwd: null pointer");
    int rc = attn_check(B, T, H, hd);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (T <= 64) {
        if (hd == 64) return launch_fwd<64, 8, 4>(qkv, out, lse, B, T, H, scale, s);
        return launch_fwd<32, 8, 4>(qkv, out, lse, B, T, H, scale, s);
    }
    if (hd == 64) return launch_fwd<64, 26, 4>(qkv, out, lse, B, T, H, scale, s);
    return launch_fwd<32, 26, 4>(qkv, out, lse, B, T, H, scale, s);
}

extern "C" int ssl4p_attention_bwd_mma(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                   int B, int T, int H, int hd, float scale, void* stream) {
    SSL4P_REQUIRE(qkv && out && dout && lse && dqkv, "attention_bwd: null pointer");
    int rc = attn_check(B, T, H, hd);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (T <= 64) {
        if (hd == 64) return launch_bwd<64, 8, 4>(qkv, out, dout, lse, dqkv, B, T, H, scale, s);
        return launch_bwd<32, 8, 4>(qkv, out, dout, lse, dqkv, B, T, H, scale, s);
    }
    if (hd == 64) return launch_bwd<64, 26, 8>(qkv, out, dout, lse, dqkv, B, T, H, scale, s);
    return launch_bwd<32, 26, 8>(qkv, out, dout, lse, dqkv, B, T, H, scale, s);
}
