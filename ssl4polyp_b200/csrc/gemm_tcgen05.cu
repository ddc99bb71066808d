// bf16 GEMM on the 5th-gen tensor cores: TMA (SWIZZLE_128B) -> shared memory -> tcgen05.mma (fp32
// accumulators in TMEM, double buffered) -> tcgen05.ld epilogue with bias / GELU / residual fusion.
//
// Replaces the cuBLAS `addmm` calls behind every nn.Linear of timm 0.4.12's Block/Attention/Mlp as used by
// /root/reference/src/ssl4polyp/models/mae/models_mae.py:39-41,53-55,47,58 (SURVEY §2d K1,K9,K11-K14,K16) and
// their autograd backward (dgrad = NN layout, wgrad = TN layout with deterministic split-K).
//
// Persistent kernel, one CTA per SM, 320 threads:
//   warp 0      : TMA producer (one elected lane)
//   warp 1      : TMEM allocator + tcgen05.mma issuer (one elected lane)
//   warps 2..9  : epilogue; warp w owns TMEM lanes 32*(w%4) .. 32*(w%4)+31  (= 32 rows of the 128-row tile); the two
//                 warps that share a lane quarter take alternate 32-column chunks.  Eight warps (two per scheduler)
//                 because at K = 512..768 the epilogue (bias/GELU/residual + stores) costs as much issue time as the
//                 main loop: it has to run on all four schedulers with a second warp to hide tcgen05.ld/erf latency.
// Pipelines: smem full/empty ring (kStages), TMEM full/empty (2 accumulators).
#include "common.cuh"
#include "ssl4p.h"
#include <mutex>
#include <stdlib.h>
#include <unordered_map>

namespace ssl4p {

constexpr int BM = 128;          // tile rows (UMMA M, cta_group::1)
constexpr int BK = 64;           // K per stage: 64 bf16 = 128 B = one SWIZZLE_128B row
constexpr int UMMA_K = 16;
// Epilogue warps: 16 for the GELU / dGELU epilogues, 8 otherwise.  At K = 512..768 every part of those two epilogues
// (tcgen05.ld, transpose, erf math or the gelu' operand load, stores) adds to a per-warp chain longer than the main loop of
// the tile (scripts/gemm_dbg_sweep.sh); splitting it over twice the warps gives 144 -> 130 us and 140 -> 119 us at the
// decoder shape.  The price is 64 KB of transpose patches (4 KB per warp), i.e. one pipeline stage less, which costs the
// plain bf16 epilogues at K >= 1536 more than they gain - they stay at 8 warps.
__host__ __device__ constexpr bool epi_out_f32(int epi) {
    return epi == SSL4P_EPI_BIAS_RESIDUAL_F32 || epi == SSL4P_EPI_BIAS_F32 || epi == SSL4P_EPI_PARTIAL_F32;
}
__host__ __device__ constexpr int epi_warps(int epi) { return (epi == SSL4P_EPI_BIAS_GELU_BF16 || epi == SSL4P_EPI_DGELU_BF16) ? 16 : 8; }
__host__ __device__ constexpr int gemm_threads(int epi) { return 64 + 32 * epi_warps(epi); }

struct GemmParams {
    int M, N, K;                 // C[M,N] = sum_k A[m,k] B[n,k]
    int ldc, ldc2, ldaux;        // leading dimensions (elements) of C, C2 and aux
    int k_per_split;             // K range per split (multiple of BK); split index = tile / (mt*nt)
    int splits;
    void* C;                     // primary output
    void* C2;                    // secondary output (GELU: activated copy)
    const float* bias;           // [N] or null
    const void* aux;             // residual (fp32 [M,ldaux]) or pre-activation (bf16 [M,ldaux])
    long long split_stride;      // elements between split-K partial planes of C
    int resid_prefetch;          // RESIDUAL epilogue: bulk-prefetch the tile's residual rows into L2 (K <= 1024; SSL4P_GEMM_RESID_PREFETCH)
    int* sched;                  // dynamic tile scheduler slot {next tile, finished units} (zero between launches) or null = static
};

template <int BN, bool A_MN, bool B_MN, int EPI, int CG>
struct SmemLayout {
    static constexpr int kABytes = BM * BK * 2;                              // this CTA's 128 rows of the (128 * CG)-row tile
    static constexpr int kBRows = BN / CG;                                   // N elements of the tile this CTA stages (CG = 2: its half)
    static constexpr int kBRowsLoad = B_MN ? ((kBRows + 63) / 64) * 64 : kBRows;   // MN-major boxes are 64 elements wide
    static constexpr int kBBytes = kBRowsLoad * BK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kEpiBytes = epi_warps(EPI) * 32 * 32 * 4;     // per-warp transpose patches of the epilogue
    static constexpr int kBudget = 224 * 1024 - kEpiBytes;
    static constexpr int kStagesRaw = kBudget / kStageBytes;
    static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
    static constexpr int kBarBytes = 1024;
    static constexpr int kTotal = kStages * kStageBytes + kBarBytes + kEpiBytes + 1024 /*align slack*/;
};

// ---- epilogue ------------------------------------------------------------------------------------
// tcgen05.ld hands every thread ONE row x 32 columns.  Written to global memory like that, a warp-wide 16-byte access
// touches 32 different 128-byte lines (one L1 tag cycle each), and at K = 512..768 those accesses cost more than the
// main loop.  So each warp first transposes its 32x32 fp32 block through a private 4 KB shared-memory patch (XOR
// swizzle, conflict free both ways) and then touches global memory in row-contiguous 512-byte warp accesses:
// fp32 outputs 4 rows x 128 B per instruction, bf16 outputs 8 rows x 64 B.
constexpr int kStageBytesPerWarp = 32 * 32 * 4;

__device__ __forceinline__ float4 f4_add(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 shfl_f4(float4 v, int src) {
    return make_float4(__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src), __shfl_sync(0xffffffffu, v.z, src),
                       __shfl_sync(0xffffffffu, v.w, src));
}
// The bias of the <= 4 column chunks a warp handles in one tile (4 x 32 floats), one float4 per lane: lane 8k + j holds
// columns [n0 + chunk0 + 64 k + 4 j, +4).  Loaded BEFORE the wait for the accumulator, so its latency is never exposed.
__device__ __forceinline__ float4 load_bias_lane(const GemmParams& p, int n0, int chunk0, int chunk_step) {
    const int lane = (int)lane_id();
    const int col = n0 + chunk0 + (lane >> 3) * chunk_step + (lane & 7) * 4;
    if (p.bias == nullptr || col >= p.N) return make_float4(0.f, 0.f, 0.f, 0.f);
    return __ldg(reinterpret_cast<const float4*>(p.bias + col));
}

// The second epilogue operand of a chunk (fp32 residual: 8 x 16 B per lane; saved gelu': 4 x 16 B per lane), in the
// transposed (row-contiguous) mapping.  Loaded one chunk ahead so that its HBM latency hides behind the previous chunk.
struct AuxRegs { uint4 v[8]; };

template <int EPI>
__device__ __forceinline__ void load_aux(const GemmParams& p, int row_base, int col0, AuxRegs& a) {
    const int lane = (int)lane_id();
    if constexpr (EPI == SSL4P_EPI_BIAS_RESIDUAL_F32) {
        const int col = col0 + (lane & 7) * 4;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int row = row_base + 4 * i + (lane >> 3);
            a.v[i] = (row < p.M && col < p.N) ? ldg_v4(reinterpret_cast<const float*>(p.aux) + (size_t)row * p.ldaux + col)
                                              : make_uint4(0u, 0u, 0u, 0u);
        }
    } else if constexpr (EPI == SSL4P_EPI_DGELU_BF16) {
        const int col = col0 + (lane & 3) * 8;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = row_base + 8 * i + (lane >> 2);
            a.v[i] = (row < p.M && col < p.N) ? ldg_v4(reinterpret_cast<const __nv_bfloat16*>(p.aux) + (size_t)row * p.ldaux + col)
                                              : make_uint4(0u, 0u, 0u, 0u);
        }
    }
}

// stg: shared-space address of this warp's 4 KB transpose patch.
template <int EPI>
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, uint32_t stg, int row_base, int col0, int split, uint32_t (&acc)[32],
                                               const AuxRegs& aux, const float4 bias_lane, const int kchunk) {
    const int lane = (int)lane_id();
    constexpr bool kHasBias = EPI == SSL4P_EPI_BIAS_BF16 || EPI == SSL4P_EPI_BIAS_GELU_BF16 || EPI == SSL4P_EPI_BIAS_RESIDUAL_F32 ||
                              EPI == SSL4P_EPI_BIAS_F32;
    constexpr bool kOutF32 = EPI == SSL4P_EPI_BIAS_RESIDUAL_F32 || EPI == SSL4P_EPI_BIAS_F32 || EPI == SSL4P_EPI_PARTIAL_F32;

    // ---- transpose through the patch: every lane writes its row (32 columns), then reads 8 x 16 B in the
    // row-contiguous mapping.  All shared-memory reads are issued before any global store.
#pragma unroll
    for (int c = 0; c < 8; ++c)
        sts_v4(stg + lane * 128 + ((c ^ (lane & 7)) << 4), acc[4 * c], acc[4 * c + 1], acc[4 * c + 2], acc[4 * c + 3]);
    __syncwarp();
    float4 t[8];
    if constexpr (kOutF32) {
        const int cc = lane & 7;                                   // 16-byte chunk = 4 columns
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int rr = 4 * i + (lane >> 3);
            t[i] = lds_v4f(stg + rr * 128 + ((cc ^ (rr & 7)) << 4));
        }
    } else {
        const int cc = lane & 3;                                   // 16-byte bf16 chunk = 8 columns
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int rr = 8 * i + (lane >> 2);
            t[2 * i] = lds_v4f(stg + rr * 128 + (((2 * cc) ^ (rr & 7)) << 4));
            t[2 * i + 1] = lds_v4f(stg + rr * 128 + (((2 * cc + 1) ^ (rr & 7)) << 4));
        }
    }
    __syncwarp();                                                  // the patch may be rewritten by the next chunk

    if constexpr (kOutF32) {
        const int col = col0 + (lane & 7) * 4;
        const bool col_ok = col < p.N;                             // N % 8 == 0 (checked on the host)
        float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
        if constexpr (kHasBias) b = shfl_f4(bias_lane, kchunk * 8 + (lane & 7));
        float* outp = reinterpret_cast<float*>(p.C) + (EPI == SSL4P_EPI_PARTIAL_F32 ? (size_t)split * p.split_stride : 0);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int row = row_base + 4 * i + (lane >> 3);
            float4 v = f4_add(t[i], b);
            if constexpr (EPI == SSL4P_EPI_BIAS_RESIDUAL_F32)
                v = f4_add(v, make_float4(__uint_as_float(aux.v[i].x), __uint_as_float(aux.v[i].y), __uint_as_float(aux.v[i].z), __uint_as_float(aux.v[i].w)));
            if (row < p.M && col_ok) stg_v4f(outp + (size_t)row * p.ldc + col, v);
        }
    } else {
        const int col = col0 + (lane & 3) * 8;
        const bool col_ok = col < p.N;
        float4 b0 = make_float4(0.f, 0.f, 0.f, 0.f), b1 = b0;
        if constexpr (kHasBias) {
            b0 = shfl_f4(bias_lane, kchunk * 8 + 2 * (lane & 3));
            b1 = shfl_f4(bias_lane, kchunk * 8 + 2 * (lane & 3) + 1);
        }
        float csum[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};   // DGELU: column sums of this warp's 32 rows (fc1 bias gradient)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = row_base + 8 * i + (lane >> 2);
            const float4 lo = t[2 * i], hi = t[2 * i + 1];
            float v[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
            if constexpr (kHasBias) {                              // (x + 0.0f is not a no-op for the compiler: -0.0f)
                v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w; v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
            }
            if constexpr (EPI == SSL4P_EPI_DGELU_BF16) {
                const uint32_t hw[4] = {aux.v[i].x, aux.v[i].y, aux.v[i].z, aux.v[i].w};
#pragma unroll
                for (int q = 0; q < 4; ++q) { const float2 hf = unpack_bf16(hw[q]); v[2 * q] *= hf.x; v[2 * q + 1] *= hf.y; }
            }
            const bool ok = row < p.M && col_ok;
            if constexpr (EPI == SSL4P_EPI_DGELU_BF16) {
#pragma unroll
                for (int j = 0; j < 8; ++j) csum[j] += ok ? v[j] : 0.f;
            }
            if constexpr (EPI == SSL4P_EPI_BIAS_GELU_BF16) {
                // C = gelu'(h) (what the backward multiplies by), C2 = gelu(h): one erf evaluation serves both
                float g[8];
#pragma unroll
                for (int j = 0; j < 8; j += 2) {
                    float a0, a1, d0, d1;
                    gelu_and_grad_f2(v[j], v[j + 1], a0, a1, d0, d1);
                    g[j] = a0; g[j + 1] = a1; v[j] = d0; v[j + 1] = d1;
                }
                if (ok) stg_v4(reinterpret_cast<__nv_bfloat16*>(p.C2) + (size_t)row * p.ldc2 + col,
                               pack_bf16(g[0], g[1]), pack_bf16(g[2], g[3]), pack_bf16(g[4], g[5]), pack_bf16(g[6], g[7]));
            }
            if (ok) stg_v4(reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col,
                           pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
        }
        if constexpr (EPI == SSL4P_EPI_DGELU_BF16) {
            if (p.C2 != nullptr) {                                 // fixed-order butterfly over the 8 lanes that share a column group
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    csum[j] += __shfl_xor_sync(0xffffffffu, csum[j], 4);
                    csum[j] += __shfl_xor_sync(0xffffffffu, csum[j], 8);
                    csum[j] += __shfl_xor_sync(0xffffffffu, csum[j], 16);
                }
                if (lane < 4 && col_ok && row_base < p.M) {
                    float* dst = reinterpret_cast<float*>(p.C2) + (size_t)(row_base >> 5) * p.N + col;
                    stg_v4f(dst, make_float4(csum[0], csum[1], csum[2], csum[3]));
                    stg_v4f(dst + 4, make_float4(csum[4], csum[5], csum[6], csum[7]));
                }
            }
        }
    }
}

// TMA store of one [32 rows x 32 bf16] box from shared memory (SWIZZLE_64B layout) to the output; rows / columns
// beyond the tensor are clipped by the hardware
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, uint32_t smem_src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_src), "r"(c0), "r"(c1) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// EPI_BIAS_BF16 through TMA: the lane keeps its row (the tcgen05.ld layout), adds the bias, packs to bf16 and writes 64
// bytes into the [32 x 64 B] patch (16-byte chunk c of row r at c ^ ((r >> 1) & 3) = SWIZZLE_64B); one lane hands the
// box to the TMA unit.  No read-back, no per-thread global stores.
__device__ __forceinline__ void epilogue_chunk_tma(const CUtensorMap* tmC, const GemmParams& p, uint32_t stg, int row_base, int col0,
                                                   uint32_t (&acc)[32], const float4 bias_lane, const int kchunk) {
    const int lane = (int)lane_id();
    uint32_t w[16];
    if (p.bias != nullptr) {
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
            const float4 b = shfl_f4(bias_lane, kchunk * 8 + j4);
            w[2 * j4] = pack_bf16(__uint_as_float(acc[4 * j4]) + b.x, __uint_as_float(acc[4 * j4 + 1]) + b.y);
            w[2 * j4 + 1] = pack_bf16(__uint_as_float(acc[4 * j4 + 2]) + b.z, __uint_as_float(acc[4 * j4 + 3]) + b.w);
        }
    } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) w[j] = pack_bf16(__uint_as_float(acc[2 * j]), __uint_as_float(acc[2 * j + 1]));
    }
    if (lane == 0) tma_store_wait_read();                          // the previous box of this patch has left shared memory
    __syncwarp();
#pragma unroll
    for (int c = 0; c < 4; ++c)
        sts_v4(stg + lane * 64 + ((c ^ ((lane >> 1) & 3)) << 4), w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
    fence_proxy_async();
    __syncwarp();
    if (lane == 0) tma_store_2d(tmC, stg, col0, row_base);
}

// ---- the kernel ---------------------------------------------------------------------------------
// CG = 1: one CTA per 128 x BN tile (cta_group::1).  CG = 2: a CTA pair (cluster of 2 = one TPC) per 256 x BN tile
// (cta_group::2): CTA r owns rows [128 r, 128 r + 128) of the tile (its A rows, its TMEM accumulator, its epilogue) and
// stages columns [r BN/2, (r+1) BN/2) of B; rank 0 issues the MMAs, the full barriers live in rank 0, every commit is
// multicast to both CTAs.
template <int BN, bool A_MN, bool B_MN, int EPI, int CG>
__global__ void __launch_bounds__(gemm_threads(EPI), 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
    using L = SmemLayout<BN, A_MN, B_MN, EPI, CG>;
    constexpr int TM = BM * CG;                               // tile rows
    constexpr int kStages = L::kStages;
    constexpr int kEpiWarps = epi_warps(EPI);
    constexpr int kChunkStep = (kEpiWarps / 4) * 32;          // columns between two chunks of the same warp
    constexpr uint32_t kTmemCols = 512;                      // 2 accumulators x up to 256 columns
    constexpr uint32_t kAccStride = 256;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_a = smem;                                  // kStages x 16 KB
    uint8_t* smem_b = smem + kStages * L::kABytes;           // kStages x BN*128 B
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * L::kStageBytes);
    uint64_t* full_bar = bars;                               // [kStages]
    uint64_t* empty_bar = bars + kStages;                    // [kStages]
    uint64_t* tmem_full = bars + 2 * kStages;                // [2]
    uint64_t* tmem_empty = bars + 2 * kStages + 2;           // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);
    // Tile queue: ONE thread per unit (the producer of the CTA / of the pair's leader) takes tiles off a global counter and
    // hands them to every other role through this ring.  A persistent grid with a static tile list stalls whenever some of
    // its CTAs cannot be placed - the NCCL all-reduce kernels of the overlapped gradient exchange hold SMs for ~150 us and a
    // CTA that needs a whole SM's shared memory cannot share one with them: the late CTAs then run their full share as a
    // second wave.  With the counter a late CTA finds the queue drained and leaves at once.
    constexpr int kTQ = 4;
    uint64_t* tq_full = bars + 2 * kStages + 5;              // [kTQ]
    uint64_t* tq_empty = bars + 2 * kStages + 5 + kTQ;       // [kTQ] (CG = 2: the leader's copy is the one in use)
    volatile int* tq_tile = reinterpret_cast<volatile int*>(bars + 2 * kStages + 5 + 2 * kTQ);   // [kTQ]
    uint8_t* epi_stage = smem + kStages * L::kStageBytes + L::kBarBytes;   // kEpiWarps x 4 KB

    const int warp = threadIdx.x >> 5;
    const int mt = (p.M + TM - 1) / TM;
    const int nt = (p.N + BN - 1) / BN;
    const int tiles_mn = mt * nt;
    const int total_tiles = tiles_mn * p.splits;
    uint32_t rank = 0;                                        // rank inside the CTA pair
    if constexpr (CG == 2) rank = cluster_ctarank();
    const int unit = CG == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;       // tile-processing unit (CTA or CTA pair)
    const int nunits = CG == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;

    int first_tile = unit;                                    // scheduler thread only
    if (warp == 0 && lane_id() == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        // the first tile comes off the counter too (a CTA that starts late must not own any tile); the round trip hides
        // behind the barrier initialisation and the TMEM allocation below
        if (p.sched != nullptr && rank == 0) first_tile = atomicAdd(p.sched, 1);
    }
    if (warp == 1) {
        if (lane_id() == 0) {
            for (int i = 0; i < kStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
            for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], kEpiWarps * CG); }
            // consumers of a queue slot: the MMA warp and the epilogue warps (CG = 2: of both CTAs, plus the peer's producer)
            for (int i = 0; i < kTQ; ++i) { mbar_init(&tq_full[i], 1); mbar_init(&tq_empty[i], CG == 2 ? 2 + 2 * kEpiWarps : 1 + kEpiWarps); }
            mbar_fence_init();
            fence_proxy_async();
        }
        __syncwarp();
        if constexpr (CG == 2) tmem_alloc_cg2<kTmemCols>(tmem_slot); else tmem_alloc<kTmemCols>(tmem_slot);
    }
    tc_fence_before();
    if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();      // barriers of BOTH CTAs initialised before any TMA / commit
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait_then_trigger();                                  // everything above overlapped the previous kernel's tail

    // next tile of this unit for a consumer role (-1 = no more); `single`: called by one thread instead of a whole warp
    int tq_slot = 0; uint32_t tq_phase = 0;
    auto next_tile = [&](bool single) -> int {
        if (CG == 2 && rank != 0) mbar_wait_cluster(&tq_full[tq_slot], tq_phase); else mbar_wait(&tq_full[tq_slot], tq_phase);
        const int t = tq_tile[tq_slot];
        if (!single) __syncwarp();
        if (single || lane_id() == 0) {
            if constexpr (CG == 2) mbar_arrive_cluster(mapa_u32(smem_u32(&tq_empty[tq_slot]), 0));
            else mbar_arrive(&tq_empty[tq_slot]);
        }
        if (++tq_slot == kTQ) { tq_slot = 0; tq_phase ^= 1; }
        return t;
    };

    if (warp == 0) {
        // ===================== TMA producer (and, in the unit's leader CTA, the tile scheduler) =====================
        if (lane_id() == 0) {
            int stage = 0; uint32_t phase = 0;
            // the barrier that counts this CTA's bytes: its own (CG = 1) or the leader's (CG = 2)
            auto load = [&](void* dst, const CUtensorMap* m, int st, int c0, int c1) {
                if constexpr (CG == 2) tma_load_2d_cg2(dst, m, mapa_u32(smem_u32(&full_bar[st]), 0), c0, c1);
                else tma_load_2d(dst, m, &full_bar[st], c0, c1);
            };
            const bool scheduler = CG == 1 || rank == 0;
            // scheduler: the request for tile i+1 goes out before the loads of tile i and its answer is published from
            // INSIDE that tile's k loop (after the third k-block), never at the tile boundary: the peer CTA's producer needs
            // the index the moment the leader's producer turns to the next tile (a hand-off at the boundary cost ~0.5 us per
            // tile).  One tile of look-ahead only: a unit that asked for two tiles up front would starve its neighbours in
            // the one-wave split-K launches.
            auto publish = [&](int t) {
                const int pub = t < total_tiles ? t : -1;
                // every consumer (of both CTAs) has read the slot's last tile
                if constexpr (CG == 2) mbar_wait_cluster(&tq_empty[tq_slot], tq_phase ^ 1); else mbar_wait(&tq_empty[tq_slot], tq_phase ^ 1);
                tq_tile[tq_slot] = pub;
                if constexpr (CG == 2) {
                    st_cluster_u32(mapa_u32(smem_u32(const_cast<int*>(&tq_tile[tq_slot])), 1), (uint32_t)pub);
                    mbar_arrive_cluster(mapa_u32(smem_u32(&tq_full[tq_slot]), 1));
                }
                mbar_arrive(&tq_full[tq_slot]);
                if (++tq_slot == kTQ) { tq_slot = 0; tq_phase ^= 1; }
            };
            int tile = first_tile;
            if (scheduler) publish(tile);
            while (true) {
                int next = 0;
                if (scheduler) {
                    if (tile >= total_tiles) break;
                    next = p.sched != nullptr ? atomicAdd(p.sched, 1) : tile + nunits;
                } else {
                    tile = next_tile(true);
                    if (tile < 0) break;
                }
                const int split = tile / tiles_mn;
                const int t = tile - split * tiles_mn;
                const int m0 = (t / nt) * TM + (int)rank * BM;                         // this CTA's A rows
                const int n0 = (t % nt) * BN + (int)rank * L::kBRows;                  // this CTA's share of B
                const int k_begin = split * p.k_per_split;
                const int k_end = min(p.K, k_begin + p.k_per_split);
                const int k_pub = min(k_begin + 2 * BK, k_end - BK);                   // k-block after which `next` is published
                bool published = !scheduler;
                for (int k = k_begin; k < k_end; k += BK) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (CG == 1 || rank == 0) mbar_expect_tx(&full_bar[stage], L::kStageBytes * CG);   // both CTAs' bytes
                    uint8_t* sa = smem_a + stage * L::kABytes;
                    uint8_t* sb = smem_b + stage * L::kBBytes;
                    if constexpr (!A_MN) {
                        load(sa, &tmA, stage, k, m0);                                // box [64 k] x [128 rows]
                    } else {
#pragma unroll
                        for (int c = 0; c < BM / 64; ++c)                            // box [64 m] x [64 k rows]
                            load(sa + c * (BK * 128), &tmA, stage, m0 + c * 64, k);
                    }
                    if constexpr (!B_MN) {
                        load(sb, &tmB, stage, k, n0);                                // box [64 k] x [kBRows rows]
                    } else {
#pragma unroll
                        for (int c = 0; c < L::kBRowsLoad / 64; ++c)
                            load(sb + c * (BK * 128), &tmB, stage, n0 + c * 64, k);
                    }
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                    if (!published && k >= k_pub) { publish(next); published = true; }
                }
                if (!published) publish(next);                                        // (empty split-K range)
                tile = next;
            }
            if (scheduler && p.sched != nullptr) {
                // the unit that finishes last puts the slot back to zero for the next launch (every other unit has made its
                // final request by then)
                __threadfence();
                if (atomicAdd(p.sched + 1, 1) == nunits - 1) { atomicExch(p.sched, 0); atomicExch(p.sched + 1, 0); }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer (CG = 2: the leader CTA only) =====================
        if (CG == 1 || rank == 0) {
        constexpr uint32_t idesc = umma_idesc_bf16(TM, BN, A_MN, B_MN);
        // K-major SW128: 8-row groups 1024 B apart (SBO), LBO unused.  MN-major SW128: 64-element MN chunks
        // BK*128 B apart (LBO), 8-k-row groups 1024 B apart (SBO).
        constexpr uint32_t a_lbo = A_MN ? BK * 128 : 16, a_sbo = 1024;
        constexpr uint32_t b_lbo = B_MN ? BK * 128 : 16, b_sbo = 1024;
        constexpr uint32_t a_kstep = A_MN ? (UMMA_K / 8) * 1024 : UMMA_K * 2;   // bytes per UMMA_K step
        constexpr uint32_t b_kstep = B_MN ? (UMMA_K / 8) * 1024 : UMMA_K * 2;
        const bool leader = lane_id() == 0;     // tcgen05.commit tracks the issuing thread: one fixed lane issues all
        int stage = 0; uint32_t phase = 0;
        int acc = 0; uint32_t acc_phase = 0;
        while (true) {
            const int tile = next_tile(false);
            if (tile < 0) break;
            const int split = tile / tiles_mn;
            const int k_begin = split * p.k_per_split;
            const int k_end = min(p.K, k_begin + p.k_per_split);
            mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc * kAccStride;
            uint32_t accumulate = 0;
            for (int k = k_begin; k < k_end; k += BK) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                if (leader) {
                    const uint32_t sa = smem_u32(smem_a + stage * L::kABytes);
                    const uint32_t sb = smem_u32(smem_b + stage * L::kBBytes);
#pragma unroll
                    for (int kk = 0; kk < BK / UMMA_K; ++kk) {
                        const uint64_t adesc = umma_smem_desc(sa + kk * a_kstep, a_lbo, a_sbo);
                        const uint64_t bdesc = umma_smem_desc(sb + kk * b_kstep, b_lbo, b_sbo);
                        if constexpr (CG == 2) umma_bf16_cg2(tmem_d, adesc, bdesc, idesc, accumulate);
                        else umma_bf16(tmem_d, adesc, bdesc, idesc, accumulate);
                        accumulate = 1;
                    }
                    // frees the smem slot (in both CTAs) when the MMAs retire
                    if constexpr (CG == 2) umma_commit_cg2(&empty_bar[stage], 3); else umma_commit(&empty_bar[stage]);
                }
                __syncwarp();
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
            if (leader) {                                        // accumulator complete -> epilogue (of both CTAs)
                if constexpr (CG == 2) umma_commit_cg2(&tmem_full[acc], 3); else umma_commit(&tmem_full[acc]);
            }
            __syncwarp();
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        }
    } else {
        // ===================== epilogue warps =====================
        const int q = warp & 3;                                       // TMEM lane quarter this warp may access
        const int chunk0 = ((warp - 2) >> 2) * 32;                    // the quarter's two warps interleave 32-column chunks
        int acc = 0; uint32_t acc_phase = 0;
        while (true) {
            const int tile = next_tile(false);
            if (tile < 0) break;
            const int split = tile / tiles_mn;
            const int t = tile - split * tiles_mn;
            const int m0 = (t / nt) * TM + (int)rank * BM, n0 = (t % nt) * BN;
            const int row_base = m0 + q * 32;
            constexpr bool kAux = EPI == SSL4P_EPI_BIAS_RESIDUAL_F32 || EPI == SSL4P_EPI_DGELU_BF16;
            AuxRegs aux_cur, aux_nxt;
            // Second epilogue operand.  dGELU (16 warps: at most kNC = 2 chunks per warp and tile): ALL of the tile's gelu'
            // values of this warp are requested before the accumulator wait - with one chunk of look-ahead the load latency
            // was the critical chain of the kernel (117 us with everything else switched off).  Residual (8 warps, 4 chunks,
            // 32 registers per chunk): one chunk ahead.
            constexpr bool kAuxAll = EPI == SSL4P_EPI_DGELU_BF16;
            constexpr int kNC = (BN + kChunkStep - 1) / kChunkStep;
            AuxRegs aux_all[kAuxAll ? kNC : 1];
            if constexpr (kAuxAll) {
#pragma unroll
                for (int ci = 0; ci < kNC; ++ci) {
                    const int c = chunk0 + ci * kChunkStep;
                    if (c < BN && n0 + c < p.N) load_aux<EPI>(p, row_base, n0 + c, aux_all[ci]);
                }
            } else if constexpr (kAux) {
                // the residual of the whole tile row (up to 1 KB, contiguous) is pulled into L2 now, one row per lane, by the
                // first warp of each lane quarter: the chunk-by-chunk register loads below (one chunk of look-ahead) then pay
                // L2 latency, not HBM latency - they were the stall of this epilogue (ncu r02d: long scoreboard on aux)
                if (chunk0 == 0 && p.resid_prefetch) {
                    const int row = row_base + (int)lane_id();
                    const int ncols = min(BN, p.N - n0);
                    if (row < p.M && ncols > 0) {
                        const float* src = reinterpret_cast<const float*>(p.aux) + (size_t)row * p.ldaux + n0;
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(ncols * 4) : "memory");
                    }
                }
                load_aux<EPI>(p, row_base, n0 + chunk0, aux_cur);      // in flight while the MMAs of the tile finish
            }
            constexpr bool kBias = EPI == SSL4P_EPI_BIAS_BF16 || EPI == SSL4P_EPI_BIAS_GELU_BF16 || EPI == SSL4P_EPI_BIAS_RESIDUAL_F32 ||
                                   EPI == SSL4P_EPI_BIAS_F32;
            float4 bias_lane = make_float4(0.f, 0.f, 0.f, 0.f);
            if constexpr (kBias) bias_lane = load_bias_lane(p, n0, chunk0, kChunkStep);
            mbar_wait(&tmem_full[acc], acc_phase);
            tc_fence_after();
            const uint32_t stg = smem_u32(epi_stage) + (warp - 2) * kStageBytesPerWarp;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * kAccStride;
            const bool empty_split = split * p.k_per_split >= p.K;   // empty split-K range: the accumulator was never written
            if constexpr (kAuxAll) {
#pragma unroll
                for (int ci = 0; ci < kNC; ++ci) {
                    const int c = chunk0 + ci * kChunkStep;
                    const int col0 = n0 + c;
                    if (c < BN && col0 < p.N) {
                        uint32_t r[32];
                        tmem_ld_32x32(taddr + c, r);
                        tmem_ld_wait();
                        if (row_base < p.M) epilogue_chunk<EPI>(p, stg, row_base, col0, split, r, aux_all[ci], bias_lane, ci);
                    }
                }
            } else
#pragma unroll 1
            for (int c = chunk0; c < BN; c += kChunkStep) {
                const int col0 = n0 + c;
                if (col0 >= p.N) break;
                if constexpr (kAux) if (c + kChunkStep < BN) load_aux<EPI>(p, row_base, col0 + kChunkStep, aux_nxt);
                uint32_t r[32];
                tmem_ld_32x32(taddr + c, r);
                tmem_ld_wait();
                if (empty_split) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] = 0u;
                }
                if constexpr (EPI == SSL4P_EPI_BIAS_BF16) {
                    if (row_base < p.M) epilogue_chunk_tma(&tmC, p, stg, row_base, col0, r, bias_lane, (c - chunk0) / kChunkStep);
                } else
                if (row_base < p.M) epilogue_chunk<EPI>(p, stg, row_base, col0, split, r, aux_cur, bias_lane, (c - chunk0) / kChunkStep);
                if constexpr (kAux) aux_cur = aux_nxt;
            }
            tc_fence_before();
            __syncwarp();
            if (lane_id() == 0) {                                     // the MMA issuer (leader CTA) may overwrite this accumulator
                if constexpr (CG == 2) mbar_arrive_cluster(mapa_u32(smem_u32(&tmem_empty[acc]), 0));
                else mbar_arrive(&tmem_empty[acc]);
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    }

    if constexpr (EPI == SSL4P_EPI_BIAS_BF16) {
        if (warp >= 2 && lane_id() == 0) tma_store_wait_all();     // shared memory must outlive the bulk stores
    }
    tc_fence_before();
    // CG = 2: the leader's MMAs read the peer's shared memory and write its TMEM: nobody leaves before both are done
    if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        if constexpr (CG == 2) tmem_dealloc_cg2<kTmemCols>(tmem_base); else tmem_dealloc<kTmemCols>(tmem_base);
    }
}

// ---- host side ----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(f);
    });
    return fn;
}

// 2-D bf16 tensor map: `inner` contiguous elements per row, `rows` rows with pitch `ld` elements;
// box = 64 x box_rows, SWIZZLE_128B, zero fill out of bounds.
static int make_tmap(CUtensorMap* out, const void* ptr, long long inner, long long rows, long long ld, int box_rows) {
    struct Key { const void* p; long long a, b, c; int r; bool operator==(const Key& o) const { return p == o.p && a == o.a && b == o.b && c == o.c && r == o.r; } };
    struct Hash { size_t operator()(const Key& k) const { return std::hash<const void*>()(k.p) ^ (size_t)(k.a * 1315423911ull) ^ (size_t)(k.b * 2654435761ull) ^ (size_t)(k.c << 7) ^ (size_t)k.r; } };
    static thread_local std::unordered_map<Key, CUtensorMap, Hash> cache;
    Key key{ptr, inner, rows, ld, box_rows};
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return SSL4P_OK; }

    EncodeTiledFn enc = get_encode_fn();
    if (!enc) { set_error("cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)"); return SSL4P_ERR_CUDA; }
    SSL4P_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA operand pointer must be 16-byte aligned");
    SSL4P_REQUIRE((ld * 2) % 16 == 0, "TMA operand row pitch must be a multiple of 16 bytes (ld=%lld)", ld);
    cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (CUresult %d) inner=%lld rows=%lld ld=%lld box_rows=%d", (int)r, inner, rows, ld, box_rows); return SSL4P_ERR_CUDA; }
    if (cache.size() > 4096) cache.clear();
    cache.emplace(key, *out);
    return SSL4P_OK;
}

// bf16 output as a 2-D tensor map for TMA stores: box = 32 columns x 32 rows (one epilogue chunk), SWIZZLE_64B
static int make_tmap_out(CUtensorMap* out, void* ptr, long long cols, long long rows, long long ld) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) { set_error("cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)"); return SSL4P_ERR_CUDA; }
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {32u, 32u};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (output) failed (CUresult %d) cols=%lld rows=%lld ld=%lld", (int)r, cols, rows, ld); return SSL4P_ERR_CUDA; }
    return SSL4P_OK;
}

// CTA pairs that can be resident at once (cudaOccupancyMaxActiveClusters of the first pair kernel launched; 0 = not queried
// yet).  Not every board places 74 two-CTA clusters: SM harvesting can leave TPCs with a single SM.
static int g_max_pairs = 0;
static int pair_units() {
    const int by_sms = num_sms() / 2;
    return g_max_pairs > 0 && g_max_pairs < by_sms ? g_max_pairs : by_sms;
}
extern "C" int ssl4p_gemm_max_cta_pairs(void) { return g_max_pairs; }

template <int BN, bool A_MN, bool B_MN, int EPI, int CG>
static int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GemmParams& p, cudaStream_t stream) {
    using L = SmemLayout<BN, A_MN, B_MN, EPI, CG>;
    auto kern = gemm_tcgen05_kernel<BN, A_MN, B_MN, EPI, CG>;
    static bool attr_set = false;
    if (!attr_set) {
        SSL4P_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
        attr_set = true;
    }
    if (CG == 2 && g_max_pairs == 0) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2 * 74); cfg.blockDim = dim3((unsigned)gemm_threads(EPI)); cfg.dynamicSmemBytes = L::kTotal;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) == cudaSuccess && n > 0) g_max_pairs = n; else { cudaGetLastError(); g_max_pairs = 1 << 20; }
    }
    const int mt = (p.M + BM * CG - 1) / (BM * CG), nt = (p.N + BN - 1) / BN;
    const int total = mt * nt * p.splits;
    const int units = CG == 2 ? pair_units() : num_sms();      // CTAs or CTA pairs (one pair = the two SMs of a TPC)
    const int grid = (total < units ? total : units) * CG;
    cudaError_t e = launch_ex(kern, dim3((unsigned)grid), dim3((unsigned)gemm_threads(EPI)), (size_t)L::kTotal, stream, CG, tmA, tmB, tmC, p);
    if (e != cudaSuccess) { set_error("gemm_tcgen05_kernel launch failed: %s", cudaGetErrorString(e)); return SSL4P_ERR_CUDA; }
    return check_launch("gemm_tcgen05_kernel");
}

// Only the (layout, epilogue) pairs the model uses are instantiated.
template <int BN, int CG>
static int dispatch(int layout, int epi, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& c, const GemmParams& p, cudaStream_t s) {
    if (layout == SSL4P_GEMM_NT) {
        switch (epi) {
            case SSL4P_EPI_BIAS_BF16: return launch<BN, false, false, SSL4P_EPI_BIAS_BF16, CG>(a, b, c, p, s);
            case SSL4P_EPI_BIAS_GELU_BF16: return launch<BN, false, false, SSL4P_EPI_BIAS_GELU_BF16, CG>(a, b, c, p, s);
            case SSL4P_EPI_BIAS_RESIDUAL_F32: return launch<BN, false, false, SSL4P_EPI_BIAS_RESIDUAL_F32, CG>(a, b, c, p, s);
            case SSL4P_EPI_BIAS_F32: return launch<BN, false, false, SSL4P_EPI_BIAS_F32, CG>(a, b, c, p, s);
        }
    } else if (layout == SSL4P_GEMM_NN) {
        switch (epi) {
            case SSL4P_EPI_BIAS_BF16: return launch<BN, false, true, SSL4P_EPI_BIAS_BF16, CG>(a, b, c, p, s);
            case SSL4P_EPI_DGELU_BF16: return launch<BN, false, true, SSL4P_EPI_DGELU_BF16, CG>(a, b, c, p, s);
        }
    } else {
        switch (epi) {
            case SSL4P_EPI_PARTIAL_F32: return launch<BN, true, true, SSL4P_EPI_PARTIAL_F32, CG>(a, b, c, p, s);
        }
    }
    set_error("gemm: epilogue %d is not available for layout %d", epi, layout);
    return SSL4P_ERR_UNSUPPORTED;
}

// ---- dynamic tile scheduler slots (ssl4p_gemm_set_scheduler_buffer) ---------------------------------------------------
// A caller-owned, zero-initialised device buffer of 32-byte slots; every launch takes one ({next tile, finished units}, put
// back to zero by the kernel itself).  Launches recorded into a CUDA graph keep their slot for the life of the graph, so
// they draw from the first half of the buffer and never hand a slot out twice (exhausted -> static tile lists); eager
// launches rotate through the second half.
struct SchedBuffer { int* base = nullptr; long long slots = 0; int device = -1; long long captured = 0; long long eager = 0; };
static SchedBuffer g_sched;
static std::mutex g_sched_mutex;
constexpr int kSchedSlotInts = 8;

static int* take_sched_slot(cudaStream_t stream) {
    std::lock_guard<std::mutex> lock(g_sched_mutex);
    if (g_sched.base == nullptr || g_sched.slots < 2) return nullptr;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev != g_sched.device) return nullptr;
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &st) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    const long long half = g_sched.slots / 2;
    long long idx;
    if (st == cudaStreamCaptureStatusActive) {
        if (g_sched.captured >= half) return nullptr;
        idx = g_sched.captured++;
    } else {
        idx = half + (g_sched.eager++ % (g_sched.slots - half));
    }
    return g_sched.base + idx * kSchedSlotInts;
}

// 0 = choose per shape, 1 = cta_group::1 only, 2 = the CTA pair wherever the shape allows it
// (ssl4p_gemm_set_cta_group; initial value from the environment variable SSL4P_GEMM_CG)
static int& cta_group_mode() {
    static int mode = [] { const char* e = getenv("SSL4P_GEMM_CG"); const int v = e ? atoi(e) : 0; return v >= 0 && v <= 2 ? v : 0; }();
    return mode;
}

// Wave-quantisation aware choice of the tile: minimise ceil(tiles / units) * per-tile cost.  Every CTA computes
// 128 x bn x K per tile in both modes, so the per-tile cost is ~ bn / rate + a fixed term; `rate` = relative tensor
// throughput of the configuration (smaller N tiles re-read A from shared memory per MMA; a CTA pair fills
// (128 + bn/2) x 64 operands per k-block and SM instead of (128 + bn) x 64).
// bn_fixed / cg_fixed != 0 restrict the search (force_bn of the C-ABI call).
// Policy 0 (per shape): both groups go through the same wave-quantisation cost, the single-CTA tile scaled by its measured
// throughput relative to the pair at equal quantisation (profiles/r02n_gemm_{mae,cls,large}.txt, one board, both groups back
// to back): split-K wgrads 0.83-0.93 (-> 0.87), K >= 1536 0.93-0.95, K = 1024 about even, K <= 768 the pair LOSES 4-28 % (the
// tile time is the epilogue's and two epilogues share one accumulator hand-off).  The quantisation term then decides cases
// like the classifier's M = 25 216 (197 single-CTA row tiles x 3 = 3.99 waves of 148, but 99 pair tiles x 3 = 4.01 waves of
// 74: the pair was 20-23 % slower there although K = 2304 / 3072).
static double single_rate(int K, int epi) {
    if (epi == SSL4P_EPI_PARTIAL_F32) return 0.87;
    if (K >= 1536) return 0.93;
    if (K >= 1024) return 0.98;
    return 1.12;
}

static void pick_config(int M, int N, int K, int epi, int splits, int bn_fixed, int cg_fixed, int& bn_out, int& cg_out) {
    const int sms = num_sms();
    static const int cands[4] = {256, 192, 128, 64};
    static const double rate1[4] = {1.00, 0.95, 0.85, 0.60};
    static const double rate2[4] = {1.00, 0.95, 0.85, 0.0};
    const int mode = cg_fixed ? cg_fixed : cta_group_mode();
    const bool pair_ok = M > BM && sms >= 2 && bn_fixed != 64;
    const double r1 = single_rate(K, epi);
    double best_cost = 1e30;
    bn_out = bn_fixed ? bn_fixed : 128; cg_out = 1;
    for (int cg = 1; cg <= 2; ++cg) {
        if (cg == 2 && (mode == 1 || !pair_ok)) continue;
        if (cg == 1 && mode == 2 && pair_ok) continue;
        for (int i = 0; i < 4; ++i) {
            const double rate = cg == 2 ? rate2[i] : rate1[i] * r1;
            const int bn = cands[i];
            if (rate <= 0.0 || (bn_fixed && bn != bn_fixed)) continue;
            const long long tiles = (long long)((M + BM * cg - 1) / (BM * cg)) * ((N + bn - 1) / bn) * splits;
            const long long units = cg == 2 ? pair_units() : sms;
            const long long waves = (tiles + units - 1) / units;
            const double cost = (double)waves * (bn / rate + 24.0);
            if (cost < best_cost) { best_cost = cost; bn_out = bn; cg_out = cg; }
        }
    }
}

}  // namespace ssl4p

using namespace ssl4p;

extern "C" size_t ssl4p_gemm_workspace_bytes(int M, int N, int K, int layout, int epilogue) {
    (void)K; (void)layout;
    if (epilogue != SSL4P_EPI_PARTIAL_F32) return 0;
    return (size_t)SSL4P_MAX_SPLITS * (size_t)M * (size_t)N * sizeof(float);
}

extern "C" int ssl4p_gemm_set_scheduler_buffer(void* zeroed_device_buffer, size_t bytes) {
    std::lock_guard<std::mutex> lock(g_sched_mutex);
    if (zeroed_device_buffer == nullptr || bytes < 2 * kSchedSlotInts * sizeof(int)) { g_sched = SchedBuffer(); return SSL4P_OK; }
    SSL4P_REQUIRE((reinterpret_cast<uintptr_t>(zeroed_device_buffer) & 31) == 0, "gemm scheduler buffer must be 32-byte aligned");
    int dev = -1;
    SSL4P_CUDA(cudaGetDevice(&dev));
    g_sched.base = (int*)zeroed_device_buffer;
    g_sched.slots = (long long)(bytes / (kSchedSlotInts * sizeof(int)));
    g_sched.device = dev; g_sched.captured = 0; g_sched.eager = 0;
    return SSL4P_OK;
}

extern "C" int ssl4p_gemm_set_cta_group(int mode) {
    const int prev = cta_group_mode();
    if (mode >= 0 && mode <= 2) cta_group_mode() = mode;
    return prev;
}

extern "C" int ssl4p_gemm_pick_splits(int M, int N, int K) {
    // ONE wave of 128x256 tiles: the largest split count whose tile total still fits the SM count (measured on B200,
    // scripts/wgrad_sweep.py: a second, partly filled wave costs more than the longer K loop of a single full one);
    // each split keeps >= 4 K-blocks.
    // With the CTA pair the unit is a 256x256 tile on one of sms/2 pairs.
    const bool pair = cta_group_mode() != 1 && M > BM;
    const int sms = pair ? pair_units() : num_sms();
    const int tiles = ((M + (pair ? 2 * BM : BM) - 1) / (pair ? 2 * BM : BM)) * ((N + 255) / 256);
    int s = sms / tiles;
    const int kb = (K + BK - 1) / BK;
    if (s > kb / 4) s = kb / 4;
    if (s > SSL4P_MAX_SPLITS) s = SSL4P_MAX_SPLITS;
    if (s < 1) s = 1;
    return s;
}

extern "C" int ssl4p_gemm_bf16(const void* A, const void* B, void* C, void* C2, const float* bias, const void* aux,
                               int M, int N, int K, int lda, int ldb, int ldc, int ldc2, int ldaux,
                               int layout, int epilogue, int splits, int force_bn, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    SSL4P_REQUIRE(A && B && C, "gemm: null operand");
    SSL4P_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: bad shape M=%d N=%d K=%d", M, N, K);
    SSL4P_REQUIRE(layout >= 0 && layout <= 2, "gemm: bad layout %d", layout);
    const bool a_mn = layout == SSL4P_GEMM_TN;
    const bool b_mn = layout != SSL4P_GEMM_NT;
    if (splits < 1) splits = 1;
    SSL4P_REQUIRE(splits == 1 || epilogue == SSL4P_EPI_PARTIAL_F32, "gemm: split-K needs the PARTIAL_F32 epilogue");
    SSL4P_REQUIRE(splits <= SSL4P_MAX_SPLITS, "gemm: too many splits");
    if (epilogue == SSL4P_EPI_BIAS_GELU_BF16) SSL4P_REQUIRE(C2 != nullptr, "gemm: GELU epilogue needs C2");
    if (epilogue == SSL4P_EPI_BIAS_RESIDUAL_F32 || epilogue == SSL4P_EPI_DGELU_BF16) SSL4P_REQUIRE(aux != nullptr, "gemm: epilogue needs aux");
    // vectorised epilogue accesses
    SSL4P_REQUIRE(ldc % 8 == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0, "gemm: C must be 16B aligned with ldc %% 8 == 0");
    SSL4P_REQUIRE(N % 8 == 0, "gemm: N must be a multiple of 8 (vectorised epilogue), got %d", N);
    if (C2) SSL4P_REQUIRE(ldc2 % 8 == 0 && (reinterpret_cast<uintptr_t>(C2) & 15) == 0, "gemm: C2 alignment");
    if (aux) SSL4P_REQUIRE(ldaux % 8 == 0 && (reinterpret_cast<uintptr_t>(aux) & 15) == 0, "gemm: aux alignment");
    if (bias) SSL4P_REQUIRE((reinterpret_cast<uintptr_t>(bias) & 15) == 0, "gemm: bias alignment");

    // force_bn: low 16 bits = N tile (0 = choose), bits 16.. = CTA group (0 = choose, 1, 2)
    int bn, cg;
    pick_config(M, N, K, epilogue, splits, force_bn & 0xffff, (force_bn >> 16) & 0xff, bn, cg);
    SSL4P_REQUIRE(bn == 64 || bn == 128 || bn == 192 || bn == 256, "gemm: unsupported N tile %d", bn);
    SSL4P_REQUIRE(cg == 1 || (cg == 2 && bn >= 128), "gemm: unsupported CTA group %d for N tile %d", cg, bn);

    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.ldc2 = ldc2; p.ldaux = ldaux;
    p.splits = splits;
    const int kb = (K + BK - 1) / BK;
    p.k_per_split = ((kb + splits - 1) / splits) * BK;
    p.C = C; p.C2 = C2; p.bias = bias; p.aux = aux;
    p.split_stride = (long long)M * ldc;
    p.sched = take_sched_slot(stream);
    // measured (scripts/r02_ab3.sh): dec proj (K = 512) 63.5 -> 58.4 us, enc proj (K = 768) unchanged, but dec fc2 (K = 2048,
    // main-loop bound) 95.7 -> 99.3 us: the prefetches compete with the operand loads -> short K only
    static const int resid_prefetch = [] { const char* e = getenv("SSL4P_GEMM_RESID_PREFETCH"); return e ? atoi(e) : -1; }();
    p.resid_prefetch = resid_prefetch >= 0 ? resid_prefetch : (K <= 1024 ? 1 : 0);

    CUtensorMap tmA, tmB;
    int rc;
    if (!a_mn) rc = make_tmap(&tmA, A, K, M, lda, BM);            // A[M,K], K contiguous
    else       rc = make_tmap(&tmA, A, M, K, lda, BK);            // A stored [K,M], M contiguous
    if (rc) return rc;
    if (!b_mn) rc = make_tmap(&tmB, B, K, N, ldb, bn / cg);       // B[N,K], K contiguous (a CTA pair: half the tile per CTA)
    else       rc = make_tmap(&tmB, B, N, K, ldb, BK);            // B stored [K,N], N contiguous
    if (rc) return rc;

    CUtensorMap tmC = tmA;                                         // placeholder unless the epilogue stores through TMA
    if (epilogue == SSL4P_EPI_BIAS_BF16) { rc = make_tmap_out(&tmC, C, N, M, ldc); if (rc) return rc; }
    if (cg == 2) {
        if (bn == 256) return dispatch<256, 2>(layout, epilogue, tmA, tmB, tmC, p, stream);
        if (bn == 192) return dispatch<192, 2>(layout, epilogue, tmA, tmB, tmC, p, stream);
        if (bn == 128) return dispatch<128, 2>(layout, epilogue, tmA, tmB, tmC, p, stream);
    }
    if (bn == 256) return dispatch<256, 1>(layout, epilogue, tmA, tmB, tmC, p, stream);
    if (bn == 192) return dispatch<192, 1>(layout, epilogue, tmA, tmB, tmC, p, stream);
    if (bn == 128) return dispatch<128, 1>(layout, epilogue, tmA, tmB, tmC, p, stream);
    if (bn == 64) return dispatch<64, 1>(layout, epilogue, tmA, tmB, tmC, p, stream);
    set_error("gemm: dispatch fell through");
    return SSL4P_ERR_ARG;
}
