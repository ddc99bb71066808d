// Pipelined tcgen05 attention forward (second generation; replaces the forward of attention_tc.cu on the hot path).
//
// Same contract as ssl4p_attention_fwd (timm 0.4.12 Attention.forward, /root/reference/src/ssl4polyp/models/mae/
// models_mae.py:166,186; models.py:205-207).  What changed against the first generation, all of it to take the
// latency chain  TMA -> S MMA -> softmax -> P V MMA -> O read-out  off the critical path (ncu: 37 % of the warp
// samples sat in mbarrier waits, tensor pipe and XU both ~30 % busy):
//   * all eight softmax warps work on ONE 128-query tile: two threads per row (alternate 32-column chunks), the row
//     is read from TMEM exactly once into registers, maximum and sum are exchanged through shared memory.  The S
//     buffer is released ~300 cycles after it was produced, so two S buffers keep the tensor pipe ahead;
//   * the O read-out of tile t-1 is deferred until after the exponentials of tile t, i.e. P V of tile t-1 runs under
//     the softmax of tile t, and one P buffer (64 KB) is enough;
//   * Q/K/V tiles are ring-buffered (2 stages at T = 197, 4 at T = 50) in compact boxes of exactly Npad token rows,
//     so the loads of the next (image, head slice) overlap the current one.
// TMEM: S0 at column 0, S1 at 224, O at 448 (T <= 208); T in (208, 256] falls back to one S buffer (S 0, O 256).
#include "attn_common.cuh"
#include <stdlib.h>

namespace ssl4p {

struct AttnFwd2Params {
    int T, H, items, slices_per_img;
    int Npad;            // keys padded to 16
    int NB;              // 128-query tiles per head
    int RB, NBOX;        // TMA box rows and boxes per operand (NBOX * RB >= Npad)
    int KC;              // 64-key chunks of P
    int stages;          // Q/K/V ring depth
    int nS;              // S buffers in TMEM (2, or 1 when Npad > 208)
    int dbg;             // timing experiments only (SSL4P_ATTN_DEBUG bit mask; results are wrong when non-zero)
    float scale;
    __nv_bfloat16* out;  // [B*T, H*hd]
    float* lse;          // [B, H, T]
    long long* prof;     // optional [16] phase clocks of CTA 0 / warp 0 (timing experiments; NULL on the hot path)
};

constexpr int kMaxStages = 4;

template <int HD>
__global__ void __launch_bounds__(kAttnThreads, 1)
attention_fwd2_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnFwd2Params p) {
    constexpr int HP = 64 / HD;                                  // heads per work item
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int opnd_bytes = p.NBOX * p.RB * 128;                  // one operand (Q, K or V) of one stage
    const int stage_bytes = 3 * opnd_bytes;
    uint8_t* sStage = smem;
    uint8_t* sP = smem + p.stages * stage_bytes;                 // KC chunks of [128 rows x 128 B]
    uint8_t* tail = sP + p.KC * kTileBytes;
    float* sMax = reinterpret_cast<float*>(tail);                // [2 parity][2 halves][128]
    float* sSum = sMax + 512;                                    // [2 parity][2 halves][128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(tail + 4096);
    uint64_t* qkv_full = bars;                                   // [kMaxStages]
    uint64_t* qkv_empty = bars + kMaxStages;                     // [kMaxStages]
    uint64_t* s_full = bars + 2 * kMaxStages;                    // [2]
    uint64_t* s_free = s_full + 2;                               // [2]
    uint64_t* p_full = s_free + 2;                               // [1]
    uint64_t* o_full = p_full + 1;                               // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = lane_id();
    const int NT = HP * p.NB;                                    // tiles per item
    const int my_items = p.items > (int)blockIdx.x ? (p.items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    const int total = my_items * NT;
    const uint32_t s_col_stride = 224;
    const uint32_t o_col = p.nS == 2 ? 448u : 256u;

    if (warp == 8 && lane == 0) tma_prefetch_desc(&tmQKV);
    if (warp == 9) {
        if (lane == 0) {
            for (int i = 0; i < kMaxStages; ++i) { mbar_init(&qkv_full[i], 1); mbar_init(&qkv_empty[i], 1); }
            for (int i = 0; i < 2; ++i) { mbar_init(&s_full[i], 1); mbar_init(&s_free[i], 256); }
            mbar_init(p_full, 256);
            mbar_init(o_full, 1);
            mbar_fence_init();
            fence_proxy_async();
        }
        __syncwarp();
        tmem_alloc<512>(tmem_slot);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 8) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            const int HHD = p.H * HD;
            for (int il = 0; il < my_items; ++il) {
                const int item = blockIdx.x + il * gridDim.x;
                const int b = item / p.slices_per_img, sl = item % p.slices_per_img;
                const int st = il % p.stages, use = il / p.stages;
                mbar_wait(&qkv_empty[st], (use & 1) ^ 1);
                if ((p.dbg & 4) && il >= p.stages) { mbar_arrive(&qkv_full[st]); continue; }
                mbar_expect_tx(&qkv_full[st], (uint32_t)stage_bytes);
                uint8_t* base = sStage + st * stage_bytes;
                for (int i = 0; i < p.NBOX; ++i) {
                    tma_load_3d(base + i * p.RB * 128, &tmQKV, &qkv_full[st], sl * 64, i * p.RB, b);
                    tma_load_3d(base + opnd_bytes + i * p.RB * 128, &tmQKV, &qkv_full[st], HHD + sl * 64, i * p.RB, b);
                    tma_load_3d(base + 2 * opnd_bytes + i * p.RB * 128, &tmQKV, &qkv_full[st], 2 * HHD + sl * 64, i * p.RB, b);
                }
            }
        }
        __syncwarp();
    } else if (warp == 9) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc_s = idesc_rt(128, p.Npad, false, false);
            const uint32_t idesc_o = idesc_rt(128, 64, false, true);
            const uint32_t stage0 = smem_u32(sStage), sPa = smem_u32(sP);
            const int ksteps = (p.dbg & 2) ? 1 : (p.Npad >> 4);
            int items_seen = 0;                                   // qkv_full phases already observed (items are visited in order)
            auto issue_s = [&](int g) {
                const int il = g / NT, t = g - il * NT;
                const int st = il % p.stages;
                if (il >= items_seen) { mbar_wait(&qkv_full[st], (il / p.stages) & 1); items_seen = il + 1; }
                const int sb = g % p.nS, use = g / p.nS;
                mbar_wait(&s_free[sb], (use & 1) ^ 1);
                tc_fence_after();
                const int hh = t / p.NB, mb = t - hh * p.NB;
                const uint32_t q0 = stage0 + st * stage_bytes + mb * kTileBytes + hh * (HD * 2);
                const uint32_t k0 = stage0 + st * stage_bytes + opnd_bytes + hh * (HD * 2);
                const uint32_t tmem_d = tmem_base + sb * s_col_stride;
#pragma unroll
                for (int kk = 0; kk < HD / 16; ++kk)
                    umma_bf16(tmem_d, umma_smem_desc(q0 + kk * 32, 16, 1024), umma_smem_desc(k0 + kk * 32, 16, 1024), idesc_s, kk > 0);
                umma_commit(&s_full[sb]);
            };
            for (int g = 0; g < p.nS && g < total; ++g) issue_s(g);
            for (int g = 0; g < total; ++g) {
                const int il = g / NT, t = g - il * NT;
                const int st = il % p.stages;
                mbar_wait(p_full, g & 1);                         // P(g) in shared memory, O(g-1) drained
                tc_fence_after();
                const uint32_t v0 = stage0 + st * stage_bytes + 2 * opnd_bytes;
                for (int ks = 0; ks < ksteps; ++ks)
                    umma_bf16(tmem_base + o_col, umma_smem_desc(sPa + (ks >> 2) * kTileBytes + (ks & 3) * 32, 16, 1024),
                              umma_smem_desc(v0 + ks * 2048, kTileBytes, 1024), idesc_o, ks > 0);
                umma_commit(o_full);
                if (t == NT - 1) umma_commit(&qkv_empty[st]);     // last tile of the item: its Q/K/V stage may be refilled
                if (g + p.nS < total) issue_s(g + p.nS);
            }
        }
        __syncwarp();
    } else {
        // ===================== softmax warps (all eight on one tile, two threads per row) =====================
        const int q4 = warp & 3, half = warp >> 2;
        const int row = q4 * 32 + lane;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q4 * 32) << 16);
        const float sl2 = p.scale * kLog2eF;
        const int HHD = p.H * HD;
        const uint32_t sMax_a = smem_u32(sMax), sSum_a = smem_u32(sSum), sP_a = smem_u32(sP);

        // O / rowsum -> bf16 rows of `out`, log-sum-exp; tile described by (gp, b, head, hh, qrow, mx)
        auto epilogue = [&](int gp, int b, int head, int hh, int qrow, float mx) {
            const int pp = gp & 1;
            const float sum = lds32f(sSum_a + 4 * (pp * 256 + row)) + lds32f(sSum_a + 4 * (pp * 256 + 128 + row));
            const float inv = 1.0f / sum;
            if constexpr (HD == 64) {
                uint32_t r[32];
                tmem_ld_32x32(lane_addr + o_col + half * 32, r);
                tmem_ld_wait();
                if (qrow < p.T && !(p.dbg & 16)) {
                    __nv_bfloat16* dst = p.out + (size_t)(b * p.T + qrow) * HHD + head * HD + half * 32;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        uint4 o;
                        o.x = pack_bf16(__uint_as_float(r[8 * i + 0]) * inv, __uint_as_float(r[8 * i + 1]) * inv);
                        o.y = pack_bf16(__uint_as_float(r[8 * i + 2]) * inv, __uint_as_float(r[8 * i + 3]) * inv);
                        o.z = pack_bf16(__uint_as_float(r[8 * i + 4]) * inv, __uint_as_float(r[8 * i + 5]) * inv);
                        o.w = pack_bf16(__uint_as_float(r[8 * i + 6]) * inv, __uint_as_float(r[8 * i + 7]) * inv);
                        *reinterpret_cast<uint4*>(dst + 8 * i) = o;
                    }
                }
            } else {
                uint32_t r[16];
                tmem_ld_32x16(lane_addr + o_col + hh * 32 + half * 16, r);
                tmem_ld_wait();
                if (qrow < p.T && !(p.dbg & 16)) {
                    __nv_bfloat16* dst = p.out + (size_t)(b * p.T + qrow) * HHD + head * HD + half * 16;
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        uint4 o;
                        o.x = pack_bf16(__uint_as_float(r[8 * i + 0]) * inv, __uint_as_float(r[8 * i + 1]) * inv);
                        o.y = pack_bf16(__uint_as_float(r[8 * i + 2]) * inv, __uint_as_float(r[8 * i + 3]) * inv);
                        o.z = pack_bf16(__uint_as_float(r[8 * i + 4]) * inv, __uint_as_float(r[8 * i + 5]) * inv);
                        o.w = pack_bf16(__uint_as_float(r[8 * i + 6]) * inv, __uint_as_float(r[8 * i + 7]) * inv);
                        *reinterpret_cast<uint4*>(dst + 8 * i) = o;
                    }
                }
            }
            if (half == 0 && qrow < p.T) p.lse[((size_t)b * p.H + head) * p.T + qrow] = mx * p.scale + __logf(sum);
            tc_fence_before();
        };

        long long pc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        const bool prof = p.prof != nullptr && blockIdx.x == 0 && threadIdx.x == 0;
#define PROF_T(k) do { if (prof) { const long long now_ = clock64(); pc[k] += now_ - tprev; tprev = now_; } } while (0)
        long long tprev = clock64();
        int g = 0;
        int pv_b = 0, pv_head = 0, pv_hh = 0, pv_qrow = 0;       // the tile whose O is still in TMEM
        float pv_mx = 0.f;
        for (int il = 0; il < my_items; ++il) {
            const int item = blockIdx.x + il * gridDim.x;
            const int b = item / p.slices_per_img, sl = item % p.slices_per_img;
            for (int t = 0; t < NT; ++t, ++g) {
                const int hh = t / p.NB, mb = t - hh * p.NB;
                const int head = sl * HP + hh;
                const int qrow = mb * 128 + row;
                const int sb = g % p.nS, use = g / p.nS;
                const int par = g & 1;
                PROF_T(8);
                mbar_wait(&s_full[sb], use & 1);
                tc_fence_after();
                PROF_T(0);
                // ---- this thread's share of the row: 32-column chunks 2i + half, read from TMEM exactly once
                uint32_t s[4][32];
                const uint32_t saddr = lane_addr + sb * s_col_stride;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int c = (2 * i + half) * 32;
                    if (c < p.Npad) tmem_ld_32x32(saddr + c, s[i]);
                }
                tmem_ld_wait();
                tc_fence_before();
                mbar_arrive(&s_free[sb]);                        // the tensor core may overwrite this S buffer
                PROF_T(1);
                float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};      // independent chains (two warps per scheduler: ILP matters)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int c = (2 * i + half) * 32;
                    if (c + 32 <= p.T) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) mx4[j & 3] = fmaxf(mx4[j & 3], __uint_as_float(s[i][j]));
                    } else if (c < p.T) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) if (c + j < p.T) mx4[j & 3] = fmaxf(mx4[j & 3], __uint_as_float(s[i][j]));
                    }
                }
                float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
                sts32f(sMax_a + 4 * (par * 256 + half * 128 + row), mx);
                PROF_T(2);
                named_bar_sync(1, 256);
                PROF_T(3);
                mx = fmaxf(mx, lds32f(sMax_a + 4 * (par * 256 + (half ^ 1) * 128 + row)));
                const float msl = mx * sl2;
                // ---- p = exp2((s - max) * scale * log2e) -> bf16 pairs (kept in registers), partial row sum
                float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int c = (2 * i + half) * 32;
                    if (c < p.Npad) {
                        if (p.dbg & 1) {
#pragma unroll
                            for (int j = 0; j < 32; j += 2) {
                                const float e0 = fmaf(__uint_as_float(s[i][j]), sl2, -msl), e1 = fmaf(__uint_as_float(s[i][j + 1]), sl2, -msl);
                                sum4[(j >> 1) & 3] += e0 + e1;
                                s[i][j >> 1] = pack_bf16(e0, e1);
                            }
                        } else if (c + 32 <= p.T) {
#pragma unroll
                            for (int j = 0; j < 32; j += 2) {
                                const float e0 = fast_exp2(fmaf(__uint_as_float(s[i][j]), sl2, -msl));
                                const float e1 = fast_exp2(fmaf(__uint_as_float(s[i][j + 1]), sl2, -msl));
                                sum4[(j >> 1) & 3] += e0 + e1;
                                s[i][j >> 1] = pack_bf16(e0, e1);
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; j += 2) {
                                const float e0 = (c + j < p.T) ? fast_exp2(fmaf(__uint_as_float(s[i][j]), sl2, -msl)) : 0.f;
                                const float e1 = (c + j + 1 < p.T) ? fast_exp2(fmaf(__uint_as_float(s[i][j + 1]), sl2, -msl)) : 0.f;
                                sum4[(j >> 1) & 3] += e0 + e1;
                                s[i][j >> 1] = pack_bf16(e0, e1);
                            }
                        }
                    }
                }
                const float sum = (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);
                sts32f(sSum_a + 4 * (par * 256 + half * 128 + row), sum);
                PROF_T(4);
                // ---- the previous tile: P V has had the whole exponential phase to finish; drain its O
                if (g > 0) {
                    mbar_wait(o_full, (g - 1) & 1);
                    tc_fence_after();
                    PROF_T(5);
                    epilogue(g - 1, pv_b, pv_head, pv_hh, pv_qrow, pv_mx);
                }
                PROF_T(6);
                // ---- P (bf16) into the K-major SW128 tile the tensor core reads
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int c = (2 * i + half) * 32;
                    if (c < p.Npad && !(p.dbg & 8)) {
                        const uint32_t tile = sP_a + (c >> 6) * kTileBytes;
                        const int sub = (c & 63) >> 3;
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            sts128(tile + sw128_off(row, sub + k), s[i][4 * k], s[i][4 * k + 1], s[i][4 * k + 2], s[i][4 * k + 3]);
                    }
                }
                fence_proxy_async();                              // generic-proxy smem writes -> visible to the tensor core
                mbar_arrive(p_full);                              // P(g) written and O(g-1) drained
                PROF_T(7);
                pv_b = b; pv_head = head; pv_hh = hh; pv_qrow = qrow; pv_mx = mx;
            }
        }
        if (total > 0) {
            named_bar_sync(1, 256);                               // partial sums of the last tile visible
            mbar_wait(o_full, (total - 1) & 1);
            tc_fence_after();
            epilogue(total - 1, pv_b, pv_head, pv_hh, pv_qrow, pv_mx);
        }
        if (prof) { for (int k = 0; k < 9; ++k) p.prof[k] = pc[k]; p.prof[9] = total; }
#undef PROF_T
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

constexpr int kFwd2SmemBudget = 232448 - 6144;                  // stages + P; 4 KB exchange + 1 KB barriers + 1 KB slack on top = the 227 KB limit

template <int HD>
static int launch_fwd2(const void* qkv, void* out, float* lse, int B, int T, int H, float scale, cudaStream_t s) {
    AttnFwd2Params p;
    p.T = T; p.H = H; p.slices_per_img = H * HD / 64; p.items = B * p.slices_per_img;
    p.Npad = (T + 15) / 16 * 16;
    p.NB = (T + 127) / 128;
    if (p.Npad <= 128) { p.RB = p.Npad; p.NBOX = 1; }
    else { p.RB = ((p.Npad / 2) + 7) / 8 * 8; p.NBOX = 2; }
    p.KC = (p.Npad + 63) / 64;
    p.nS = p.Npad <= 208 ? 2 : 1;
    const int stage_bytes = 3 * p.NBOX * p.RB * 128;
    int stages = (kFwd2SmemBudget - p.KC * kTileBytes) / stage_bytes;
    if (stages > kMaxStages) stages = kMaxStages;
    SSL4P_REQUIRE(stages >= p.nS || stages >= 1 && p.nS == 1, "attention_fwd: shared memory plan failed (T=%d)", T);
    p.stages = stages;
    p.scale = scale; p.out = (__nv_bfloat16*)out; p.lse = lse;
    { const char* e = getenv("SSL4P_ATTN_DEBUG"); p.dbg = e ? atoi(e) : 0; }
    p.prof = (p.dbg & 64) ? reinterpret_cast<long long*>(lse) : nullptr;      // experiment: phase clocks overwrite the head of lse
    CUtensorMap tm;
    int rc = make_tmap_tokens(&tm, qkv, 3LL * H * HD, T, B, p.RB);
    if (rc) return rc;
    const int smem = stages * stage_bytes + p.KC * kTileBytes + 4096 + 1024 + 1024;
    auto kern = attention_fwd2_kernel<HD>;
    static bool attr = false;
    if (!attr) { SSL4P_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwd2SmemBudget + 6144)); attr = true; }
    const int grid = p.items < num_sms() ? p.items : num_sms();
    kern<<<grid, kAttnThreads, smem, s>>>(tm, p);
    return check_launch("attention_fwd2_kernel");
}

}  // namespace ssl4p

using namespace ssl4p;

extern "C" int ssl4p_attention_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int hd, float scale, void* stream) {
    SSL4P_REQUIRE(qkv && out && lse, "attention_fwd: null pointer");
    SSL4P_REQUIRE(B > 0 && H > 0 && T > 0, "attention: bad shape");
    SSL4P_REQUIRE(hd == 32 || hd == 64, "attention: head_dim %d not supported (32 or 64)", hd);
    SSL4P_REQUIRE((H * hd) % 64 == 0, "attention: H*head_dim (%d) must be a multiple of 64", H * hd);
    SSL4P_REQUIRE(T <= 256, "attention: sequence length %d not supported (<= 256)", T);
    cudaStream_t s = (cudaStream_t)stream;
    // Measured on B200 (scripts/attn_bench.py, B=256): the ring-buffered kernel wins for one-tile sequences (T=50: 62 vs
    // 76 us); with two query tiles per head the two-warpgroup kernel of attention_tc.cu is ahead (T=197: 199 vs 211 us).
    // For T <= 64 (the 50-token encoder of the pretraining step) nothing is tensor bound: 0.09 GFLOP/img against 79 MB
    // of qkv/out traffic per layer; the register-resident mma.sync kernel of attention.cu (one CTA per (image, head),
    // no TMEM round trips) finishes in 51 us where the tcgen05 pipelines need 76-84 us.
    const char* force = getenv("SSL4P_ATTN_FWD");
    if (!force && T <= 64) return ssl4p_attention_fwd_mma(qkv, out, lse, B, T, H, hd, scale, stream);
    const bool use2 = force ? (force[0] == '2') : (T <= 128);
    if (!use2) return ssl4p_attention_fwd_tc1(qkv, out, lse, B, T, H, hd, scale, stream);
    if (hd == 64) return launch_fwd2<64>(qkv, out, lse, B, T, H, scale, s);
    return launch_fwd2<32>(qkv, out, lse, B, T, H, scale, s);
}
