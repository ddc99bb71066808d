// Fused multi-head self-attention forward/backward on the 5th-gen tensor cores (tcgen05 + TMEM), TMA-fed.
//
// Replaces timm 0.4.12 Attention.forward as used by /root/reference/src/ssl4polyp/models/mae/models_mae.py:166,186 and
// models.py:205-207:  softmax(q k^T * hd^-0.5) v  on qkv [B, T, 3, H, hd]  (T = 50 encoder, 197 decoder/classifier;
// hd = 64 / 32), and its autograd backward.  Nothing of size [B,H,T,T] ever touches HBM.
//
// Work item = (image b, 64-column slice of the head dimension) = one head (hd 64) or a PAIR of heads (hd 32): every
// operand tile is then a [rows x 64 bf16] TMA box in the SWIZZLE_128B layout the GEMM kernel already uses, and a head
// inside a pair is selected by the K offset of the UMMA descriptor (K-major operands) or by the half of the 64
// accumulator columns that is kept (MN-major operands).  Rows t >= T are zero-filled by TMA (3-D tensor map
// {column, token, image}), so padding needs no HBM traffic and no masking of the operands.
//
// Forward  (persistent, 1 CTA / SM, 320 threads):
//   warps 0-3 / 4-7 : two softmax warpgroups; thread r owns row r of a 128-query tile (TMEM lane r)
//   warp 8          : TMA producer (Q, K, V tiles of the next item)
//   warp 9          : tcgen05.mma issuer, event driven: per warpgroup a chain  S = Q K^T -> [softmax] -> O = P V
//   S (fp32, <= 256 columns) lives in TMEM; P (bf16) goes through shared memory in the K-major SW128 layout; O is
//   accumulated in the TMEM columns S occupied (S is dead once P has been written).
// Backward (persistent, 1 CTA / SM, 320 threads): per head, blocks of 128 keys x 128 queries:
//   S^T = K Q^T and dP^T = V dO^T into TMEM; 8 warps turn them into P^T and dS^T (bf16, shared memory) using the saved
//   log-sum-exp and D = rowsum(dO * O); then dV += P^T dO, dK += dS^T Q, dQ += dS K accumulate in TMEM.
//   Deterministic (no atomics).
#include "attn_common.cuh"
#include <mutex>
#include <stdlib.h>
#include <type_traits>
#include <unordered_map>

namespace ssl4p {

// Timing ablations of the backward (SSL4P_ATTN_DEBUG, scripts/r02_attn_dbg.sh) are compiled in only with -DSSL4P_ATTN_ABLATION:
// in the shipped kernel the switches are constant-false and cost nothing (the forward lost 18 us to such tests).
#ifdef SSL4P_ATTN_ABLATION
#define ABLATE(bit) ((p.debug & (bit)) != 0)
#else
#define ABLATE(bit) false
#endif

struct AttnFwdParams {
    int B, pair;                               // pair: two images share one 128-row tile (T <= 64), item = (image pair, slice)
    int T, H, items, slices_per_img;           // slices_per_img = H*hd/64
    int Npad, NB, NKB;                         // keys padded to 16; query tiles of 128; key boxes of 128 rows
    int sep;                                   // 1: O has its own TMEM columns (deferred read-out), 0: O overwrites S
    int backoff_ns;                            // MMA issuer: sleep this long after a polling sweep without progress (0: spin)
    int s_stride;                              // sep: TMEM columns between the two S buffers
    float scale;
    __nv_bfloat16* out;                        // [B*T, H*hd]
    float* lse;                                // [B, H, T]
};

// =========================================================================================================
// forward
// =========================================================================================================
// Two softmax warpgroups, each a chain  S = Q K^T -> softmax -> P -> O = P V  over its own tiles (g = wg, wg+2, ...).
// TMEM plan `sep` (2 * round32(Npad) + 2 * hd <= 512, e.g. the 197-token decoder at hd 32): S_w at w * s_stride, O_w at
// 2 * s_stride + w * hd.  O no longer aliases S, so the issuer starts S of the chain's NEXT tile the moment P of the
// current one is complete (before P V), and the warpgroup reads O of tile g out only after the exponentials of tile
// g+2: both round trips (commit -> mbarrier -> wake-up) that used to sit between two tiles are hidden.
// Otherwise (`sep` = 0) O is accumulated over the dead S columns and read out before the next S may be issued.
// PAIR (two images per tile, T <= 64) and SEP (O in its own TMEM columns) are compile-time: the softmax loop is sensitive to
// every extra live value and branch (a handful of runtime switches cost the backward 5 %, the forward 14 %).
template <int HD, bool PAIR, bool SEP>
__global__ void __launch_bounds__(kAttnThreads, 1)
attention_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmQKV64, const AttnFwdParams p) {
    constexpr int HP = 64 / HD;                                  // heads per work item
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;                                          // 2 tiles
    uint8_t* sK = smem + 2 * kTileBytes;                         // 2 tiles
    uint8_t* sV = smem + 4 * kTileBytes;                         // 2 tiles
    uint8_t* sP = smem + 6 * kTileBytes;                         // 2 warpgroups x 4 key chunks x 16 KB
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 14 * kTileBytes);
    uint64_t* qk_full = bars;                                    // Q and K of the item landed
    uint64_t* qk_empty = bars + 1;                               // all S MMAs of the item retired: Q/K may be overwritten
    uint64_t* v_full = bars + 10;                                // V of the item landed
    uint64_t* v_empty = bars + 11;                               // all P V MMAs of the item retired
    // one-tile sequences (T <= 128): the two tiles of each operand buffer form a 2-deep ring of items; stage 1 barriers:
    uint64_t* qk_full1 = bars + 12; uint64_t* qk_empty1 = bars + 13; uint64_t* v_full1 = bars + 14; uint64_t* v_empty1 = bars + 15;
    const bool ring = PAIR || (p.NB == 1 && p.NKB == 1);          // (pair mode: always one tile per operand)
    uint64_t* s_full = bars + 2;                                 // [2]
    uint64_t* p_full = bars + 4;                                 // [2]
    uint64_t* o_full = bars + 6;                                 // [2]
    uint64_t* s_empty = bars + 8;                                // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

    const int warp = threadIdx.x >> 5;
    const int lane = lane_id();
    const int NT = HP * p.NB;                                    // tiles per item
    const int my_items = p.items > (int)blockIdx.x ? (p.items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    constexpr bool sep = SEP;

    if (warp == 8 && lane == 0) { tma_prefetch_desc(&tmQKV); tma_prefetch_desc(&tmQKV64); }
    if (warp == 9) {
        if (lane == 0) {
            mbar_init(qk_full, 1); mbar_init(qk_empty, NT);
            mbar_init(v_full, 1); mbar_init(v_empty, NT);
            mbar_init(qk_full1, 1); mbar_init(qk_empty1, NT);
            mbar_init(v_full1, 1); mbar_init(v_empty1, NT);
            for (int i = 0; i < 2; ++i) {
                mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 128); mbar_init(&o_full[i], 1); mbar_init(&s_empty[i], 128);
            }
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
    pdl_wait_then_trigger();

    if (warp == 8) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            const int HHD = p.H * HD;
            for (int il = 0; il < my_items; ++il) {
                const int item = blockIdx.x + il * gridDim.x;
                const int b = item / p.slices_per_img, sl = item % p.slices_per_img;
                // Q/K are dead once the item's S MMAs have retired (about half way through the item), V after its last
                // P V: two barrier pairs, so the next item's Q/K arrive while this item is still in its softmax
                const int stage = ring ? (il & 1) : 0, use = ring ? (il >> 1) : il;
                uint64_t* qf = stage ? qk_full1 : qk_full; uint64_t* qe = stage ? qk_empty1 : qk_empty;
                uint64_t* vf = stage ? v_full1 : v_full;   uint64_t* ve = stage ? v_empty1 : v_empty;
                const int so = stage * kTileBytes;
                mbar_wait(qe, (use & 1) ^ 1);
                mbar_expect_tx(qf, (uint32_t)(p.NB + p.NKB) * kTileBytes);
                if (PAIR) {
                    // images 2b and 2b+1 in rows [0,64) and [64,128) of every operand tile (64-row boxes; rows t >= T and a
                    // second image beyond the batch are zero-filled by TMA)
                    for (int im = 0; im < 2; ++im) {
                        tma_load_3d(sQ + so + im * (kTileBytes / 2), &tmQKV64, qf, sl * 64, 0, 2 * b + im);
                        tma_load_3d(sK + so + im * (kTileBytes / 2), &tmQKV64, qf, HHD + sl * 64, 0, 2 * b + im);
                    }
                    mbar_wait(ve, (use & 1) ^ 1);
                    mbar_expect_tx(vf, (uint32_t)kTileBytes);
                    for (int im = 0; im < 2; ++im)
                        tma_load_3d(sV + so + im * (kTileBytes / 2), &tmQKV64, vf, 2 * HHD + sl * 64, 0, 2 * b + im);
                    continue;
                }
                for (int i = 0; i < p.NB; ++i) tma_load_3d(sQ + so + i * kTileBytes, &tmQKV, qf, sl * 64, i * 128, b);
                for (int i = 0; i < p.NKB; ++i) tma_load_3d(sK + so + i * kTileBytes, &tmQKV, qf, HHD + sl * 64, i * 128, b);
                mbar_wait(ve, (use & 1) ^ 1);
                mbar_expect_tx(vf, (uint32_t)p.NKB * kTileBytes);
                for (int i = 0; i < p.NKB; ++i) tma_load_3d(sV + so + i * kTileBytes, &tmQKV, vf, 2 * HHD + sl * 64, i * 128, b);
                if (!ring && il + 1 < my_items) {                 // single-buffered operands: warm L2 with the next item's boxes
                    const int item2 = item + gridDim.x;
                    const int b2 = item2 / p.slices_per_img, s2 = item2 % p.slices_per_img;
                    for (int i = 0; i < p.NKB; ++i) {
                        tma_prefetch_3d(&tmQKV, s2 * 64, i * 128, b2);
                        tma_prefetch_3d(&tmQKV, HHD + s2 * 64, i * 128, b2);
                        tma_prefetch_3d(&tmQKV, 2 * HHD + s2 * 64, i * 128, b2);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 9) {
        // ===================== MMA issuer (one thread, two chains) =====================
        if (lane == 0) {
            const uint32_t idesc_s = idesc_rt(128, p.Npad, false, false);
            const uint32_t idesc_o = idesc_rt(128, HD, false, true);   // hd 32: N = 32, half of a 64-wide MN-major swizzle atom
            const int total = my_items * NT;
            int nxt[2] = {0, 1};                                  // tile whose P V is pending, per chain
            int s_done[2] = {0, 0};                               // tiles of the chain whose S has been issued
            int loaded = 0, loaded_v = 0;                         // completions of qk_full / v_full observed so far (in order)
            const uint32_t sQa = smem_u32(sQ), sKa = smem_u32(sK), sVa = smem_u32(sV), sPa = smem_u32(sP);
            const uint32_t s_col[2] = {0u, sep ? (uint32_t)p.s_stride : 256u};
            const uint32_t o_col[2] = {sep ? 2u * p.s_stride : 0u, sep ? 2u * p.s_stride + HD : 256u};
            // S of tile u (index inside chain w) if its operands have landed and the S columns are free
            auto try_issue_s = [&](int w) -> bool {
                const int u = s_done[w], g = 2 * u + w;
                if (g >= total) return false;
                const int il = g / NT, t = g - il * NT;
                if (loaded <= il) {                                                   // items land in order; a barrier is at most one phase ahead
                    const bool ok = ring ? mbar_test_wait((loaded & 1) ? qk_full1 : qk_full, (loaded >> 1) & 1) : mbar_test_wait(qk_full, loaded & 1);
                    if (ok) ++loaded;
                }
                if (loaded <= il) return false;
                if (!sep && !mbar_test_wait(&s_empty[w], (u & 1) ^ 1)) return false;  // alias plan: O of the previous tile read out
                tc_fence_after();
                const int hh = t / p.NB, mb = t - hh * p.NB;
                const uint32_t so = ring ? (uint32_t)(il & 1) * kTileBytes : 0u;
                const uint32_t a0 = sQa + so + mb * kTileBytes + hh * (HD * 2);
                const uint32_t b0 = sKa + so + hh * (HD * 2);
#pragma unroll
                for (int kk = 0; kk < HD / 16; ++kk)
                    umma_bf16(tmem_base + s_col[w], umma_smem_desc(a0 + kk * 32, 16, 1024), umma_smem_desc(b0 + kk * 32, 16, 1024), idesc_s, kk > 0);
                umma_commit(&s_full[w]);
                umma_commit((ring && (il & 1)) ? qk_empty1 : qk_empty);
                ++s_done[w];
                return true;
            };
            while (nxt[0] < total || nxt[1] < total) {
                bool progress = false;
#pragma unroll
                for (int w = 0; w < 2; ++w) {
                    const int g = nxt[w];
                    if (g >= total) continue;
                    const int u = g >> 1, il = g / NT, t = g - il * NT;
                    if (s_done[w] <= u) { progress |= try_issue_s(w); continue; }     // S(g) itself not issued yet
                    if (!mbar_test_wait(&p_full[w], u & 1)) continue;
                    tc_fence_after();
                    // sep plan: P(g) complete means S(g) has been consumed -> the chain's next S goes first, then P V
                    if (sep && s_done[w] == u + 1) try_issue_s(w);
                    if (loaded_v <= il) {
                        const bool ok = ring ? mbar_test_wait((loaded_v & 1) ? v_full1 : v_full, (loaded_v >> 1) & 1) : mbar_test_wait(v_full, loaded_v & 1);
                        if (ok) ++loaded_v;
                    }
                    if (loaded_v <= il) continue;                 // V not there yet: come back (p_full stays complete)
                    tc_fence_after();
                    const uint32_t pa = sPa + (uint32_t)w * 4u * kTileBytes;
                    const int ksteps = p.Npad >> 4;
                    const uint32_t vso = ring ? (uint32_t)(il & 1) * kTileBytes : 0u;
                    const uint32_t voff = (HD == 32) ? (uint32_t)(t / p.NB) * 64u : 0u;   // head of the pair: +64 B inside the 128 B row
                    for (int ks = 0; ks < ksteps; ++ks)
                        umma_bf16(tmem_base + o_col[w], umma_smem_desc(pa + (ks >> 2) * kTileBytes + (ks & 3) * 32, 16, 1024),
                                  umma_smem_desc(sVa + vso + ks * 2048 + voff, kTileBytes, 1024), idesc_o, ks > 0);
                    umma_commit(&o_full[w]);
                    umma_commit((ring && (il & 1)) ? v_empty1 : v_empty);
                    nxt[w] = g + 2;
                    progress = true;
                }
                // This thread shares its scheduler with two softmax warps: polling flat out takes a third of that
                // scheduler's issue slots away from the warps everybody is waiting for.
                if (!progress && p.backoff_ns) __nanosleep(p.backoff_ns);
            }
        }
        __syncwarp();
    } else {
        // ===================== softmax warpgroups =====================
        const int wg = warp >> 2, q4 = warp & 3;
        const int row = q4 * 32 + lane;
        const uint32_t tlanes = tmem_base + ((uint32_t)(q4 * 32) << 16);
        // pair mode: rows [0,64) are image 0 and see key columns [0,64), rows [64,128) image 1 / columns [64,128)
        const int img = PAIR ? (q4 >> 1) : 0;
        const int kbase = img * 64;                              // first key column of this row's image
        const int klen = PAIR ? 64 : p.Npad;                   // key columns this row reads
        const int lrow = PAIR ? (row & 63) : row;              // query index inside the image (before the tile offset)
        const uint32_t taddr = tlanes + (sep ? (uint32_t)(wg * p.s_stride) : (uint32_t)wg * 256u) + (uint32_t)kbase;   // S of this chain
        const uint32_t oaddr = tlanes + (sep ? (uint32_t)(2 * p.s_stride + wg * HD) : (uint32_t)wg * 256u);            // O of this chain
        const uint32_t pw = smem_u32(sP) + wg * 4 * kTileBytes;
        const float sl2 = p.scale * kLog2eF;
        if (PAIR) {
            // P is block diagonal: the key chunk of the OTHER image is never written by this row and must read as zero
            const uint32_t other = pw + (uint32_t)(1 - img) * kTileBytes;
#pragma unroll
            for (int i = 0; i < 8; ++i) sts_v4(other + sw128_off(row, i), 0u, 0u, 0u, 0u);
        }
        // O / rowsum -> bf16 row of `out`, log-sum-exp of the row
        auto read_out = [&](int b, int head, int qrow, float mx, float sum) {
            const float inv = 1.0f / sum;
            const bool ok = qrow < p.T && b < p.B;
#pragma unroll
            for (int hf = 0; hf < HD / 32; ++hf) {
                uint32_t r[32];
                tmem_ld_32x32(oaddr + hf * 32, r);
                tmem_ld_wait();
                if (ok) {
                    __nv_bfloat16* dst = p.out + ((size_t)(b * p.T + qrow) * p.H + head) * HD + hf * 32;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        stg_v4(dst + 8 * i,
                               pack_bf16(__uint_as_float(r[8 * i + 0]) * inv, __uint_as_float(r[8 * i + 1]) * inv),
                               pack_bf16(__uint_as_float(r[8 * i + 2]) * inv, __uint_as_float(r[8 * i + 3]) * inv),
                               pack_bf16(__uint_as_float(r[8 * i + 4]) * inv, __uint_as_float(r[8 * i + 5]) * inv),
                               pack_bf16(__uint_as_float(r[8 * i + 6]) * inv, __uint_as_float(r[8 * i + 7]) * inv));
                }
            }
            if (ok) p.lse[((size_t)b * p.H + head) * p.T + qrow] = mx * p.scale + __logf(sum);
        };
        // running maximum over one 32-column chunk held in registers (columns c .. c+31 of this row's key range)
        auto chunk_max = [&](const uint32_t (&r)[32], int c, float mx) -> float {
            if (c + 32 <= p.T) {
                float m0 = mx, m1 = mx, m2 = mx, m3 = mx;
#pragma unroll
                for (int jj = 0; jj < 32; jj += 4) {
                    m0 = fmaxf(m0, __uint_as_float(r[jj])); m1 = fmaxf(m1, __uint_as_float(r[jj + 1]));
                    m2 = fmaxf(m2, __uint_as_float(r[jj + 2])); m3 = fmaxf(m3, __uint_as_float(r[jj + 3]));
                }
                return fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
            }
#pragma unroll
            for (int jj = 0; jj < 32; ++jj) if (c + jj < p.T) mx = fmaxf(mx, __uint_as_float(r[jj]));
            return mx;
        };
        int pb = 0, phead = 0, pqrow = 0, pu = -1;               // tile whose O is still in TMEM (sep plan)
        float pmx = 0.f, psum = 1.f;
        for (int il = 0; il < my_items; ++il) {
            const int item = blockIdx.x + il * gridDim.x;
            const int b0 = item / p.slices_per_img, sl = item % p.slices_per_img;
            const int b = PAIR ? 2 * b0 + img : b0;
            for (int t = 0; t < NT; ++t) {
                const int g = il * NT + t;
                if ((g & 1) != wg) continue;
                const int u = g >> 1;
                const int hh = t / p.NB, mb = t - hh * p.NB;
                const int head = sl * HP + hh;
                const int qrow = mb * 128 + lrow;
                mbar_wait(&s_full[wg], u & 1);
                tc_fence_after();
                // ---- pass 1: row maximum over the valid keys (the TMEM load of chunk c+1 is in flight during the max of chunk c)
                float mx = -INFINITY;
                {
                    uint32_t ra[32], rb[32];
                    tmem_ld_32x32(taddr, ra);
                    for (int c = 0; c < klen; c += 64) {
                        tmem_ld_wait();
                        if (c + 32 < klen) tmem_ld_32x32(taddr + c + 32, rb);
                        mx = chunk_max(ra, c, mx);
                        if (c + 32 >= klen) break;
                        tmem_ld_wait();
                        if (c + 64 < klen) tmem_ld_32x32(taddr + c + 64, ra);
                        mx = chunk_max(rb, c + 32, mx);
                    }
                }
                // sep plan: the P buffer of this chain is still being read by P V of the previous tile until o_full
                if (sep && pu >= 0) { mbar_wait(&o_full[wg], pu & 1); tc_fence_after(); }
                // ---- pass 2: p = exp2((s - max) * scale * log2e), row sum, bf16 P into the SW128 K-major tile; the scale /
                // subtract and the row sum run two elements per instruction (FFMA2 / FADD2), the TMEM load of the next
                // chunk is in flight during the exponentials of this one
                const float msl = mx * sl2;
                const f32x2 sl2v = f2_splat(sl2), mslv = f2_splat(-msl);
                f32x2 acc0 = f2_splat(0.f), acc1 = f2_splat(0.f);
                auto chunk_p = [&](const uint32_t (&r)[32], int c) {
                    const uint32_t tile = pw + (uint32_t)((kbase + c) >> 6) * kTileBytes;
                    const int sub = ((kbase + c) & 63) >> 3;
                    const bool tail = c + 32 > p.T;
                    const bool last16 = c + 16 >= klen;
#pragma unroll
                    for (int h16 = 0; h16 < 2; ++h16) {           // 16 columns at a time keeps the live exponentials at 16
                        // key columns past the padded length (208 of 224 at T = 197) are never read by P V: no exponentials for
                        // them.  head_dim 64 only: at head_dim 32 the branch costs more than the 7 % of exponentials it saves
                        // (decoder shape 127 -> 131 us; classifier shape 82 -> 76 us, scripts/r02_attn2.sh)
                        if (HD == 64 && h16 == 1 && last16) break;
                        float e[16];
#pragma unroll
                        for (int jj = 0; jj < 16; jj += 2) {
                            float x0, x1;
                            f2_unpack(f2_fma(f2_pack(__uint_as_float(r[h16 * 16 + jj]), __uint_as_float(r[h16 * 16 + jj + 1])), sl2v, mslv), x0, x1);
                            e[jj] = fast_exp2(x0); e[jj + 1] = fast_exp2(x1);
                        }
                        if (tail) {
#pragma unroll
                            for (int jj = 0; jj < 16; ++jj) if (c + h16 * 16 + jj >= p.T) e[jj] = 0.f;
                        }
#pragma unroll
                        for (int jj = 0; jj < 16; jj += 4) {
                            acc0 = f2_add(acc0, f2_pack(e[jj], e[jj + 1]));
                            acc1 = f2_add(acc1, f2_pack(e[jj + 2], e[jj + 3]));
                        }
#pragma unroll
                        for (int i = 0; i < 2; ++i)
                            sts_v4(tile + sw128_off(row, sub + h16 * 2 + i), pack_bf16(e[8 * i + 0], e[8 * i + 1]), pack_bf16(e[8 * i + 2], e[8 * i + 3]),
                                   pack_bf16(e[8 * i + 4], e[8 * i + 5]), pack_bf16(e[8 * i + 6], e[8 * i + 7]));
                    }
                };
                {
                    uint32_t ra[32], rb[32];
                    tmem_ld_32x32(taddr, ra);
                    for (int c = 0; c < klen; c += 64) {
                        tmem_ld_wait();
                        if (c + 32 < klen) tmem_ld_32x32(taddr + c + 32, rb);
                        chunk_p(ra, c);
                        if (c + 32 >= klen) break;
                        tmem_ld_wait();
                        if (c + 64 < klen) tmem_ld_32x32(taddr + c + 64, ra);
                        chunk_p(rb, c + 32);
                    }
                }
                float sum;
                {
                    float a, b2, c2, d2;
                    f2_unpack(acc0, a, b2); f2_unpack(acc1, c2, d2);
                    sum = (a + b2) + (c2 + d2);
                }
                if (sep) {
                    // O of the chain's previous tile (complete: o_full was observed above) leaves TMEM before P V of this
                    // tile may overwrite it, i.e. before p_full
                    if (pu >= 0) read_out(pb, phead, pqrow, pmx, psum);
                    fence_proxy_async();                          // generic-proxy smem writes -> visible to the tensor core
                    tc_fence_before();
                    mbar_arrive(&p_full[wg]);
                    pb = b; phead = head; pqrow = qrow; pmx = mx; psum = sum; pu = u;
                } else {
                    fence_proxy_async();
                    tc_fence_before();
                    mbar_arrive(&p_full[wg]);
                    mbar_wait(&o_full[wg], u & 1);
                    tc_fence_after();
                    read_out(b, head, qrow, mx, sum);
                    tc_fence_before();
                    mbar_arrive(&s_empty[wg]);
                }
            }
        }
        if (sep && pu >= 0) {                                     // O of the chain's last tile
            mbar_wait(&o_full[wg], pu & 1);
            tc_fence_after();
            read_out(pb, phead, pqrow, pmx, psum);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

// =========================================================================================================
// backward
// =========================================================================================================
struct AttnBwdParams {
    int B, pair;                               // pair: two images per 128-row tile (T <= 64); half block h = queries of image h
    int T, H, items, slices_per_img;
    int Npad, NB;
    int debug;                                 // timing experiments only (SSL4P_ATTN_DEBUG; results are WRONG with a bit set): 1 = no blk_done wait, 2 = no shared-memory stores of P^T / dS^T, 4 = no exponentials, 8 = no dV / dK MMAs, 16 = no dQ MMAs, 32 = no score MMAs, 64 = no per-item statistics loads
    float scale;
    const __nv_bfloat16* out;                  // [B*T, H*hd]
    const __nv_bfloat16* dout;                 // [B*T, H*hd]
    const float* lse;                          // [B, H, T]
    __nv_bfloat16* dqkv;                       // [B*T, 3*H*hd]
};

// 32 fp32 TMEM columns of this thread's row -> 32 bf16 at dst (16-byte stores)
__device__ __forceinline__ void store_row32_bf16(__nv_bfloat16* dst, const uint32_t (&r)[32]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        uint4 o;
        o.x = pack_bf16(__uint_as_float(r[8 * i + 0]), __uint_as_float(r[8 * i + 1]));
        o.y = pack_bf16(__uint_as_float(r[8 * i + 2]), __uint_as_float(r[8 * i + 3]));
        o.z = pack_bf16(__uint_as_float(r[8 * i + 4]), __uint_as_float(r[8 * i + 5]));
        o.w = pack_bf16(__uint_as_float(r[8 * i + 6]), __uint_as_float(r[8 * i + 7]));
        *reinterpret_cast<uint4*>(dst + 8 * i) = o;
    }
}

// TMEM column map (512 columns): two score buffers of a HALF block (128 keys x 64 queries), buffer u at 128 u:
// S^T [0,64) dP^T [64,128);  then dV [256,320)  dK [320,384)  dQ[qt] [384 + 64 qt, +64).
// Half blocks are numbered j = 0,1,2,... in issue order and use buffer j % kNBuf, so the tensor pipe computes the scores of
// half block j+1 (and the dV/dK/dQ updates of j-1) while the element-wise warps turn half block j into P^T / dS^T.
constexpr uint32_t kColST = 0, kColDP = 64, kColBuf = 128;
// Gradient accumulators behind the first two score buffers: dV, dK (HD columns each) and dQ per query tile (HD columns).  At
// head_dim 32 the gradient MMAs run with N = 32 (one head: the B operand starts 64 bytes into the 128-byte swizzle row), so
// they end at column 384 and a THIRD score buffer fits at [384, 512): the tensor pipe can then be two half blocks ahead of the
// element-wise warps, which used to wait ~1/3 of the time for scores (MMA -> commit -> mbarrier -> wake-up is ~450 clk and
// sat between every two half blocks).  At head_dim 64 the accumulators take [256, 512) and there are two buffers.
template <int HD> struct BwdCols {
    static constexpr int kNBuf = HD == 32 ? 3 : 2;
    static constexpr uint32_t kDV = 256, kDK = 256 + HD, kDQ = 256 + 2 * HD;       // dQ of query tile qt at kDQ + qt * HD
    __device__ __forceinline__ static uint32_t buf(int i) { return i < 2 ? (uint32_t)i * kColBuf : 384u; }
};

// Position of a half block inside one work item: head of the pair, key tile, query tile, half of the query tile.
struct HalfBlk {
    int hh, kt, qt, h;
    __device__ __forceinline__ static int nq(int Npad, int qt) { return min(128, Npad - qt * 128); }
    __device__ __forceinline__ static int halves(int Npad, int qt) { return nq(Npad, qt) > 64 ? 2 : 1; }
    __device__ __forceinline__ int nqh(int Npad) const { const int n = nq(Npad, qt); return h == 0 ? min(64, n) : n - 64; }
    __device__ __forceinline__ void next(int Npad, int NB) {
        if (h + 1 < halves(Npad, qt)) { ++h; return; }
        h = 0;
        if (++qt < NB) return;
        qt = 0;
        if (++kt < NB) return;
        kt = 0; ++hh;
    }
};

// Backward: one more warp than the forward - the MMA issue is split over TWO threads (warps 9 and 10).  With a single issuer the
// kernel was bound by that thread's instruction stream (about 20 descriptor-building MMA issues and 3-4 commits per half block:
// with the element-wise work switched off entirely the decoder shape still took 329 of 349 us, SSL4P_ATTN_DEBUG ablation).
constexpr int kAttnBwdThreads = kAttnThreads + 32;

template <int HD, bool PAIR>
__global__ void __launch_bounds__(kAttnBwdThreads, 1)
attention_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                        const __grid_constant__ CUtensorMap tmQKV64, const __grid_constant__ CUtensorMap tmDO64, const AttnBwdParams p) {
    constexpr int HP = 64 / HD;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;                                          // 2 tiles each
    uint8_t* sK = smem + 2 * kTileBytes;
    uint8_t* sV = smem + 4 * kTileBytes;
    uint8_t* sDO = smem + 6 * kTileBytes;
    uint8_t* sPT = smem + 8 * kTileBytes;                        // [128 keys] x 2 chunks of 64 queries
    uint8_t* sDS = smem + 10 * kTileBytes;
    float* sL = reinterpret_cast<float*>(smem + 12 * kTileBytes);   // [HP][256] -lse * log2e (-inf for q >= T)
    float* sD = sL + HP * 256;                                      // [HP][256] -D * scale
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 12 * kTileBytes + 4096);
    uint64_t* in_full = bars;                                    // [2] operand stages: one-tile sequences (T <= 128) use the two
    uint64_t* in_empty = bars + 2;                               // [2] tiles of every operand buffer as a 2-deep ring of items
    using Cols = BwdCols<HD>;
    constexpr int kNBuf = Cols::kNBuf;                           // score buffers in TMEM: half block j uses buffer j % kNBuf
    uint64_t* st_full = bars + 4;                                // [3] scores of half block j in its TMEM buffer
    uint64_t* pt_full = bars + 7;                                // [3] P^T / dS^T of half block j in shared memory, buffer drained
    uint64_t* blk_done = bars + 10;                              // all tensor reads of a block's P^T / dS^T retired
    uint64_t* dvk_full = bars + 11;
    uint64_t* dvk_empty = bars + 12;
    uint64_t* dq_full = bars + 13;
    uint64_t* dq_empty = bars + 14;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);
    const bool ring = PAIR || p.NB == 1;                         // item il lives in stage il & 1, otherwise always stage 0

    const int warp = threadIdx.x >> 5;
    const int lane = lane_id();
    const int my_items = p.items > (int)blockIdx.x ? (p.items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    const int HHD = p.H * HD;

    if (warp == 8 && lane == 0) { tma_prefetch_desc(&tmQKV); tma_prefetch_desc(&tmDO); tma_prefetch_desc(&tmQKV64); tma_prefetch_desc(&tmDO64); }
    if (warp == 9) {
        if (lane == 0) {
            for (int i = 0; i < 2; ++i) { mbar_init(&in_full[i], 1); mbar_init(&in_empty[i], 2); }   // both issuers release a stage
            // one arrival per element-wise WARP (its lanes meet at __syncwarp first): 8 arrivals on a barrier instead of 256
            for (int i = 0; i < kNBuf; ++i) { mbar_init(&st_full[i], 1); mbar_init(&pt_full[i], 8); }
            mbar_init(blk_done, 2);                              // dV / dK issuer + dQ issuer
            mbar_init(dvk_full, 1); mbar_init(dvk_empty, 8);
            mbar_init(dq_full, 1); mbar_init(dq_empty, 8);
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
    pdl_wait_then_trigger();

    if (warp == 8) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            for (int il = 0; il < my_items; ++il) {
                const int item = blockIdx.x + il * gridDim.x;
                const int b = item / p.slices_per_img, sl = item % p.slices_per_img;
                const int stage = ring ? (il & 1) : 0, use = ring ? (il >> 1) : il;
                mbar_wait(&in_empty[stage], (use & 1) ^ 1);
                mbar_expect_tx(&in_full[stage], (uint32_t)(4 * p.NB) * kTileBytes);
                if (PAIR) {                                     // images 2b, 2b+1 -> rows [0,64), [64,128) of every operand tile
                    for (int im = 0; im < 2; ++im) {
                        const int slot = stage * kTileBytes + im * (kTileBytes / 2);
                        tma_load_3d(sQ + slot, &tmQKV64, &in_full[stage], sl * 64, 0, 2 * b + im);
                        tma_load_3d(sK + slot, &tmQKV64, &in_full[stage], HHD + sl * 64, 0, 2 * b + im);
                        tma_load_3d(sV + slot, &tmQKV64, &in_full[stage], 2 * HHD + sl * 64, 0, 2 * b + im);
                        tma_load_3d(sDO + slot, &tmDO64, &in_full[stage], sl * 64, 0, 2 * b + im);
                    }
                    continue;
                }
                for (int i = 0; i < p.NB; ++i) {
                    const int slot = (stage + i) * kTileBytes;
                    tma_load_3d(sQ + slot, &tmQKV, &in_full[stage], sl * 64, i * 128, b);
                    tma_load_3d(sK + slot, &tmQKV, &in_full[stage], HHD + sl * 64, i * 128, b);
                    tma_load_3d(sV + slot, &tmQKV, &in_full[stage], 2 * HHD + sl * 64, i * 128, b);
                    tma_load_3d(sDO + slot, &tmDO, &in_full[stage], sl * 64, i * 128, b);
                }
                // the operands are single buffered: pull the next item's boxes into L2 now, so that the loads issued
                // after this item's last MMA pay L2 latency instead of HBM latency
                if (!ring && il + 1 < my_items) {
                    const int item2 = item + gridDim.x;
                    const int b2 = item2 / p.slices_per_img, sl2_ = item2 % p.slices_per_img;
                    for (int i = 0; i < p.NB; ++i) {
                        tma_prefetch_3d(&tmQKV, sl2_ * 64, i * 128, b2);
                        tma_prefetch_3d(&tmQKV, HHD + sl2_ * 64, i * 128, b2);
                        tma_prefetch_3d(&tmQKV, 2 * HHD + sl2_ * 64, i * 128, b2);
                        tma_prefetch_3d(&tmDO, sl2_ * 64, i * 128, b2);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 9) {
        // ===================== MMA issuer A: dV += P^T dO, dK += dS^T Q =====================
        // The whole warp runs the loop (uniform bookkeeping, every lane waits on the barriers); the tcgen05 instructions sit
        // under elect.sync, which the compiler knows to be single-threaded: with `if (lane == 0)` around the loop it wrapped
        // every UTCHMMA / UTCBAR in an ELECT + BRA.U.ANY retry loop (the operands live in uniform registers).
        {
            auto mma = [&](uint32_t d, uint64_t ad, uint64_t bd, uint32_t id, uint32_t acc) { if (elect_one()) umma_bf16(d, ad, bd, id, acc); };
            auto commit = [&](uint64_t* bar) { if (elect_one()) umma_commit(bar); };
            const uint32_t sQ0 = smem_u32(sQ), sDO0 = smem_u32(sDO);
            const uint32_t sPTa = smem_u32(sPT), sDSa = smem_u32(sDS);
            const uint32_t idesc_acc = idesc_rt(128, HD, false, true);     // A K-major (P^T / dS^T rows = keys), B MN-major (N = one head)
            int j = 0, nkt = 0;                                  // half blocks, key tiles issued so far (all items)
            for (int il = 0; il < my_items; ++il) {
                const int stage = ring ? (il & 1) : 0, use = ring ? (il >> 1) : il;
                mbar_wait(&in_full[stage], use & 1);
                tc_fence_after();
                const uint32_t sQa = sQ0 + stage * kTileBytes, sDOa = sDO0 + stage * kTileBytes;
                for (HalfBlk c{0, 0, 0, 0}; c.hh < HP; c.next(p.Npad, p.NB), ++j) {
                    if (c.qt == 0 && c.h == 0) mbar_wait(dvk_empty, (nkt & 1) ^ 1);   // previous key tile's dV/dK have been read out
                    mbar_wait(&pt_full[j % kNBuf], (j / kNBuf) & 1);   // P^T / dS^T of this half block written
                    tc_fence_after();
                    const uint32_t hoff = HD == 32 ? (uint32_t)c.hh * 64u : 0u;   // hd 32: the head's 32 columns inside the 64-wide MN-major row
                    const int ks0 = c.h * 4, ks1 = ks0 + (c.nqh(p.Npad) >> 4);
                    for (int ks = ks0; ks < ks1 && !ABLATE(8); ++ks) {
                        const uint32_t aoff = (ks >> 2) * kTileBytes + (ks & 3) * 32;
                        mma(tmem_base + Cols::kDV, umma_smem_desc(sPTa + aoff, 16, 1024),
                                  umma_smem_desc(sDOa + c.qt * kTileBytes + ks * 2048 + hoff, kTileBytes, 1024), idesc_acc, (c.qt > 0) || (ks > 0));
                        mma(tmem_base + Cols::kDK, umma_smem_desc(sDSa + aoff, 16, 1024),
                                  umma_smem_desc(sQa + c.qt * kTileBytes + ks * 2048 + hoff, kTileBytes, 1024), idesc_acc, (c.qt > 0) || (ks > 0));
                    }
                    if (c.h + 1 == HalfBlk::halves(p.Npad, c.qt)) {
                        commit(blk_done);                   // (one of two arrivals: issuer B commits its dQ MMAs)
                        if (c.qt == p.NB - 1) { commit(dvk_full); ++nkt; }
                    }
                }
                commit(&in_empty[stage]);
            }
        }
        __syncwarp();
    } else if (warp == 10) {
        // ===================== MMA issuer B: scores (S^T, dP^T) and dQ += dS K =====================
        {
            auto mma = [&](uint32_t d, uint64_t ad, uint64_t bd, uint32_t id, uint32_t acc) { if (elect_one()) umma_bf16(d, ad, bd, id, acc); };
            auto commit = [&](uint64_t* bar) { if (elect_one()) umma_commit(bar); };
            uint32_t sQa = smem_u32(sQ), sKa = smem_u32(sK), sVa = smem_u32(sV), sDOa = smem_u32(sDO);   // + stage offset per item
            const uint32_t sQ0 = sQa, sK0 = sKa, sV0 = sVa, sDO0 = sDOa;
            const uint32_t sDSa = smem_u32(sDS);
            const uint32_t idesc_dq = idesc_rt(128, HD, true, true);       // A MN-major (dS^T read as [keys][queries]), B MN-major
            int j = 0, nhead = 0;                                // half blocks, heads issued so far (all items)
            // scores of one half block: S^T = K[kt] Q[qt,h]^T and dP^T = V[kt] dO[qt,h]^T into TMEM buffer jj % kNBuf
            auto issue_scores = [&](const HalfBlk& x, int jj) {
                const uint32_t koff = x.hh * (HD * 2);
                const uint32_t idesc_s = idesc_rt(128, x.nqh(p.Npad), false, false);
                const uint32_t tb = tmem_base + Cols::buf(jj % kNBuf);
                const uint32_t qoff = x.qt * kTileBytes + x.h * (64 * 128) + koff;
                if (ABLATE(32)) { commit(&st_full[jj % kNBuf]); return; }
#pragma unroll
                for (int kk = 0; kk < HD / 16; ++kk)
                    mma(tb + kColST, umma_smem_desc(sKa + x.kt * kTileBytes + koff + kk * 32, 16, 1024),
                              umma_smem_desc(sQa + qoff + kk * 32, 16, 1024), idesc_s, kk > 0);
#pragma unroll
                for (int kk = 0; kk < HD / 16; ++kk)
                    mma(tb + kColDP, umma_smem_desc(sVa + x.kt * kTileBytes + koff + kk * 32, 16, 1024),
                              umma_smem_desc(sDOa + qoff + kk * 32, 16, 1024), idesc_s, kk > 0);
                commit(&st_full[jj % kNBuf]);
            };
            for (int il = 0; il < my_items; ++il) {
                const int stage = ring ? (il & 1) : 0, use = ring ? (il >> 1) : il;
                mbar_wait(&in_full[stage], use & 1);
                tc_fence_after();
                sQa = sQ0 + stage * kTileBytes; sKa = sK0 + stage * kTileBytes; sVa = sV0 + stage * kTileBytes; sDOa = sDO0 + stage * kTileBytes;
                HalfBlk ahead{0, 0, 0, 0};                        // next half block whose scores have not been issued
                int ja = j;
                for (int i = 0; i < kNBuf && ahead.hh < HP; ++i) { issue_scores(ahead, ja++); ahead.next(p.Npad, p.NB); }
                for (HalfBlk c{0, 0, 0, 0}; c.hh < HP; c.next(p.Npad, p.NB), ++j) {
                    if (c.qt == 0 && c.h == 0 && c.kt == 0) mbar_wait(dq_empty, (nhead & 1) ^ 1);   // previous head's dQ has been read out
                    mbar_wait(&pt_full[j % kNBuf], (j / kNBuf) & 1);   // dS^T of this half block written, its TMEM buffer drained
                    tc_fence_after();
                    if (c.h + 1 == HalfBlk::halves(p.Npad, c.qt)) {
                        // dQ[qt] += dS K[kt] : A = dS^T tile read MN-major (M = queries: 2 chunks of 64, 16 KB apart; K = keys)
                        const uint32_t hoff = HD == 32 ? (uint32_t)c.hh * 64u : 0u;
                        if (!ABLATE(16))
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks)
                            mma(tmem_base + Cols::kDQ + c.qt * HD, umma_smem_desc(sDSa + ks * 2048, kTileBytes, 1024),
                                      umma_smem_desc(sKa + c.kt * kTileBytes + ks * 2048 + hoff, kTileBytes, 1024), idesc_dq, (c.kt > 0) || (ks > 0));
                        commit(blk_done);
                        if (c.qt == p.NB - 1 && c.kt == p.NB - 1) { commit(dq_full); ++nhead; }
                    }
                    // refill the buffer just drained (after the dQ MMAs: they gate blk_done, which the element-wise warps wait
                    // for first)
                    if (ahead.hh < HP) { issue_scores(ahead, ja++); ahead.next(p.Npad, p.NB); }
                }
                commit(&in_empty[stage]);
            }
        }
        __syncwarp();
    } else {
        // ===================== element-wise warps (256 threads) =====================
        const int q4 = warp & 3, half = warp >> 2;
        const int row = q4 * 32 + lane;                          // key row inside the 128-key tile (TMEM lane)
        const uint32_t tlane = tmem_base + ((uint32_t)(q4 * 32) << 16);
        const float sl2 = p.scale * kLog2eF;
        const int tid = threadIdx.x;                             // 0..255
        const uint32_t sLa = smem_u32(sL), sDa = smem_u32(sD), sPTa = smem_u32(sPT), sDSa = smem_u32(sDS);
        int j = 0, nblk = 0;                                     // half blocks / blocks processed (all items)
        int wkt = 0, whd = 0;                                    // key tiles / heads whose gradients have been written out
        // Deferred read-out of the TMEM gradient accumulators: it is done inside the NEXT half block, between its
        // arithmetic and its shared-memory stores, so the wait for the last dV/dK (dQ) MMAs hides behind useful work.
        int pend_kt = -1, pend_kt_hh = 0, pend_dq_hh = -1, pend_b = 0, pend_sl = 0;
        const int img = PAIR ? (q4 >> 1) : 0;                  // pair mode: image of this thread's tile row
        const int lrow = PAIR ? (row & 63) : row;              // token index of the row inside its image (before the tile offset)
        if (PAIR) {
            // P^T / dS^T are block diagonal: chunk h (queries of image h) is only ever written by the key rows of image h; the
            // rows of the other image must read as zero.  Warps 0-3 clear P^T, warps 4-7 dS^T.
            const uint32_t base = (half == 0 ? sPTa : sDSa) + (uint32_t)(1 - img) * kTileBytes;
#pragma unroll
            for (int i = 0; i < 8; ++i) sts_v4(base + sw128_off(row, i), 0u, 0u, 0u, 0u);
        }
        auto write_dvk = [&]() {                                 // dV / dK of key tile pend_kt: warps 0-3 write dV, warps 4-7 dK
            const int key = pend_kt * 128 + lrow;
            const int pb_ = PAIR ? 2 * pend_b + img : pend_b;
            const bool ok = key < p.T && pb_ < p.B;
            const int head = pend_sl * HP + pend_kt_hh;
            mbar_wait(dvk_full, wkt & 1);
            tc_fence_after();
            const uint32_t col = half == 0 ? Cols::kDV : Cols::kDK;
            __nv_bfloat16* dst = p.dqkv + (size_t)(pb_ * p.T + key) * (3 * HHD) + (half == 0 ? 2 * HHD : HHD) + head * HD;
            if constexpr (HD == 64) {
#pragma unroll
                for (int h2 = 0; h2 < 2; ++h2) {
                    uint32_t r[32];
                    tmem_ld_32x32(tlane + col + h2 * 32, r);
                    tmem_ld_wait();
                    if (ok) store_row32_bf16(dst + h2 * 32, r);
                }
            } else {
                uint32_t r[32];
                tmem_ld_32x32(tlane + col, r);
                tmem_ld_wait();
                if (ok) store_row32_bf16(dst, r);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(dvk_empty);
            ++wkt; pend_kt = -1;
        };
        auto write_dq = [&]() {                                  // dQ of head pend_dq_hh: warps 0-3 take query tile 0, warps 4-7 tile 1
            const int head = pend_sl * HP + pend_dq_hh;
            mbar_wait(dq_full, whd & 1);
            tc_fence_after();
            if (half < p.NB) {
                const int q = half * 128 + lrow;
                const int pb_ = PAIR ? 2 * pend_b + img : pend_b;
                const bool ok = q < p.T && pb_ < p.B;
                __nv_bfloat16* dst = p.dqkv + (size_t)(pb_ * p.T + q) * (3 * HHD) + head * HD;
                if constexpr (HD == 64) {
#pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2) {
                        uint32_t r[32];
                        tmem_ld_32x32(tlane + Cols::kDQ + half * 64 + h2 * 32, r);
                        tmem_ld_wait();
                        if (ok) store_row32_bf16(dst + h2 * 32, r);
                    }
                } else {
                    uint32_t r[32];
                    tmem_ld_32x32(tlane + Cols::kDQ + half * 32, r);
                    tmem_ld_wait();
                    if (ok) store_row32_bf16(dst, r);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(dq_empty);
            ++whd; pend_dq_hh = -1;
        };
        for (int il = 0; il < my_items; ++il) {
            const int item = blockIdx.x + il * gridDim.x;
            const int b = item / p.slices_per_img, sl = item % p.slices_per_img;
            // ---- per-query statistics of both heads: lse*log2e and D*scale, D = sum_d dO*O.  All global loads of the
            // thread are issued before the first use (one memory round trip per item).
            {
                const int q = tid;                               // slot in the statistics arrays = query row of the tile(s)
                const int qb = PAIR ? 2 * b + (q >> 6) : b;     // image and token of that slot
                const int qt_ = PAIR ? (q & 63) : q;
                uint4 ra[HP][HD / 8], rd[HP][HD / 8];
                float lv[HP], dv[HP];
                if (qt_ < p.T && qb < p.B && (!PAIR || q < 128) && !ABLATE(64)) {
#pragma unroll
                    for (int hh = 0; hh < HP; ++hh) {
                        const size_t off = ((size_t)(qb * p.T + qt_) * p.H + sl * HP + hh) * HD;
#pragma unroll
                        for (int c = 0; c < HD / 8; ++c) { ra[hh][c] = ldg_v4(p.out + off + c * 8); rd[hh][c] = ldg_v4(p.dout + off + c * 8); }
                        lv[hh] = __ldg(p.lse + ((size_t)qb * p.H + sl * HP + hh) * p.T + qt_) * kLog2eF;
                    }
#pragma unroll
                    for (int hh = 0; hh < HP; ++hh) {
                        float acc = 0.f;
#pragma unroll
                        for (int c = 0; c < HD / 8; ++c) {
                            const uint32_t aw[4] = {ra[hh][c].x, ra[hh][c].y, ra[hh][c].z, ra[hh][c].w};
                            const uint32_t dw[4] = {rd[hh][c].x, rd[hh][c].y, rd[hh][c].z, rd[hh][c].w};
#pragma unroll
                            for (int u = 0; u < 4; ++u) { const float2 x = unpack_bf16(aw[u]), y = unpack_bf16(dw[u]); acc += x.x * y.x + x.y * y.y; }
                        }
                        dv[hh] = acc * p.scale;
                    }
                } else {
#pragma unroll
                    for (int hh = 0; hh < HP; ++hh) { lv[hh] = INFINITY; dv[hh] = 0.f; }
                }
                // the NEXT item's O rows (128 contiguous bytes per query: both heads) go to L2 now; dO is already covered by the
                // producer's tile prefetch.  The statistics loads of the next item then cost an L2 round trip, not HBM's.
                if (il + 1 < my_items) {
                    const int item2 = item + gridDim.x;
                    const int b2 = item2 / p.slices_per_img, s2 = item2 % p.slices_per_img;
                    const int qb2 = PAIR ? 2 * b2 + (q >> 6) : b2;
                    if (qt_ < p.T && qb2 < p.B && (!PAIR || q < 128)) {
                        const __nv_bfloat16* nxt = p.out + ((size_t)(qb2 * p.T + qt_) * p.H + s2 * HP) * HD;
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(nxt));
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(p.lse + ((size_t)qb2 * p.H + s2 * HP) * p.T + qt_));
                    }
                }
                named_bar_sync(1, 256);                          // every thread is done reading the previous item's statistics
#pragma unroll
                for (int hh = 0; hh < HP; ++hh) { sts32f(sLa + (hh * 256 + q) * 4, -lv[hh]); sts32f(sDa + (hh * 256 + q) * 4, -dv[hh]); }   // stored NEGATED: addends of the packed FMAs below
                named_bar_sync(1, 256);
            }

            for (int hh = 0; hh < HP; ++hh) {
                for (int kt = 0; kt < p.NB; ++kt) {
                    const int key = kt * 128 + lrow;
                    const bool key_ok = key < p.T && (!PAIR || 2 * b + img < p.B);
                    const bool keys_all_ok = __all_sync(0xffffffffu, key_ok);
                    for (int qt = 0; qt < p.NB; ++qt, ++nblk) {
                        const int nq = min(128, p.Npad - qt * 128);
                        const int nh = nq > 64 ? 2 : 1;
                        for (int h = 0; h < nh; ++h, ++j) {
                            const int nqh = h == 0 ? min(64, nq) : nq - 64;
                            const int c = half * 32;                            // this warp's 32 query columns of the half block
                            // pair mode: half block h holds the queries of image h; key rows of the other image stay zero
                            const bool active = c < nqh && (!PAIR || img == h);
                            uint32_t pw[16], dw[16];
                            mbar_wait(&st_full[j % kNBuf], (j / kNBuf) & 1);
                            tc_fence_after();
                            if (active) {
                                uint32_t rs[32], rd[32];
                                const uint32_t tb = tlane + Cols::buf(j % kNBuf);
                                tmem_ld_32x32(tb + kColST + c, rs);
                                tmem_ld_32x32(tb + kColDP + c, rd);
                                const int q0 = qt * 128 + h * 64 + c;                // slot of the first query in the statistics arrays
                                const uint32_t la = sLa + (hh * 256 + q0) * 4, da = sDa + (hh * 256 + q0) * 4;
                                tmem_ld_wait();
                                // Padding QUERIES need no masking: their statistics slot holds -lse = -inf, so p = exp2(-inf) = 0
                                // and dS = 0 * finite = 0.  Padding KEYS (zero K rows: s = 0, p = exp2(-lse) != 0) do; the select
                                // is only compiled into the path taken by warps that own such rows.
                                // two elements per FFMA2 / FMUL2: (s * scale*log2e - lse*log2e), (dP * scale - D*scale) and their product
                                const f32x2 sl2v = f2_splat(sl2), scv = f2_splat(p.scale);
                                auto body = [&](auto mask_keys) {
#pragma unroll
                                    for (int jj = 0; jj < 32; jj += 4) {
                                        const float4 l4 = lds_v4f(la + jj * 4);      // warp-uniform address: one broadcast wavefront
                                        const float4 d4 = lds_v4f(da + jj * 4);      // (both arrays hold the NEGATED statistics)
                                        const float nl[4] = {l4.x, l4.y, l4.z, l4.w}, nd[4] = {d4.x, d4.y, d4.z, d4.w};
                                        float pv[4], ev[4];
#pragma unroll
                                        for (int u = 0; u < 4; u += 2) {
                                            float a0, a1;
                                            f2_unpack(f2_fma(f2_pack(__uint_as_float(rs[jj + u]), __uint_as_float(rs[jj + u + 1])), sl2v, f2_pack(nl[u], nl[u + 1])), a0, a1);
                                            pv[u] = ABLATE(4) ? a0 : fast_exp2(a0);
                                            pv[u + 1] = ABLATE(4) ? a1 : fast_exp2(a1);
                                            const f32x2 t = f2_fma(f2_pack(__uint_as_float(rd[jj + u]), __uint_as_float(rd[jj + u + 1])), scv, f2_pack(nd[u], nd[u + 1]));
                                            f2_unpack(f2_mul(f2_pack(pv[u], pv[u + 1]), t), ev[u], ev[u + 1]);
                                            if (decltype(mask_keys)::value && !key_ok) { pv[u] = 0.f; pv[u + 1] = 0.f; ev[u] = 0.f; ev[u + 1] = 0.f; }
                                        }
                                        pw[(jj >> 1)] = pack_bf16(pv[0], pv[1]); pw[(jj >> 1) + 1] = pack_bf16(pv[2], pv[3]);
                                        dw[(jj >> 1)] = pack_bf16(ev[0], ev[1]); dw[(jj >> 1) + 1] = pack_bf16(ev[2], ev[3]);
                                    }
                                };
                                if (keys_all_ok) body(std::false_type{}); else body(std::true_type{});
                            }
                            if (pend_kt >= 0) write_dvk();
                            if (pend_dq_hh >= 0) write_dq();
                            // the P^T / dS^T tiles are single buffered: the tensor reads of the previous block must have retired
                            if (h == 0 && nblk > 0 && !ABLATE(1)) mbar_wait(blk_done, (nblk - 1) & 1);
                            if (active && !ABLATE(2)) {
                                const uint32_t tile = (uint32_t)h * kTileBytes;
                                const int sub = c >> 3;
#pragma unroll
                                for (int i = 0; i < 4; ++i) {
                                    const uint32_t off = tile + sw128_off(row, sub + i);
                                    sts_v4(sPTa + off, pw[4 * i], pw[4 * i + 1], pw[4 * i + 2], pw[4 * i + 3]);
                                    sts_v4(sDSa + off, dw[4 * i], dw[4 * i + 1], dw[4 * i + 2], dw[4 * i + 3]);
                                }
                            }
                            fence_proxy_async();
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&pt_full[j % kNBuf]);
                        }
                    }
                    pend_kt = kt; pend_kt_hh = hh; pend_b = b; pend_sl = sl;
                }
                pend_dq_hh = hh;                                 // read out inside the next half block (of the next item, if any)
            }
        }
        if (pend_kt >= 0) write_dvk();                            // the CTA's last item: nothing left to hide behind
        if (pend_dq_hh >= 0) write_dq();
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

// ---- host side ----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn3)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn3 attn_encode_fn() {
    static EncodeTiledFn3 fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn3>(f);
    });
    return fn;
}

// bf16 [B, T, cols] (cols contiguous): box = 64 columns x 128 tokens x 1 image, SWIZZLE_128B, zero fill for t >= T.
int make_tmap_tokens(CUtensorMap* out, const void* ptr, long long cols, long long T, long long B, int box_rows) {
    struct Key { const void* p; long long c, t, b; int r; bool operator==(const Key& o) const { return p == o.p && c == o.c && t == o.t && b == o.b && r == o.r; } };
    struct Hash { size_t operator()(const Key& k) const { return std::hash<const void*>()(k.p) ^ (size_t)(k.c * 1315423911ull) ^ (size_t)(k.t * 2654435761ull) ^ (size_t)(k.b << 9) ^ (size_t)k.r; } };
    static thread_local std::unordered_map<Key, CUtensorMap, Hash> cache;
    Key key{ptr, cols, T, B, box_rows};
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return SSL4P_OK; }
    EncodeTiledFn3 enc = attn_encode_fn();
    if (!enc) { set_error("cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)"); return SSL4P_ERR_CUDA; }
    SSL4P_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "attention: operand pointer must be 16-byte aligned");
    SSL4P_REQUIRE(cols % 64 == 0, "attention: H*head_dim must be a multiple of 64 (got row of %lld columns)", cols);
    cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)T, (cuuint64_t)B};
    cuuint64_t strides[2] = {(cuuint64_t)cols * 2, (cuuint64_t)cols * 2 * (cuuint64_t)T};
    SSL4P_REQUIRE(box_rows >= 8 && box_rows <= 256 && box_rows % 8 == 0, "attention: bad TMA box (%d rows)", box_rows);
    cuuint32_t box[3] = {64u, (cuuint32_t)box_rows, 1u};
    cuuint32_t estr[3] = {1u, 1u, 1u};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("attention: cuTensorMapEncodeTiled failed (CUresult %d) cols=%lld T=%lld B=%lld", (int)r, cols, T, B); return SSL4P_ERR_CUDA; }
    if (cache.size() > 1024) cache.clear();
    cache.emplace(key, *out);
    return SSL4P_OK;
}

constexpr int kFwdSmem = 14 * kTileBytes + 256 + 1024;
constexpr int kBwdSmem = 12 * kTileBytes + 4096 + 256 + 1024;

template <int HD>
static int launch_fwd_tc(const void* qkv, void* out, float* lse, int B, int T, int H, float scale, cudaStream_t s) {
    CUtensorMap tm, tm64;
    int rc = make_tmap_tokens(&tm, qkv, 3LL * H * HD, T, B, 128);
    if (rc) return rc;
    rc = make_tmap_tokens(&tm64, qkv, 3LL * H * HD, T, B, 64);
    if (rc) return rc;
    AttnFwdParams p;
    p.B = B; p.T = T; p.H = H; p.slices_per_img = H * HD / 64;
    // Short sequences (T <= 64, the masked MAE encoder: 50 tokens): two images share every 128-row tile, the score matrix is
    // 128 x 128 with the two diagonal 64 x 64 blocks used - half the tiles, MMAs and softmax rows of one image per tile.
    p.pair = (T <= 64) ? 1 : 0;
    if (const char* e = getenv("SSL4P_ATTN_PAIR")) p.pair = (T <= 64 && e[0] != '0') ? 1 : 0;   // tests: either mapping
    p.items = (p.pair ? (B + 1) / 2 : B) * p.slices_per_img;
    p.Npad = p.pair ? 128 : (T + 15) / 16 * 16; p.NB = (T + 127) / 128; p.NKB = (p.Npad + 127) / 128;
    p.scale = scale; p.out = (__nv_bfloat16*)out; p.lse = lse;
    p.backoff_ns = 48;
    if (const char* e = getenv("SSL4P_ATTN_BACKOFF")) p.backoff_ns = atoi(e);
    p.s_stride = (p.Npad + 31) / 32 * 32;
    // Separate O columns pay off when a chain has several tiles per item (T > 128: 175 -> 135 us at the decoder shape); for
    // one-tile items (T = 50: 49 vs 55 us) the read-out that is deferred by a whole item costs more than it hides.
    const bool fits = 2 * p.s_stride + 2 * HD <= 512;
    p.sep = (fits && p.NB >= 2) ? 1 : 0;
    if (const char* e = getenv("SSL4P_ATTN_SEP")) p.sep = (fits && e[0] != '0') ? 1 : 0;      // tests: force either TMEM plan
    auto kern = p.pair ? (p.sep ? attention_fwd_tc_kernel<HD, true, true> : attention_fwd_tc_kernel<HD, true, false>)
                       : (p.sep ? attention_fwd_tc_kernel<HD, false, true> : attention_fwd_tc_kernel<HD, false, false>);
    static bool attr[4] = {false, false, false, false};
    const int ki = (p.pair ? 2 : 0) + (p.sep ? 1 : 0);
    if (!attr[ki]) { SSL4P_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem)); attr[ki] = true; }
    const int grid = p.items < num_sms() ? p.items : num_sms();
    cudaError_t e = launch_ex(kern, dim3((unsigned)grid), dim3(kAttnThreads), (size_t)kFwdSmem, s, 1, tm, tm64, p);
    if (e != cudaSuccess) { set_error("attention_fwd_tc_kernel launch failed: %s", cudaGetErrorString(e)); return SSL4P_ERR_CUDA; }
    return check_launch("attention_fwd_tc_kernel");
}

template <int HD>
static int launch_bwd_tc(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int B, int T, int H,
                         float scale, cudaStream_t s) {
    CUtensorMap tmQ, tmD, tmQ64, tmD64;
    int rc = make_tmap_tokens(&tmQ, qkv, 3LL * H * HD, T, B, 128);
    if (rc) return rc;
    rc = make_tmap_tokens(&tmD, dout, (long long)H * HD, T, B, 128);
    if (rc) return rc;
    rc = make_tmap_tokens(&tmQ64, qkv, 3LL * H * HD, T, B, 64);
    if (rc) return rc;
    rc = make_tmap_tokens(&tmD64, dout, (long long)H * HD, T, B, 64);
    if (rc) return rc;
    AttnBwdParams p;
    p.B = B; p.T = T; p.H = H; p.slices_per_img = H * HD / 64;
    p.pair = (T <= 64) ? 1 : 0;                                  // see launch_fwd_tc
    if (const char* e = getenv("SSL4P_ATTN_PAIR")) p.pair = (T <= 64 && e[0] != '0') ? 1 : 0;
    p.items = (p.pair ? (B + 1) / 2 : B) * p.slices_per_img;
    p.Npad = p.pair ? 128 : (T + 15) / 16 * 16; p.NB = (T + 127) / 128;
    p.scale = scale; p.out = (const __nv_bfloat16*)out; p.dout = (const __nv_bfloat16*)dout; p.lse = lse; p.dqkv = (__nv_bfloat16*)dqkv;
    p.debug = 0;
    if (const char* e = getenv("SSL4P_ATTN_DEBUG")) p.debug = atoi(e);
    auto kern = p.pair ? attention_bwd_tc_kernel<HD, true> : attention_bwd_tc_kernel<HD, false>;
    static bool attr[2] = {false, false};
    if (!attr[p.pair]) { SSL4P_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem)); attr[p.pair] = true; }
    const int grid = p.items < num_sms() ? p.items : num_sms();
    cudaError_t e = launch_ex(kern, dim3((unsigned)grid), dim3(kAttnBwdThreads), (size_t)kBwdSmem, s, 1, tmQ, tmD, tmQ64, tmD64, p);
    if (e != cudaSuccess) { set_error("attention_bwd_tc_kernel launch failed: %s", cudaGetErrorString(e)); return SSL4P_ERR_CUDA; }
    return check_launch("attention_bwd_tc_kernel");
}

}  // namespace ssl4p

using namespace ssl4p;

static int attn_tc_check(int B, int T, int H, int hd) {
    SSL4P_REQUIRE(B > 0 && H > 0 && T > 0, "attention: bad shape");
    SSL4P_REQUIRE(hd == 32 || hd == 64, "attention: head_dim %d not supported (32 or 64)", hd);
    SSL4P_REQUIRE((H * hd) % 64 == 0, "attention: H*head_dim (%d) must be a multiple of 64", H * hd);
    SSL4P_REQUIRE(T <= 256, "attention: sequence length %d not supported (<= 256)", T);
    return SSL4P_OK;
}

extern "C" int ssl4p_attention_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int hd, float scale, void* stream) {
    SSL4P_REQUIRE(qkv && out && lse, "attention_fwd: null pointer");
    int rc = attn_tc_check(B, T, H, hd);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (hd == 64) return launch_fwd_tc<64>(qkv, out, lse, B, T, H, scale, s);
    return launch_fwd_tc<32>(qkv, out, lse, B, T, H, scale, s);
}

extern "C" int ssl4p_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                   int B, int T, int H, int hd, float scale, void* stream) {
    SSL4P_REQUIRE(qkv && out && dout && lse && dqkv, "attention_bwd: null pointer");
    int rc = attn_tc_check(B, T, H, hd);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (hd == 64) return launch_bwd_tc<64>(qkv, out, dout, lse, dqkv, B, T, H, scale, s);
    return launch_bwd_tc<32>(qkv, out, dout, lse, dqkv, B, T, H, scale, s);
}
