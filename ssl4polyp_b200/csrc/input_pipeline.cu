// Pretraining input transform on the device: uint8 HWC frames -> RandomResizedCrop (PIL bicubic, bit-exact) ->
// horizontal flip -> ToTensor -> Normalize -> fp32 NCHW, the tensor MaskedAutoencoderViT.forward takes.
// Reference call site: /root/reference/src/ssl4polyp/models/mae/main_pretrain.py:156-160 (a torchvision Compose that
// runs per image on DataLoader workers).  The arithmetic restated here is Pillow's libImaging/Resample.c for 8-bit
// channels (ImagingResample: precompute_coeffs, normalize_coeffs_8bpc, ImagingResampleHorizontal_8bpc,
// ImagingResampleVertical_8bpc) and torchvision's to_tensor / normalize; oracle/input_oracle.py is the CPU
// restatement, pinned bit for bit against the installed Pillow / torchvision (tests/golden/input_golden.npz).
//
// Byte / integer work, HBM- and LSU-bound (no tensor cores): the crop parameters are drawn on the host (a few RNG
// calls per image, the reference's own sequence), the host sends the raw uint8 crop rectangles in one copy per batch
// (a memcpy per image on the host instead of a bicubic resample: 2-6 ms of a core per image), and two kernels do the rest:
//   1. resample_coeffs_kernel   per (image, axis): the 224 windows and their 22-bit fixed-point weights, in fp64
//      with explicitly rounded operations (no FMA contraction: Pillow is compiled for baseline x86-64).
//   2. crop_resize_normalize_kernel   per (band of output rows, image): horizontal pass of the source rows the band
//      needs into a uint8 shared-memory tile (Pillow's intermediate image is uint8 too), vertical pass out of the
//      tile, then flip / 255 / mean / std and coalesced fp32 stores.
#include "common.cuh"
#include "ssl4p.h"

namespace ssl4p {

constexpr int kPrecBits = 32 - 8 - 2;        // Resample.c PRECISION_BITS
constexpr int kMaxTaps = SSL4P_RESAMPLE_MAX_TAPS;
constexpr int kDescStride = 8;               // int64 per image: offset, pitch, top, left, height, width, flip, 0

__device__ __forceinline__ double bicubic_weight(double x) {        // Resample.c bicubic_filter, a = -0.5
    if (x < 0.0) x = -x;
    if (x < 1.0) {                                                   // ((a + 2) x - (a + 3)) x x + 1
        double t = __dmul_rn(1.5, x);
        t = __dadd_rn(t, -2.5);
        t = __dmul_rn(t, x);
        t = __dmul_rn(t, x);
        return __dadd_rn(t, 1.0);
    }
    if (x < 2.0) {                                                   // (((x - 5) x + 8) x - 4) a
        double t = __dadd_rn(x, -5.0);
        t = __dmul_rn(t, x);
        t = __dadd_rn(t, 8.0);
        t = __dmul_rn(t, x);
        t = __dadd_rn(t, -4.0);
        return __dmul_rn(t, -0.5);
    }
    return 0.0;
}

// Workspace per (image n, axis a; a = 0 horizontal, 1 vertical): int32 bounds[S][2] = (first source index, taps;
// taps == 0 marks a skipped pass) followed by int32 weights[kMaxTaps][S] (tap-major: lanes = adjacent outputs).
__device__ __forceinline__ size_t axis_words(int S) { return (size_t)S * (kMaxTaps + 2); }

__device__ __forceinline__ double bilinear_weight(double x) {       // Resample.c bilinear_filter (support 1)
    if (x < 0.0) x = -x;
    return x < 1.0 ? __dadd_rn(1.0, -x) : 0.0;
}

__device__ __forceinline__ double filter_weight(int filter, double x) {
    return filter == SSL4P_RESAMPLE_BILINEAR ? bilinear_weight(x) : bicubic_weight(x);
}

__global__ void resample_coeffs_kernel(const long long* __restrict__ desc, int S, int filter, int* __restrict__ ws) {
    const int n = blockIdx.x, axis = blockIdx.y;
    const long long* d = desc + (size_t)n * kDescStride;
    const int in_size = (int)(axis == 0 ? d[5] : d[4]);
    int* bounds = ws + ((size_t)n * 2 + axis) * axis_words(S);
    int* weights = bounds + 2 * S;
    for (int xx = threadIdx.x; xx < S; xx += blockDim.x) {
        if (in_size == S) {                                          // ImagingResample skips a pass that keeps the size
            bounds[2 * xx] = xx;
            bounds[2 * xx + 1] = 0;
            continue;
        }
        const double scale = __ddiv_rn((double)in_size, (double)S);  // (in1 - in0) / outSize, box = the whole crop
        const double filterscale = scale < 1.0 ? 1.0 : scale;
        const double support = __dmul_rn(filter == SSL4P_RESAMPLE_BILINEAR ? 1.0 : 2.0, filterscale);
        const double ss = __ddiv_rn(1.0, filterscale);
        const double center = __dmul_rn((double)xx + 0.5, scale);    // in0 (= 0.0) + (xx + 0.5) * scale
        int xmin = __double2int_rz(__dadd_rn(__dadd_rn(center, -support), 0.5));
        if (xmin < 0) xmin = 0;
        int xmax = __double2int_rz(__dadd_rn(__dadd_rn(center, support), 0.5));
        if (xmax > in_size) xmax = in_size;
        xmax -= xmin;
        if (xmax > kMaxTaps) xmax = kMaxTaps;                        // unreachable: the host checks the largest crop
        double ww = 0.0;
        for (int x = 0; x < xmax; ++x)
            ww = __dadd_rn(ww, filter_weight(filter, __dmul_rn(__dadd_rn(__dadd_rn((double)(x + xmin), -center), 0.5), ss)));
        for (int x = 0; x < xmax; ++x) {
            double w = filter_weight(filter, __dmul_rn(__dadd_rn(__dadd_rn((double)(x + xmin), -center), 0.5), ss));
            if (ww != 0.0) w = __ddiv_rn(w, ww);
            const double f = __dmul_rn(w, (double)(1 << kPrecBits));
            weights[(size_t)x * S + xx] = w < 0.0 ? __double2int_rz(__dadd_rn(-0.5, f)) : __double2int_rz(__dadd_rn(0.5, f));
        }
        bounds[2 * xx] = xmin;
        bounds[2 * xx + 1] = xmax;
    }
}

__device__ __forceinline__ int clip8(int acc) {
    acc >>= kPrecBits;
    return acc < 0 ? 0 : (acc > 255 ? 255 : acc);
}

// kS > 0: the output size is a compile-time constant (224: the index divisions become multiplies); kS == 0: runtime S
template <int kS>
__global__ void __launch_bounds__(256)
crop_resize_normalize_kernel(const uint8_t* __restrict__ frames, const long long* __restrict__ desc, const int* __restrict__ ws,
                             int S_rt, int band, int tile_rows, float m0, float m1, float m2, float s0, float s1, float s2,
                             float* __restrict__ out, uint8_t* __restrict__ out_u8) {
    extern __shared__ uint8_t tile[];                                // [tile_rows][3 S] uint8: rows of the horizontally resized crop
    const int S = kS > 0 ? kS : S_rt;
    const int n = blockIdx.y;
    const int r0 = blockIdx.x * band;
    const int rows = min(band, S - r0);
    const long long* d = desc + (size_t)n * kDescStride;
    const long long off = d[0], pitch = d[1];
    const int top = (int)d[2], left = (int)d[3], ch = (int)d[4], cw = (int)d[5];
    const bool flip = d[6] != 0;
    const int* bx = ws + ((size_t)n * 2) * axis_words(S);
    const int* wx = bx + 2 * S;
    const int* by = bx + axis_words(S);
    const int* wy = by + 2 * S;
    const bool hpass = cw != S, vpass = ch != S;
    const int row_bytes = 3 * S;

    int ylo = r0, yhi = r0 + rows;
    if (vpass) {
        ylo = by[2 * r0];
        yhi = 0;
        for (int r = 0; r < rows; ++r) yhi = max(yhi, by[2 * (r0 + r)] + by[2 * (r0 + r) + 1]);
    }
    const int nrows = min(yhi - ylo, tile_rows);                     // the host sizes tile_rows from the largest crop

    // horizontal pass (or a plain copy) of source rows [ylo, ylo + nrows) of the crop
    const uint8_t* base = frames + off + (long long)(top + ylo) * pitch + (long long)left * 3;
    for (int idx = threadIdx.x; idx < nrows * S; idx += blockDim.x) {
        const int y = idx / S, xx = idx - y * S;
        const uint8_t* row = base + (long long)y * pitch;
        uint8_t* dst = tile + y * row_bytes + xx * 3;
        if (hpass) {
            const int xmin = bx[2 * xx], taps = bx[2 * xx + 1];
            const uint8_t* p = row + xmin * 3;
            int a0 = 1 << (kPrecBits - 1), a1 = a0, a2 = a0;
            const int* kp = wx + xx;
#pragma unroll 4
            for (int x = 0; x < taps; ++x, p += 3, kp += S) {
                const int k = __ldg(kp);
                a0 += (int)p[0] * k;
                a1 += (int)p[1] * k;
                a2 += (int)p[2] * k;
            }
            dst[0] = (uint8_t)clip8(a0); dst[1] = (uint8_t)clip8(a1); dst[2] = (uint8_t)clip8(a2);
        } else {
            const uint8_t* p = row + xx * 3;
            dst[0] = p[0]; dst[1] = p[1]; dst[2] = p[2];
        }
    }
    __syncthreads();

    // vertical pass out of the tile (one thread = one output pixel, three channel accumulators sharing each weight
    // load), then hflip -> /255 -> (x - mean) / std (torchvision to_tensor + normalize, fp32)
    for (int idx = threadIdx.x; idx < rows * S; idx += blockDim.x) {
        const int r = idx / S, xx = idx - r * S;
        const int rr = r0 + r;
        int v0, v1, v2;
        if (vpass) {
            const int ymin = by[2 * rr], taps = by[2 * rr + 1];
            const uint8_t* p = tile + (ymin - ylo) * row_bytes + xx * 3;
            const int* kp = wy + rr;
            int a0 = 1 << (kPrecBits - 1), a1 = a0, a2 = a0;
#pragma unroll 4
            for (int t = 0; t < taps; ++t, p += row_bytes, kp += S) {
                const int k = __ldg(kp);
                a0 += (int)p[0] * k;
                a1 += (int)p[1] * k;
                a2 += (int)p[2] * k;
            }
            v0 = clip8(a0); v1 = clip8(a1); v2 = clip8(a2);
        } else {
            const uint8_t* p = tile + r * row_bytes + xx * 3;
            v0 = p[0]; v1 = p[1]; v2 = p[2];
        }
        if (out_u8 != nullptr) {                                     // the resized image itself (uint8 HWC), for the augmentation kernels
            uint8_t* o8 = out_u8 + (((size_t)n * S + rr) * S + xx) * 3;
            o8[0] = (uint8_t)v0; o8[1] = (uint8_t)v1; o8[2] = (uint8_t)v2;
            continue;
        }
        const int xo = flip ? S - 1 - xx : xx;
        float* o = out + (((size_t)n * 3) * S + rr) * S + xo;
        const size_t plane = (size_t)S * S;
        o[0] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v0, 255.0f), m0), s0);
        o[plane] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v1, 255.0f), m1), s1);
        o[2 * plane] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v2, 255.0f), m2), s2);
    }
}

// rows of the horizontally resized crop a band of `band` output rows can touch, for the largest crop of the batch
static int band_tile_rows(int band, int max_crop, int S) {
    const double scale = max_crop > S ? (double)max_crop / S : 1.0;
    return (int)((band - 1) * scale) + 2 * (int)ceil(2.0 * scale) + 3;
}

}  // namespace ssl4p

using namespace ssl4p;
#define STREAM ((cudaStream_t)stream)

extern "C" size_t ssl4p_crop_resize_workspace_bytes(int N, int S) {
    return (size_t)N * 2 * (size_t)S * (kMaxTaps + 2) * sizeof(int);
}

static int crop_resize_impl(const void* frames, const long long* desc, int N, int S, int max_crop, int filter,
                            const float* mean3, const float* std3, float* out, uint8_t* out_u8,
                            void* workspace, size_t workspace_bytes, void* stream) {
    SSL4P_REQUIRE(frames && desc && (out || out_u8) && mean3 && std3, "crop_resize: null argument");
    SSL4P_REQUIRE(N >= 0 && S > 0 && S <= 1024 && max_crop > 0, "crop_resize: bad shape N=%d S=%d max_crop=%d", N, S, max_crop);
    SSL4P_REQUIRE(filter == SSL4P_RESAMPLE_BICUBIC || filter == SSL4P_RESAMPLE_BILINEAR, "crop_resize: unknown filter %d", filter);
    const double scale = max_crop > S ? (double)max_crop / S : 1.0;
    const int ksize = (int)ceil((filter == SSL4P_RESAMPLE_BILINEAR ? 1.0 : 2.0) * scale) * 2 + 1;   // Resample.c: ceil(support) * 2 + 1
    SSL4P_REQUIRE(ksize <= kMaxTaps, "crop_resize: a %d-pixel crop to %d needs %d taps per output (limit %d)", max_crop, S, ksize, kMaxTaps);
    SSL4P_REQUIRE(workspace && workspace_bytes >= ssl4p_crop_resize_workspace_bytes(N, S), "crop_resize: workspace too small");
    if (N == 0) return SSL4P_OK;
    // the tallest band whose tile fits the budget (56 KB: four CTAs per SM hide the byte-load latency of the horizontal pass)
    static const int g_tile_kb = getenv("SSL4P_RESIZE_TILE_KB") ? atoi(getenv("SSL4P_RESIZE_TILE_KB")) : 56;   // measured: 96 -> 0.81 / 2.39 ms, 56 -> 0.70 / 2.17 ms (576x720 / 1072x1920 sources)
    int band = 1;
    const int cands[6] = {32, 16, 8, 4, 2, 1};
    size_t smem = 0;
    for (int i = 0; i < 6; ++i) {
        smem = (size_t)band_tile_rows(cands[i], max_crop, S) * 3 * S;
        if (smem <= (size_t)g_tile_kb * 1024 || cands[i] == 1) { band = cands[i]; break; }
    }
    SSL4P_REQUIRE(smem <= 227 * 1024, "crop_resize: crop %d -> %d does not fit shared memory", max_crop, S);
    const int tile_rows = band_tile_rows(band, max_crop, S);
    auto kernel = S == 224 ? crop_resize_normalize_kernel<224> : crop_resize_normalize_kernel<0>;
    SSL4P_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    resample_coeffs_kernel<<<dim3(N, 2), 256, 0, STREAM>>>(desc, S, filter, (int*)workspace);
    int rc = check_launch("resample_coeffs_kernel");
    if (rc) return rc;
    kernel<<<dim3((S + band - 1) / band, N), 256, smem, STREAM>>>(
        (const uint8_t*)frames, desc, (const int*)workspace, S, band, tile_rows,
        mean3[0], mean3[1], mean3[2], std3[0], std3[1], std3[2], out, out_u8);
    return check_launch("crop_resize_normalize_kernel");
}

extern "C" int ssl4p_crop_resize_normalize_u8(const void* frames, const long long* desc, int N, int S, int max_crop, int filter,
                                              const float* mean3, const float* std3, float* out,
                                              void* workspace, size_t workspace_bytes, void* stream) {
    SSL4P_REQUIRE(out, "crop_resize: null output");
    return crop_resize_impl(frames, desc, N, S, max_crop, filter, mean3, std3, out, nullptr, workspace, workspace_bytes, stream);
}

extern "C" int ssl4p_crop_resize_u8(const void* frames, const long long* desc, int N, int S, int max_crop, int filter,
                                    void* out_u8, void* workspace, size_t workspace_bytes, void* stream) {
    static const float unit[3] = {0.f, 0.f, 0.f}, one[3] = {1.f, 1.f, 1.f};
    SSL4P_REQUIRE(out_u8, "crop_resize_u8: null output");
    return crop_resize_impl(frames, desc, N, S, max_crop, filter, unit, one, nullptr, (uint8_t*)out_u8, workspace, workspace_bytes, stream);
}
