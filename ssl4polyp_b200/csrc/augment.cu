// Fine-tune TRAIN augmentations on the device, applied to the resized uint8 image (HWC):
//   ColorJitter(0.4, 0.5, 0.25, 0.01) -> GaussianBlur((25, 25), sigma in (0.001, 2)) -> RandomHorizontalFlip ->
//   RandomVerticalFlip -> RandomRotation(180) -> ToTensor -> Normalize
// Reference call site: /root/reference/src/ssl4polyp/classification/data/transforms.py:233-258 (stage 'train'; a torchvision
// Compose on PIL images, one image at a time on DataLoader workers).  The arithmetic restated here is Pillow's
// (ImageEnhance = Image.blend with a degenerate image: libImaging/Blend.c; Convert.c rgb2l, rgb2hsv_row, hsv2rgb;
// Image.rotate NEAREST = Geometry.c affine_fixed) and torchvision's (adjust_hue's uint8 wrap-around, gaussian_blur with a
// reflect pad and round-half-even).  oracle/input_oracle.py is the CPU restatement, pinned against the installed libraries
// by oracle/make_augment_golden.py (the HSV conversions over all 2^24 inputs); these kernels are bit-identical to it.
// All random parameters are drawn on the host with the reference's own torch RNG calls (input_pipeline.TrainAugment).
//
// Byte / integer work: one CTA per image for the colour chain (the whole 150 KB image lives in shared memory; the contrast
// step needs the image-wide mean of the CURRENT image in the chain), HBM-bound streaming kernels for the rest.
#include "common.cuh"
#include "ssl4p.h"

namespace ssl4p {

constexpr int kAugInts = SSL4P_AUG_PARAM_INTS;       // int32 words per image
constexpr int kAugTaps = SSL4P_AUG_MAX_TAPS;

// Blend.c: out = (UINT8)(a + alpha * (b - a)) in float (separate multiply and add); clipped when alpha extrapolates
__device__ __forceinline__ int blend1(int a, int b, float alpha, bool interp) {
    const float t = __fadd_rn((float)a, __fmul_rn(alpha, (float)(b - a)));
    if (interp) return (int)t;
    return t <= 0.0f ? 0 : (t >= 255.0f ? 255 : (int)t);
}
__device__ __forceinline__ int rgb2l(int r, int g, int b) { return (r * 19595 + g * 38470 + b * 7471 + 0x8000) >> 16; }
__device__ __forceinline__ int clip255(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }

// Convert.c rgb2hsv_row (float quotients; hue assembled in double, stored to float, scaled in double, truncated)
__device__ __forceinline__ void rgb2hsv(int r, int g, int b, int& uh, int& us, int& uv) {
    const int maxc = max(r, max(g, b)), minc = min(r, min(g, b));
    uv = maxc;
    if (minc == maxc) { uh = 0; us = 0; return; }
    const float cr = (float)(maxc - minc);
    const float s = __fdiv_rn(cr, (float)maxc);
    const float rc = __fdiv_rn((float)(maxc - r), cr), gc = __fdiv_rn((float)(maxc - g), cr), bc = __fdiv_rn((float)(maxc - b), cr);
    double hd;
    if (r == maxc) hd = (double)__fsub_rn(bc, gc);
    else if (g == maxc) hd = __dsub_rn(__dadd_rn(2.0, (double)rc), (double)bc);
    else hd = __dsub_rn(__dadd_rn(4.0, (double)gc), (double)rc);
    float h = __double2float_rn(hd);
    double t = __dadd_rn(__ddiv_rn((double)h, 6.0), 1.0);        // in [5/6, 11/6]: fmod(t, 1) = t - 1 (exact) or t
    if (t >= 1.0) t = __dsub_rn(t, 1.0);
    h = __double2float_rn(t);
    uh = clip255(__double2int_rz(__dmul_rn((double)h, 255.0)));
    us = clip255(__double2int_rz(__dmul_rn((double)s, 255.0)));
}

// Convert.c hsv2rgb
__device__ __forceinline__ void hsv2rgb(int h, int s, int v, int& r, int& g, int& b) {
    if (s == 0) { r = g = b = v; return; }
    const double h6 = __ddiv_rn(__dmul_rn((double)h, 6.0), 255.0);
    const int i = (int)floor(h6);
    const double f = (double)__double2float_rn(__dsub_rn(h6, (double)i));
    const double fs = (double)__double2float_rn(__ddiv_rn((double)s, 255.0));
    const double vd = (double)v;
    const int p = clip255((int)floor(__dadd_rn(__dmul_rn(vd, __dsub_rn(1.0, fs)), 0.5)));
    const int q = clip255((int)floor(__dadd_rn(__dmul_rn(vd, __dsub_rn(1.0, __dmul_rn(fs, f))), 0.5)));
    const int t = clip255((int)floor(__dadd_rn(__dmul_rn(vd, __dsub_rn(1.0, __dmul_rn(fs, __dsub_rn(1.0, f)))), 0.5)));
    switch (i % 6) {
        case 0: r = v; g = t; b = p; break;
        case 1: r = q; g = v; b = p; break;
        case 2: r = p; g = v; b = t; break;
        case 3: r = p; g = q; b = v; break;
        case 4: r = t; g = p; b = v; break;
        default: r = v; g = p; b = q; break;
    }
}

// One CTA per image; img [N, S*S, 3] uint8 in place or to `out`.  params int32 [N, kAugInts]: [0..3] order of the four
// operations (0 brightness, 1 contrast, 2 saturation, 3 hue), [4] hue shift (uint8), [16..18] float bits of the brightness /
// contrast / saturation factors.
__global__ void __launch_bounds__(1024)
color_jitter_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, const int* __restrict__ params, int npix) {
    extern __shared__ uint8_t simg[];
    __shared__ unsigned long long s_sum;
    const int n = blockIdx.x;
    const int* prm = params + (size_t)n * kAugInts;
    const int nbytes = npix * 3;
    const uint8_t* src = in + (size_t)n * nbytes;
    for (int i = threadIdx.x * 4; i < nbytes; i += blockDim.x * 4)           // npix % 4 == 0 is required by the host
        *reinterpret_cast<uint32_t*>(simg + i) = *reinterpret_cast<const uint32_t*>(src + i);
    __syncthreads();
    for (int k = 0; k < 4; ++k) {
        const int op = prm[k];
        if (op == 0) {                                            // ImageEnhance.Brightness: blend(black, img, f)
            const float f = __int_as_float(prm[16]);
            const bool interp = f >= 0.0f && f <= 1.0f;
            for (int i = threadIdx.x; i < nbytes; i += blockDim.x) simg[i] = (uint8_t)blend1(0, simg[i], f, interp);
        } else if (op == 1) {                                     // ImageEnhance.Contrast: blend(int(mean(L) + 0.5), img, f)
            const float f = __int_as_float(prm[17]);
            const bool interp = f >= 0.0f && f <= 1.0f;
            if (threadIdx.x == 0) s_sum = 0ull;
            __syncthreads();
            unsigned int part = 0;
            for (int i = threadIdx.x; i < npix; i += blockDim.x) part += (unsigned)rgb2l(simg[3 * i], simg[3 * i + 1], simg[3 * i + 2]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
            if ((threadIdx.x & 31) == 0) atomicAdd(&s_sum, (unsigned long long)part);      // integer sum: order-independent
            __syncthreads();
            const int mean = __double2int_rz(__dadd_rn(__ddiv_rn((double)s_sum, (double)npix), 0.5));
            for (int i = threadIdx.x; i < nbytes; i += blockDim.x) simg[i] = (uint8_t)blend1(mean, simg[i], f, interp);
        } else if (op == 2) {                                     // ImageEnhance.Color: blend(L, img, f)
            const float f = __int_as_float(prm[18]);
            const bool interp = f >= 0.0f && f <= 1.0f;
            for (int i = threadIdx.x; i < npix; i += blockDim.x) {
                const int r = simg[3 * i], g = simg[3 * i + 1], b = simg[3 * i + 2];
                const int L = rgb2l(r, g, b);
                simg[3 * i] = (uint8_t)blend1(L, r, f, interp);
                simg[3 * i + 1] = (uint8_t)blend1(L, g, f, interp);
                simg[3 * i + 2] = (uint8_t)blend1(L, b, f, interp);
            }
        } else if (op == 3) {                                     // adjust_hue: RGB -> HSV, h += shift (mod 256), HSV -> RGB
            const int shift = prm[4];
            for (int i = threadIdx.x; i < npix; i += blockDim.x) {
                int h, s, v, r, g, b;
                rgb2hsv(simg[3 * i], simg[3 * i + 1], simg[3 * i + 2], h, s, v);
                hsv2rgb((h + shift) & 255, s, v, r, g, b);
                simg[3 * i] = (uint8_t)r; simg[3 * i + 1] = (uint8_t)g; simg[3 * i + 2] = (uint8_t)b;
            }
        }
        __syncthreads();
    }
    uint8_t* dst = out + (size_t)n * nbytes;
    for (int i = threadIdx.x * 4; i < nbytes; i += blockDim.x * 4)
        *reinterpret_cast<uint32_t*>(dst + i) = *reinterpret_cast<const uint32_t*>(simg + i);
}

__device__ __forceinline__ int reflect(int i, int n) { return i < 0 ? -i : (i >= n ? 2 * (n - 1) - i : i); }

// Separable Gaussian in float64, the oracle's order of operations (tap 0 first, separate multiply and add).
// pass 0: uint8 [N,S,S,3] -> double tmp (rows);  taps float [N, kAugTaps] (ntaps used)
__global__ void __launch_bounds__(256)
blur_rows_kernel(const uint8_t* __restrict__ in, const float* __restrict__ taps, int ntaps, int S, double* __restrict__ tmp) {
    const int n = blockIdx.z, y = blockIdx.y;
    const int row = S * 3;
    const uint8_t* src = in + ((size_t)n * S + y) * row;
    const float* k = taps + (size_t)n * kAugTaps;
    const int r = ntaps >> 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < row; i += gridDim.x * blockDim.x) {
        const int x = i / 3, c = i - 3 * x;
        double acc = 0.0;
        for (int j = 0; j < ntaps; ++j)
            acc = __dadd_rn(acc, __dmul_rn((double)__ldg(k + j), (double)src[reflect(x + j - r, S) * 3 + c]));
        tmp[((size_t)n * S + y) * row + i] = acc;
    }
}
// pass 1: columns of tmp -> round half to even -> uint8
__global__ void __launch_bounds__(256)
blur_cols_kernel(const double* __restrict__ tmp, const float* __restrict__ taps, int ntaps, int S, uint8_t* __restrict__ out) {
    const int n = blockIdx.z, y = blockIdx.y;
    const int row = S * 3;
    const float* k = taps + (size_t)n * kAugTaps;
    const int r = ntaps >> 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < row; i += gridDim.x * blockDim.x) {
        double acc = 0.0;
        for (int j = 0; j < ntaps; ++j)
            acc = __dadd_rn(acc, __dmul_rn((double)__ldg(k + j), tmp[((size_t)n * S + reflect(y + j - r, S)) * row + i]));
        out[((size_t)n * S + y) * row + i] = (uint8_t)clip255(__double2int_rn(acc));
    }
}

// flips + Image.rotate(NEAREST, expand=False, fill 0) as ONE gather, then ToTensor + Normalize -> fp32 [N,3,S,S]
// (and, optionally, the uint8 HWC result).  params: [5] hflip, [6] vflip, [7] rotation mode (0 = 16.16 fixed-point affine,
// 1 identity, 2 = 180, 3 = 90 ccw, 4 = 270), [8..13] a0..a5.
__global__ void __launch_bounds__(256)
flip_rotate_normalize_kernel(const uint8_t* __restrict__ in, const int* __restrict__ params, int S, float m0, float m1, float m2,
                             float s0, float s1, float s2, float* __restrict__ out, uint8_t* __restrict__ out_u8) {
    const int n = blockIdx.z, y = blockIdx.y;
    const int* prm = params + (size_t)n * kAugInts;
    const int hflip = prm[5], vflip = prm[6], mode = prm[7];
    const long long a0 = prm[8], a1 = prm[9], a2 = prm[10], a3 = prm[11], a4 = prm[12], a5 = prm[13];
    const uint8_t* src = in + (size_t)n * S * S * 3;
    const size_t plane = (size_t)S * S;
    for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < S; x += gridDim.x * blockDim.x) {
        int xin, yin;
        if (mode == 0) { xin = (int)((a2 + a0 * x + a1 * y) >> 16); yin = (int)((a5 + a3 * x + a4 * y) >> 16); }
        else if (mode == 1) { xin = x; yin = y; }
        else if (mode == 2) { xin = S - 1 - x; yin = S - 1 - y; }
        else if (mode == 3) { xin = S - 1 - y; yin = x; }
        else { xin = y; yin = S - 1 - x; }
        int v0 = 0, v1 = 0, v2 = 0;
        if (xin >= 0 && xin < S && yin >= 0 && yin < S) {
            const int xs = hflip ? S - 1 - xin : xin, ys = vflip ? S - 1 - yin : yin;
            const uint8_t* p = src + ((size_t)ys * S + xs) * 3;
            v0 = p[0]; v1 = p[1]; v2 = p[2];
        }
        if (out_u8 != nullptr) {
            uint8_t* o8 = out_u8 + (((size_t)n * S + y) * S + x) * 3;
            o8[0] = (uint8_t)v0; o8[1] = (uint8_t)v1; o8[2] = (uint8_t)v2;
        }
        if (out != nullptr) {
            float* o = out + ((size_t)n * 3 * S + y) * S + x;
            o[0] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v0, 255.0f), m0), s0);
            o[plane] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v1, 255.0f), m1), s1);
            o[2 * plane] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v2, 255.0f), m2), s2);
        }
    }
}

}  // namespace ssl4p

using namespace ssl4p;
#define STREAM ((cudaStream_t)stream)

extern "C" int ssl4p_color_jitter_u8(const void* img, void* out, const int* params, int N, int S, void* stream) {
    SSL4P_REQUIRE(img && out && params, "color_jitter: null argument");
    SSL4P_REQUIRE(N >= 0 && S > 0 && (S * S) % 4 == 0, "color_jitter: bad shape N=%d S=%d (S*S must be a multiple of 4)", N, S);
    const size_t smem = (size_t)S * S * 3;
    SSL4P_REQUIRE(smem <= 200 * 1024, "color_jitter: a %dx%d image does not fit shared memory (limit 258x258)", S, S);
    if (N == 0) return SSL4P_OK;
    SSL4P_CUDA(cudaFuncSetAttribute(color_jitter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    color_jitter_kernel<<<N, 1024, smem, STREAM>>>((const uint8_t*)img, (uint8_t*)out, params, S * S);
    return check_launch("color_jitter_kernel");
}

extern "C" size_t ssl4p_gaussian_blur_workspace_bytes(int N, int S) { return (size_t)N * S * S * 3 * sizeof(double); }

extern "C" int ssl4p_gaussian_blur_u8(const void* img, void* out, const float* taps, int ntaps, int N, int S,
                                      void* workspace, size_t workspace_bytes, void* stream) {
    SSL4P_REQUIRE(img && out && taps, "gaussian_blur: null argument");
    SSL4P_REQUIRE(N >= 0 && S > 0 && ntaps > 0 && (ntaps & 1) && ntaps <= kAugTaps, "gaussian_blur: bad shape N=%d S=%d taps=%d", N, S, ntaps);
    SSL4P_REQUIRE(ntaps / 2 < S, "gaussian_blur: the reflect pad (%d) needs an image larger than %d", ntaps / 2, S);
    SSL4P_REQUIRE(workspace && workspace_bytes >= ssl4p_gaussian_blur_workspace_bytes(N, S), "gaussian_blur: workspace too small");
    if (N == 0) return SSL4P_OK;
    const dim3 grid((S * 3 + 255) / 256, S, N);
    blur_rows_kernel<<<grid, 256, 0, STREAM>>>((const uint8_t*)img, taps, ntaps, S, (double*)workspace);
    int rc = check_launch("blur_rows_kernel");
    if (rc) return rc;
    blur_cols_kernel<<<grid, 256, 0, STREAM>>>((const double*)workspace, taps, ntaps, S, (uint8_t*)out);
    return check_launch("blur_cols_kernel");
}

extern "C" int ssl4p_flip_rotate_normalize_u8(const void* img, const int* params, int N, int S, const float* mean3, const float* std3,
                                              float* out, void* out_u8, void* stream) {
    SSL4P_REQUIRE(img && params && mean3 && std3 && (out || out_u8), "flip_rotate_normalize: null argument");
    SSL4P_REQUIRE(N >= 0 && S > 0, "flip_rotate_normalize: bad shape N=%d S=%d", N, S);
    if (N == 0) return SSL4P_OK;
    const dim3 grid((S + 255) / 256, S, N);
    flip_rotate_normalize_kernel<<<grid, 256, 0, STREAM>>>((const uint8_t*)img, params, S, mean3[0], mean3[1], mean3[2],
                                                           std3[0], std3[1], std3[2], out, (uint8_t*)out_u8);
    return check_launch("flip_rotate_normalize_kernel");
}
