// Microbenchmark: tcgen05.ld throughput per SM (how many bytes/clk can the epilogue/softmax warps pull out of TMEM?)
#include "../../ssl4polyp_b200/csrc/common.cuh"
#include <cstdio>
using namespace ssl4p;

__device__ __forceinline__ void ld_x64(uint32_t taddr, uint32_t (&r)[64]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
        "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]),
          "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]),
          "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
        : "r"(taddr) : "memory");
}

// mode 0: x32 load + wait per iteration;  mode 1: two x32 loads in flight + wait;  mode 2: x64 + wait
template <int MODE>
__global__ void k(int iters, int nwarps_active, long long* out, float* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tmem_alloc<512>(&slot);
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 256;
    float acc = 0.f;
    __syncthreads();
    long long t0 = clock64();
    if (warp < nwarps_active) {
        for (int i = 0; i < iters; ++i) {
            if (MODE == 0) {
                uint32_t r[32]; tmem_ld_32x32(base + (i & 3) * 32, r); tmem_ld_wait();
                acc += __uint_as_float(r[0] ^ r[9] ^ r[18] ^ r[31]);
            } else if (MODE == 1) {
                uint32_t r[32], q[32]; tmem_ld_32x32(base + (i & 3) * 32, r); tmem_ld_32x32(base + 128 + (i & 3) * 32, q); tmem_ld_wait();
                acc += __uint_as_float(r[0] ^ r[9] ^ q[18] ^ q[31]);
            } else {
                uint32_t r[64]; ld_x64(base + (i & 1) * 64, r); tmem_ld_wait();
                acc += __uint_as_float(r[0] ^ r[9] ^ r[40] ^ r[63]);
            }
        }
    }
    __syncthreads();
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    if (acc == 12345.678f) sink[0] = acc;
    tc_fence_before(); __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(slot); }
}

int main() {
    long long* d; float* s; cudaMalloc(&d, 8 * 148); cudaMalloc(&s, 4);
    const int iters = 2000;
    for (int mode = 0; mode < 3; ++mode)
        for (int nw = 4; nw <= 8; nw += 4) {
            if (mode == 0) k<0><<<148, 256>>>(iters, nw, d, s);
            if (mode == 1) k<1><<<148, 256>>>(iters, nw, d, s);
            if (mode == 2) k<2><<<148, 256>>>(iters, nw, d, s);
            cudaError_t e = cudaDeviceSynchronize();
            long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
            const double cols = (mode == 0 ? 32 : 64);
            const double bytes = (double)iters * nw * 32 * cols * 4;
            printf("mode %d warps %d: %lld clk, %.1f B/clk/SM, %.1f clk per iteration (%s)\n", mode, nw, h, bytes / h, (double)h / iters, cudaGetErrorString(e));
        }
    return 0;
}
