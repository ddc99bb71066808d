// Microbenchmark: latency of the synchronisation hops the attention kernels are built from.
//  A: one thread: tcgen05.mma (128x64x16, SS) + commit -> mbarrier wait                      (MMA round trip)
//  B: same with 13 MMAs per commit
//  C: 256 threads arrive on an mbarrier -> one other thread waits, arrives back -> 256 wait   (ping-pong, no MMA)
//  D: ping-pong with an MMA round trip in the middle (the softmax <-> MMA-issuer handshake)
//  E: fence.proxy.async + st.shared before the arrive (what the P hand-off adds)
#include "../../ssl4polyp_b200/csrc/common.cuh"
#include <cstdio>
using namespace ssl4p;

template <int MODE>
__global__ void __launch_bounds__(288, 1) k(int iters, long long* out) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 65536);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bars + 8);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (warp == 8) {
        if (lane == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 256); mbar_init(&bars[2], 1); mbar_fence_init(); fence_proxy_async(); }
        __syncwarp();
        tmem_alloc<512>(slot);
    }
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tm = *slot;
    const uint32_t idesc = umma_idesc_bf16(128, 64, false, true);
    const uint32_t sa = smem_u32(smem), sb = smem_u32(smem + 32768);
    long long t0 = clock64();
    if (MODE <= 1) {
        if (warp == 8 && lane == 0) {
            const int n = MODE == 0 ? 1 : 13;
            for (int i = 0; i < iters; ++i) {
                for (int j = 0; j < n; ++j) umma_bf16(tm, umma_smem_desc(sa + (j & 3) * 32, 16, 1024), umma_smem_desc(sb + (j & 7) * 2048, 16384, 1024), idesc, j > 0);
                umma_commit(&bars[0]);
                mbar_wait(&bars[0], i & 1);
                tc_fence_after();
            }
        }
    } else {
        if (warp < 8) {
            for (int i = 0; i < iters; ++i) {
                if (MODE == 4) { reinterpret_cast<volatile uint32_t*>(smem)[threadIdx.x * 4] = i; fence_proxy_async(); }
                tc_fence_before();
                mbar_arrive(&bars[1]);
                mbar_wait(&bars[2], i & 1);
                tc_fence_after();
            }
        } else if (lane == 0) {
            for (int i = 0; i < iters; ++i) {
                mbar_wait(&bars[1], i & 1);
                tc_fence_after();
                if (MODE >= 3) {
                    umma_bf16(tm, umma_smem_desc(sa, 16, 1024), umma_smem_desc(sb, 16384, 1024), idesc, 0);
                    umma_commit(&bars[2]);
                } else {
                    mbar_arrive(&bars[2]);
                }
            }
        }
    }
    __syncthreads();
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    tc_fence_before(); __syncthreads();
    if (warp == 8) { tc_fence_after(); tmem_dealloc<512>(tm); }
}

int main() {
    long long* d; cudaMalloc(&d, 8 * 148);
    const int iters = 2000, smem = 65536 + 1024 + 256;
    const char* names[5] = {"A  1 MMA + commit -> wait (same thread)", "B 13 MMA + commit -> wait (same thread)", "C 256-thread arrive <-> 1-thread arrive ping-pong",
                            "D ping-pong through 1 MMA + commit", "E = D + st.shared + fence.proxy.async"};
    for (int m = 0; m < 5; ++m) {
        auto run = [&](auto kern) { cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); kern<<<148, 288, smem>>>(iters, d); };
        if (m == 0) run(k<0>); if (m == 1) run(k<1>); if (m == 2) run(k<2>); if (m == 3) run(k<3>); if (m == 4) run(k<4>);
        cudaError_t e = cudaDeviceSynchronize();
        long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        printf("%-52s %8.1f clk / iteration (%s)\n", names[m], (double)h / iters, cudaGetErrorString(e));
    }
    return 0;
}
