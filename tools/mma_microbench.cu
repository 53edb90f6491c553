// Microbenchmark: tcgen05.mma kind::f16 issue/execute cost per instruction on sm_100a for the
// shapes the scan kernel can use (M=128; N in {64,128,256}; A from TMEM or shared memory).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o mma_microbench mma_microbench.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../learnedmetricindex_b200/csrc/lmi_ptx.cuh"
using namespace lmi;

template <int N, bool A_TMEM, int NACC>
__global__ void __launch_bounds__(320, 1) k_mma(int reps, int per_commit, long long* out, const uint8_t* src, int copy,
                                               volatile int* stop_flag_unused) {
    extern __shared__ uint8_t raw[];
    uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar;
    __shared__ uint64_t cbar[4];
    __shared__ volatile int done;
    __shared__ long long copied;
    __shared__ uint32_t tbase;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += 320) reinterpret_cast<uint32_t*>(sm)[i] = 0;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); for (int i = 0; i < 4; ++i) ptx::mbar_init(&cbar[i], 1); done = 0; copied = 0; ptx::fence_barrier_init(); }
    if (warp == 0) { ptx::tmem_alloc(&tbase, 512); ptx::tmem_relinquish(); }
    ptx::fence_proxy_async();
    ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
    const uint32_t tmem = tbase;
    __shared__ uint64_t never;
    if (threadIdx.x == 0) ptx::mbar_init(&never, 1);
    __syncthreads();
    const int side = copy >> 4;  // 1: other 8 warps spin on an mbarrier; 2: they stream tcgen05.ld
    if (warp >= 2 && warp < 10 && side) {
        const uint32_t lane_field = (uint32_t)((warp & 3) * 32) << 16;
        uint32_t acc = 0;
        while (!done) {
            if (side == 1) { acc += ptx::mbar_try_wait(&never, 0); }
            else { uint32_t v[32]; ptx::tmem_ld32(tmem + lane_field + 256 + (warp >> 2) * 32, v); ptx::tmem_ld_wait(); acc += v[lane]; }
        }
        if (acc == 0x12345678u) out[0] = 1;
    }
    copy &= 15;
    if (warp == 1) {
        const uint32_t idesc = ptx::umma_idesc_f16(128, N);
        const uint64_t bdesc = ptx::umma_desc_sw128_kmajor(ptx::smem_u32(sm));            // B: N rows x 64 fp16
        const uint64_t adesc = ptx::umma_desc_sw128_kmajor(ptx::smem_u32(sm + 64 * 1024)); // A: 128 rows x 64 fp16
        uint32_t phase = 0;
        long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            if (ptx::elect_one()) {
                for (int i = 0; i < per_commit; i += 4) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t d = tmem + 256 + ((i / 4) % NACC) * N % 256;
                        if (A_TMEM) ptx::mma_f16_ts(d, tmem + k * 8, bdesc + k * 2, idesc, 1u);
                        else ptx::mma_f16_ss(d, adesc + k * 2, bdesc + k * 2, idesc, 1u);
                    }
                }
                ptx::mma_commit(&bar);
            }
            __syncwarp();
            ptx::mbar_wait(&bar, phase);
            phase ^= 1;
        }
        long long t1 = clock64();
        if (lane == 0) { out[blockIdx.x] = t1 - t0; done = 1; }
    } else if (warp == 0 && copy) {
        // streams 8 KB bulk copies (L2-resident source) into a 4 x 8 KB ring at 32 KB.. of shared memory
        if (lane == 0) {
            uint32_t ph[4] = {0, 0, 0, 0};
            long long n = 0;
            for (int i = 0; i < 4; ++i) {
                ptx::mbar_arrive_expect_tx(&cbar[i], 8192);
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(ptx::smem_u32(sm + 24 * 1024 + i * 8192)), "l"(src + ((blockIdx.x * 4 + i) % 64) * 8192), "r"(8192), "r"(ptx::smem_u32(&cbar[i])) : "memory");
            }
            while (!done) {
                for (int i = 0; i < 4; ++i) {
                    ptx::mbar_wait(&cbar[i], ph[i]); ph[i] ^= 1; ++n;
                    ptx::mbar_arrive_expect_tx(&cbar[i], 8192);
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(ptx::smem_u32(sm + 24 * 1024 + i * 8192)), "l"(src + ((blockIdx.x * 4 + i + n) % 64) * 8192), "r"(8192), "r"(ptx::smem_u32(&cbar[i])) : "memory");
                }
            }
            for (int i = 0; i < 4; ++i) ptx::mbar_wait(&cbar[i], ph[i]);
            out[148 + blockIdx.x] = n * 8192;
        }
    }
    ptx::tc_fence_before(); __syncthreads();
    if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}

template <int N, bool A_TMEM, int NACC>
void run(const char* name, int grid, int copy = 0) {
    long long* d; cudaMalloc(&d, sizeof(long long) * 296); cudaMemset(d, 0, sizeof(long long) * 296);
    static uint8_t* src = nullptr; if (!src) { cudaMalloc(&src, 64 * 8192); cudaMemset(src, 0, 64 * 8192); }
    auto kern = k_mma<N, A_TMEM, NACC>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    const int reps = 200, per = 64;
    kern<<<grid, 320, 100 * 1024>>>(10, per, d, src, copy, nullptr);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<<<grid, 320, 100 * 1024>>>(reps, per, d, src, copy, nullptr);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[296]; cudaMemcpy(h, d, sizeof(long long) * 296, cudaMemcpyDeviceToHost);
    const double n = (double)reps * per;
    printf("%-28s grid %3d: %6.1f cycles/MMA (SM0)  %7.2f ms  %7.1f TFLOP/s  %s\n", name, grid, h[0] / n, ms,
           2.0 * 128 * N * 16 * n * grid / (ms * 1e-3) / 1e12, cudaGetErrorString(e));
    if (copy) printf("    concurrent bulk copy into smem: %.1f B/cycle on SM0\n", (double)h[148] / (double)h[0]);
    cudaFree(d);
}

int main() {
    run<64, true, 2>("N=64 tmem, 8 warps spin", 148, 16);
    run<64, true, 2>("N=64 tmem, 8 warps LDTM", 148, 32);
    run<64, true, 2>("N=64 tmem, spin + copy", 148, 17);
    run<64, true, 2>("N=64 tmem, LDTM + copy", 148, 33);
    run<64, true, 2>("N=64 A=tmem +copy", 1, 1);
    run<64, true, 2>("N=64 A=tmem +copy", 148, 1);
    run<128, true, 2>("N=128 A=tmem +copy", 148, 1);
    run<256, true, 1>("N=256 A=tmem +copy", 148, 1);
    for (int grid : {148}) {
        run<64, true, 1>("N=64  A=tmem 1 acc", grid);
        run<64, true, 2>("N=64  A=tmem 2 acc", grid);
        run<128, true, 1>("N=128 A=tmem 1 acc", grid);
        run<128, true, 2>("N=128 A=tmem 2 acc", grid);
        run<256, true, 1>("N=256 A=tmem 1 acc", grid);
        run<64, false, 2>("N=64  A=smem 2 acc", grid);
        run<128, false, 2>("N=128 A=smem 2 acc", grid);
        run<256, false, 1>("N=256 A=smem 1 acc", grid);
    }
    return 0;
}
