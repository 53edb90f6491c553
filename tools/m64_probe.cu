// Probe: which TMEM lanes does tcgen05.mma kind::f16 cta_group::1 read (A in TMEM) and write (D)
// for M = 64 vs M = 128?  A[l][k] = (k == 0) ? l + 1 : 0 for every TMEM lane l, B = all ones, so
// D[m][n] = (A lane that fed row m) + 1; D is pre-filled with -7 to show the lanes left untouched.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -o m64_probe m64_probe.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../learnedmetricindex_b200/csrc/lmi_ptx.cuh"

using namespace lmi;

__global__ void __launch_bounds__(128) probe(float* out, int M, int N) {
    __shared__ alignas(1024) uint8_t bsm[16384];
    __shared__ alignas(8) uint64_t bar;
    __shared__ uint32_t tbase;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 16384 / 2; i += 128) reinterpret_cast<__half*>(bsm)[i] = __float2half(1.0f);
    if (threadIdx.x == 0) {
        ptx::mbar_init(&bar, 1);
        ptx::fence_barrier_init();
    }
    if (warp == 0) {
        ptx::tmem_alloc(&tbase, 512);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = tbase;
    const uint32_t lane_field = static_cast<uint32_t>(warp * 32) << 16;
    uint32_t a[32], dinit[32];
    for (int j = 0; j < 32; ++j) {
        a[j] = 0u;
        dinit[j] = __float_as_uint(-7.f);
    }
    const __half2 h = __halves2half2(__float2half((float)(threadIdx.x + 1)), __float2half(0.f));
    a[0] = *reinterpret_cast<const uint32_t*>(&h);
    ptx::tmem_st32(tmem + lane_field + 0, a);
    for (int c = 0; c < 4; ++c) ptx::tmem_st32(tmem + lane_field + 256 + 32 * c, dinit);
    ptx::tmem_st_wait();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    if (threadIdx.x == 0) {
        const uint64_t bdesc = ptx::umma_desc_sw128_kmajor(ptx::smem_u32(bsm));
        const uint32_t idesc = ptx::umma_idesc_f16(M, N);
        ptx::mma_f16_ts(tmem + 256, tmem + 0, bdesc, idesc, 0u);
        ptx::mma_commit(&bar);
    }
    ptx::mbar_wait(&bar, 0);
    ptx::tc_fence_after();
    for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        ptx::tmem_ld32(tmem + lane_field + 256 + 32 * c, v);
        ptx::tmem_ld_wait();
        for (int j = 0; j < 32; ++j) out[(size_t)threadIdx.x * 128 + 32 * c + j] = __uint_as_float(v[j]);
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, 512);
    }
}

int main() {
    float* d;
    cudaMalloc(&d, 128 * 128 * sizeof(float));
    static float h[128 * 128];
    const int cfg[][2] = {{128, 64}, {64, 64}, {64, 128}};
    for (auto& c : cfg) {
        cudaMemset(d, 0, sizeof(h));
        probe<<<1, 128>>>(d, c[0], c[1]);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
            printf("M=%d N=%d: %s\n", c[0], c[1], cudaGetErrorString(e));
            return 1;
        }
        cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        printf("M=%d N=%d  (per TMEM lane: D[lane][col 0], col N-1, col N, col 127; -7 = untouched)\n", c[0], c[1]);
        for (int l = 0; l < 128; ++l) {
            // is the row uniform over columns [0, N)?
            bool uni = true;
            for (int j = 1; j < c[1] && j < 128; ++j) uni &= (h[l * 128 + j] == h[l * 128]);
            printf("  lane %3d: %6.1f %6.1f | %6.1f %6.1f %s\n", l, h[l * 128], h[l * 128 + c[1] - 1],
                   c[1] < 128 ? h[l * 128 + c[1]] : -1.f, h[l * 128 + 127], uni ? "" : "(non-uniform)");
        }
    }
    return 0;
}
