// Power microbenchmark: SUSTAINED tcgen05.mma kind::f16 throughput under the board's power cap for the
// operand arrangements the scan kernel can use (M = 128; N in {64,128,256}; A from TMEM or shared memory;
// random fp16 operands -- zeros draw far less power), optionally with the B-operand traffic of the real
// kernel streamed L2 -> shared memory (or HBM -> shared memory for every third copy) at the MMA's pace,
// and with eight "epilogue" warps draining the accumulator (tcgen05.ld) once per 48 MMAs.
// Every configuration runs ~1.2 s twice (the first run settles the power state); the second is reported.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o mma_power mma_power.cu
#include <chrono>
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include "../learnedmetricindex_b200/csrc/lmi_ptx.cuh"
using namespace lmi;

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ uint32_t rnd_half2(uint32_t seed, int zero) {
    if (zero) return 0u;
    const uint32_t h = hash32(seed);
    const __half a = __float2half(((int)(h & 0xffff) - 32768) * (0.1f / 32768.f));
    const __half b = __float2half(((int)(h >> 16) - 32768) * (0.1f / 32768.f));
    return (uint32_t)__half_as_ushort(a) | ((uint32_t)__half_as_ushort(b) << 16);
}

// copy: 0 none, 1 L2-resident source, 2 every third copy streams from a large (HBM) source
// side: 0 none, 1 eight warps drain N accumulator columns once per 48 MMAs
template <int N, bool A_TMEM, int NACC>
__global__ void __launch_bounds__(320, 1) k_power(long long reps, int per_commit, long long* out, const uint8_t* src,
                                                 size_t src_bytes, int copy, int side, int zero) {
    extern __shared__ uint8_t raw[];
    uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar;
    __shared__ uint64_t cbar[4];
    __shared__ volatile int done;
    __shared__ volatile long long progress;  // MMAs committed so far
    __shared__ uint32_t tbase;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // B at 0 (N rows x 128 B <= 32 KB), A at 64 KB (128 rows x 128 B), copy ring at 96 KB (4 x 8 KB)
    for (int i = threadIdx.x; i < 128 * 1024 / 4; i += 320)
        reinterpret_cast<uint32_t*>(sm)[i] = rnd_half2(i * 2654435761u + blockIdx.x, zero);
    if (threadIdx.x == 0) {
        ptx::mbar_init(&bar, 1);
        for (int i = 0; i < 4; ++i) ptx::mbar_init(&cbar[i], 1);
        done = 0;
        progress = 0;
        ptx::fence_barrier_init();
    }
    if (warp == 0) { ptx::tmem_alloc(&tbase, 512); ptx::tmem_relinquish(); }
    ptx::fence_proxy_async();
    ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
    const uint32_t tmem = tbase;
    if (warp >= 2 && warp < 6) {  // A operand in TMEM: columns [0,256) of the warp's lane quarter
        const uint32_t lane_field = (uint32_t)((warp & 3) * 32) << 16;
        for (int c = 0; c < 256; c += 32) {
            uint32_t v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = rnd_half2((threadIdx.x * 256 + c + j) * 40503u + blockIdx.x, zero);
            ptx::tmem_st32(tmem + lane_field + c, v);
        }
        ptx::tmem_st_wait();
    }
    ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
    if (warp >= 2 && warp < 10 && side) {
        // "epilogue": once per 48 MMAs read this warp's share of an accumulator of N columns
        const uint32_t lane_field = (uint32_t)((warp & 3) * 32) << 16;
        uint32_t acc = 0;
        long long seen = 0;
        while (!done) {
            const long long p = progress;
            if (p - seen >= 48) {
                seen += 48;
                for (int c = ((warp - 2) >> 2) * (N / 2); c < ((warp - 2) >> 2) * (N / 2) + N / 2; c += 32) {
                    uint32_t v[32];
                    ptx::tmem_ld32(tmem + lane_field + 256 + c, v);
                    ptx::tmem_ld_wait();
                    float m = __uint_as_float(v[0]);
#pragma unroll
                    for (int j = 1; j < 32; ++j) m = fmaxf(m, __uint_as_float(v[j]));
                    acc += __float_as_uint(m);
                }
            }
        }
        if (acc == 0x12345678u) out[0] = 1;
    }
    if (warp == 1) {
        const uint32_t idesc = ptx::umma_idesc_f16(128, N);
        const uint64_t bdesc = ptx::umma_desc_sw128_kmajor(ptx::smem_u32(sm));
        const uint64_t adesc = ptx::umma_desc_sw128_kmajor(ptx::smem_u32(sm + 64 * 1024));
        uint32_t phase = 0;
        long long t0 = clock64();
        for (long long r = 0; r < reps; ++r) {
            if (ptx::elect_one()) {
                for (int i = 0; i < per_commit; i += 4) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t d = tmem + 256 + ((i / 4) % NACC) * N % 256;
                        // A: 8 columns per k-step, walking over the 256 columns so that every step reads fresh data
                        if (A_TMEM) ptx::mma_f16_ts(d, tmem + ((i + k) * 8) % 256, bdesc + k * 2, idesc, 1u);
                        else ptx::mma_f16_ss(d, adesc + k * 2, bdesc + k * 2, idesc, 1u);
                    }
                }
                ptx::mma_commit(&bar);
            }
            __syncwarp();
            ptx::mbar_wait(&bar, phase);
            phase ^= 1;
            if (lane == 0) progress = (r + 1) * per_commit;
        }
        long long t1 = clock64();
        if (lane == 0) { out[blockIdx.x] = t1 - t0; done = 1; }
    } else if (warp == 0 && copy) {
        // B-operand traffic at the MMA's pace: N x 32 bytes per MMA, in 8 KB bulk copies into a 4-slot ring
        if (lane == 0) {
            uint32_t ph[4] = {0, 0, 0, 0};
            long long n = 0;
            const size_t l2_span = 64 * 8192;
            const size_t my_hbm = ((size_t)blockIdx.x * (src_bytes / 148)) & ~(size_t)8191;
            const size_t hbm_span = (src_bytes / 148) & ~(size_t)8191;
            auto issue = [&](int i) {
                size_t off = ((blockIdx.x * 4 + n) % 64) * 8192;
                if (copy == 2 && n % 3 == 0) off = l2_span + my_hbm + ((size_t)(n / 3) * 8192) % (hbm_span - l2_span - 8192);
                ptx::mbar_arrive_expect_tx(&cbar[i], 8192);
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(ptx::smem_u32(sm + 96 * 1024 + i * 8192)), "l"(src + off), "r"(8192),
                             "r"(ptx::smem_u32(&cbar[i])) : "memory");
                ++n;
            };
            int inflight = 0, head = 0, tail = 0;
            while (!done) {
                if (inflight < 4 && n * 8192 <= progress * (long long)(N * 32)) {
                    issue(head);
                    head = (head + 1) & 3;
                    ++inflight;
                } else if (inflight > 0 && ptx::mbar_try_wait(&cbar[tail], ph[tail])) {
                    ph[tail] ^= 1;
                    tail = (tail + 1) & 3;
                    --inflight;
                }
            }
            while (inflight > 0) {
                ptx::mbar_wait(&cbar[tail], ph[tail]);
                ph[tail] ^= 1;
                tail = (tail + 1) & 3;
                --inflight;
            }
            out[148 + blockIdx.x] = n * 8192;
        }
    }
    ptx::tc_fence_before(); __syncthreads();
    if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}

static uint8_t* g_src = nullptr;
static size_t g_src_bytes = (size_t)16 << 30;
static double now_s() {
    return std::chrono::duration<double>(std::chrono::system_clock::now().time_since_epoch()).count();
}

template <int N, bool A_TMEM, int NACC>
void run(const char* name, int copy = 0, int side = 0, int zero = 0, double seconds = 1.2) {
    long long* d; cudaMalloc(&d, sizeof(long long) * 296); cudaMemset(d, 0, sizeof(long long) * 296);
    auto kern = k_power<N, A_TMEM, NACC>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 130 * 1024);
    const int per = 48;
    const long long reps = (long long)(seconds / (per * (N / 2.0) / 1.4e9));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms = 0; cudaError_t e = cudaSuccess; double t_a = 0, t_b = 0;
    for (int pass = 0; pass < 2; ++pass) {
        t_a = now_s();
        cudaEventRecord(e0);
        kern<<<148, 320, 130 * 1024>>>(reps, per, d, g_src, g_src_bytes, copy, side, zero);
        cudaEventRecord(e1);
        e = cudaDeviceSynchronize();
        t_b = now_s();
        cudaEventElapsedTime(&ms, e0, e1);
    }
    long long h[296]; cudaMemcpy(h, d, sizeof(long long) * 296, cudaMemcpyDeviceToHost);
    const double n = (double)reps * per;
    const double tf = 2.0 * 128 * N * 16 * n * 148 / (ms * 1e-3) / 1e12;
    printf("%-34s %7.1f TFLOP/s  %6.1f cyc/MMA  ~%4.0f MHz  copied %5.1f GB/s/SM-avg  [%.3f, %.3f] %s\n", name, tf,
           h[0] / n, h[0] / (ms * 1e3), copy ? (double)h[148] / (ms * 1e-3) / 1e9 : 0.0, t_a, t_b, cudaGetErrorString(e));
    fflush(stdout);
    cudaFree(d);
}

int main() {
    if (cudaMalloc(&g_src, g_src_bytes) != cudaSuccess) { g_src_bytes = (size_t)1 << 30; cudaMalloc(&g_src, g_src_bytes); }
    cudaMemset(g_src, 0x3c, g_src_bytes);
    run<64, true, 2>("N=64  A=tmem 2acc zeros", 0, 0, 1);
    run<64, true, 2>("N=64  A=tmem 2acc");
    run<128, true, 1>("N=128 A=tmem 1acc");
    run<128, true, 2>("N=128 A=tmem 2acc");
    run<256, true, 1>("N=256 A=tmem 1acc");
    run<64, false, 2>("N=64  A=smem 2acc");
    run<128, false, 2>("N=128 A=smem 2acc");
    run<256, false, 1>("N=256 A=smem 1acc");
    run<64, true, 2>("N=64  A=tmem +L2copy", 1);
    run<128, true, 1>("N=128 A=tmem +L2copy", 1);
    run<64, true, 2>("N=64  A=tmem +L2/HBM copy", 2);
    run<128, true, 1>("N=128 A=tmem +L2/HBM copy", 2);
    run<64, true, 2>("N=64  A=tmem +L2/HBM copy +LDTM", 2, 1);
    run<128, true, 1>("N=128 A=tmem +L2/HBM copy +LDTM", 2, 1);
    run<64, true, 2>("N=64  A=tmem +LDTM", 0, 1);
    run<128, true, 1>("N=128 A=tmem +LDTM", 0, 1);
    return 0;
}
