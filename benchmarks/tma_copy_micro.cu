// Micro-benchmark behind the design of csrc/kernels/shuffle_tma.cu: how fast can ONE warp per SM move bytes with
// cp.async.bulk (global -> shared ring -> global), and which leg limits it?
//   mode 0: load + store (the copy), mode 1: loads only, mode 2: stores only (ring content is whatever is there)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/bin/tma_copy_micro benchmarks/tma_copy_micro.cu
// Run:   build/bin/tma_copy_micro            (prints one line per configuration: GB/s of payload moved)
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <unistd.h>
#include <cuda_runtime.h>

#include "../csrc/kernels/common.cuh"

using namespace uda;

template <int CHUNK, int STAGES, int AHEAD, int WARPS>
__global__ void __launch_bounds__(32 * WARPS, 1) copy_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, long long nchunks,
                                                             int mode) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t mbar[WARPS][STAGES];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint8_t* ring = smem + (size_t)w * STAGES * CHUNK;
    if (lane == 0) {
        for (int s = 0; s < STAGES; ++s) mbar_init(&mbar[w][s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (lane != 0) return;
    const long long stride = (long long)gridDim.x * WARPS, first = (long long)blockIdx.x * WARPS + w;
    const long long n_my = first < nchunks ? (nchunks - first + stride - 1) / stride : 0;
    for (long long it = 0; it < n_my + AHEAD; ++it) {
        if (it < n_my) {
            const int s = (int)(it % STAGES);
            if (it >= STAGES && mode != 1) tma_store_wait_read<STAGES - 1 - AHEAD>();
            const long long c = first + it * stride;
            if (mode != 2) {
                mbar_expect_tx(&mbar[w][s], CHUNK);
                tma_load_1d(ring + (size_t)s * CHUNK, src + c * CHUNK, CHUNK, &mbar[w][s]);
            }
        }
        const long long jt = it - AHEAD;
        if (jt >= 0 && jt < n_my) {
            const int s2 = (int)(jt % STAGES);
            if (mode != 2) mbar_wait(&mbar[w][s2], (uint32_t)(jt / STAGES) & 1u);
            if (mode != 1) {
                const long long c = first + jt * stride;
                tma_store_1d(dst + c * CHUNK, ring + (size_t)s2 * CHUNK, CHUNK);
                tma_store_commit();
            }
        }
    }
    if (mode != 1) tma_store_wait<0>();
}

// LSU reference: plain 16-byte loads and stores, `threads` per CTA, 4 in flight per thread
__global__ void lsu_copy_kernel(const int4* __restrict__ src, int4* __restrict__ dst, long long n16) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n16; i += 4 * stride) {
        int4 v0 = ld_stream(src + i), v1 = ld_stream(src + i + stride), v2 = ld_stream(src + i + 2 * stride), v3 = ld_stream(src + i + 3 * stride);
        st_stream(dst + i, v0); st_stream(dst + i + stride, v1); st_stream(dst + i + 2 * stride, v2); st_stream(dst + i + 3 * stride, v3);
    }
    for (; i < n16; i += stride) st_stream(dst + i, ld_stream(src + i));
}

void run_lsu(const uint8_t* src, uint8_t* dst, long long bytes, int ctas, int threads) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    for (int i = 0; i < 2; ++i) lsu_copy_kernel<<<ctas, threads>>>((const int4*)src, (int4*)dst, bytes / 16);
    cudaEventRecord(a);
    for (int i = 0; i < 3; ++i) lsu_copy_kernel<<<ctas, threads>>>((const int4*)src, (int4*)dst, bytes / 16);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    printf("LSU copy ctas=%3d threads=%4d : %8.1f GB/s  (%.1f GB/s per CTA)\n", ctas, threads, bytes * 3.0 / ms / 1e6, bytes * 3.0 / ms / 1e6 / ctas);
}

template <int CHUNK, int STAGES, int AHEAD, int WARPS>
void run(const uint8_t* src, uint8_t* dst, long long bytes, int ctas, int mode) {
    auto k = copy_kernel<CHUNK, STAGES, AHEAD, WARPS>;
    const int smem = CHUNK * STAGES * WARPS;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const long long nchunks = bytes / CHUNK;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    for (int i = 0; i < 2; ++i) k<<<ctas, 32 * WARPS, smem>>>(src, dst, nchunks, mode);
    cudaEventRecord(a);
    for (int i = 0; i < 3; ++i) k<<<ctas, 32 * WARPS, smem>>>(src, dst, nchunks, mode);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    cudaError_t e = cudaGetLastError();
    printf("chunk=%6d stages=%2d ahead=%d warps=%d ctas=%3d mode=%d : %8.1f GB/s  (%.1f GB/s per CTA)%s\n", CHUNK, STAGES, AHEAD, WARPS,
           ctas, mode, nchunks * (double)CHUNK * 3 / ms / 1e6, nchunks * (double)CHUNK * 3 / ms / 1e6 / ctas,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
}

// Bidirectional mode ("b <start epoch>"): two processes, started with swapped CUDA_VISIBLE_DEVICES, write to each other at the
// same time; every case starts at an agreed wall-clock second so that the two directions overlap.
static void wait_until(double t) {
    timespec ts;
    for (;;) {
        clock_gettime(CLOCK_REALTIME, &ts);
        if ((double)ts.tv_sec + 1e-9 * ts.tv_nsec >= t) return;
        usleep(200);
    }
}
template <int CHUNK, int STAGES, int AHEAD, int WARPS>
void run_at(double t, const uint8_t* src, uint8_t* dst, long long bytes, int ctas) {
    auto k = copy_kernel<CHUNK, STAGES, AHEAD, WARPS>;
    const int smem = CHUNK * STAGES * WARPS;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const long long nchunks = bytes / CHUNK;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    k<<<ctas, 32 * WARPS, smem>>>(src, dst, nchunks, 0);
    cudaDeviceSynchronize();
    wait_until(t);
    cudaEventRecord(a);
    for (int i = 0; i < 20; ++i) k<<<ctas, 32 * WARPS, smem>>>(src, dst, nchunks, 0);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    printf("bidirectional tma chunk=%6d stages=%2d ahead=%d warps=%d ctas=%3d : %8.1f GB/s per direction (%.1f GB/s per CTA)\n", CHUNK, STAGES,
           AHEAD, WARPS, ctas, nchunks * (double)CHUNK * 20 / ms / 1e6, nchunks * (double)CHUNK * 20 / ms / 1e6 / ctas);
}
void run_lsu_at(double t, const uint8_t* src, uint8_t* dst, long long bytes, int ctas) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    lsu_copy_kernel<<<ctas, 1024>>>((const int4*)src, (int4*)dst, bytes / 16);
    cudaDeviceSynchronize();
    wait_until(t);
    cudaEventRecord(a);
    for (int i = 0; i < 20; ++i) lsu_copy_kernel<<<ctas, 1024>>>((const int4*)src, (int4*)dst, bytes / 16);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    printf("bidirectional lsu ctas=%3d : %8.1f GB/s per direction\n", ctas, bytes * 20.0 / ms / 1e6);
}

int main(int argc, char** argv) {
    const long long bytes = 2LL << 30;
    uint8_t *src, *dst;
    cudaMalloc(&src, bytes);
    if (argc > 2 && argv[1][0] == 'b') {
        double t = atof(argv[2]);
        int can = 0;
        cudaDeviceCanAccessPeer(&can, 0, 1);
        if (!can) { printf("no peer access 0->1\n"); return 1; }
        cudaDeviceEnablePeerAccess(1, 0);
        cudaSetDevice(1);
        cudaMalloc(&dst, bytes);
        cudaMemset(dst, 2, bytes);
        cudaDeviceSynchronize();
        cudaSetDevice(0);
        cudaMemset(src, 1, bytes);
        cudaDeviceSynchronize();
        run_at<16384, 3, 2, 4>(t, src, dst, bytes, 16); t += 1.0;
        run_at<16384, 3, 2, 4>(t, src, dst, bytes, 32); t += 1.0;
        run_at<16384, 3, 2, 4>(t, src, dst, bytes, 64); t += 1.0;
        run_at<32768, 6, 4, 1>(t, src, dst, bytes, 32); t += 1.0;
        run_at<8192, 6, 4, 4>(t, src, dst, bytes, 32); t += 1.0;
        run_lsu_at(t, src, dst, bytes, 148); t += 1.0;
        run_lsu_at(t, src, dst, bytes, 296);
        return 0;
    }
    const bool peer = argc > 1 && argv[1][0] == 'p';
    if (peer) {   // destination on GPU 1, written from GPU 0 over NVLink
        int can = 0;
        cudaDeviceCanAccessPeer(&can, 0, 1);
        if (!can) { printf("no peer access 0->1\n"); return 1; }
        cudaDeviceEnablePeerAccess(1, 0);
        cudaSetDevice(1);
        cudaMalloc(&dst, bytes);
        cudaMemset(dst, 2, bytes);
        cudaDeviceSynchronize();
        cudaSetDevice(0);
        cudaMemset(src, 1, bytes);
        printf("== destination on the peer GPU ==\n");
        for (int ctas : {16, 32, 48, 64}) {
            run<32768, 6, 4, 1>(src, dst, bytes, ctas, 0);
            run<16384, 3, 2, 4>(src, dst, bytes, ctas, 0);
            run<8192, 3, 2, 8>(src, dst, bytes, ctas, 0);
            run<4096, 6, 4, 8>(src, dst, bytes, ctas, 0);
            run<4096, 3, 2, 16>(src, dst, bytes, ctas, 0);
            run_lsu(src, dst, bytes, ctas, 1024);
        }
        run<16384, 3, 2, 4>(src, dst, bytes, 32, 2);
        run<16384, 3, 2, 4>(src, dst, bytes, 148, 0);
        run_lsu(src, dst, bytes, 148, 1024);
        run_lsu(src, dst, bytes, 296, 1024);
        return 0;
    }
    cudaMalloc(&dst, bytes);
    cudaMemset(src, 1, bytes);
    cudaMemset(dst, 2, bytes);
    for (int mode = 0; mode < 3; ++mode) {
        run<32768, 6, 4, 1>(src, dst, bytes, 32, mode);
        run<16384, 12, 8, 1>(src, dst, bytes, 32, mode);
        run<8192, 24, 16, 1>(src, dst, bytes, 32, mode);
        run<4096, 48, 32, 1>(src, dst, bytes, 32, mode);
        run<16384, 3, 2, 4>(src, dst, bytes, 32, mode);
        run<8192, 6, 4, 4>(src, dst, bytes, 32, mode);
        run<8192, 3, 2, 8>(src, dst, bytes, 32, mode);
        run<4096, 6, 4, 8>(src, dst, bytes, 32, mode);
    }
    run<8192, 6, 4, 4>(src, dst, bytes, 16, 0);
    run<8192, 6, 4, 4>(src, dst, bytes, 64, 0);
    run<8192, 6, 4, 4>(src, dst, bytes, 148, 0);
    run<32768, 6, 4, 1>(src, dst, bytes, 148, 0);
    // correctness of mode 0
    uint8_t h[64];
    cudaMemcpy(h, dst + 12345 * 4096, 64, cudaMemcpyDeviceToHost);
    printf("dst sample byte = %d (expect 1)\n", (int)h[7]);
    return 0;
}
