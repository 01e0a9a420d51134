// K1 — the map-output all-to-all shuffle as a push over NVLink 5 / NVSwitch peer memory.
//
// Reference behaviour being replaced (pull model, three hops per 1 MiB chunk):
//   src/DataNet/RDMAClient.cc:559-600  reducer SENDs a text fetch request
//   src/MOFServer/IndexInfo.cc:237-335 provider looks up the index, O_DIRECT-reads the chunk
//   src/DataNet/RDMAServer.cc:537-631  provider RDMA-WRITEs the chunk, then SENDs an ack
//   src/DataNet/RDMAServer.cc:44-136   credit flow control + backlog
// Here the provider side *pushes*: one persistent kernel walks the device-resident MOF index, and
// for every (map, reducer) segment stores the bytes straight into the reducer rank's receive arena
// (a peer-mapped pointer; plain st.global over NVLink), then publishes "bytes landed" with a
// system-scope release reduction on a counter that lives in the destination rank's memory. The
// reducer side acquires those counters before it touches a segment. No request, no ack message, no
// credits: the arena is sized for the whole partition, so a push never blocks, which also makes the
// protocol deadlock free when every rank is producer and consumer at once.
#include "common.cuh"
#include "../../include/uda_b200.h"

namespace uda {

constexpr int PUSH_THREADS = 256;
constexpr int PUSH_CHUNK = 64 * 1024;   // bytes copied by one CTA between two flag publications
constexpr int PUSH_UNROLL = 4;

// chunk_prefix[j] = number of chunks in segments < j ; chunk_prefix[nseg] = total
__global__ void __launch_bounds__(1024) push_plan_kernel(const UdaPushDesc* __restrict__ d, int nseg,
                                                         unsigned* __restrict__ chunk_prefix) {
    __shared__ unsigned ws[32];
    __shared__ unsigned carry;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nseg; base += 1024) {
        int j = base + tid;
        unsigned c = j < nseg ? (unsigned)((d[j].bytes + PUSH_CHUNK - 1) / PUSH_CHUNK) : 0u, x = c;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) ws[w] = x;
        __syncthreads();
        if (w == 0) {
            unsigned t = ws[lane], z = t;
            for (int o = 1; o < 32; o <<= 1) {
                unsigned y = __shfl_up_sync(0xffffffffu, z, o);
                if (lane >= o) z += y;
            }
            ws[lane] = z - t;
        }
        __syncthreads();
        unsigned excl = carry + ws[w] + x - c;
        if (j < nseg) chunk_prefix[j] = excl;
        __syncthreads();
        if (tid == 1023) carry = excl + c;
        __syncthreads();
    }
    if (tid == 0) chunk_prefix[nseg] = carry;
}

__device__ __forceinline__ void copy_chunk(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, long long bytes) {
    const int tid = threadIdx.x;
    if ((((uintptr_t)src ^ (uintptr_t)dst) & 15) != 0) {  // incompatible phases: byte path (never hit by our layouts)
        for (long long i = tid; i < bytes; i += PUSH_THREADS) dst[i] = src[i];
        return;
    }
    long long head = (16 - ((uintptr_t)src & 15)) & 15;
    if (head > bytes) head = bytes;
    if (tid < head) dst[tid] = src[tid];
    const long long body = (bytes - head) >> 4;
    const int4* s = reinterpret_cast<const int4*>(src + head);
    int4* d = reinterpret_cast<int4*>(dst + head);
    long long i = tid;
    for (; i + (long long)(PUSH_UNROLL - 1) * PUSH_THREADS < body; i += (long long)PUSH_UNROLL * PUSH_THREADS) {
        int4 v[PUSH_UNROLL];
#pragma unroll
        for (int u = 0; u < PUSH_UNROLL; ++u) v[u] = ld_stream(s + i + (long long)u * PUSH_THREADS);
#pragma unroll
        for (int u = 0; u < PUSH_UNROLL; ++u) st_stream(d + i + (long long)u * PUSH_THREADS, v[u]);
    }
    for (; i < body; i += PUSH_THREADS) st_stream(d + i, ld_stream(s + i));
    const long long tail0 = head + (body << 4);
    if (tail0 + tid < bytes) dst[tail0 + tid] = src[tail0 + tid];
}

// nway > 1 interleaves the chunk order nway-fold: with peer-major descriptors of similar size per peer,
// consecutive CTAs then work on different destinations at the same time, so the HBM-bound local copy and the
// NVLink-bound remote pushes overlap instead of running one peer after the other.
__global__ void __launch_bounds__(PUSH_THREADS) push_kernel(const UdaPushDesc* __restrict__ d, int nseg,
                                                            const unsigned* __restrict__ chunk_prefix, unsigned nway) {
    const unsigned total = chunk_prefix[nseg];
    const unsigned per = (total + nway - 1) / nway;
    for (unsigned ci = blockIdx.x; ci < per * nway; ci += gridDim.x) {
        const unsigned c = (ci % nway) * per + ci / nway;
        if (c >= total) continue;
        int lo = 0, hi = nseg - 1;  // last segment with chunk_prefix[j] <= c
        while (lo < hi) {
            int mid = (lo + hi + 1) >> 1;
            if (chunk_prefix[mid] <= c) lo = mid; else hi = mid - 1;
        }
        const UdaPushDesc sd = d[lo];
        const long long off = (long long)(c - chunk_prefix[lo]) * PUSH_CHUNK;
        const long long len = sd.bytes - off < PUSH_CHUNK ? sd.bytes - off : PUSH_CHUNK;
        copy_chunk((const uint8_t*)sd.src + off, (uint8_t*)sd.dst + off, len);
        __syncthreads();  // all of this CTA's stores are issued and ordered before thread 0's release
        if (threadIdx.x == 0 && sd.flag) red_add_release_sys(sd.flag, (unsigned long long)len);
    }
}

// Reducer side: block until every expected segment has fully landed. flags[j] >= expect[j].
// status[0] is set to 1 on timeout so the host can raise the failure up-call instead of hanging.
__global__ void wait_flags_kernel(const unsigned long long* __restrict__ flags, const unsigned long long* __restrict__ expect,
                                  int n, unsigned long long max_spins, unsigned* __restrict__ status) {
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        unsigned long long spins = 0;
        while (ld_acquire_sys(flags + j) < expect[j]) {
            if (++spins > max_spins) { atomicExch(status, 1u); break; }
            __nanosleep(200);
        }
    }
}

// All-rank barrier over peer memory: slot[r] of every rank's barrier pad is written by rank r.
__global__ void barrier_kernel(unsigned long long* const* __restrict__ pads /* [world] peer pointers */, int rank,
                               int world, unsigned long long epoch, unsigned long long max_spins,
                               unsigned* __restrict__ status) {
    const int p = threadIdx.x;
    if (p < world) {
        __threadfence_system();
        st_release_sys(pads[p] + rank, epoch);
        unsigned long long spins = 0;
        while (ld_acquire_sys(pads[rank] + p) < epoch) {
            if (++spins > max_spins) { atomicExch(status, 2u); break; }
            __nanosleep(200);
        }
    }
}

// Same barrier, epoch taken from (and advanced in) a counter in this rank's own memory: the launch has no
// host-side state, so it can sit inside a captured CUDA graph that is replayed once per step. Every rank runs the
// same sequence of barriers, so the counters stay in step.
__global__ void barrier_counted_kernel(unsigned long long* const* __restrict__ pads, int rank, int world,
                                       unsigned long long* __restrict__ counter, unsigned long long max_spins,
                                       unsigned* __restrict__ status) {
    __shared__ unsigned long long s_epoch;
    if (threadIdx.x == 0) { s_epoch = *counter + 1ull; *counter = s_epoch; }
    __syncthreads();
    const unsigned long long epoch = s_epoch;
    const int p = threadIdx.x;
    if (p < world) {
        __threadfence_system();
        st_release_sys(pads[p] + rank, epoch);
        unsigned long long spins = 0;
        while (ld_acquire_sys(pads[rank] + p) < epoch) {
            if (++spins > max_spins) { atomicExch(status, 2u); break; }
            __nanosleep(200);
        }
    }
}

}  // namespace uda

extern "C" long long uda_push_scratch_bytes(int nseg) { return ((long long)nseg + 1) * (long long)sizeof(unsigned); }

// `nway` interleaves the chunk order over the peer-major descriptor table (number of destination ranks: the local
// copy and the remote pushes then run concurrently); 1 = chunks in descriptor order. It is a call parameter: the
// library keeps no per-process push state, so several jobs / devices in one process cannot disturb each other.
extern "C" int uda_shuffle_push_nway(const UdaPushDesc* descs_dev, int nseg, void* scratch, int max_ctas, int nway,
                                     cudaStream_t stream) {
    using namespace uda;
    if (nway < 1 || nway > 4096) return UDA_E_INVALID;
    if (nseg <= 0) return 0;
    unsigned* prefix = (unsigned*)scratch;
    push_plan_kernel<<<1, 1024, 0, stream>>>(descs_dev, nseg, prefix);
    int grid = max_ctas > 0 ? max_ctas : kNumSMs * 8;
    push_kernel<<<grid, PUSH_THREADS, 0, stream>>>(descs_dev, nseg, prefix, (unsigned)nway);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_shuffle_push(const UdaPushDesc* descs_dev, int nseg, void* scratch, int max_ctas, cudaStream_t stream) {
    return uda_shuffle_push_nway(descs_dev, nseg, scratch, max_ctas, 1, stream);
}

extern "C" int uda_wait_flags(const unsigned long long* flags_dev, const unsigned long long* expect_dev, int n,
                              unsigned long long max_spins, unsigned* status_dev, cudaStream_t stream) {
    if (n <= 0) return 0;
    int b = (n + 127) / 128;
    if (b > 148) b = 148;
    uda::wait_flags_kernel<<<b, 128, 0, stream>>>(flags_dev, expect_dev, n, max_spins, status_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_barrier_all(unsigned long long* const* pads_dev, int rank, int world, unsigned long long epoch,
                               unsigned long long max_spins, unsigned* status_dev, cudaStream_t stream) {
    if (world < 1 || world > 64) return UDA_E_INVALID;
    uda::barrier_kernel<<<1, 64, 0, stream>>>(pads_dev, rank, world, epoch, max_spins, status_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_barrier_all_counted(unsigned long long* const* pads_dev, int rank, int world, unsigned long long* counter_dev,
                                       unsigned long long max_spins, unsigned* status_dev, cudaStream_t stream) {
    if (world < 1 || world > 64 || !counter_dev) return UDA_E_INVALID;
    uda::barrier_counted_kernel<<<1, 64, 0, stream>>>(pads_dev, rank, world, counter_dev, max_spins, status_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// uda_shuffle_push_static — the same push as uda_shuffle_push, for callers that may not allocate.
//
// Who needs it: a shuffle provider embedded in a process whose device memory is owned by someone
// else. Typical cases are a JNI plugin inside a Hadoop/YARN container where another runtime has
// already reserved the GPU's free memory, a service that sizes all of its buffers once at start-up and
// treats any later cudaMalloc as a bug (allocation synchronises the device and can fail under
// fragmentation), and code that runs between other people's kernels where an implicit device
// synchronisation is not acceptable. Such a caller still owns the MOF bytes, the arena and the
// descriptor table (inputs and outputs), but has nowhere to get the kernel's chunk-prefix scratch
// from. Here that scratch is a fixed array inside the library image.
//
// Limit: at most UDA_PUSH_STATIC_MAX_SEGMENTS descriptors per call ((max+1) * 4 bytes = 64 KiB + 4 B of
// device memory, held by every GPU that loads this library until the process exits).
//
// Set-up: the first call on each device looks up the array's device address
// (cudaGetSymbolAddress, with that device current) and caches it; make that first call outside any
// stream capture. Later calls on the same device only launch kernels.
//
// Sharing: every call on a device uses the same array and finds the previous call's prefix table in
// it. Nothing in it has to start at zero (the plan kernel rewrites entries [0, nseg] before the push
// kernel reads them), so there is nothing to clear; but two calls whose kernels overlap on the GPU
// (different streams, no dependency between them) race on the table. Issue them on one stream or
// order them with events.
constexpr int PUSH_STATIC_MAX_SEG = UDA_PUSH_STATIC_MAX_SEGMENTS;
static __device__ __align__(256) unsigned g_push_static_prefix[PUSH_STATIC_MAX_SEG + 1];

#include <mutex>
namespace {
constexpr int kMaxDevices = 64;
std::mutex g_push_static_mu;
unsigned* g_push_static_ptr[kMaxDevices] = {nullptr};
}  // namespace

extern "C" int uda_shuffle_push_static(const UdaPushDesc* descs_dev, int nseg, int max_ctas, cudaStream_t stream) {
    using namespace uda;
    if (nseg < 0) return UDA_E_INVALID;
    if (nseg > PUSH_STATIC_MAX_SEG) return UDA_E_TOO_LARGE;  // "nseg exceeds UDA_PUSH_STATIC_MAX_SEGMENTS"
    if (nseg == 0) return 0;
    int dev = -1;
    UDA_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return UDA_E_INVALID;
    unsigned* prefix = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_push_static_mu);
        if (!g_push_static_ptr[dev]) {
            // In host code the array's name is only a handle for the symbol, not a device address:
            // resolve it once per device, with that device current.
            unsigned* p = nullptr;
            cudaError_t e = cudaGetSymbolAddress((void**)&p, g_push_static_prefix);
            if (e != cudaSuccess) return (int)e;
            if (!p) return UDA_E_INVALID;
            g_push_static_ptr[dev] = p;
        }
        prefix = g_push_static_ptr[dev];
    }
    push_plan_kernel<<<1, 1024, 0, stream>>>(descs_dev, nseg, prefix);
    int grid = max_ctas > 0 ? max_ctas : kNumSMs * 8;
    push_kernel<<<grid, PUSH_THREADS, 0, stream>>>(descs_dev, nseg, prefix, 1u);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_shuffle_push_static_max_segments(void) { return PUSH_STATIC_MAX_SEG; }
