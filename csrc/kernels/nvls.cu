// K10 — broadcast of small side tables (partition splitters, bucket boundary keys, configuration) to every rank.
//
// Reference context: there is no counterpart — Hadoop ships such tables as files in the distributed cache. With
// NVSwitch the natural mechanism is an NVLS multicast store: one `multimem.st` on a multicast address lands the
// data in every rank's copy of a symmetric buffer inside the switch, so the root pays for one store stream
// instead of P-1. The multicast mapping comes from torch's symmetric memory rendezvous
// (uda_b200/parallel/nvls.py); when the fabric offers no multicast the same API falls back to unicast peer
// stores from this file.
#include "common.cuh"
#include "../../include/uda_b200.h"

namespace uda {

__global__ void multicast_store_kernel(const int4* __restrict__ src, unsigned long long mc_addr, long long nvec) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
        const int4 v = src[i];
        asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr + (unsigned long long)i * 16),
                     "f"(__int_as_float(v.x)), "f"(__int_as_float(v.y)), "f"(__int_as_float(v.z)), "f"(__int_as_float(v.w))
                     : "memory");
    }
    __threadfence_system();
}

__global__ void unicast_store_kernel(const int4* __restrict__ src, const unsigned long long* __restrict__ peers, int world,
                                     long long nvec) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
        const int4 v = src[i];
        for (int p = 0; p < world; ++p) st_stream(reinterpret_cast<int4*>(peers[p]) + i, v);
    }
    __threadfence_system();
}

}  // namespace uda

// bytes must be a multiple of 16; src and the multicast address 16-byte aligned.
extern "C" int uda_multicast_store(const void* src, unsigned long long multicast_ptr, long long bytes, cudaStream_t stream) {
    if (bytes <= 0) return 0;
    if ((bytes & 15) || (multicast_ptr & 15) || ((uintptr_t)src & 15) || multicast_ptr == 0) return UDA_E_INVALID;
    long long nvec = bytes >> 4;
    int b = (int)((nvec + 255) / 256 < 148LL * 4 ? (nvec + 255) / 256 : 148LL * 4);
    uda::multicast_store_kernel<<<b, 256, 0, stream>>>((const int4*)src, multicast_ptr, nvec);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_unicast_store(const void* src, const unsigned long long* peer_ptrs_dev, int world, long long bytes,
                                 cudaStream_t stream) {
    if (bytes <= 0) return 0;
    if ((bytes & 15) || ((uintptr_t)src & 15) || world < 1) return UDA_E_INVALID;
    long long nvec = bytes >> 4;
    int b = (int)((nvec + 255) / 256 < 148LL * 4 ? (nvec + 255) / 256 : 148LL * 4);
    uda::unicast_store_kernel<<<b, 256, 0, stream>>>((const int4*)src, peer_ptrs_dev, world, nvec);
    UDA_CHECK_LAUNCH();
    return 0;
}
