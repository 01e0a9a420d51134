// In-shared-memory pairwise merge-path tree shared by the entry merge (merge.cu) and the fused record
// merge (merge_records.cu).
#pragma once
#include "common.cuh"

namespace uda {

template <int THREADS>
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total_out) {
    __shared__ uint32_t ws[THREADS / 32];
    __shared__ uint32_t tot;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t x = v;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    __syncthreads();  // protect ws/tot from the previous call
    if (lane == 31) ws[w] = x;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t acc = 0;
        for (int i = 0; i < THREADS / 32; ++i) { uint32_t t = ws[i]; ws[i] = acc; acc += t; }
        tot = acc;
    }
    __syncthreads();
    *total_out = tot;
    return ws[w] + x - v;
}

// one level of the in-shared-memory merge tree: lists (soff, L) in src -> pairs merged into dst
template <int THREADS, int VT>
__device__ __forceinline__ void merge_level(const Entry* __restrict__ src, const uint32_t* __restrict__ soff, int L,
                                            Entry* __restrict__ dst, uint32_t* __restrict__ doff, uint32_t total) {
    const int Lp = (L + 1) >> 1;
    for (int p = threadIdx.x; p <= Lp; p += THREADS) doff[p] = p < Lp ? soff[2 * p] : total;
    const uint32_t o = threadIdx.x * VT;
    if (o < total) {
        // pair containing output position o: largest p with soff[2p] <= o
        int plo = 0, phi = Lp - 1;
        while (plo < phi) {
            int mid = (plo + phi + 1) >> 1;
            if (soff[2 * mid] <= o) plo = mid; else phi = mid - 1;
        }
        int p = plo;
        uint32_t a0 = soff[2 * p];
        uint32_t a1 = soff[min(2 * p + 1, L)];
        uint32_t b1 = soff[min(2 * p + 2, L)];
        // merge path on diagonal d
        uint32_t d = o - a0, la = a1 - a0, lb = b1 - a1;
        uint32_t lo = d > lb ? d - lb : 0u, hi = d < la ? d : la;
        while (lo < hi) {
            uint32_t mid = (lo + hi) >> 1;
            Entry ea = src[a0 + mid], eb = src[a1 + (d - 1 - mid)];
            if (!entry_less(eb, ea)) lo = mid + 1; else hi = mid;  // A[mid] <= B[d-1-mid]
        }
        uint32_t ia = a0 + lo, ib = a1 + (d - lo);
        uint32_t ea_end = a1, eb_end = b1;
        bool ha = ia < ea_end, hb = ib < eb_end;
        Entry ea, eb;
        if (ha) ea = src[ia];
        if (hb) eb = src[ib];
#pragma unroll
        for (int t = 0; t < VT; ++t) {
            uint32_t pos = o + t;
            if (pos >= total) break;
            if (!ha && !hb) {  // pair exhausted: the next non-empty pair starts exactly at pos
                do { ++p; } while (soff[min(2 * p + 2, L)] <= pos);
                ia = soff[2 * p];
                ea_end = soff[min(2 * p + 1, L)];
                ib = ea_end;
                eb_end = soff[min(2 * p + 2, L)];
                ha = ia < ea_end; hb = ib < eb_end;
                if (ha) ea = src[ia];
                if (hb) eb = src[ib];
            }
            bool takeA = ha && (!hb || !entry_less(eb, ea));
            if (takeA) {
                dst[pos] = ea;
                ++ia; ha = ia < ea_end;
                if (ha) ea = src[ia];
            } else {
                dst[pos] = eb;
                ++ib; hb = ib < eb_end;
                if (hb) eb = src[ib];
            }
        }
    }
}

}  // namespace uda
