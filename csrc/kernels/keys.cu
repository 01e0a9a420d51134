// K3 — key normalise: turn serialized Hadoop keys into fixed-width, order-preserving 16-byte sort
// entries so that every later kernel (radix sort, merge-path partition, smem merge) compares
// integers instead of calling a byte comparator per heap sift.
//
// Reference behaviour being replaced: src/Merger/CompareFunc.cc:70-91 (memcmp with a type specific
// header skip: Text = leading VInt, BytesWritable = 4-byte length, fixed-width Writables = none, then
// shorter-key-first), invoked per compare from the heap at src/Merger/StreamRW.h:163.
//
// Entry layout: see common.cuh. The first 10 *compared* key bytes become the 80-bit prefix. Keys
// longer than 10 compared bytes (or shorter: zero padded) may tie on the prefix; the tie-break kernel
// in merge.cu re-orders equal-prefix groups with the exact comparator.
#include "common.cuh"
#include "../../include/uda_b200.h"

namespace uda {

// Fixed-stride records (TeraSort fast path). One launch covers many runs: blockIdx.y = run.
__global__ void extract_keys_fixed_kernel(const UdaRunDesc* __restrict__ runs, int stride, int key_off, int key_len,
                                          Entry* __restrict__ out) {
    const UdaRunDesc rd = runs[blockIdx.y];
    const uint8_t* base = (const uint8_t*)rd.base;
    Entry* o = out + rd.out_off;
    const unsigned long long hi_bits = ((unsigned long long)rd.run_id & 0xFFFFull) << 32;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < rd.nrec; i += (long long)gridDim.x * blockDim.x) {
        unsigned long long w0;
        uint32_t t;
        key_prefix(base + i * stride + key_off, key_len, &w0, &t);
        Entry e;
        e.w0 = w0;
        e.w1 = ((unsigned long long)t << 48) | hi_bits | (unsigned long long)(uint32_t)(rd.idx0 + i);
        o[i] = e;
    }
}

// Variable-length IFile records addressed through a record-offset array (general path).
// rec_off[i] points at the record header (VInt klen, VInt vlen). key_kind: 0 raw bytes, 1 Text
// (skip leading VInt), 2 BytesWritable (skip 4 bytes).
__global__ void extract_keys_var_kernel(const uint8_t* __restrict__ data, const long long* __restrict__ rec_off,
                                        long long n, int key_kind, uint32_t run_id, uint32_t idx0,
                                        Entry* __restrict__ out, uint32_t* __restrict__ long_key_flag) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const uint8_t* p = data + rec_off[i];
        long long klen, vlen;
        int n1 = vint_decode(p, &klen);
        int n2 = vint_decode(p + n1, &vlen);
        const uint8_t* k = p + n1 + n2;
        int len = (int)klen;
        if (key_kind == 1 && len > 0) {
            int s = vint_size_from_first((int8_t)k[0]);
            k += s; len -= s;
        } else if (key_kind == 2) {
            k += 4; len -= 4;
        }
        if (len < 0) len = 0;
        if (len != 10) atomicOr(long_key_flag, 1u);  // prefix ties are possible -> exact tie-break needed
        unsigned long long w0;
        uint32_t t;
        key_prefix(k, len, &w0, &t);
        Entry e;
        e.w0 = w0;
        e.w1 = ((unsigned long long)t << 48) | (((unsigned long long)run_id & 0xFFFFull) << 32) | (unsigned long long)(uint32_t)(idx0 + i);
        out[i] = e;
    }
}

}  // namespace uda

extern "C" int uda_extract_keys_fixed(const UdaRunDesc* runs_dev, int nruns, long long max_nrec, int stride, int key_off,
                                      int key_len, void* entries_out, cudaStream_t stream) {
    if (nruns <= 0 || max_nrec <= 0) return 0;
    long long bx = (max_nrec + 255) / 256;
    long long cap = (148LL * 16 + nruns - 1) / nruns;
    if (cap < 1) cap = 1;
    if (bx > cap) bx = cap;
    dim3 grid((unsigned)bx, (unsigned)nruns);
    uda::extract_keys_fixed_kernel<<<grid, 256, 0, stream>>>(runs_dev, stride, key_off, key_len, (uda::Entry*)entries_out);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_extract_keys_var(const void* data, const long long* rec_off_dev, long long n, int key_kind,
                                    unsigned run_id, unsigned idx0, void* entries_out, unsigned* long_key_flag_dev,
                                    cudaStream_t stream) {
    if (n <= 0) return 0;
    int blocks = (int)((n + 255) / 256 < 148LL * 16 ? (n + 255) / 256 : 148LL * 16);
    uda::extract_keys_var_kernel<<<blocks, 256, 0, stream>>>((const uint8_t*)data, rec_off_dev, n, key_kind, run_id, idx0,
                                                           (uda::Entry*)entries_out, long_key_flag_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}
