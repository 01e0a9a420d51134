// K5 / K6 — IFile record decode (boundary discovery) and encode (framing) on the device.
//
// Reference behaviour being replaced:
//   decode: BaseSegment::nextKVInternal (src/Merger/StreamRW.cc:334-404) — two VInts, EOF = (-1,-1),
//           bounds check; the two-buffer straddle join (src/Merger/StreamRW.cc:592-662) disappears
//           because a fetched segment is contiguous in the receive arena.
//   encode: write_kv_to_stream (src/Merger/StreamRW.cc:184-201, EOF at :216-218) with
//           StreamUtility::serializeLong (src/CommUtils/IOUtility.cc:167-196).
//
// Fast path: records of constant (klen, vlen) have a constant stride; frame/unframe are pure strided
// copies. General path: boundary discovery is a pointer chase (record i+1 starts where record i ends),
// parallelised as  (1) every byte position speculatively decodes "if a record started here, where does
// the next one start"  (2) pointer doubling inside a shared-memory chunk turns that into per-chunk
// entry->exit maps  (3) a tiny serial pass threads the true entry offset through the chunks
// (4) each chunk walks from its true entry and emits offsets at positions given by a prefix sum of the
// per-chunk record counts.
#include "common.cuh"
#include "../../include/uda_b200.h"

namespace uda {

// ---------------------------------------------------------------------------------- fixed stride
__global__ void frame_fixed_kernel(const uint8_t* __restrict__ raw, long long n, int klen, int vlen,
                                   uint8_t* __restrict__ out, int hdr) {
    const int rl = klen + vlen, stride = hdr + rl;
    uint8_t h[20];
    int h1 = vint_encode(h, klen);
    vint_encode(h + h1, vlen);
    const long long total = n * stride;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        long long i = t / stride;
        int o = (int)(t - i * stride);
        out[t] = o < hdr ? h[o] : raw[i * rl + (o - hdr)];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { out[total] = 0xFF; out[total + 1] = 0xFF; }
}

__global__ void unframe_fixed_kernel(const uint8_t* __restrict__ framed, long long n, int stride, int hdr,
                                     uint8_t* __restrict__ raw) {
    const int rl = stride - hdr;
    const long long total = n * rl;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        long long i = t / rl;
        int o = (int)(t - i * rl);
        raw[t] = framed[i * stride + hdr + o];
    }
}

// ---------------------------------------------------------------------------------- general decode
#ifndef UDA_DC_DOUBLING
#define UDA_DC_DOUBLING 0   // 1 = the earlier pointer-doubling resolution of the chunk maps
#endif
constexpr int DC_CHUNK = 4096;           // bytes per chunk (one CTA)
constexpr int DC_THREADS = 256;
constexpr unsigned DC_EXIT = 0x80000000u;  // high bit: the chain left the chunk; low bits = overshoot

// "next record start" if a record header began at absolute position p; returns p for EOF / invalid.
// `byte_at(q)` returns the segment byte at absolute offset q (global memory, or a shared-memory copy of the chunk).
template <class ByteAt>
__device__ __forceinline__ long long vint_at(ByteAt byte_at, long long p, int n) {
    // n = total encoded size (2..9) of a multi-byte VInt whose first byte is at p
    const int8_t b = (int8_t)byte_at(p);
    const bool neg = b < -120;
    unsigned long long t = 0;
    for (int i = 1; i < n; ++i) t = (t << 8) | byte_at(p + i);
    return neg ? (long long)~t : (long long)t;
}
template <class ByteAt>
__device__ __forceinline__ long long next_from_rd(ByteAt byte_at, long long len, long long p, bool* eof) {
    *eof = false;
    if (p >= len) { *eof = true; return p; }
    const int8_t b0 = (int8_t)byte_at(p);
    const int n1 = vint_size_from_first(b0);
    if (p + n1 >= len) { *eof = true; return p; }
    const long long klen = n1 == 1 ? (long long)b0 : vint_at(byte_at, p, n1);
    const int8_t b1 = (int8_t)byte_at(p + n1);
    const int n2 = vint_size_from_first(b1);
    if (p + n1 + n2 > len) { *eof = true; return p; }
    const long long vlen = n2 == 1 ? (long long)b1 : vint_at(byte_at, p + n1, n2);
    // EOF marker, or garbage on a speculative path (bounded so the sum below cannot overflow)
    if (klen < 0 || vlen < 0 || klen > len || vlen > len) { *eof = true; return p; }
    const long long nx = p + n1 + n2 + klen + vlen;
    if (nx > len) { *eof = true; return p; }
    return nx;
}
__device__ __forceinline__ long long next_from(const uint8_t* __restrict__ seg, long long len, long long p, bool* eof) {
    return next_from_rd([seg](long long q) { return (unsigned)seg[q]; }, len, p, eof);
}

// Phase 1+2: per chunk, for every byte offset e in the chunk: exit[e] = where the chain started at e
// leaves the chunk (offset relative to the next chunk's start, or EOF), and cnt[e] = records on the way.
// Only the entries that can be true chain entries are ever read, but all are computed (speculation).
__global__ void __launch_bounds__(DC_THREADS) decode_chunk_maps_kernel(const uint8_t* __restrict__ seg, long long len,
                                                                      long long chunk0, unsigned* __restrict__ exit_map,
                                                                      unsigned* __restrict__ cnt_map) {
    __shared__ unsigned nxt[DC_CHUNK];  // in-chunk target, or DC_EXIT|overshoot, or 0xFFFFFFFF for EOF
    __shared__ unsigned cnt[DC_CHUNK];
    __shared__ __align__(16) uint8_t sb[DC_CHUNK + 32];  // the chunk + the longest header that can start in it (2 x 9 bytes)
    const long long base = (chunk0 + blockIdx.x) * DC_CHUNK;
    const long long wbase = (long long)blockIdx.x * DC_CHUNK;  // position inside the scratch window
    for (int i = threadIdx.x; i < DC_CHUNK + 32; i += DC_THREADS) sb[i] = base + i < len ? seg[base + i] : (uint8_t)0;
    __syncthreads();
    auto byte_at = [&](long long q) { return (unsigned)sb[q - base]; };
    for (int e = threadIdx.x; e < DC_CHUNK; e += DC_THREADS) {
        long long p = base + e;
        bool eof;
        long long nx = next_from_rd(byte_at, len, p, &eof);
        if (eof) { nxt[e] = 0xFFFFFFFFu; cnt[e] = 0; }
        else if (nx - base >= DC_CHUNK) { nxt[e] = DC_EXIT | (unsigned)(nx - base - DC_CHUNK); cnt[e] = 1; }
        else { nxt[e] = (unsigned)(nx - base); cnt[e] = 1; }
    }
    __syncthreads();
#if UDA_DC_DOUBLING
    // pointer doubling: a record is >= 2 bytes so at most DC_CHUNK/2 hops = 12 rounds; stop as soon as every chain
    // has left the chunk (records of ~100 bytes need 6 rounds)
    for (int round = 0; round < 12; ++round) {
        unsigned nn[DC_CHUNK / DC_THREADS], cc[DC_CHUNK / DC_THREADS];
        bool all_out = true;
#pragma unroll
        for (int i = 0; i < DC_CHUNK / DC_THREADS; ++i) {
            int e = threadIdx.x + i * DC_THREADS;
            unsigned t = nxt[e];
            if (t & DC_EXIT) { nn[i] = t; cc[i] = cnt[e]; }
            else { nn[i] = nxt[t]; cc[i] = cnt[e] + cnt[t]; all_out = false; }
        }
        if (__syncthreads_and(all_out)) break;
#pragma unroll
        for (int i = 0; i < DC_CHUNK / DC_THREADS; ++i) {
            int e = threadIdx.x + i * DC_THREADS;
            nxt[e] = nn[i]; cnt[e] = cc[i];
        }
        __syncthreads();
    }
#else
    // Chains only move forward (a record is >= 2 bytes), so "where does the chain from e leave the chunk" resolves in
    // ONE backward sweep: when offset e is visited, everything its hop can reach has been resolved already. Pointer
    // doubling did the same with log2(hops) rounds over all 4096 entries and was bound by shared-memory wavefronts
    // (~18 k per chunk); the sweep needs ~1.5 k.
    //   pass A: every warp sweeps its own strip of 512 offsets from the top, 32 offsets at a time. A hop that stays
    //           inside the group of 32 (< 32 bytes: rare) is resolved with shuffles, highest offset first.
    //           After the pass an entry holds its final exit, or an offset inside a LATER strip.
    //   pass B: strips from the second-to-last down to the first pick up the (already final) value of that offset.
    {
        constexpr int STRIP = DC_CHUNK / (DC_THREADS / 32);
        const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
        const int s0 = wrp * STRIP, s1 = s0 + STRIP;
        for (int g = s1 - 32; g >= s0; g -= 32) {
            const int e = g + lane;
            const unsigned t = nxt[e], c = cnt[e];
            unsigned ex = t, cc = c;
            const bool ingroup = !(t & DC_EXIT) && t < (unsigned)(g + 32);
            bool resolved = true;
            if (!(t & DC_EXIT) && t < (unsigned)s1) {
                if (!ingroup) { ex = nxt[t]; cc = c + cnt[t]; }   // a later group of this strip: already swept
                else resolved = false;
            }
            unsigned pending = __ballot_sync(0xffffffffu, !resolved);
            while (pending) {
                const int tl = ingroup ? (int)(t - (unsigned)g) : lane;
                const unsigned vex = __shfl_sync(0xffffffffu, ex, tl), vcc = __shfl_sync(0xffffffffu, cc, tl);
                const int vres = __shfl_sync(0xffffffffu, (int)resolved, tl);
                if (!resolved && vres) { ex = vex; cc = c + vcc; resolved = true; }
                pending = __ballot_sync(0xffffffffu, !resolved);
            }
            nxt[e] = ex;
            cnt[e] = cc;
            __syncwarp();
        }
        for (int s = DC_THREADS / 32 - 2; s >= 0; --s) {
            __syncthreads();
            for (int e = s * STRIP + threadIdx.x; e < (s + 1) * STRIP; e += DC_THREADS) {
                const unsigned t = nxt[e];
                if (!(t & DC_EXIT)) { nxt[e] = nxt[t]; cnt[e] += cnt[t]; }   // t lies in a later strip: final by now
            }
        }
        __syncthreads();
    }
#endif
    for (int e = threadIdx.x; e < DC_CHUNK; e += DC_THREADS) {
        exit_map[wbase + e] = nxt[e];
        cnt_map[wbase + e] = cnt[e];
    }
}

// Phase 3: thread the true entry offset through the chunks of one scratch window. The chain
//     entry(c+1) = exit_map[c][entry(c)]
// is serial in c, and one dependent load per 4 KiB chunk is what a first version paid (7.5 GB/s). It is done in two
// levels instead: groups of DC_GROUP chunks get their own entry -> exit table, built speculatively for all 4096
// possible entry offsets in parallel (chains that start at different offsets fall into step after a few records,
// so the loads of a group's threads coalesce); one thread then hops from group to group, and one thread per group
// fills in the per-chunk entries.
// entry[c] = offset inside chunk c where the first record header of that chunk begins (0xFFFFFFFE: chunk
// lies inside a long record, 0xFFFFFFFF: the chain ended earlier), first[c] = index of that record.
constexpr int DC_GROUP = 64;
struct DecodeCarry { unsigned long long entry; unsigned long long total; unsigned long long done; };

// walks chunks [w0, w1) of the window starting with overshoot/entry e; optionally records entry[]/first[]
__device__ __forceinline__ void walk_chunks(const unsigned* __restrict__ exit_map, const unsigned* __restrict__ cnt_map, long long w0,
                                            long long w1, long long chunk0, unsigned long long& e, unsigned long long& total, bool& done,
                                            unsigned* __restrict__ entry, unsigned long long* __restrict__ first) {
    for (long long w = w0; w < w1; ++w) {
        if (first) first[chunk0 + w] = total;
        if (done) { if (entry) entry[chunk0 + w] = 0xFFFFFFFFu; continue; }
        if (e >= (unsigned long long)DC_CHUNK) { if (entry) entry[chunk0 + w] = 0xFFFFFFFEu; e -= DC_CHUNK; continue; }  // inside a long record
        if (entry) entry[chunk0 + w] = (unsigned)e;
        const unsigned x = exit_map[w * DC_CHUNK + e];
        total += cnt_map[w * DC_CHUNK + e];
        if (x == 0xFFFFFFFFu) done = true;
        else e = x & ~DC_EXIT;
    }
}

__global__ void __launch_bounds__(DC_THREADS) decode_group_maps_kernel(const unsigned* __restrict__ exit_map, const unsigned* __restrict__ cnt_map,
                                                                      long long nwin, unsigned* __restrict__ gexit, unsigned* __restrict__ gcnt) {
    const long long w0 = (long long)blockIdx.x * DC_GROUP;
    const long long w1 = w0 + DC_GROUP < nwin ? w0 + DC_GROUP : nwin;
    for (int e0 = threadIdx.x; e0 < DC_CHUNK; e0 += DC_THREADS) {
        unsigned long long e = (unsigned long long)e0, total = 0;
        bool done = false;
        walk_chunks(exit_map, cnt_map, w0, w1, 0, e, total, done, nullptr, nullptr);
        // a record longer than 2 GiB cannot be represented (nor fetched by the reference: int32 lengths)
        gexit[(long long)blockIdx.x * DC_CHUNK + e0] = done ? 0xFFFFFFFFu : (unsigned)(e > 0x7FFFFFFFull ? 0x7FFFFFFFull : e);
        gcnt[(long long)blockIdx.x * DC_CHUNK + e0] = (unsigned)total;
    }
}

__global__ void decode_thread_groups_kernel(const unsigned* __restrict__ exit_map, const unsigned* __restrict__ cnt_map,
                                            const unsigned* __restrict__ gexit, const unsigned* __restrict__ gcnt, long long nwin,
                                            unsigned long long* __restrict__ gentry, unsigned long long* __restrict__ gfirst,
                                            unsigned long long* __restrict__ gdone, DecodeCarry* __restrict__ carry,
                                            unsigned long long* __restrict__ nrec_out) {
    unsigned long long e = carry->entry, total = carry->total;
    bool done = carry->done != 0;
    const long long ng = (nwin + DC_GROUP - 1) / DC_GROUP;
    for (long long g = 0; g < ng; ++g) {
        gentry[g] = e; gfirst[g] = total; gdone[g] = done ? 1ull : 0ull;
        if (done) continue;
        const long long w0 = g * DC_GROUP, w1 = w0 + DC_GROUP < nwin ? w0 + DC_GROUP : nwin;
        const unsigned long long span = (unsigned long long)(w1 - w0) * DC_CHUNK;
        if (e >= span) { e -= span; continue; }                      // a long record covers the whole group
        if (e >= (unsigned long long)DC_CHUNK) {                      // the chain enters in the middle of the group: walk it
            walk_chunks(exit_map, cnt_map, w0, w1, 0, e, total, done, nullptr, nullptr);
            continue;
        }
        const unsigned x = gexit[g * DC_CHUNK + e];
        total += gcnt[g * DC_CHUNK + e];
        if (x == 0xFFFFFFFFu) done = true;
        else e = x;
    }
    carry->entry = e; carry->total = total; carry->done = done ? 1ull : 0ull;
    *nrec_out = total;
}

__global__ void decode_fill_chunks_kernel(const unsigned* __restrict__ exit_map, const unsigned* __restrict__ cnt_map, long long chunk0,
                                          long long nwin, const unsigned long long* __restrict__ gentry,
                                          const unsigned long long* __restrict__ gfirst, const unsigned long long* __restrict__ gdone,
                                          unsigned* __restrict__ entry, unsigned long long* __restrict__ first) {
    const long long ng = (nwin + DC_GROUP - 1) / DC_GROUP;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < ng; g += (long long)gridDim.x * blockDim.x) {
        unsigned long long e = gentry[g], total = gfirst[g];
        bool done = gdone[g] != 0;
        const long long w0 = g * DC_GROUP, w1 = w0 + DC_GROUP < nwin ? w0 + DC_GROUP : nwin;
        walk_chunks(exit_map, cnt_map, w0, w1, chunk0, e, total, done, entry, first);
    }
}

// Phase 4: every chunk walks its (short) true chain and writes the offsets.
__global__ void decode_emit_kernel(const uint8_t* __restrict__ seg, long long len, long long nchunks,
                                   const unsigned* __restrict__ entry, const unsigned long long* __restrict__ first,
                                   long long* __restrict__ rec_off, const unsigned long long* __restrict__ nrec) {
    for (long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x; c < nchunks; c += (long long)gridDim.x * blockDim.x) {
        unsigned e = entry[c];
        if (e >= 0xFFFFFFFEu) continue;
        long long p = c * DC_CHUNK + e;
        const long long lim = (c + 1) * DC_CHUNK;
        unsigned long long i = first[c];
        while (p < lim) {
            bool eof;
            long long nx = next_from(seg, len, p, &eof);
            if (eof) { rec_off[i] = p; return; }  // rec_off[nrec] = position of the EOF marker
            rec_off[i++] = p;
            p = nx;
        }
        if (c == nchunks - 1 || entry[c + 1] == 0xFFFFFFFFu) {
            if (i == *nrec) rec_off[i] = p;
        }
    }
}

// ---------------------------------------------------------------------------------- general encode
// Gathers (key,val) pairs addressed by offset arrays into IFile framing. out_off = exclusive prefix of
// the encoded sizes (computed by the caller with uda_ifile_encoded_sizes + a scan).
__global__ void encoded_sizes_kernel(const int* __restrict__ klen, const int* __restrict__ vlen, long long n,
                                     unsigned* __restrict__ sizes) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        sizes[i] = (unsigned)(vint_size(klen[i]) + vint_size(vlen[i]) + klen[i] + vlen[i]);
}

__global__ void encode_var_kernel(const uint8_t* __restrict__ keys, const long long* __restrict__ key_off,
                                  const int* __restrict__ klen, const uint8_t* __restrict__ vals,
                                  const long long* __restrict__ val_off, const int* __restrict__ vlen, long long n,
                                  const unsigned long long* __restrict__ out_off, uint8_t* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = warp; i < n; i += nwarps) {
        uint8_t* o = out + out_off[i];
        uint8_t h[20];
        int h1 = vint_encode(h, klen[i]);
        int h2 = vint_encode(h + h1, vlen[i]);
        const int hdr = h1 + h2;
        if (lane < hdr) o[lane] = h[lane];
        const uint8_t* k = keys + key_off[i];
        for (int b = lane; b < klen[i]; b += 32) o[hdr + b] = k[b];
        const uint8_t* v = vals + val_off[i];
        for (int b = lane; b < vlen[i]; b += 32) o[hdr + klen[i] + b] = v[b];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long end = out_off[n];
        out[end] = 0xFF; out[end + 1] = 0xFF;
    }
}

}  // namespace uda

extern "C" int uda_ifile_frame_fixed(const void* raw, long long n, int klen, int vlen, void* out, cudaStream_t stream) {
    using namespace uda;
    if (klen < 0 || vlen < 0) return UDA_E_INVALID;
    int hdr = vint_size(klen) + vint_size(vlen);
    long long total = n * (hdr + klen + vlen);
    long long b = (total + 255) / 256;
    if (b > kNumSMs * 16) b = kNumSMs * 16;
    if (b < 1) b = 1;
    frame_fixed_kernel<<<(int)b, 256, 0, stream>>>((const uint8_t*)raw, n, klen, vlen, (uint8_t*)out, hdr);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_ifile_unframe_fixed(const void* framed, long long n, int klen, int vlen, void* raw, cudaStream_t stream) {
    using namespace uda;
    int hdr = vint_size(klen) + vint_size(vlen);
    long long total = n * (klen + vlen);
    if (total <= 0) return 0;
    long long b = (total + 255) / 256;
    if (b > kNumSMs * 16) b = kNumSMs * 16;
    unframe_fixed_kernel<<<(int)b, 256, 0, stream>>>((const uint8_t*)framed, n, hdr + klen + vlen, hdr, (uint8_t*)raw);
    UDA_CHECK_LAUNCH();
    return 0;
}

constexpr long long DC_WINDOW_CHUNKS = 16384;  // 64 MiB of segment per scratch window

extern "C" long long uda_ifile_decode_scratch_bytes(long long seg_len) {
    long long nchunks = (seg_len + uda::DC_CHUNK - 1) / uda::DC_CHUNK;
    if (nchunks < 1) nchunks = 1;
    long long win = nchunks < DC_WINDOW_CHUNKS ? nchunks : DC_WINDOW_CHUNKS;
    // window maps (4+4 B per byte) + per-chunk entry (4 B) and first (8 B) + per-group tables + carry
    const long long ng = (win + uda::DC_GROUP - 1) / uda::DC_GROUP;
    return win * uda::DC_CHUNK * 8 + nchunks * 12 + ng * (uda::DC_CHUNK * 8 + 24) + 1024;
}

// rec_off_out must hold (max_records + 1) entries; *nrec_dev receives the record count and
// rec_off_out[nrec] the offset of the EOF marker.
extern "C" int uda_ifile_decode_offsets(const void* seg, long long seg_len, long long* rec_off_out,
                                        unsigned long long* nrec_dev, void* scratch, cudaStream_t stream) {
    using namespace uda;
    if (seg_len <= 0) return UDA_E_INVALID;
    const long long nchunks = (seg_len + DC_CHUNK - 1) / DC_CHUNK;
    const long long win = nchunks < DC_WINDOW_CHUNKS ? nchunks : DC_WINDOW_CHUNKS;
    uint8_t* s = (uint8_t*)scratch;
    DecodeCarry* carry = (DecodeCarry*)s; s += 256;
    unsigned* exit_map = (unsigned*)s; s += win * DC_CHUNK * 4;
    unsigned* cnt_map = (unsigned*)s; s += win * DC_CHUNK * 4;
    unsigned long long* first = (unsigned long long*)s; s += nchunks * 8;
    const long long ngw = (win + DC_GROUP - 1) / DC_GROUP;
    unsigned long long* gentry = (unsigned long long*)s; s += ngw * 8;
    unsigned long long* gfirst = (unsigned long long*)s; s += ngw * 8;
    unsigned long long* gdone = (unsigned long long*)s; s += ngw * 8;
    unsigned* gexit = (unsigned*)s; s += ngw * DC_CHUNK * 4;
    unsigned* gcnt = (unsigned*)s; s += ngw * DC_CHUNK * 4;
    unsigned* entry = (unsigned*)s;
    UDA_CUDA(cudaMemsetAsync(carry, 0, sizeof(DecodeCarry), stream));
    for (long long c0 = 0; c0 < nchunks; c0 += win) {
        const long long nw = nchunks - c0 < win ? nchunks - c0 : win;
        decode_chunk_maps_kernel<<<(unsigned)nw, DC_THREADS, 0, stream>>>((const uint8_t*)seg, seg_len, c0, exit_map, cnt_map);
        const long long ng = (nw + DC_GROUP - 1) / DC_GROUP;
        decode_group_maps_kernel<<<(unsigned)ng, DC_THREADS, 0, stream>>>(exit_map, cnt_map, nw, gexit, gcnt);
        decode_thread_groups_kernel<<<1, 1, 0, stream>>>(exit_map, cnt_map, gexit, gcnt, nw, gentry, gfirst, gdone, carry, nrec_dev);
        decode_fill_chunks_kernel<<<(unsigned)((ng + 127) / 128), 128, 0, stream>>>(exit_map, cnt_map, c0, nw, gentry, gfirst, gdone, entry, first);
    }
    long long b = (nchunks + 127) / 128;
    if (b > kNumSMs * 8) b = kNumSMs * 8;
    decode_emit_kernel<<<(int)b, 128, 0, stream>>>((const uint8_t*)seg, seg_len, nchunks, entry, first, rec_off_out, nrec_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_ifile_encoded_sizes(const int* klen_dev, const int* vlen_dev, long long n, unsigned* sizes_out,
                                       cudaStream_t stream) {
    if (n <= 0) return 0;
    int b = (int)((n + 255) / 256 < 148LL * 16 ? (n + 255) / 256 : 148LL * 16);
    uda::encoded_sizes_kernel<<<b, 256, 0, stream>>>(klen_dev, vlen_dev, n, sizes_out);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_ifile_encode_var(const void* keys, const long long* key_off, const int* klen, const void* vals,
                                    const long long* val_off, const int* vlen, long long n,
                                    const unsigned long long* out_off_dev, void* out, cudaStream_t stream) {
    long long b = n > 0 ? (n * 32 + 255) / 256 : 1;
    if (b > 148LL * 16) b = 148LL * 16;
    uda::encode_var_kernel<<<(int)b, 256, 0, stream>>>((const uint8_t*)keys, key_off, klen, (const uint8_t*)vals, val_off, vlen,
                                                     n, out_off_dev, (uint8_t*)out);
    UDA_CHECK_LAUNCH();
    return 0;
}
