/* uda_b200 — C ABI of the sm_100a kernel library (libuda_kernels.so).
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on and returns 0 or a
 * cudaError_t value (negative values are argument errors raised by this library, see UDA_E_*).
 * Unless stated otherwise an entry point allocates nothing: scratch is passed in by the caller and
 * its required size is returned by the matching *_scratch_bytes() function.
 *
 * Reference citations name the CPU code path of chaoyang223/UDA whose behaviour the kernel replaces.
 */
#ifndef UDA_B200_H
#define UDA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __DRIVER_TYPES_H__
typedef struct CUstream_st* cudaStream_t;
#endif

#define UDA_E_INVALID (-1)   /* bad argument */
#define UDA_E_TOO_LARGE (-2) /* size exceeds a documented maximum */

/* 16-byte sort entry, compared as the 128-bit integer (w0:w1); see csrc/kernels/common.cuh. */
typedef struct {
    unsigned long long w0; /* key bytes 0..7 big-endian */
    unsigned long long w1; /* key bytes 8..9 << 48 | run << 32 | record index in run */
} UdaEntry;

/* One sorted run (= one fetched map-output segment) as seen by the reducer-side kernels. */
typedef struct {
    const void* base;   /* first record of the run (device pointer, may be a peer/arena address) */
    long long nrec;     /* records in the run */
    long long out_off;  /* first entry of this run in the concatenated entry array */
    unsigned run_id;    /* stable-order rank of the run (map index) */
    unsigned idx0;      /* record index assigned to the first record */
} UdaRunDesc;

/* One contiguous copy of the fused shuffle: MOF partition segment -> peer receive arena. */
typedef struct {
    const void* src;              /* local MOF bytes (segment start) */
    void* dst;                    /* destination in the (peer-mapped) receive arena */
    long long bytes;              /* segment length (partLength) */
    unsigned long long* flag;     /* landed-bytes counter in the destination rank's arena */
} UdaPushDesc;

/* ---- synthetic data + validation (no reference counterpart: Hadoop teragen/teravalidate,
 *      scripts/regression/sortcountRunner.sh:123-241) ------------------------------------------ */
int uda_teragen(void* out, long long n, unsigned long long seed, unsigned long long first_row, int framed,
                int key_bits, cudaStream_t stream);
int uda_validate_fixed(const void* recs, long long n, int stride, int key_off, unsigned long long* result3,
                       cudaStream_t stream);

/* ---- K3 key normalise (reference: src/Merger/CompareFunc.cc:70-113) ---------------------------- */
int uda_extract_keys_fixed(const UdaRunDesc* runs_dev, int nruns, long long max_nrec, int stride, int key_off,
                           int key_len, void* entries_out, cudaStream_t stream);
int uda_extract_keys_var(const void* data, const long long* rec_off_dev, long long n, int key_kind, unsigned run_id,
                         unsigned idx0, void* entries_out, unsigned* long_key_flag_dev, cudaStream_t stream);

/* ---- K2 map-side LSD radix sort of entries (reference: Hadoop MapTask sort, implied by the sorted
 *      MOFs consumed at src/Merger/MergeQueue.h:299-321) ------------------------------------------ */
long long uda_radix_sort_scratch_bytes(long long n);
/* Sorts by the top key_bytes (1..10) bytes of the 80-bit key prefix; stable. Result ends in `a`
 * (b is the ping-pong buffer of the same size). */
int uda_radix_sort_entries(void* a, void* b, long long n, int key_bytes, void* scratch, cudaStream_t stream);

/* ---- K4 reducer-side k-way merge of sorted runs (reference: src/Merger/MergeQueue.h:126-321 heap +
 *      src/Merger/StreamRW.cc:151-225 write_kv_to_stream hot loop) ------------------------------- */
long long uda_merge_scratch_bytes(long long n, int nruns);
/* entries: runs concatenated, run r = [run_off[r], run_off[r+1]); each run sorted. sorted_out gets all
 * n entries in (key, run, index) order. nruns <= 4096, n < 2^32. */
int uda_kway_merge_entries(const void* entries, const long long* run_off_dev, int nruns, long long n,
                           void* sorted_out, void* scratch, cudaStream_t stream);
/* Fused front end for fixed-stride records: key normalise + histogram in one pass over the landed
 * segments (described by runs_dev), then the same merge. entries_tmp / sorted_out: n entries each. */
int uda_merge_runs_fixed(const UdaRunDesc* runs_dev, const long long* run_off_dev, int nruns, long long n,
                         long long max_nrec, int stride, int key_off, int key_len, void* entries_tmp,
                         void* sorted_out, void* scratch, cudaStream_t stream);
/* Fused fixed-stride merge that reads and writes every record exactly once (sampled k-way merge-path
 * partition, TMA-staged record tiles ordered in shared memory by key bins + exact in-bin ranking, independent of
 * the run count; merge tree for duplicate-heavy tiles; records stored straight to their merged position). out: n*stride bytes (+2 for the EOF marker). align2 = 1 when stride, key_off, every run base
 * and `out` are even (TeraSort). Runs need 32 readable bytes after their last record. ctas_per_sm caps the
 * resident merge CTAs per SM (0 = occupancy limit; lower it when a push kernel runs concurrently). */
long long uda_merge_records_scratch_bytes(long long n, int nruns);
int uda_merge_records_supported(int nruns, long long n, int stride, int key_len);
int uda_merge_records_fixed(const UdaRunDesc* runs_dev, int nruns, long long n, long long max_nrec, int stride,
                            int key_off, int key_len, int align2, void* out, int write_eof, void* scratch,
                            int ctas_per_sm, cudaStream_t stream);
/* key-range cuts of a sorted fixed-stride run (spill path): equi-depth splitters, and positions of splitters */
int uda_sample_splitters(const void* base, long long n, int stride, int key_off, int key_len, int nkeys,
                         unsigned long long* keys_out_dev, cudaStream_t stream);
int uda_lower_bound_records(const void* base, long long n, int stride, int key_off, int key_len,
                            const unsigned long long* keys_dev, int nkeys, long long* pos_out_dev, cudaStream_t stream);
int uda_merge_records_stats(const void* scratch, unsigned long long* out2_host, cudaStream_t stream);
/* {oversize buckets, bisection iterations, tiles ordered by the merge tree instead of the key bins}; synchronises */
int uda_merge_records_stats3(const void* scratch, unsigned long long* out3_host, cudaStream_t stream);
/* tile capacity (record slots) and dynamic shared memory per CTA the fused merge uses for this shape */
int uda_merge_records_tile(int nruns, int stride, int* cap_out, long long* smem_out);
/* synchronises the stream; out2_host = {oversize buckets, bisection iterations} of the last merge */
int uda_merge_stats(const void* scratch, unsigned long long* out2_host, cudaStream_t stream);

/* ---- merged-stream materialisation (reference: src/Merger/StreamRW.cc:184-201,216-218) ---------- */
/* dst must be 16-byte aligned and hold n*stride (+2 if write_eof) bytes. run_base_dev[run] = device
 * address of record 0 of that run; source buffers need 32 readable bytes after their last record. */
int uda_gather_fixed(const void* sorted_entries, long long n, const unsigned long long* run_base_dev, int stride,
                     void* dst, int write_eof, cudaStream_t stream);
int uda_gather_var_lengths(const void* sorted_entries, long long n, const unsigned long long* run_recoff_dev,
                           unsigned* rec_len_out, cudaStream_t stream);
int uda_gather_var(const void* sorted_entries, long long n, const unsigned long long* run_base_dev,
                   const unsigned long long* run_recoff_dev, const unsigned long long* dst_off_dev, void* dst,
                   cudaStream_t stream);

/* ---- general (variable-length) reduce path: exact comparator tie-break + 64-bit stream offsets
 *      (reference: src/Merger/CompareFunc.cc:70-91 called from the segment heap, src/Merger/StreamRW.h:163) ---- */
/* Re-orders, inside `sorted_entries`, every group of entries that share the 80-bit key prefix but are not in the
 * order of the exact comparator (key_kind 0 raw bytes, 1 Text, 2 BytesWritable; equal keys keep (run, index)
 * order). run_base_dev[run] / run_recoff_dev[run] locate record idx of a run as base + recoff[idx].
 * tmp_entries: n entries of scratch (the merge's input array is free at this point). n < 2^31-1. */
long long uda_tiebreak_scratch_bytes(long long n);
int uda_tiebreak_var(void* sorted_entries, long long n, const unsigned long long* run_base_dev,
                     const unsigned long long* run_recoff_dev, int key_kind, void* tmp_entries, void* scratch,
                     cudaStream_t stream);
/* out3_host: {groups re-ordered, largest such group, entries moved}; synchronises the stream */
int uda_tiebreak_stats(const void* scratch, long long n, unsigned long long* out3_host, cudaStream_t stream);
/* out[i] = sum of v[0..i) as 64-bit, out[m] = total (record lengths -> byte offsets of the merged stream) */
long long uda_scan_u32_to_u64_scratch_bytes(long long m);
int uda_scan_u32_to_u64(const unsigned* v, long long m, unsigned long long* out, void* scratch, cudaStream_t stream);
/* writes the IFile EOF marker (VInt -1, VInt -1) at dst + *total_dev (dst + 0 when total_dev is null) */
int uda_write_eof_at(void* dst, const unsigned long long* total_dev, cudaStream_t stream);

/* ---- K1 fused shuffle push over peer memory (reference: src/DataNet/RDMAClient.cc:559-600 request,
 *      src/MOFServer/IndexInfo.cc:237-335 chunk read, src/DataNet/RDMAServer.cc:537-631 write+ack,
 *      :44-136 credits) --------------------------------------------------------------------------- */
long long uda_push_scratch_bytes(int nseg);
int uda_shuffle_push(const UdaPushDesc* descs_dev, int nseg, void* scratch, int max_ctas, cudaStream_t stream);
/* the same push with the chunk order interleaved over `nway` destination ranks (peer-major descriptor table), so the
 * local copy and the remote pushes run concurrently; uda_shuffle_push is nway = 1 (chunks in descriptor order) */
int uda_shuffle_push_nway(const UdaPushDesc* descs_dev, int nseg, void* scratch, int max_ctas, int nway,
                          cudaStream_t stream);
/* Copy-engine variant of the push: `nsub` descriptor tables of nseg/nsub entries each, concatenated in the order
 * they must land (sub-partition 0 first). `num_ctas` single-warp CTAs (default 32) each claim one SM through a
 * 192 KiB shared-memory ring and move 32 KiB chunks with cp.async.bulk in both directions; the landed-byte counters
 * are advanced per chunk exactly as by uda_shuffle_push. `nway` interleaves the chunk order inside a sub-partition
 * over the peer-major descriptor order. scratch: uda_push_tma_scratch_bytes(nseg). */
long long uda_push_tma_scratch_bytes(int nseg);
/* flush_chunks: chunks (16 KiB each) a copy warp completes between two advances of the landed-byte counters
 * (1..64, 0 = default 8): each advance costs one system-scope fence. */
int uda_shuffle_push_tma(const UdaPushDesc* descs_dev, int nseg, int nsub, int nway, void* scratch, int num_ctas,
                         int flush_chunks, cudaStream_t stream);
/* Receive ring with credits (reference: two 1 MiB buffers per map refilled on consumption, src/Merger/StreamRW.cc:462-486,
 * sender throttled by credits, src/DataNet/RDMAServer.cc:44-136). The destination arenas hold fewer sub-partitions than the
 * step has; before a copy warp touches sub-partition j it publishes what it has in flight and waits until
 * credits_dev[p] >= need0 + j for every rank p < world (credits_dev: this rank's credit slots, written by the reducers).
 * A wait that exceeds max_spins polls sets *status_dev = 3 and the warp stops pushing. */
int uda_shuffle_push_tma_ring(const UdaPushDesc* descs_dev, int nseg, int nsub, int nway, void* scratch, int num_ctas,
                              int flush_chunks, const unsigned long long* credits_dev, int world, long long need0,
                              unsigned long long max_spins, unsigned* status_dev, cudaStream_t stream);
/* reducer side: ++*counter_dev, then the new count is stored (system-scope release) to credit_slots_dev[p] for p < world */
int uda_credit_publish(unsigned long long* const* credit_slots_dev, int world, unsigned long long* counter_dev,
                       cudaStream_t stream);
/* No counterpart in the reference (it has no device code and nothing that must avoid allocation).
 * Same push as uda_shuffle_push (chunks in descriptor order), but the chunk-prefix scratch is a fixed array inside the library,
 * so the call never allocates: for providers embedded in a process that does not own the GPU's free
 * memory (JNI plugin next to another runtime, fixed-budget services) or that must not trigger the
 * implicit device synchronisation of cudaMalloc. At most UDA_PUSH_STATIC_MAX_SEGMENTS descriptors per
 * call (returns UDA_E_TOO_LARGE beyond that). The first call on each device resolves the array's
 * address and must be made outside stream capture. All calls on a device share the array: calls
 * whose kernels overlap on the GPU race; keep them on one stream or order them with events. */
#define UDA_PUSH_STATIC_MAX_SEGMENTS 16384
int uda_shuffle_push_static(const UdaPushDesc* descs_dev, int nseg, int max_ctas, cudaStream_t stream);
int uda_shuffle_push_static_max_segments(void);
int uda_wait_flags(const unsigned long long* flags_dev, const unsigned long long* expect_dev, int n,
                   unsigned long long max_spins, unsigned* status_dev, cudaStream_t stream);
int uda_barrier_all(unsigned long long* const* pads_dev, int rank, int world, unsigned long long epoch,
                    unsigned long long max_spins, unsigned* status_dev, cudaStream_t stream);
/* the same barrier with the epoch kept in *counter_dev (this rank's memory, advanced by the kernel): no host-side
 * state, so the launch can be part of a captured CUDA graph that is replayed every step */
int uda_barrier_all_counted(unsigned long long* const* pads_dev, int rank, int world, unsigned long long* counter_dev,
                            unsigned long long max_spins, unsigned* status_dev, cudaStream_t stream);

/* ---- K7 CRC32, zlib polynomial (no reference counterpart: the reference transfers Hadoop's IFile
 *      checksum and ignores it, src/MOFServer/IndexInfo.cc:308-310, src/Merger/StreamRW.cc:361-367) - */
int uda_crc32(const void* data, long long n, unsigned* acc_scratch, unsigned* out_dev, void* out_be4_dev,
              cudaStream_t stream);

/* ---- K5/K6 IFile decode / encode (reference: src/Merger/StreamRW.cc:334-404 nextKVInternal,
 *      :184-201 serialise, src/CommUtils/IOUtility.cc:167-196 VInt) ------------------------------- */
int uda_ifile_frame_fixed(const void* raw, long long n, int klen, int vlen, void* out, cudaStream_t stream);
int uda_ifile_unframe_fixed(const void* framed, long long n, int klen, int vlen, void* raw, cudaStream_t stream);
long long uda_ifile_decode_scratch_bytes(long long seg_len);
int uda_ifile_decode_offsets(const void* seg, long long seg_len, long long* rec_off_out, unsigned long long* nrec_dev,
                             void* scratch, cudaStream_t stream);
int uda_ifile_encoded_sizes(const int* klen_dev, const int* vlen_dev, long long n, unsigned* sizes_out,
                            cudaStream_t stream);
int uda_ifile_encode_var(const void* keys, const long long* key_off, const int* klen, const void* vals,
                         const long long* val_off, const int* vlen, long long n, const unsigned long long* out_off_dev,
                         void* out, cudaStream_t stream);

/* ---- K9 block decompression on the device (reference: decompress thread + dlopen'ed codecs,
 *      src/Merger/DecompressorWrapper.cc:168-197, src/Merger/SnappyDecompressor.cc:89-105). A compressed segment is
 *      Hadoop block framing: [u32 BE rawLen][u32 BE compLen][compLen bytes of raw Snappy] repeated. -------------- */
long long uda_snappy_table_bytes(int max_blocks);
/* out2_dev = {blocks found, total plain bytes}; status_dev: 0 ok, 1 more than max_blocks, 2 corrupt block header */
int uda_snappy_scan_blocks(const void* seg, long long seg_len, int max_blocks, void* table, unsigned long long* out2_dev,
                           unsigned* status_dev, cudaStream_t stream);
/* one warp per block; `out` holds the total plain bytes; status_dev becomes 3 when a block is corrupt */
int uda_snappy_decode_blocks(const void* seg, const void* table, const unsigned long long* nblocks_dev, int max_blocks, void* out,
                             unsigned* status_dev, cudaStream_t stream);
/* LZO1X blocks (reference: src/Merger/LzoDecompressor.cc:121-167) in the same framing: the table comes from
 * uda_snappy_scan_blocks (the framing is codec independent); one warp per block, status_dev becomes 3 on a corrupt block */
int uda_lzo_decode_blocks(const void* seg, const void* table, const unsigned long long* nblocks_dev, int max_blocks, void* out,
                          unsigned* status_dev, cudaStream_t stream);

/* ---- symmetric receive arena plumbing (reference: registered MR + rkey exchange,
 *      src/DataNet/RDMAComm.cc:67-154,178-218, src/DataNet/RDMAComm.h:35-39). These allocate. ---- */
int uda_dev_malloc(void** ptr, long long bytes);
int uda_dev_free(void* ptr);
int uda_ipc_handle_bytes(void);
int uda_ipc_get_handle(void* ptr, void* handle_out);
int uda_ipc_open_handle(const void* handle_in, void** ptr_out);
int uda_ipc_close_handle(void* ptr);
int uda_can_access_peer(int dev, int peer, int* out);
int uda_memset_async(void* ptr, int value, long long bytes, cudaStream_t stream);
int uda_host_alloc_pinned(void** ptr, long long bytes);
int uda_host_free_pinned(void* ptr);
/* runtime services for hosts that use this ABI without a CUDA runtime of their own (csrc/runtime/device_merge.cc) */
int uda_device_count(int* out);
int uda_set_device(int dev);
int uda_mem_info(long long* free_bytes, long long* total_bytes);
int uda_memcpy_h2d_async(void* dst_dev, const void* src_host, long long bytes, cudaStream_t stream);
int uda_memcpy_d2h_async(void* dst_host, const void* src_dev, long long bytes, cudaStream_t stream);
int uda_stream_synchronize(cudaStream_t stream);
/* text for a return code of this library (negative: UDA_E_*, positive: cudaError_t) */
const char* uda_error_string(int rc);

/* ---- K1 second mode: fused key partition + record scatter for map outputs not yet grouped by reducer
 *      (no reference counterpart: Hadoop's MapTask partitions; the transfer it fuses with replaces
 *      src/DataNet/RDMAServer.cc:537-631). mode 0 = Hadoop HashPartitioner over key[key_skip..key_len),
 *      mode 1 = total-order range over P-1 big-endian 8-byte splitters. stride even, <= 104; P <= 256. ---- */
int uda_partition_count(const void* recs, long long n, int stride, int key_off, int key_len, int key_skip, int mode,
                        int P, const unsigned long long* splitters_dev, unsigned long long* counts_dev,
                        cudaStream_t stream);
int uda_partition_scatter(const void* recs, long long n, int stride, int key_off, int key_len, int key_skip, int mode,
                          int P, const unsigned long long* splitters_dev, const unsigned long long* dst_dev,
                          unsigned long long* cursor_dev, cudaStream_t stream);
/* ---- combiner (WordCount-style sum of 4-byte big-endian values over equal adjacent keys of a sorted
 *      fixed-stride stream; the reference leaves combining to Hadoop's reducer) ---------------------- */
int uda_combine_heads(const void* recs, long long n, int stride, int key_off, int key_len, unsigned* head_dev,
                      cudaStream_t stream);
int uda_combine_emit(const void* recs, long long n, long long groups, int stride, int key_off, int key_len, int val_off,
                     const unsigned* gid_dev, void* out, unsigned long long* sums_dev, cudaStream_t stream);
/* device-wide exclusive scan of m < 2^31 u32 values (out may alias v; out[m] = total) */
long long uda_scan_u32_scratch_bytes(long long m);
int uda_scan_u32(const unsigned* v, long long m, unsigned* out, unsigned* sums_scratch, cudaStream_t stream);

/* ---- K10 broadcast of side tables: NVLS multicast store (multimem.st) or unicast peer stores
 *      (no reference counterpart: Hadoop distributes such tables as files) ------------------------ */
int uda_multicast_store(const void* src, unsigned long long multicast_ptr, long long bytes, cudaStream_t stream);
int uda_unicast_store(const void* src, const unsigned long long* peer_ptrs_dev, int world, long long bytes,
                      cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* UDA_B200_H */
