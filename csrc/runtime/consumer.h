// Consumer role ("NetMerger"): fetches one partition from every map output and streams the merged
// key/value sequence to the reducer.
//
// Reference behaviour covered:
//   * MergeManager_main / reduce_downcall_handler / finalize  (src/Merger/NetMergerMain.cc:44-77,
//     src/Merger/reducer.cc:56-217,354-410): INIT -> comparator, codec, buffer sizing with page alignment
//     and minimum check, memory pool; FETCH -> queue a map output; FINAL -> flag
//   * calculateMemPool (src/Merger/reducer.cc:453-496): (num_kv_bufs + 10) buffer pairs; with
//     compression the pair is split by mapred.rdma.compression.buffer.ratio
//   * MergeManager online and hybrid modes (src/Merger/MergeManager.cc:47-365): fetching phase with a
//     fetchOverMessage up-call every 20 segments, merging phase that emits blocks of whole records through
//     dataFromUda; hybrid = sqrt(N) bounded groups (LPQs) merged to local spill files
//     <dir>/uda.<task>.lpq-NNN, then a final merge (RPQ) over those files
//   * MergeQueue / PriorityQueue (src/Merger/MergeQueue.h:126-427): binary min-heap on the current key
//   * Segment double buffering, straddling-record join, next-chunk prefetch (src/Merger/StreamRW.cc:240-662)
//   * block framing of compressed segments and the decompress thread
//     (src/Merger/DecompressorWrapper.cc:27-383; Snappy/LZO via dlopen, SnappyDecompressor.cc, LzoDecompressor.cc)
// The data structures are new: a segment is a pull-based byte source (network double buffer, file, or a
// decompressing adaptor) under a record reader that joins straddling records in a carry buffer.
#pragma once
#include <atomic>
#include <deque>
#include <condition_variable>
#include <functional>
#include <list>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "compare.h"
#include "protocol.h"
#include "queues.h"
#include "transport.h"
#include "uda_rt.h"

namespace uda {

constexpr size_t kStagingBytes = 1u << 20;  // NETLEV_KV_POOL_EXPO 20: merged blocks handed to the host
constexpr int kExtraBuffers = 10;           // EXTRA_RDMA_BUFFERS (src/Merger/reducer.cc:50)
constexpr int kProgressEvery = 20;          // fetchOverMessage cadence (src/Merger/MergeManager.cc:124-130)

// ---- byte sources -------------------------------------------------------------------------------------
struct HostSegment;   // device_merge.h
long long fetch_stall_us();   // process-wide: microseconds merge threads waited for chunks that had not landed yet
long long fetch_stalls();

class ByteSource {
public:
    virtual ~ByteSource() = default;
    // Next block of the logical stream; the previous block becomes invalid. false = end of stream.
    // Throws UdaError when the underlying fetch failed.
    virtual bool next_block(const uint8_t** data, size_t* len) = 0;
    // Total bytes this source will deliver when that is known after the first block (-1 otherwise)
    virtual long long size_hint() const { return -1; }
};

struct BufferPair {
    uint8_t* buf[2];
    size_t size[2];
};

// (num_kv_bufs + 10) pairs carved from one region (the reference registers it as one MR)
class BufferPool {
public:
    BufferPool(int pairs, size_t size0, size_t size1);
    ~BufferPool();
    BufferPair borrow();  // blocks until a pair is free
    void give_back(const BufferPair& p);
    int pairs() const { return pairs_; }
    size_t total_bytes() const { return total_; }

private:
    uint8_t* region_ = nullptr;
    size_t total_ = 0;
    int pairs_;
    std::vector<BufferPair> free_;
    std::mutex mu_;
    std::condition_variable cv_;
};

enum class Codec { NONE, SNAPPY, LZO };

// Decompress one Hadoop codec block; returns false on corrupt input. Uses the system library when it
// can be dlopen()ed (like the reference) and a built-in decoder otherwise (Snappy only).
bool codec_decompress(Codec c, const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len, const HostCallbacks* host);
Codec codec_for_class(const std::string& java_class);  // "null" -> NONE; unknown -> throws (fallback contract)
bool snappy_uncompress_builtin(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len);
size_t snappy_compress(const uint8_t* in, size_t n, std::vector<uint8_t>* out);            // greedy LZ77 matcher, real copies
size_t snappy_compress_literal(const uint8_t* in, size_t n, std::vector<uint8_t>* out);  // test helper: valid, uncompressed stream
// LZO1X (lzo1x.cc): built-in block decoder used when liblzo2 cannot be loaded, and a greedy compressor for tests / tools
bool lzo1x_decompress_builtin(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len);
bool lzo1x_decompress_lanes(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len, unsigned lanes);
size_t lzo1x_compress(const uint8_t* in, size_t n, std::vector<uint8_t>* out);

// One map output partition being fetched chunk by chunk into a borrowed buffer pair.
class FetchSource : public ByteSource {
public:
    FetchSource(InputClient* client, BufferPool* pool, std::string host, FetchRequest first, size_t chunk_size);
    ~FetchSource() override;
    void start();                       // issue the first fetch
    bool wait_first_chunk();            // true when the first chunk has landed (false: failed)
    bool next_block(const uint8_t** data, size_t* len) override;
    long long raw_length() const { return raw_len_.load(std::memory_order_relaxed); }
    long long part_length() const { return part_len_.load(std::memory_order_relaxed); }
    long long size_hint() const override { return part_length(); }
    const std::string& map_id() const { return req_.map_id; }

private:
    enum State { EMPTY, FETCHING, READY, BUSY };
    FetchRequest prepare(int slot);
    void send(int slot, const FetchRequest& r);
    InputClient* client_;
    BufferPool* pool_;
    std::string host_;
    FetchRequest req_;
    size_t chunk_;
    BufferPair pair_{};
    bool have_pair_ = false;
    State state_[2] = {EMPTY, EMPTY};
    size_t filled_[2] = {0, 0};
    long long fetched_ = 0, requested_ = 0;
    // written by the completion of every chunk (under mu_, always the same values after the first ack), read by the
    // merge thread through the getters without the lock
    std::atomic<long long> part_len_{-1}, raw_len_{-1};
    int next_slot_ = 0;   // slot the consumer reads next
    int busy_slot_ = -1;
    bool failed_ = false;
    std::mutex mu_;
    std::condition_variable cv_;
};

// Hadoop block-compressed stream -> plain bytes: [u32 BE rawLen][u32 BE compLen][compLen bytes]...
// The reduce task's decompress thread (reference: DecompressorWrapper's thread, src/Merger/DecompressorWrapper.cc:168-197):
// every compressed segment keeps its next block being decoded here while the merge consumes the current one.
class DecompressWorker {
public:
    DecompressWorker();
    ~DecompressWorker();   // finishes the queued jobs, then joins
    void submit(std::function<void()> job);

private:
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<std::function<void()>> q_;
    bool stop_ = false;
    std::thread th_;
};

class DecompressingSource : public ByteSource {
public:
    // worker == nullptr: every block is decoded inside next_block() on the caller's thread
    DecompressingSource(std::unique_ptr<ByteSource> inner, Codec codec, long long raw_total, const HostCallbacks* host,
                        DecompressWorker* worker = nullptr);
    ~DecompressingSource() override;
    bool next_block(const uint8_t** data, size_t* len) override;
    long long size_hint() const override {   // rawLength of the index record, learnt by the inner fetch from the first ack
        auto* fs = dynamic_cast<const FetchSource*>(inner_.get());
        return fs ? fs->raw_length() : -1;
    }

private:
    bool read_exact(uint8_t* dst, size_t n);
    std::unique_ptr<ByteSource> inner_;
    Codec codec_;
    long long raw_left_;
    const HostCallbacks* host_;
    const uint8_t* cur_ = nullptr;
    size_t cur_len_ = 0;
    std::vector<uint8_t> comp_, plain_;
    // decode-ahead (worker_ != nullptr): two plain buffers; the inner source, comp_, cur_ are touched by one job at a time
    enum SlotState { EMPTY, PENDING, READY, END, ERROR };
    bool decode_into(std::vector<uint8_t>* plain);   // one block: false at the clean end, throws on corruption
    void schedule(int slot);                         // caller holds mu_
    DecompressWorker* worker_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::vector<uint8_t> ahead_[2];
    SlotState state_[2] = {EMPTY, EMPTY};
    std::string error_;
    int next_ = 0;          // slot the consumer takes next
    bool started_ = false;
};

// Spill file written by an LPQ merge (plain merged-stream framing, EOF marker at the end)
class FileSource : public ByteSource {
public:
    FileSource(std::string path, bool delete_on_close);
    ~FileSource() override;
    bool next_block(const uint8_t** data, size_t* len) override;

private:
    std::string path_;
    bool del_;
    FILE* f_ = nullptr;
    std::vector<uint8_t> buf_;
};

// ---- records ------------------------------------------------------------------------------------------
struct Record {
    const uint8_t* key = nullptr; int klen = 0;
    const uint8_t* val = nullptr; int vlen = 0;
    // the record as it lies in its source: VInt(klen) VInt(vlen) key value, `size` bytes from `raw` (the merged stream
    // carries exactly these bytes, so emitting a record is one copy)
    const uint8_t* raw = nullptr; size_t size = 0;
};

// Iterates the records of one source; a record that straddles two blocks is joined in a carry buffer.
class Segment {
public:
    Segment(std::unique_ptr<ByteSource> src, int id) : src_(std::move(src)), id_(id) {}
    bool next();                      // advance; false at the EOF marker / end of stream
    const Record& cur() const { return rec_; }
    int id() const { return id_; }
    ByteSource* source() { return src_.get(); }

private:
    bool refill();
    std::unique_ptr<ByteSource> src_;
    int id_;
    const uint8_t* p_ = nullptr;
    const uint8_t* end_ = nullptr;
    std::vector<uint8_t> carry_;
    Record rec_;
    bool done_ = false;
    bool seen_bytes_ = false;   // any byte of this segment was ever received
};

// Binary min-heap of segments ordered by (current key, segment id)
class MergeHeap {
public:
    explicit MergeHeap(KeyKind kind) : kind_(kind) {}
    void insert(std::unique_ptr<Segment> s);   // primes the segment; empty segments are dropped
    bool empty() const { return heap_.empty(); }
    size_t size() const { return heap_.size(); }
    const Record& top() const { return heap_.front().seg->cur(); }
    void advance();                            // pop-or-sift after the top record has been consumed

private:
    // The first 8 compared key bytes (big-endian, zero padded) ride in the heap array: most comparisons of a sift are
    // decided by two integers that are already in cache instead of two pointer chases and a memcmp. Equal prefixes fall
    // through to the exact comparator (a shorter key that is a prefix of a longer one has the smaller or equal padded
    // prefix, so the integer order never contradicts the comparator's).
    struct Node {
        uint64_t prefix;
        std::unique_ptr<Segment> seg;
    };
    uint64_t prefix_of(const Record& r) const;
    bool less(const Node& a, const Node& b) const;
    void sift_down(size_t i);
    void sift_up(size_t i);
    KeyKind kind_;
    std::vector<Node> heap_;
};

// ---- the reduce task ----------------------------------------------------------------------------------
struct MemoryPlan {
    int num_kv_bufs = 0;
    size_t buffer_size = 0;    // page aligned chunk size used for fetches
    size_t pair_size[2] = {0, 0};
    int pairs = 0;
};
// throws UdaError("Not enough memory for rdma buffers") / ("RDMA Buffer is too small")
MemoryPlan plan_memory(const InitParams& init, int num_kv_bufs, Codec codec, double compression_ratio, long long max_buf_cfg);

class Consumer {
public:
    Consumer(const Options& opt, const HostCallbacks* host);
    ~Consumer();
    void start();                                  // MergeManager_main
    void handle_command(const std::string& cmd);   // reduce_downcall_handler
    void exit_msg();                               // reduce_exit_msg_handler -> finalize
    // statistics for tests
    uint64_t records_emitted() const { return emitted_.load(); }
    int lpqs_spilled() const { return spilled_.load(); }
    int device_merges() const { return device_merges_.load(); }
    int device_groups() const { return device_groups_.load(); }
    int fetch_ms() const { return fetch_ms_.load(); }
    int device_ms() const { return device_ms_.load(); }

private:
    struct FetchOrder { std::string host, job, map; int reduce; };
    void on_init(const HadoopCmd& cmd);
    void merge_thread_main();
    void merge_online();
    void merge_hybrid();
    void merge_device();   // approach 3: fetch to host memory, merge on the GPU (device_merge.h), CPU heap merge as fallback
    // a partition that does not fit the device's free memory: groups of segments are merged on the device into host runs
    // (the hybrid merge's LPQs, src/Merger/MergeManager.cc:202-288, with the device as the LPQ merger), a heap merges the runs
    bool merge_device_grouped(std::vector<HostSegment>* segs, const std::function<void(const uint8_t*, size_t)>& deliver);
    // fetch `count` map outputs into `heap`; reports progress every kProgressEvery segments
    void fetching_phase(MergeHeap* heap, int count, int* progress);
    void merging_phase(MergeHeap* heap, const std::function<void(const uint8_t*, size_t)>& sink, bool final_stream);
    std::unique_ptr<ByteSource> make_source(const FetchOrder& o);
    void fail(const UdaError& e);

    Options opt_;
    const HostCallbacks* host_;
    InitParams init_;
    KeyKind key_kind_ = KeyKind::RAW_BYTES;
    Codec codec_ = Codec::NONE;
    MemoryPlan plan_;
    std::unique_ptr<InputClient> client_;
    std::unique_ptr<DecompressWorker> decomp_;   // only for compressed map outputs; outlives every source (reset after the merge thread)
    std::unique_ptr<BufferPool> pool_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::list<FetchOrder> fetch_list_;
    std::atomic<bool> stop_{false}, failed_{false}, initialised_{false};
    std::thread merge_thread_, warm_thread_;
    std::vector<uint8_t> staging_[2];
    std::atomic<uint64_t> emitted_{0};
    std::atomic<int> spilled_{0};
    std::atomic<int> device_merges_{0}, device_groups_{0}, fetch_ms_{0}, device_ms_{0};
    int segment_seq_ = 0;
    long long fetch_wait_us_ = 0, fetch_copy_us_ = 0;   // merge_device(): time blocked in next_block / appending to host memory
};

}  // namespace uda
