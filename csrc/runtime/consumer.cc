#include "consumer.h"
#include "threads.h"
#include "device_merge.h"

#include <dlfcn.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <random>

namespace uda {

// ========================================================================================== buffers
BufferPool::BufferPool(int pairs, size_t s0, size_t s1) : pairs_(pairs) {
    const size_t page = (size_t)getpagesize();
    auto up = [&](size_t x) { return (x + page - 1) / page * page; };
    const size_t a0 = up(s0), a1 = up(s1);
    total_ = (a0 + a1) * (size_t)pairs;
    void* p = nullptr;
    if (posix_memalign(&p, page, total_ ? total_ : page) != 0) throw UdaError("failed to allocate the shuffle buffer region");
    region_ = (uint8_t*)p;
    for (int i = 0; i < pairs; ++i) {
        BufferPair bp;
        bp.buf[0] = region_ + (a0 + a1) * (size_t)i;
        bp.buf[1] = bp.buf[0] + a0;
        bp.size[0] = s0;
        bp.size[1] = s1;
        free_.push_back(bp);
    }
}
BufferPool::~BufferPool() { free(region_); }
BufferPair BufferPool::borrow() {
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [&] { return !free_.empty(); });
    BufferPair p = free_.back();
    free_.pop_back();
    return p;
}
void BufferPool::give_back(const BufferPair& p) {
    {
        std::lock_guard<std::mutex> lk(mu_);
        free_.push_back(p);
    }
    cv_.notify_one();
}

// ========================================================================================== codecs
Codec codec_for_class(const std::string& c) {
    if (c.empty() || c == "null") return Codec::NONE;
    if (c.find("SnappyCodec") != std::string::npos) return Codec::SNAPPY;
    if (c.find("lzo.LzoCodec") != std::string::npos || c.find("LzoCodec") != std::string::npos) return Codec::LZO;
    UDA_LOG(lsERROR, "unsupported compression codec '%s'", c.c_str());
    throw UdaError("unsupported compression codec");
}

// Raw Snappy format: varint uncompressed length, then literal / copy elements.
bool snappy_uncompress_builtin(const uint8_t* in, size_t n, uint8_t* out, size_t out_len) {
    size_t ip = 0, op = 0;
    uint64_t ulen = 0;
    int shift = 0;
    for (;;) {
        if (ip >= n || shift > 35) return false;
        uint8_t b = in[ip++];
        ulen |= (uint64_t)(b & 0x7F) << shift;
        if (!(b & 0x80)) break;
        shift += 7;
    }
    if (ulen != out_len) return false;
    while (ip < n) {
        const uint8_t tag = in[ip++];
        size_t len, off;
        switch (tag & 3) {
            case 0: {  // literal
                len = (tag >> 2) + 1;
                if (len > 60) {
                    const size_t nb = len - 60;
                    if (ip + nb > n) return false;
                    len = 0;
                    for (size_t i = 0; i < nb; ++i) len |= (size_t)in[ip + i] << (8 * i);
                    len += 1;
                    ip += nb;
                }
                if (ip + len > n || op + len > out_len) return false;
                memcpy(out + op, in + ip, len);
                ip += len;
                op += len;
                continue;
            }
            case 1:
                if (ip + 1 > n) return false;
                len = ((tag >> 2) & 7) + 4;
                off = ((size_t)(tag >> 5) << 8) | in[ip];
                ip += 1;
                break;
            case 2:
                if (ip + 2 > n) return false;
                len = (tag >> 2) + 1;
                off = in[ip] | ((size_t)in[ip + 1] << 8);
                ip += 2;
                break;
            default:
                if (ip + 4 > n) return false;
                len = (tag >> 2) + 1;
                off = in[ip] | ((size_t)in[ip + 1] << 8) | ((size_t)in[ip + 2] << 16) | ((size_t)in[ip + 3] << 24);
                ip += 4;
                break;
        }
        if (off == 0 || off > op || op + len > out_len) return false;
        for (size_t i = 0; i < len; ++i) out[op + i] = out[op + i - off];  // overlapping copies are byte-serial by definition
        op += len;
    }
    return op == out_len;
}

// Greedy single-pass Snappy compressor (hash of 4 bytes -> last position, copies with 2-byte offsets). The
// reference only decompresses (the map side compresses inside Hadoop); this exists so that tests and benchmarks can
// produce real compressed segments without a libsnappy on the box.
namespace {
inline uint32_t load32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
void emit_literal(const uint8_t* p, size_t len, std::vector<uint8_t>* out) {
    while (len) {
        const size_t l = std::min<size_t>(len, 65536);
        if (l <= 60) out->push_back((uint8_t)((l - 1) << 2));
        else if (l <= 256) { out->push_back(60 << 2); out->push_back((uint8_t)(l - 1)); }
        else { out->push_back(61 << 2); out->push_back((uint8_t)((l - 1) & 0xFF)); out->push_back((uint8_t)((l - 1) >> 8)); }
        out->insert(out->end(), p, p + l);
        p += l; len -= l;
    }
}
void emit_copy(size_t off, size_t len, std::vector<uint8_t>* out) {
    while (len) {
        size_t l = std::min<size_t>(len, 64);
        if (len - l > 0 && len - l < 4) l = len - 4;             // never leave a tail shorter than a copy may be
        if (l >= 4 && l <= 11 && off < 2048) {                    // 1-byte-offset form
            out->push_back((uint8_t)(1 | ((l - 4) << 2) | ((off >> 8) << 5)));
            out->push_back((uint8_t)(off & 0xFF));
        } else {
            out->push_back((uint8_t)(2 | ((l - 1) << 2)));
            out->push_back((uint8_t)(off & 0xFF));
            out->push_back((uint8_t)(off >> 8));
        }
        len -= l;
    }
}
}  // namespace

size_t snappy_compress(const uint8_t* in, size_t n, std::vector<uint8_t>* out) {
    out->clear();
    uint64_t v = n;
    do {
        uint8_t b = v & 0x7F;
        v >>= 7;
        out->push_back(b | (v ? 0x80 : 0));
    } while (v);
    constexpr int kBits = 14;
    std::vector<uint32_t> table(1u << kBits, 0xFFFFFFFFu);
    size_t ip = 0, lit = 0;
    while (ip + 4 <= n) {
        const uint32_t h = (load32(in + ip) * 0x1E35A7BDu) >> (32 - kBits);
        const uint32_t cand = table[h];
        table[h] = (uint32_t)ip;
        if (cand != 0xFFFFFFFFu && ip - cand <= 65535 && load32(in + cand) == load32(in + ip)) {
            size_t len = 4;
            while (ip + len < n && in[cand + len] == in[ip + len]) ++len;
            emit_literal(in + lit, ip - lit, out);
            emit_copy(ip - cand, len, out);
            ip += len;
            lit = ip;
        } else {
            ++ip;
        }
    }
    emit_literal(in + lit, n - lit, out);
    return out->size();
}

size_t snappy_compress_literal(const uint8_t* in, size_t n, std::vector<uint8_t>* out) {
    out->clear();
    uint64_t v = n;
    do {
        uint8_t b = v & 0x7F;
        v >>= 7;
        out->push_back(b | (v ? 0x80 : 0));
    } while (v);
    size_t pos = 0;
    while (pos < n) {
        size_t len = std::min<size_t>(n - pos, 65536);
        if (len <= 60) out->push_back((uint8_t)((len - 1) << 2));
        else if (len <= 256) { out->push_back(60 << 2); out->push_back((uint8_t)(len - 1)); }
        else { out->push_back(61 << 2); out->push_back((uint8_t)((len - 1) & 0xFF)); out->push_back((uint8_t)((len - 1) >> 8)); }
        out->insert(out->end(), in + pos, in + pos + len);
        pos += len;
    }
    return out->size();
}

namespace {
typedef int (*snappy_uncompress_fn)(const char*, size_t, char*, size_t*);
typedef int (*lzo_decompress_fn)(const unsigned char*, unsigned long, unsigned char*, unsigned long*, void*);
typedef int (*lzo_init_fn)(unsigned, int, int, int, int, int, int, int, int, int);

snappy_uncompress_fn load_snappy() {
    static snappy_uncompress_fn fn = [] {
        for (const char* name : {"libsnappy.so.1", "libsnappy.so"}) {
            if (void* h = dlopen(name, RTLD_LAZY)) {
                if (void* s = dlsym(h, "snappy_uncompress")) return (snappy_uncompress_fn)s;
            }
        }
        return (snappy_uncompress_fn) nullptr;
    }();
    return fn;
}

lzo_decompress_fn resolve_lzo(const HostCallbacks* host) {
    lzo_decompress_fn fn = nullptr;
    void* h = nullptr;
    for (const char* name : {"liblzo2.so.2", "liblzo2.so"}) if ((h = dlopen(name, RTLD_LAZY))) break;
    if (!h) return nullptr;
    // decompressor chosen by io.compression.codec.lzo.decompressor (default LZO1X), safe variant
    std::string which = conf_get(host, "io.compression.codec.lzo.decompressor", "LZO1X");
    std::string sym = "lzo1x_decompress_safe";
    if (which.size() >= 5 && (which[3] == '1' || which[3] == '2')) {
        std::string base = which.substr(0, 5);
        std::transform(base.begin(), base.end(), base.begin(), ::tolower);
        sym = base + "_decompress" + (which.find("SAFE") != std::string::npos || which == which.substr(0, 5) ? "_safe" : "");
    }
    fn = (lzo_decompress_fn)dlsym(h, sym.c_str());
    if (!fn) fn = (lzo_decompress_fn)dlsym(h, "lzo1x_decompress_safe");
    return fn;
}
// resolved once per process by whichever reduce task decompresses first (several merge / decompress threads may race here:
// the initialisation of a function-local static is the synchronisation)
lzo_decompress_fn load_lzo(const HostCallbacks* host) {
    static const lzo_decompress_fn fn = resolve_lzo(host);
    return fn;
}
}  // namespace

bool codec_decompress(Codec c, const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len, const HostCallbacks* host) {
    if (c == Codec::SNAPPY) {
        if (auto fn = load_snappy()) {
            size_t ol = out_len;
            return fn((const char*)in, in_len, (char*)out, &ol) == 0 && ol == out_len;
        }
        return snappy_uncompress_builtin(in, in_len, out, out_len);
    }
    if (c == Codec::LZO) {
        if (auto fn = load_lzo(host)) {
            unsigned long ol = out_len;
            return fn(in, in_len, out, &ol, nullptr) == 0 && ol == out_len;
        }
        // no liblzo2 on this machine: the built-in decoder handles the LZO1X family's stream format (LZO1X-1, -999)
        std::string which = conf_get(host, "io.compression.codec.lzo.decompressor", "LZO1X");
        if (which.compare(0, 5, "LZO1X") != 0)
            throw UdaError("LZO decompressor " + which + " needs liblzo2, which cannot be loaded (built in: LZO1X only)");
        return lzo1x_decompress_builtin(in, in_len, out, out_len);
    }
    if (in_len != out_len) return false;
    memcpy(out, in, in_len);
    return true;
}

// ========================================================================================== sources
// process-wide: time merge threads spent waiting for a chunk that had not landed yet (the prefetch did not hide the fetch)
static std::atomic<long long> g_fetch_stall_us{0}, g_fetch_stalls{0};
long long fetch_stall_us() { return g_fetch_stall_us.load(); }
long long fetch_stalls() { return g_fetch_stalls.load(); }

FetchSource::FetchSource(InputClient* client, BufferPool* pool, std::string host, FetchRequest first, size_t chunk)
    : client_(client), pool_(pool), host_(std::move(host)), req_(std::move(first)), chunk_(chunk) {}

FetchSource::~FetchSource() {
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [&] { return state_[0] != FETCHING && state_[1] != FETCHING; });  // never free a buffer a fetch may still write
    if (have_pair_) pool_->give_back(pair_);
}

void FetchSource::start() {
    pair_ = pool_->borrow();
    have_pair_ = true;
    FetchRequest r;
    {
        std::lock_guard<std::mutex> lk(mu_);
        r = prepare(0);
    }
    send(0, r);
}

// caller holds mu_: marks the slot as being fetched and returns the request to send once the lock is dropped
FetchRequest FetchSource::prepare(int slot) {
    state_[slot] = FETCHING;
    FetchRequest r = req_;
    r.map_offset = fetched_;
    r.chunk_size = (int)std::min(chunk_, pair_.size[slot]);
    return r;
}

// caller must NOT hold mu_: a failing connection runs completion callbacks synchronously
void FetchSource::send(int slot, const FetchRequest& r) {
    try {
        client_->fetch(host_, r, pair_.buf[slot], pair_.size[slot], [this, slot](bool ok, const FetchAck& ack) {
            bool more = false;
            FetchRequest next;
            const int other = slot ^ 1;
            {
                std::lock_guard<std::mutex> lk(mu_);
                // whatever transport delivered it, an ack is only believed when its numbers are possible: a chunk fits its
                // buffer, lengths are not negative, the partition does not shrink below what has been fetched
                const bool sane = ack.sent >= 0 && (size_t)ack.sent <= pair_.size[slot] && ack.part_length >= 0 && ack.raw_length >= 0 &&
                                  fetched_ + ack.sent <= ack.part_length;
                if (ok && !sane && ack.path != kPathTooLong && ack.path != kFetchRejected)
                    UDA_LOG(lsERROR, "impossible fetch ack for map %s: sent=%lld part=%lld raw=%lld after %lld bytes", req_.map_id.c_str(), ack.sent,
                            ack.part_length, ack.raw_length, fetched_);
                if (!ok || !sane || ack.path == kPathTooLong || ack.path == kFetchRejected) {
                    if (ok && ack.path == kPathTooLong) UDA_LOG(lsERROR, "Mof path is too long, UDA supports path of max %zu chars", kMaxMofPath);
                    else if (ok) UDA_LOG(lsERROR, "the provider rejected the fetch of map %s", req_.map_id.c_str());
                    failed_ = true;
                    state_[slot] = EMPTY;
                } else {
                    filled_[slot] = (size_t)ack.sent;
                    fetched_ += ack.sent;
                    part_len_ = ack.part_length;
                    raw_len_ = ack.raw_length;
                    req_.path = ack.path;  // later requests carry the resolved index record
                    req_.file_offset = ack.file_offset;
                    req_.raw_length = ack.raw_length;
                    req_.part_length = ack.part_length;
                    state_[slot] = READY;
                    // prefetch the next chunk into the other buffer
                    if (state_[other] == EMPTY && fetched_ < part_len_ && ack.sent > 0) { next = prepare(other); more = true; }
                }
                // notify under the lock: the destructor waits on cv_ for "no fetch in flight" and frees this object as
                // soon as it can re-acquire mu_, so nothing of *this may be touched after the unlock unless a new fetch
                // (more == true, state FETCHING) keeps the object alive
                cv_.notify_all();
            }
            if (more) send(other, next);
        });
    } catch (const UdaError& e) {
        UDA_LOG(lsERROR, "fetch of map %s failed: %s", req_.map_id.c_str(), e.info().c_str());
        {
            std::lock_guard<std::mutex> lk(mu_);
            failed_ = true;
            state_[slot] = EMPTY;
            cv_.notify_all();  // under the lock, see above
        }
    }
}

bool FetchSource::wait_first_chunk() {
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [&] { return failed_ || state_[0] == READY || state_[0] == BUSY; });
    return !failed_;
}

bool FetchSource::next_block(const uint8_t** data, size_t* len) {
    std::unique_lock<std::mutex> lk(mu_);
    if (busy_slot_ >= 0) {  // the consumer is done with the previous block
        state_[busy_slot_] = EMPTY;
        const int freed = busy_slot_;
        busy_slot_ = -1;
        const bool in_flight = state_[freed ^ 1] == FETCHING;
        if (!failed_ && !in_flight && part_len_ >= 0 && fetched_ < part_len_) {
            FetchRequest r = prepare(freed);
            lk.unlock();
            send(freed, r);
            lk.lock();
        }
    }
    for (;;) {
        if (failed_) throw UdaError("fetching a map output failed");
        if (state_[next_slot_] == READY) break;
        const bool more = state_[0] == FETCHING || state_[1] == FETCHING;
        if (!more) return false;  // everything fetched and consumed
        const auto t0 = std::chrono::steady_clock::now();
        cv_.wait(lk);
        g_fetch_stall_us += std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count();
        ++g_fetch_stalls;
    }
    const int s = next_slot_;
    state_[s] = BUSY;
    busy_slot_ = s;
    next_slot_ ^= 1;
    *data = pair_.buf[s];
    *len = filled_[s];
    if (*len == 0) return false;
    return true;
}

DecompressWorker::DecompressWorker() {
    th_ = std::thread([this] {
        for (;;) {
            std::function<void()> job;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return stop_ || !q_.empty(); });
                if (q_.empty()) return;   // stop_ and nothing left
                job = std::move(q_.front());
                q_.pop_front();
            }
            job();   // jobs catch their own errors and park them in their source
        }
    });
}
DecompressWorker::~DecompressWorker() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        stop_ = true;
    }
    cv_.notify_all();
    if (th_.joinable()) th_.join();
}
void DecompressWorker::submit(std::function<void()> job) {
    {
        std::lock_guard<std::mutex> lk(mu_);
        q_.push_back(std::move(job));
    }
    cv_.notify_one();
}

DecompressingSource::DecompressingSource(std::unique_ptr<ByteSource> inner, Codec codec, long long raw_total, const HostCallbacks* host,
                                         DecompressWorker* worker)
    : inner_(std::move(inner)), codec_(codec), raw_left_(raw_total), host_(host), worker_(worker) {}

DecompressingSource::~DecompressingSource() {
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [&] { return state_[0] != PENDING && state_[1] != PENDING; });   // a job still holds `this`
}

bool DecompressingSource::read_exact(uint8_t* dst, size_t n) {
    while (n) {
        if (cur_len_ == 0 && !inner_->next_block(&cur_, &cur_len_)) return false;
        size_t k = std::min(n, cur_len_);
        memcpy(dst, cur_, k);
        dst += k; n -= k; cur_ += k; cur_len_ -= k;
    }
    return true;
}

bool DecompressingSource::decode_into(std::vector<uint8_t>* plain) {
    uint8_t hdr[8];
    if (!read_exact(hdr, 8)) return false;  // clean end (or only the trailing IFile checksum was left)
    const uint32_t raw = (uint32_t)hdr[0] << 24 | (uint32_t)hdr[1] << 16 | (uint32_t)hdr[2] << 8 | hdr[3];
    const uint32_t comp = (uint32_t)hdr[4] << 24 | (uint32_t)hdr[5] << 16 | (uint32_t)hdr[6] << 8 | hdr[7];
    if (raw == 0 && comp == 0) return false;
    if (raw > (1u << 30) || comp > (1u << 30)) throw UdaError("corrupt compressed block header");
    comp_.resize(comp);
    if (!read_exact(comp_.data(), comp)) throw UdaError("compressed block is truncated");
    plain->resize(raw);
    if (!codec_decompress(codec_, comp_.data(), comp, plain->data(), raw, host_)) throw UdaError("decompression failed");
    return true;
}

void DecompressingSource::schedule(int slot) {
    state_[slot] = PENDING;
    worker_->submit([this, slot] {
        SlotState result = READY;
        std::string err;
        try {
            if (!decode_into(&ahead_[slot])) result = END;
        } catch (const UdaError& e) {
            result = ERROR;
            err = e.info();
        } catch (const std::exception& e) {
            result = ERROR;
            err = e.what();
        }
        std::lock_guard<std::mutex> lk(mu_);
        if (result == ERROR) error_ = err;
        state_[slot] = result;
        cv_.notify_all();   // under the lock: the destructor may free this object as soon as it sees no PENDING slot
    });
}

bool DecompressingSource::next_block(const uint8_t** data, size_t* len) {
    if (!worker_) {
        if (!decode_into(&plain_)) return false;
        *data = plain_.data();
        *len = plain_.size();
        return true;
    }
    std::unique_lock<std::mutex> lk(mu_);
    if (!started_) {
        started_ = true;
        schedule(0);
    }
    const int s = next_;
    cv_.wait(lk, [&] { return state_[s] != PENDING; });
    if (state_[s] == ERROR) throw UdaError(error_);
    if (state_[s] == END || state_[s] == EMPTY) return false;
    // the block the caller held until now (the other slot) is free again: decode the one after this into it
    state_[s ^ 1] = EMPTY;
    schedule(s ^ 1);
    state_[s] = EMPTY;      // READY consumed; the buffer stays untouched until the caller comes back (only s ^ 1 is being written)
    next_ = s ^ 1;
    *data = ahead_[s].data();
    *len = ahead_[s].size();
    return true;
}

FileSource::FileSource(std::string path, bool del) : path_(std::move(path)), del_(del) {
    f_ = fopen(path_.c_str(), "rb");
    if (!f_) throw UdaError("cannot open spill file " + path_);
    buf_.resize(kStagingBytes);
}
FileSource::~FileSource() {
    if (f_) fclose(f_);
    if (del_) ::unlink(path_.c_str());
}
bool FileSource::next_block(const uint8_t** data, size_t* len) {
    size_t n = fread(buf_.data(), 1, buf_.size(), f_);
    if (n == 0) return false;
    *data = buf_.data();
    *len = n;
    return true;
}

// ========================================================================================== records
// Hadoop serialises key and value lengths as ints: a longer length in a header is corruption (or an attack), and the
// arithmetic below (header + key + value as one size) must never see two lengths whose sum wraps.
static constexpr int64_t kMaxRecordPart = 0x7FFFFFFF;

bool Segment::next() {
    if (done_) return false;
    for (;;) {
        if (p_ != nullptr && p_ < end_) {
            int64_t kl = 0, vl = 0;
            int n1 = vint_decode(p_, end_, &kl);
            int n2 = n1 ? vint_decode(p_ + n1, end_, &vl) : 0;
            if (n1 && n2) {
                if (kl == -1 && vl == -1) { done_ = true; return false; }
                if (kl < 0 || vl < 0 || kl > kMaxRecordPart || vl > kMaxRecordPart) throw UdaError("corrupt IFile record header");
                const size_t total = (size_t)n1 + n2 + (size_t)kl + (size_t)vl;
                if ((size_t)(end_ - p_) >= total) {
                    rec_.key = p_ + n1 + n2; rec_.klen = (int)kl;
                    rec_.val = rec_.key + kl; rec_.vlen = (int)vl;
                    rec_.raw = p_; rec_.size = total;
                    p_ += total;
                    return true;
                }
            }
            carry_.assign(p_, end_);  // the record straddles the end of this block
        } else {
            carry_.clear();
        }
        // pull blocks until a whole record sits in carry_, or a fresh block can be parsed in place
        for (;;) {
            const uint8_t* d;
            size_t len;
            if (!src_->next_block(&d, &len)) {
                // every segment, fetched or spilled, closes with the EOF marker: a stream that just stops -- even at a
                // record boundary -- has lost its tail, and delivering the shorter stream would lose records silently
                // (a partition of zero bytes -- an index record with partLength 0 -- is simply empty)
                if (carry_.empty() && !seen_bytes_) { done_ = true; return false; }
                if (carry_.empty()) throw UdaError("segment ended without its EOF marker");
                throw UdaError("segment ended inside a record");
            }
            seen_bytes_ = seen_bytes_ || len > 0;
            if (carry_.empty()) { p_ = d; end_ = d + len; break; }
            SplitCursor c{carry_.data(), carry_.size(), d, len};
            int64_t kl = 0, vl = 0;
            int n1 = c.read_vint(&kl);
            int n2 = n1 ? c.read_vint(&vl) : 0;
            if (!n1 || !n2) { carry_.insert(carry_.end(), d, d + len); continue; }
            if (kl == -1 && vl == -1) { done_ = true; return false; }
            if (kl < 0 || vl < 0 || kl > kMaxRecordPart || vl > kMaxRecordPart) throw UdaError("corrupt IFile record header");
            const size_t total = (size_t)n1 + n2 + (size_t)kl + (size_t)vl;
            if (carry_.size() + len < total) { carry_.insert(carry_.end(), d, d + len); continue; }
            const size_t take = total > carry_.size() ? total - carry_.size() : 0;
            carry_.insert(carry_.end(), d, d + take);
            p_ = d + take;
            end_ = d + len;
            const uint8_t* b = carry_.data();
            rec_.key = b + n1 + n2; rec_.klen = (int)kl;
            rec_.val = rec_.key + kl; rec_.vlen = (int)vl;
            rec_.raw = b; rec_.size = total;
            return true;
        }
    }
}

uint64_t MergeHeap::prefix_of(const Record& r) const {
    const uint8_t* body;
    int n;
    key_body(kind_, r.key, r.klen, &body, &n);
    uint64_t v = 0;
    if (n >= 8) {
        memcpy(&v, body, 8);
        return __builtin_bswap64(v);
    }
    for (int i = 0; i < n; ++i) v |= (uint64_t)body[i] << (56 - 8 * i);
    return v;
}
bool MergeHeap::less(const Node& a, const Node& b) const {
    if (a.prefix != b.prefix) return a.prefix < b.prefix;
    const Record &x = a.seg->cur(), &y = b.seg->cur();
    int c = compare_keys(kind_, x.key, x.klen, y.key, y.klen);
    return c ? c < 0 : a.seg->id() < b.seg->id();
}
void MergeHeap::sift_up(size_t i) {
    while (i > 0) {
        size_t p = (i - 1) / 2;
        if (!less(heap_[i], heap_[p])) break;
        std::swap(heap_[i], heap_[p]);
        i = p;
    }
}
void MergeHeap::sift_down(size_t i) {
    // the sinking node travels in a local: children move up into the hole, one store per level instead of a swap
    const size_t n = heap_.size();
    Node x = std::move(heap_[i]);
    for (;;) {
        const size_t l = 2 * i + 1;
        if (l >= n) break;
        const size_t r = l + 1;
        const size_t c = (r < n && less(heap_[r], heap_[l])) ? r : l;
        if (!less(heap_[c], x)) break;
        heap_[i] = std::move(heap_[c]);
        i = c;
    }
    heap_[i] = std::move(x);
}
void MergeHeap::insert(std::unique_ptr<Segment> s) {
    if (!s->next()) return;  // empty partition: nothing to merge, buffers go back to the pool
    const uint64_t pre = prefix_of(s->cur());
    heap_.push_back(Node{pre, std::move(s)});
    sift_up(heap_.size() - 1);
}
void MergeHeap::advance() {
    if (heap_.front().seg->next()) {
        heap_.front().prefix = prefix_of(heap_.front().seg->cur());
        sift_down(0);
    } else {
        std::swap(heap_.front(), heap_.back());
        heap_.pop_back();
        if (!heap_.empty()) sift_down(0);
    }
}

// ========================================================================================== planning
MemoryPlan plan_memory(const InitParams& init, int num_kv_bufs, Codec codec, double ratio, long long max_buf_cfg) {
    MemoryPlan p;
    p.num_kv_bufs = num_kv_bufs;
    long long max_buf = init.max_buffer_bytes;
    const long long min_buf = init.min_buffer_bytes;
    if (init.shuffle_memory < (long long)num_kv_bufs * max_buf * 2) {  // 2: double buffer
        const long long orig = max_buf;
        max_buf = init.shuffle_memory / ((long long)num_kv_bufs * 2);
        if (max_buf < min_buf) {
            UDA_LOG(lsERROR, "Not enough memory for rdma buffers: shuffleMemorySize=%lldB; mapred.rdma.buf.size.min=%lldB",
                    init.shuffle_memory, min_buf);
            throw UdaError("Not enough memory for rdma buffers");
        }
        UDA_LOG(lsWARN, "UDA: using calculated RDMA buffer size=%lldB (not aligned yet) instead of max size=%lldB", max_buf, orig);
    }
    const long long page = getpagesize();
    const long long aligned = max_buf - max_buf % page;
    if (aligned <= 0 || aligned < min_buf) {
        UDA_LOG(lsERROR, "RDMA Buffer is too small: maxRdmaBufferSize=%lldB, pagesize=%lld, aligned_buffer_size=%lldB, min_buffer=%lldB",
                max_buf, page, aligned, min_buf);
        throw UdaError("RDMA Buffer is too small");
    }
    p.buffer_size = (size_t)aligned;
    p.pairs = num_kv_bufs + kExtraBuffers;
    if (codec == Codec::NONE) {
        p.pair_size[0] = p.pair_size[1] = p.buffer_size;
    } else {
        // the pair's 2*buffer_size is split: a small landing buffer for compressed bytes (capped at the
        // configured chunk size) and the rest for decompressed data
        long long total = 2 * aligned;
        long long comp = (long long)(total * ratio);
        comp -= comp % page;
        if (max_buf_cfg > 0 && comp > max_buf_cfg) comp = max_buf_cfg - max_buf_cfg % page;
        if (comp < page) comp = page;
        p.pair_size[0] = p.pair_size[1] = (size_t)comp;  // both slots receive compressed chunks; plain data lives in the adaptor
        p.buffer_size = (size_t)comp;
    }
    return p;
}

// ========================================================================================== consumer
Consumer::Consumer(const Options& opt, const HostCallbacks* host) : opt_(opt), host_(host) {}
Consumer::~Consumer() { exit_msg(); }

void Consumer::start() {
    log_open_netmerger(opt_.log_dir);
    /* the next line is parsed by deployment automation: keep its wording */
    UDA_LOG(lsINFO, "UDA version is %s", VERSION_UDA);
    staging_[0].resize(kStagingBytes);
    staging_[1].resize(kStagingBytes);
}

void Consumer::handle_command(const std::string& s) {
    HadoopCmd cmd;
    if (!parse_hadoop_cmd(s, &cmd)) { UDA_LOG(lsWARN, "Hadoop's command '%s' cannot be parsed", s.c_str()); return; }
    switch (cmd.header) {
        case INIT_MSG: on_init(cmd); break;
        case FETCH_MSG: {
            if (cmd.params.size() < 4) { UDA_LOG(lsERROR, "FETCH needs host, job, map and partition"); return; }
            {
                std::lock_guard<std::mutex> lk(mu_);
                fetch_list_.push_back(FetchOrder{cmd.params[0], cmd.params[1], cmd.params[2], atoi(cmd.params[3].c_str())});
            }
            cv_.notify_all();
            break;
        }
        case FINAL_MSG: UDA_LOG(lsINFO, "final merge requested"); break;
        case EXIT_MSG: exit_msg(); break;
        default: UDA_LOG(lsWARN, "unknown command %d", cmd.header); break;
    }
}

void Consumer::on_init(const HadoopCmd& cmd) {
    if (!parse_init_params(cmd, &init_)) throw UdaError("malformed INIT command");
    UDA_LOG(lsINFO, "got params from java: num_maps=%d, job_id=%s, reduce_task_id=%s, lpq_size=%d, rdma.buf.size=%lldB, minBuffer=%lldB, "
            "cmp_func=%s, comp_alg=%s, comp_block_size=%d, shuffleMemorySize=%lldB", init_.num_maps, init_.job_id.c_str(),
            init_.reduce_task_id.c_str(), init_.lpq_size, init_.max_buffer_bytes, init_.min_buffer_bytes, init_.key_class.c_str(),
            init_.codec_class.c_str(), init_.codec_block_size, init_.shuffle_memory);
    key_kind_ = key_kind_for_class(init_.key_class);  // unsupported -> UdaError -> Java falls back to vanilla
    codec_ = codec_for_class(init_.codec_class);
    int num_kv_bufs = init_.num_maps;
    if (opt_.merge_approach == 2) {
        int lpqs = init_.lpq_size > 0 ? (init_.num_maps + init_.lpq_size - 1) / init_.lpq_size : (int)std::sqrt((double)init_.num_maps);
        if (lpqs < 1) lpqs = 1;
        int parallel = atoi(conf_get(host_, "mapred.rdma.num.parallel.lpqs", "0").c_str());
        if (parallel < 3) parallel = 3;
        int per = (init_.num_maps + lpqs - 1) / lpqs;
        if (init_.num_maps >= lpqs && lpqs > 1) num_kv_bufs = std::min(init_.num_maps, per * parallel);
    }
    if (opt_.merge_approach == 3) num_kv_bufs = std::min(init_.num_maps, 8);   // the device merge drains 8 sources at a time
    if (num_kv_bufs < 1) num_kv_bufs = 1;
    double ratio = atof(conf_get(host_, "mapred.rdma.compression.buffer.ratio", "0.20").c_str());
    long long cfg_kb = atoll(conf_get(host_, "mapred.rdma.buf.size", "1024").c_str());
    plan_ = plan_memory(init_, num_kv_bufs, codec_, ratio, cfg_kb * 1024);
    client_ = make_input_client(opt_.data_port, opt_.wqes_per_conn);
    client_->start();
    pool_ = std::make_unique<BufferPool>(plan_.pairs, plan_.pair_size[0], plan_.pair_size[1]);
    /* the next line is parsed by the regression's buffer-size checks: keep its wording */
    UDA_LOG(lsINFO, " After RDMA buffers registration: buffer1 = %d bytes , buffer2 = %d bytes , buffers count = %d , total = %lld bytes)",
            (int)plan_.pair_size[0], (int)plan_.pair_size[1], plan_.pairs, (long long)pool_->total_bytes());
    initialised_ = true;
    if (opt_.merge_approach == 3)
        warm_thread_ = spawn_guarded("device-warm-up", nullptr, nullptr, [] { DeviceMerge::instance().warm_up(); });
    // mapred.uda.decompress.ahead=false decodes every block inside the merge thread instead (the A/B switch of the benchmark)
    if (codec_ != Codec::NONE && conf_get(host_, "mapred.uda.decompress.ahead", "true") != "false") decomp_ = std::make_unique<DecompressWorker>();
    merge_thread_ = std::thread([this] { merge_thread_main(); });
}

void Consumer::fail(const UdaError& e) {
    UDA_LOG(lsERROR, "UDA has encountered a critical error and will try to fallback to vanilla MSG=%s", e.full_message().c_str());
    failed_ = true;
    if (host_ && host_->failure_in_uda) host_->failure_in_uda(host_->user);
}

void Consumer::merge_thread_main() {
    try {
        switch (opt_.merge_approach) {
            case 1: merge_online(); break;
            case 2: merge_hybrid(); break;
            case 3: merge_device(); break;
            default: throw UdaError("unsupported merge approach (1=online, 2=hybrid, 3=device exist)");
        }
        UDA_LOG(lsINFO, "merge thread exit");
    } catch (const UdaError& e) {
        if (!stop_) fail(e);
    } catch (const std::exception& e) {
        if (!stop_) fail(UdaError(std::string("unexpected: ") + e.what()));
    }
}

std::unique_ptr<ByteSource> Consumer::make_source(const FetchOrder& o) {
    FetchRequest r;
    r.job_id = o.job;
    r.map_id = o.map;
    r.reduce_id = o.reduce;
    auto fs = std::make_unique<FetchSource>(client_.get(), pool_.get(), o.host, r, plan_.buffer_size);
    fs->start();
    if (codec_ == Codec::NONE) return fs;
    return std::make_unique<DecompressingSource>(std::move(fs), codec_, -1, host_, decomp_.get());
}

void Consumer::fetching_phase(MergeHeap* heap, int count, int* progress) {
    int got = 0;
    ShuffleRng rng(12345u + (uint64_t)segment_seq_);
    while (got < count) {
        std::vector<FetchOrder> batch;
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [&] { return stop_ || !fetch_list_.empty(); });
            if (stop_) throw UdaError("stopped while fetching");
            while (!fetch_list_.empty() && (int)batch.size() < count - got) {
                batch.push_back(fetch_list_.front());
                fetch_list_.pop_front();
            }
        }
        // random order, interleaved by host: the first requests of all reducers must not pile up on one provider
        batch = shuffle_interleave_hosts(std::move(batch), rng, [](const FetchOrder& o) { return o.host; });
        std::vector<std::pair<std::unique_ptr<ByteSource>, FetchSource*>> started;
        for (auto& o : batch) {
            auto src = make_source(o);
            FetchSource* fs = dynamic_cast<FetchSource*>(src.get());
            started.emplace_back(std::move(src), fs);
        }
        for (auto& st : started) {
            if (st.second && !st.second->wait_first_chunk()) throw UdaError("first fetch of a map output failed");
            heap->insert(std::make_unique<Segment>(std::move(st.first), segment_seq_++));
            ++got;
            ++*progress;
            if ((*progress % kProgressEvery == 0 || *progress == init_.num_maps) && host_ && host_->fetch_over_message)
                host_->fetch_over_message(host_->user);
        }
    }
}

void Consumer::merging_phase(MergeHeap* heap, const std::function<void(const uint8_t*, size_t)>& sink, bool final_stream) {
    int cur = 0;
    size_t used = 0;
    uint64_t pending = 0;   // records in the open block: the shared counter moves once per block, not once per record
    auto flush = [&] {
        if (used) sink(staging_[cur].data(), used);
        if (final_stream && pending) emitted_ += pending;  // counted once delivered; spill passes of the hybrid merge are not deliveries
        pending = 0;
        cur ^= 1;
        used = 0;
        if (stop_) throw UdaError("stopped while merging");
    };
    while (!heap->empty()) {
        const Record& r = heap->top();
        const size_t need = r.size;   // the record as it lies in its source: header + key + value, one copy
        if (need > kStagingBytes) {   // a single record larger than a block travels alone, straight from where it lies
            flush();
            sink(r.raw, need);
            if (final_stream) ++emitted_;
        } else {
            if (used + need > kStagingBytes) flush();  // blocks always hold whole records
            memcpy(staging_[cur].data() + used, r.raw, need);
            used += need;
            ++pending;
        }
        heap->advance();
    }
    if (used + 2 > kStagingBytes) flush();
    staging_[cur][used++] = 0xFF;  // VInt(-1) VInt(-1)
    staging_[cur][used++] = 0xFF;
    flush();
}

void Consumer::merge_online() {
    MergeHeap heap(key_kind_);
    int progress = 0;
    fetching_phase(&heap, init_.num_maps, &progress);
    merging_phase(&heap, [this](const uint8_t* b, size_t n) {
        if (host_ && host_->data_from_uda) host_->data_from_uda(host_->user, b, (int)n);
    }, true);
}

namespace {
// a fetched segment served again as a ByteSource (CPU fallback of the device merge)
class MemorySource : public ByteSource {
public:
    MemorySource(const uint8_t* p, size_t n) : p_(p), n_(n) {}
    bool next_block(const uint8_t** data, size_t* len) override {
        if (done_) return false;
        done_ = true;
        *data = p_;
        *len = n_;
        return n_ > 0;
    }
private:
    const uint8_t* p_;
    size_t n_;
    bool done_ = false;
};
}  // namespace

void Consumer::merge_device() {
    const int n = init_.num_maps;
    const auto t_start = std::chrono::steady_clock::now();
    // (the CUDA context is created by the warm-up thread that on_init started: about a second on a cold process,
    //  overlapped with the wait for the first map-completion events and with the fetches)
    std::vector<HostSegment> segs;
    segs.reserve((size_t)n);
    int progress = 0;
    ShuffleRng rng(12345u);
    while ((int)segs.size() < n) {
        std::vector<FetchOrder> batch;
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [&] { return stop_ || !fetch_list_.empty(); });
            if (stop_) throw UdaError("stopped while fetching");
            while (!fetch_list_.empty() && (int)(segs.size() + batch.size()) < n) {
                batch.push_back(fetch_list_.front());
                fetch_list_.pop_front();
            }
        }
        batch = shuffle_interleave_hosts(std::move(batch), rng, [](const FetchOrder& o) { return o.host; });
        // at most `parallel` sources hold buffer pairs at a time; they are drained round-robin so that every one of them
        // keeps its two chunks in flight (draining one source after the other leaves the rest idle)
        const size_t parallel = 8;
        for (size_t b0 = 0; b0 < batch.size(); b0 += parallel) {
            std::vector<std::unique_ptr<ByteSource>> started;
            for (size_t i = b0; i < std::min(batch.size(), b0 + parallel); ++i) started.push_back(make_source(batch[i]));
            std::vector<HostSegment> part(started.size());
            std::vector<bool> open(started.size(), true);
            size_t left = started.size();
            while (left) {
                for (size_t i = 0; i < started.size(); ++i) {
                    if (!open[i]) continue;
                    if (stop_) throw UdaError("stopped while fetching");
                    const uint8_t* d;
                    size_t len;
                    auto ta = std::chrono::steady_clock::now();
                    bool got = started[i]->next_block(&d, &len);
                    auto tb = std::chrono::steady_clock::now();
                    fetch_wait_us_ += std::chrono::duration_cast<std::chrono::microseconds>(tb - ta).count();
                    if (got) {
                        if (part[i].bytes.empty()) {
                            const long long hint = started[i]->size_hint();
                            if (hint > 0) part[i].bytes.reserve((size_t)hint + 16);
                        }
                        part[i].bytes.insert(part[i].bytes.end(), d, d + len);
                        fetch_copy_us_ += std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - tb).count();
                    } else {
                        open[i] = false;
                        --left;
                    }
                }
            }
            for (auto& hs : part) {
                index_segment(&hs);
                segs.push_back(std::move(hs));
                ++progress;
                if ((progress % kProgressEvery == 0 || progress == n) && host_ && host_->fetch_over_message)
                    host_->fetch_over_message(host_->user);
            }
        }
    }
    auto deliver = [this](const uint8_t* b, size_t len) {
        if (host_ && host_->data_from_uda) host_->data_from_uda(host_->user, b, (int)len);
    };
    long long records = 0;
    std::string why;
    const auto t_fetched = std::chrono::steady_clock::now();
    fetch_ms_ = (int)std::chrono::duration_cast<std::chrono::milliseconds>(t_fetched - t_start).count();
    UDA_LOG(lsINFO, "device merge fetch: %d ms total, %lld ms waiting for chunks, %lld ms copying", fetch_ms_.load(), fetch_wait_us_ / 1000, fetch_copy_us_ / 1000);
    DeviceMerge::Refusal refusal = DeviceMerge::Refusal::NONE;
    if (DeviceMerge::instance().merge(segs, key_kind_, kStagingBytes, deliver, &records, &why, &refusal)) {
        emitted_ += (uint64_t)records;
        device_merges_++;
        device_ms_ = (int)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t_fetched).count();
        UDA_LOG(lsINFO, "device merge of %d segments, %lld records done: fetch %d ms, device merge + delivery %d ms", n, records,
                fetch_ms_.load(), device_ms_.load());
        return;
    }
    if (refusal == DeviceMerge::Refusal::MEMORY && segs.size() >= 2) {
        UDA_LOG(lsINFO, "device merge: %s; merging in groups", why.c_str());
        if (merge_device_grouped(&segs, deliver)) {
            device_ms_ = (int)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t_fetched).count();
            return;
        }
        why += "; no grouping fits either";
    }
    UDA_LOG(lsWARN, "device merge not possible (%s): merging on the CPU", why.c_str());
    MergeHeap heap(key_kind_);
    for (auto& hs : segs)
        heap.insert(std::make_unique<Segment>(std::make_unique<MemorySource>(hs.bytes.data(), (size_t)hs.raw_len()), segment_seq_++));
    merging_phase(&heap, deliver, true);
}

bool Consumer::merge_device_grouped(std::vector<HostSegment>* segs, const std::function<void(const uint8_t*, size_t)>& deliver) {
    DeviceMerge& dm = DeviceMerge::instance();
    const size_t k = segs->size();
    std::vector<std::vector<uint8_t>> runs;
    size_t groups = 2;
    for (;; groups *= 2) {
        if (groups > k) return false;   // even single segments do not fit: the caller merges on the CPU
        // consecutive segments per group, cut at equal shares of the bytes (the arrival order is already shuffled)
        long long total = 0;
        for (const auto& sg : *segs) total += sg.raw_len();
        std::vector<size_t> first(groups + 1, k);
        first[0] = 0;
        {
            long long acc = 0;
            size_t g = 1;
            for (size_t i = 0; i < k && g < groups; ++i) {
                acc += (*segs)[i].raw_len();
                // the cut after segment i closes group g when the share is reached, keeping one segment for every later group
                while (g < groups && (acc * (long long)groups >= total * (long long)g || k - (i + 1) <= groups - g)) first[g++] = i + 1;
            }
        }
        runs.clear();
        bool ok = true, memory = false;
        for (size_t g = 0; g < groups && ok; ++g) {
            if (stop_) throw UdaError("stopped while merging");
            const size_t a = first[g], b = first[g + 1];
            if (a >= b) { runs.emplace_back(std::vector<uint8_t>{0xFF, 0xFF}); continue; }   // an empty group is an empty run
            std::vector<HostSegment> part;
            part.reserve(b - a);
            long long bytes = 2;
            for (size_t i = a; i < b; ++i) { bytes += (*segs)[i].raw_len() - 2; part.push_back(std::move((*segs)[i])); }
            std::vector<uint8_t> run;
            run.reserve((size_t)bytes);
            std::string why;
            long long recs = 0;
            DeviceMerge::Refusal refusal = DeviceMerge::Refusal::NONE;
            try {
                ok = dm.merge(part, key_kind_, kStagingBytes, [&](const uint8_t* p, size_t len) { run.insert(run.end(), p, p + len); }, &recs, &why,
                              &refusal);
            } catch (...) {
                for (size_t i = a; i < b; ++i) (*segs)[i] = std::move(part[i - a]);
                throw;
            }
            for (size_t i = a; i < b; ++i) (*segs)[i] = std::move(part[i - a]);   // the caller may still need them (CPU merge)
            if (!ok) {
                memory = refusal == DeviceMerge::Refusal::MEMORY;
                UDA_LOG(lsINFO, "device merge of group %zu/%zu refused (%s)", g + 1, groups, why.c_str());
                break;
            }
            runs.push_back(std::move(run));
        }
        if (ok) break;
        if (!memory) return false;
    }
    UDA_LOG(lsINFO, "device merge in %zu groups of segments (device memory bounds the group size); final merge of the %zu runs on the CPU", groups,
            runs.size());
    for (auto& sg : *segs) { std::vector<uint8_t>().swap(sg.bytes); std::vector<long long>().swap(sg.rec_off); }   // the runs hold the data now
    MergeHeap heap(key_kind_);
    for (auto& r : runs) heap.insert(std::make_unique<Segment>(std::make_unique<MemorySource>(r.data(), r.size()), segment_seq_++));
    device_groups_ = (int)groups;
    device_merges_ += (int)groups;
    merging_phase(&heap, deliver, true);
    return true;
}

void Consumer::merge_hybrid() {
    const int n = init_.num_maps;
    int lpqs = init_.lpq_size > 0 ? (n + init_.lpq_size - 1) / init_.lpq_size : (int)std::sqrt((double)n);
    if (lpqs < 1) lpqs = 1;
    if (n < lpqs || lpqs == 1) { merge_online(); return; }
    int parallel = atoi(conf_get(host_, "mapred.rdma.num.parallel.lpqs", "0").c_str());
    if (parallel < 3) parallel = 3;
    const int per = (n + lpqs - 1) / lpqs;
    std::vector<std::string> dirs = init_.local_dirs;
    if (dirs.empty()) dirs.push_back("/tmp");
    UDA_LOG(lsINFO, "hybrid merge: %d maps in %d LPQs of <=%d, %d in parallel", n, lpqs, per, parallel);

    ExternalQuotaQueue<std::shared_ptr<MergeHeap>> q((size_t)parallel);
    std::atomic<bool> fetch_failed{false};
    std::string fetch_err;
    std::thread fetcher([&] {
        int progress = 0, left = n;
        try {
            for (int l = 0; l < lpqs; ++l) {
                q.wait_and_reserve();
                auto h = std::make_shared<MergeHeap>(key_kind_);
                const int cnt = std::min(per, left);
                fetching_phase(h.get(), cnt, &progress);
                left -= cnt;
                q.push_reserved(h);
            }
        } catch (const std::exception& e) {
            fetch_err = e.what();
            fetch_failed = true;
            for (int l = 0; l < lpqs; ++l) q.push_reserved(nullptr);  // unblock the merger
        }
    });
    std::vector<std::string> files;
    try {
        for (int l = 0; l < lpqs; ++l) {
            auto h = q.wait_and_pop_without_dereserve();
            if (!h || fetch_failed) throw UdaError("LPQ fetch failed: " + fetch_err);
            char name[64];
            snprintf(name, sizeof(name), "/uda.%s.lpq-%03d", init_.reduce_task_id.c_str(), l);
            std::string path = dirs[(size_t)l % dirs.size()] + name;
            FILE* f = fopen(path.c_str(), "wb");
            if (!f) throw UdaError("cannot create spill file " + path);
            files.push_back(path);
            bool werr = false;
            merging_phase(h.get(), [&](const uint8_t* b, size_t len) { if (fwrite(b, 1, len, f) != len) werr = true; }, false);
            fclose(f);
            if (werr) throw UdaError("short write to spill file " + path);
            h.reset();       // returns the LPQ's buffers to the pool
            q.dereserve();   // only now may the fetcher start another LPQ
            ++spilled_;
        }
    } catch (...) {
        stop_ = true;
        cv_.notify_all();
        for (int i = 0; i < parallel + 1; ++i) q.dereserve();
        fetcher.join();
        for (auto& p : files) ::unlink(p.c_str());
        stop_ = false;
        throw;
    }
    fetcher.join();
    // final merge over the spilled runs (never compressed)
    MergeHeap rpq(key_kind_);
    for (size_t i = 0; i < files.size(); ++i)
        rpq.insert(std::make_unique<Segment>(std::make_unique<FileSource>(files[i], true), (int)i));
    merging_phase(&rpq, [this](const uint8_t* b, size_t len) {
        if (host_ && host_->data_from_uda) host_->data_from_uda(host_->user, b, (int)len);
    }, true);
}

void Consumer::exit_msg() {
    if (stop_.exchange(true)) return;
    cv_.notify_all();
    if (client_) client_->stop();  // fails in-flight fetches so that a blocked merge thread wakes up
    if (merge_thread_.joinable()) merge_thread_.join();
    if (warm_thread_.joinable()) warm_thread_.join();
    decomp_.reset();   // every source is gone with the merge thread: nothing is queued any more
    pool_.reset();
    client_.reset();
    log_close();
}

}  // namespace uda
