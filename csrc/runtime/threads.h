// Thread wrapper and fetch-order shuffling helpers of the host runtime.
//
// Reference behaviour covered here: uda_thread_create = pthread_create + JVM attach/detach + "an
// exception that escapes a native thread becomes a fallback request" (src/CommUtils/UdaUtil.cc:39-94), and the
// Fisher-Yates helpers that randomise the order in which map outputs are fetched so that the first requests of all
// reducers do not land on the same provider (src/CommUtils/UdaUtil.h vector_shuffle / list_shuffle_in_vector, used
// at src/Merger/MergeManager.cc:60-75).
//
// Here JVM attachment is lazy (csrc/jni/uda_bridge.cc attaches a thread on its first up-call and detaches it from
// a thread_local destructor), so the wrapper only has to name the thread for the log, contain exceptions and raise
// the failure up-call exactly once.
#pragma once
#include <atomic>
#include <cstdint>
#include <exception>
#include <functional>
#include <list>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "uda_rt.h"

namespace uda {

// Raises failure_in_uda at most once per process role; later failures are only logged
// (reference: src/UdaBridge.cc:75-86 ignores everything after the first failure).
class FailureLatch {
public:
    bool raise(const HostCallbacks* host, const std::string& why) {
        bool first = !failed_.exchange(true);
        UDA_LOG(lsERROR, "%s failure in native thread: %s", first ? "first" : "subsequent", why.c_str());
        if (first && host && host->failure_in_uda) host->failure_in_uda(host->user);
        return first;
    }
    bool failed() const { return failed_.load(); }
    void reset() { failed_ = false; }
private:
    std::atomic<bool> failed_{false};
};

// Starts `body` on a new thread. Anything it throws is logged with its backtrace (UdaError) and turned into a
// failure up-call through `latch`; the thread then ends normally so that join() never sees a terminate().
template <class Fn>
std::thread spawn_guarded(const char* name, const HostCallbacks* host, FailureLatch* latch, Fn body) {
    std::string tname(name ? name : "uda");
    return std::thread([tname, host, latch, body]() mutable {
        UDA_LOG(lsDEBUG, "thread %s started", tname.c_str());
        try {
            body();
        } catch (const UdaError& e) {
            if (latch) latch->raise(host, tname + ": " + e.full_message());
            else UDA_LOG(lsERROR, "thread %s: %s", tname.c_str(), e.full_message().c_str());
        } catch (const std::exception& e) {
            if (latch) latch->raise(host, tname + ": " + e.what());
            else UDA_LOG(lsERROR, "thread %s: %s", tname.c_str(), e.what());
        } catch (...) {
            if (latch) latch->raise(host, tname + ": unknown exception");
            else UDA_LOG(lsERROR, "thread %s: unknown exception", tname.c_str());
        }
        UDA_LOG(lsDEBUG, "thread %s finished", tname.c_str());
    });
}

// xorshift64*: deterministic across platforms (std::shuffle's use of the engine is implementation defined, and
// the golden tests pin the order).
struct ShuffleRng {
    uint64_t s;
    explicit ShuffleRng(uint64_t seed) : s(seed ? seed : 0x9E3779B97F4A7C15ull) {}
    uint64_t next() {
        s ^= s >> 12; s ^= s << 25; s ^= s >> 27;
        return s * 0x2545F4914F6CDD1Dull;
    }
    // unbiased integer in [0, n)
    uint64_t below(uint64_t n) {
        uint64_t lim = UINT64_MAX - UINT64_MAX % n;
        uint64_t v;
        do v = next(); while (v >= lim);
        return v % n;
    }
};

// Fisher-Yates, in place.
template <class T>
void vector_shuffle(std::vector<T>& v, ShuffleRng& rng) {
    for (size_t i = v.size(); i > 1; --i) std::swap(v[i - 1], v[rng.below(i)]);
}

// Moves a list into a vector in random order (the reference drains its pending-fetch list this way).
template <class T>
std::vector<T> list_shuffle_in_vector(std::list<T>& l, ShuffleRng& rng) {
    std::vector<T> v;
    v.reserve(l.size());
    for (auto& x : l) v.push_back(std::move(x));
    l.clear();
    vector_shuffle(v, rng);
    return v;
}

// Random order that additionally never asks the same host twice in a row while another host still has work:
// shuffle, bucket by host, then deal round-robin starting from a random host. `host_of(x)` names the bucket.
template <class T, class HostOf>
std::vector<T> shuffle_interleave_hosts(std::vector<T> v, ShuffleRng& rng, HostOf host_of) {
    vector_shuffle(v, rng);
    std::vector<std::pair<std::string, std::vector<T>>> buckets;
    for (auto& x : v) {
        std::string h = host_of(x);
        size_t b = 0;
        while (b < buckets.size() && buckets[b].first != h) ++b;
        if (b == buckets.size()) buckets.emplace_back(h, std::vector<T>());
        buckets[b].second.push_back(std::move(x));
    }
    std::vector<T> out;
    out.reserve(v.size());
    if (buckets.empty()) return out;
    size_t start = rng.below(buckets.size()), left = v.size();
    for (size_t round = 0; left; ++round)
        for (size_t i = 0; i < buckets.size(); ++i) {
            auto& bk = buckets[(start + i) % buckets.size()].second;
            if (round < bk.size()) { out.push_back(std::move(bk[round])); --left; }
        }
    return out;
}

}  // namespace uda
