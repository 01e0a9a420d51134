// Thread-safe queues used by the provider (request queue, reader queues) and the hybrid merger
// (bounded number of runs in flight).
//
// Reference behaviour: src/include/concurrent_queue.h:49-99 (MPMC queue), :116-178 (quota queue),
// :196-272 (external quota queue: reserve / push_reserved / pop_without_dereserve / dereserve, which
// bounds the number of LPQs being built). The reference's try_pop forgets to return true
// (concurrent_queue.h:75-85); this one reports success properly.
#pragma once
#include <condition_variable>
#include <deque>
#include <mutex>
#include <utility>

namespace uda {

template <typename T>
class ConcurrentQueue {
public:
    void push(T v) {
        {
            std::lock_guard<std::mutex> lk(mu_);
            q_.push_back(std::move(v));
        }
        cv_.notify_one();
    }
    bool try_pop(T& out) {
        std::lock_guard<std::mutex> lk(mu_);
        if (q_.empty()) return false;
        out = std::move(q_.front());
        q_.pop_front();
        return true;
    }
    // Blocks until an element arrives or close() is called; false means closed and drained.
    bool wait_and_pop(T& out) {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return !q_.empty() || closed_; });
        if (q_.empty()) return false;
        out = std::move(q_.front());
        q_.pop_front();
        return true;
    }
    void close() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            closed_ = true;
        }
        cv_.notify_all();
    }
    bool empty() const {
        std::lock_guard<std::mutex> lk(mu_);
        return q_.empty();
    }
    size_t size() const {
        std::lock_guard<std::mutex> lk(mu_);
        return q_.size();
    }

private:
    mutable std::mutex mu_;
    std::condition_variable cv_;
    std::deque<T> q_;
    bool closed_ = false;
};

// push() blocks while `quota` elements are queued.
template <typename T>
class QuotaQueue {
public:
    explicit QuotaQueue(size_t quota) : quota_(quota) {}
    void push(T v) {
        std::unique_lock<std::mutex> lk(mu_);
        not_full_.wait(lk, [&] { return q_.size() < quota_; });
        q_.push_back(std::move(v));
        lk.unlock();
        not_empty_.notify_one();
    }
    T wait_and_pop() {
        std::unique_lock<std::mutex> lk(mu_);
        not_empty_.wait(lk, [&] { return !q_.empty(); });
        T v = std::move(q_.front());
        q_.pop_front();
        lk.unlock();
        not_full_.notify_one();
        return v;
    }
    size_t size() const {
        std::lock_guard<std::mutex> lk(mu_);
        return q_.size();
    }

private:
    mutable std::mutex mu_;
    std::condition_variable not_full_, not_empty_;
    std::deque<T> q_;
    size_t quota_;
};

// The quota counts *reservations*, which outlive the queue residency of an element: a producer
// reserves before it starts building an element, and the consumer gives the slot back only when it has
// finished processing the popped element. That bounds elements being built + queued + being consumed.
template <typename T>
class ExternalQuotaQueue {
public:
    explicit ExternalQuotaQueue(size_t quota) : quota_(quota) {}
    void wait_and_reserve() {
        std::unique_lock<std::mutex> lk(mu_);
        slot_free_.wait(lk, [&] { return reserved_ < quota_; });
        ++reserved_;
    }
    void push_reserved(T v) {
        {
            std::lock_guard<std::mutex> lk(mu_);
            q_.push_back(std::move(v));
        }
        not_empty_.notify_one();
    }
    T wait_and_pop_without_dereserve() {
        std::unique_lock<std::mutex> lk(mu_);
        not_empty_.wait(lk, [&] { return !q_.empty(); });
        T v = std::move(q_.front());
        q_.pop_front();
        return v;
    }
    void dereserve() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            if (reserved_) --reserved_;
        }
        slot_free_.notify_one();
    }
    size_t reserved() const {
        std::lock_guard<std::mutex> lk(mu_);
        return reserved_;
    }

private:
    mutable std::mutex mu_;
    std::condition_variable slot_free_, not_empty_;
    std::deque<T> q_;
    size_t quota_, reserved_ = 0;
};

}  // namespace uda
