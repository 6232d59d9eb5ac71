// Bounded lock-free multi-producer / single-consumer ring of fixed-size POD
// items (Vyukov-style per-slot sequence numbers).  Replaces the reference's
// mutex-protected growing FIFO used for the daemon outbox (reference
// src/queue.c:26-152, inc/util/queue.h:20-46), including its unlocked q_empty
// read and leaking q_free.
#pragma once
#include <atomic>
#include <cstddef>
#include <cstdint>
#include <memory>
#include <new>
#include <type_traits>

namespace ocm {

template <typename T>
class MpscRing {
    static_assert(std::is_trivially_copyable<T>::value, "ring items are copied bytewise");
    struct Slot {
        std::atomic<uint64_t> seq;
        T item;
    };

  public:
    explicit MpscRing(size_t capacity_pow2 = 1024) {
        size_t c = 2;
        while (c < capacity_pow2) c <<= 1;
        cap_ = c;
        mask_ = c - 1;
        slots_.reset(new Slot[c]);
        for (size_t i = 0; i < c; ++i) slots_[i].seq.store(i, std::memory_order_relaxed);
    }
    MpscRing(const MpscRing &) = delete;
    MpscRing &operator=(const MpscRing &) = delete;

    // Any thread.  false when full.
    bool push(const T &v) {
        uint64_t pos = tail_.load(std::memory_order_relaxed);
        for (;;) {
            Slot &s = slots_[pos & mask_];
            uint64_t seq = s.seq.load(std::memory_order_acquire);
            int64_t dif = (int64_t)seq - (int64_t)pos;
            if (dif == 0) {
                if (tail_.compare_exchange_weak(pos, pos + 1, std::memory_order_relaxed)) {
                    s.item = v;
                    s.seq.store(pos + 1, std::memory_order_release);
                    return true;
                }
            } else if (dif < 0) {
                return false; // full
            } else {
                pos = tail_.load(std::memory_order_relaxed);
            }
        }
    }
    // Consumer thread only.  false when empty.
    bool pop(T *out) {
        Slot &s = slots_[head_ & mask_];
        uint64_t seq = s.seq.load(std::memory_order_acquire);
        if ((int64_t)seq - (int64_t)(head_ + 1) < 0) return false;
        *out = s.item;
        s.seq.store(head_ + cap_, std::memory_order_release);
        ++head_;
        return true;
    }
    bool empty() const {
        const Slot &s = slots_[head_ & mask_];
        return (int64_t)s.seq.load(std::memory_order_acquire) - (int64_t)(head_ + 1) < 0;
    }
    size_t capacity() const { return cap_; }
    size_t size_approx() const { return (size_t)(tail_.load(std::memory_order_relaxed) - head_); }

  private:
    size_t cap_, mask_;
    std::unique_ptr<Slot[]> slots_;
    alignas(64) std::atomic<uint64_t> tail_{0};
    alignas(64) uint64_t head_ = 0;
};

} // namespace ocm
