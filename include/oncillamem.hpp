// oncillamem.hpp - header-only C++17 convenience layer over the C ABI of oncillamem.h.
//
// The reference is C only (reference inc/oncillamem.h); its callers keep ocm_init/ocm_tini and
// ocm_alloc/ocm_free balanced by hand (reference test/ocm_test.c frees on every exit path).  This wrapper adds
// nothing but ownership: a Session that attaches on construction and detaches on destruction, a move-only
// Allocation that frees itself, and exceptions carrying enum ocm_err instead of -1 returns.
//
//   ocm::Session s;                                             // ocm_init / ocm_tini
//   ocm::Allocation pair(OCM_REMOTE_GPU, 1 << 20, 1 << 20);     // window + far side on the neighbour rank
//   pair.put(1 << 20);  pair.get(1 << 20);
//   float *far = static_cast<float *>(pair.remote_ptr());      // peer HBM, usable by your own kernels
#ifndef ONCILLAMEM_HPP_
#define ONCILLAMEM_HPP_

#include "oncillamem.h"
#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>

namespace ocm {

class Error : public std::runtime_error {
  public:
    Error(const std::string &what, int err) : std::runtime_error(what + ": " + ocm_strerror(err)), err_(err) {}
    int code() const { return err_; }   // enum ocm_err

  private:
    int err_;
};

namespace detail {
inline void check(int rc, const char *what) {
    if (rc == 0) return;
    // the reference-compatible calls return -1 and leave the reason in ocm_last_error(); the extensions return -err
    throw Error(what, rc == -1 ? ocm_last_error() : -rc);
}
} // namespace detail

// Attaches this process to its rank's daemon for the lifetime of the object.
class Session {
  public:
    Session() { detail::check(ocm_init(), "ocm_init"); }
    ~Session() { ocm_tini(); }
    Session(const Session &) = delete;
    Session &operator=(const Session &) = delete;
    int rank() const { return ocm_rank(); }
    int num_ranks() const { return ocm_num_ranks(); }
    int device() const { return ocm_device(); }
    void quiet() const { detail::check(ocm_quiet(), "ocm_quiet"); }
};

// One allocation; frees itself.  Move-only.
class Allocation {
  public:
    Allocation() = default;
    Allocation(enum ocm_kind kind, uint64_t local_bytes, uint64_t remote_bytes = 0, int policy = OCM_POLICY_DEFAULT,
               int rank = -1, uint32_t flags = 0) {
        struct ocm_alloc_params_ex p = {};
        p.base.kind = kind;
        p.base.local_alloc_bytes = local_bytes;
        p.base.rem_alloc_bytes = remote_bytes;
        p.policy = policy;
        p.rank = rank;
        p.flags = flags;
        h_ = ocm_alloc_ex(&p);
        if (!h_) throw Error("ocm_alloc", ocm_last_error());
    }
    // Open a block another application shared (ocm_share / ocm_attach).
    static Allocation attach(const struct ocm_share_token &tok) {
        Allocation a;
        detail::check(ocm_attach(&tok, &a.h_), "ocm_attach");
        return a;
    }
    ~Allocation() { reset(); }
    Allocation(Allocation &&o) noexcept : h_(std::exchange(o.h_, nullptr)) {}
    Allocation &operator=(Allocation &&o) noexcept {
        if (this != &o) {
            reset();
            h_ = std::exchange(o.h_, nullptr);
        }
        return *this;
    }
    Allocation(const Allocation &) = delete;
    Allocation &operator=(const Allocation &) = delete;

    void reset() {
        if (h_) ocm_free(std::exchange(h_, nullptr));
    }
    explicit operator bool() const { return h_ != nullptr; }
    ocm_alloc_t handle() const { return h_; }

    enum ocm_kind kind() const { return ocm_alloc_kind(h_); }
    bool is_remote() const { return ocm_is_remote(h_); }
    int remote_rank() const { return ocm_remote_rank(h_); }
    void *remote_ptr() const { return ocm_remote_ptr(h_); }
    void *local_ptr() const {
        void *p = nullptr;
        size_t n = 0;
        return ocm_localbuf(h_, &p, &n) == 0 ? p : nullptr;
    }
    size_t local_bytes() const {
        void *p = nullptr;
        size_t n = 0;
        return ocm_localbuf(h_, &p, &n) == 0 ? n : 0;
    }
    size_t remote_bytes() const {
        size_t n = 0;
        return ocm_remote_sz(h_, &n) == 0 ? n : 0;
    }

    // one-sided: window[src_offset] -> far[dest_offset] and back
    void put(uint64_t bytes, uint64_t src_offset = 0, uint64_t dest_offset = 0) { onesided(1, bytes, src_offset, dest_offset); }
    void get(uint64_t bytes, uint64_t src_offset = 0, uint64_t dest_offset = 0) { onesided(0, bytes, src_offset, dest_offset); }
    void put_async(uint64_t bytes, void *stream, uint64_t src_offset = 0, uint64_t dest_offset = 0) {
        struct ocm_params p = {};
        p.src_offset = src_offset, p.dest_offset = dest_offset, p.bytes = bytes, p.op_flag = 1;
        detail::check(ocm_copy_onesided_async(h_, &p, stream), "ocm_copy_onesided_async");
    }

    void get_async(uint64_t bytes, void *stream, uint64_t src_offset = 0, uint64_t dest_offset = 0) {
        struct ocm_params p = {};
        p.src_offset = src_offset, p.dest_offset = dest_offset, p.bytes = bytes, p.op_flag = 0;
        detail::check(ocm_copy_onesided_async(h_, &p, stream), "ocm_copy_onesided_async");
    }
    // n puts / gets of this allocation in ONE kernel launch (entry.op_flag: 1 = put, 0 = get); stream-ordered
    void batch(const struct ocm_params *entries, int n, void *stream = nullptr) {
        detail::check(ocm_copy_onesided_batch(h_, entries, n, stream), "ocm_copy_onesided_batch");
    }

    uint64_t signal() const {
        uint64_t v = 0;
        detail::check(ocm_signal_read(h_, &v), "ocm_signal_read");
        return v;
    }
    void wait_signal(uint64_t at_least, int timeout_ms = 10000) const {
        detail::check(ocm_signal_wait(h_, at_least, timeout_ms), "ocm_signal_wait");
    }
    struct ocm_share_token share() const {
        struct ocm_share_token t;
        detail::check(ocm_share(h_, &t), "ocm_share");
        return t;
    }

  private:
    void onesided(int op, uint64_t bytes, uint64_t so, uint64_t dof) {
        struct ocm_params p = {};
        p.src_offset = so, p.dest_offset = dof, p.bytes = bytes, p.op_flag = op;
        detail::check(ocm_copy_onesided(h_, &p), op ? "ocm_copy_onesided(put)" : "ocm_copy_onesided(get)");
    }
    ocm_alloc_t h_ = nullptr;
};

// Two-sided copy dst <- src (op_flag 1) with the offset semantics of DESIGN.md section 4.
inline void copy(Allocation &dst, Allocation &src, uint64_t bytes, uint64_t src_offset = 0, uint64_t dest_offset = 0,
                 uint64_t src_offset_2 = 0, uint64_t dest_offset_2 = 0) {
    struct ocm_params p = {};
    p.src_offset = src_offset, p.dest_offset = dest_offset;
    p.src_offset_2 = src_offset_2, p.dest_offset_2 = dest_offset_2;
    p.bytes = bytes, p.op_flag = 1;
    detail::check(ocm_copy(dst.handle(), src.handle(), &p), "ocm_copy");
}

// Map every rank's pools into this process now (set-up) instead of at first use; returns the ranks fully mapped.
inline int prefetch_views(bool gpu = true, bool host = true) {
    int r = ocm_prefetch_views((gpu ? 1u : 0u) | (host ? 2u : 0u));
    if (r < 0) throw Error("ocm_prefetch_views", -r);
    return r;
}

} // namespace ocm
#endif // ONCILLAMEM_HPP_
