#include "tables.hpp"
#include "log.hpp"
#include "../../include/oncillamem.h"
#include <algorithm>
#include <cerrno>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace ocm {

namespace {
inline uint64_t pack_head(uint32_t tag, uint32_t idx_plus1) { return ((uint64_t)tag << 32) | idx_plus1; }
inline uint32_t head_idx(uint64_t h) { return (uint32_t)h; }
inline uint32_t head_tag(uint64_t h) { return (uint32_t)(h >> 32); }

void stack_push(std::atomic<uint64_t> &head, Region *regs, uint32_t idx) {
    uint64_t h = head.load(std::memory_order_acquire);
    for (;;) {
        regs[idx].next.store(head_idx(h), std::memory_order_relaxed);
        uint64_t nh = pack_head(head_tag(h) + 1, idx + 1);
        if (head.compare_exchange_weak(h, nh, std::memory_order_acq_rel, std::memory_order_acquire)) return;
    }
}

// mask of bits [lo, hi) inside one 64-bit word (0 <= lo < hi <= 64)
inline uint64_t bits(unsigned lo, unsigned hi) {
    uint64_t m = hi >= 64 ? ~0ull : ((1ull << hi) - 1);
    return m & ~((1ull << lo) - 1);
}

// p.hint must never stay above a granule that has just become free
inline void hint_lower(std::atomic<uint64_t> &hint, uint64_t region) {
    uint64_t h = hint.load(std::memory_order_relaxed);
    while (region < h && !hint.compare_exchange_weak(h, region, std::memory_order_acq_rel, std::memory_order_relaxed)) {
    }
}

void bitmap_clear(std::atomic<uint64_t> *bm, uint64_t start, uint64_t n) {
    uint64_t i = start, end = start + n;
    while (i < end) {
        uint64_t w = i >> 6;
        unsigned lo = (unsigned)(i & 63);
        unsigned hi = (unsigned)std::min<uint64_t>(64, lo + (end - i));
        bm[w].fetch_and(~bits(lo, hi), std::memory_order_release);
        i += hi - lo;
    }
}

// Try to take [start, start+n).  On conflict everything taken so far is given
// back and *conflict receives the first granule found busy.
bool bitmap_claim(std::atomic<uint64_t> *bm, uint64_t start, uint64_t n, uint64_t *conflict) {
    uint64_t i = start, end = start + n;
    while (i < end) {
        uint64_t w = i >> 6;
        unsigned lo = (unsigned)(i & 63);
        unsigned hi = (unsigned)std::min<uint64_t>(64, lo + (end - i));
        uint64_t m = bits(lo, hi);
        uint64_t old = bm[w].load(std::memory_order_relaxed);
        for (;;) {
            if (old & m) {
                if (i > start) bitmap_clear(bm, start, i - start);
                *conflict = (w << 6) + (unsigned)__builtin_ctzll(old & m);
                return false;
            }
            if (bm[w].compare_exchange_weak(old, old | m, std::memory_order_acq_rel, std::memory_order_relaxed)) break;
        }
        i += hi - lo;
    }
    return true;
}

// First granule >= from that is free, or N.
uint64_t next_free(std::atomic<uint64_t> *bm, uint64_t from, uint64_t N) {
    uint64_t i = from;
    while (i < N) {
        uint64_t w = i >> 6;
        uint64_t word = bm[w].load(std::memory_order_relaxed) | ((i & 63) ? ((1ull << (i & 63)) - 1) : 0);
        if (word != ~0ull) return std::min<uint64_t>((w << 6) + (unsigned)__builtin_ctzll(~word), N);
        i = (w + 1) << 6;
    }
    return N;
}

// First busy granule in [start, start+n), or start+n if the run is free (read-only probe).
uint64_t first_busy(std::atomic<uint64_t> *bm, uint64_t start, uint64_t n) {
    uint64_t i = start, end = start + n;
    while (i < end) {
        uint64_t w = i >> 6;
        unsigned lo = (unsigned)(i & 63);
        unsigned hi = (unsigned)std::min<uint64_t>(64, lo + (end - i));
        uint64_t busy = bm[w].load(std::memory_order_relaxed) & bits(lo, hi);
        if (busy) return (w << 6) + (unsigned)__builtin_ctzll(busy);
        i += hi - lo;
    }
    return end;
}

// Find and claim a free run of n granules in [lo, hi).  Returns its start or UINT64_MAX.
uint64_t find_run(std::atomic<uint64_t> *bm, uint64_t lo, uint64_t hi, uint64_t n) {
    if (n == 0 || n > hi) return UINT64_MAX;   // an empty run would "succeed" without claiming anything
    uint64_t i = lo;
    while (i + n <= hi) {
        i = next_free(bm, i, hi);
        if (i + n > hi) break;
        uint64_t b = first_busy(bm, i, n);
        if (b < i + n) {
            i = b + 1;
            continue;
        }
        uint64_t conflict = 0;
        if (bitmap_claim(bm, i, n, &conflict)) return i;
        i = conflict + 1;
    }
    return UINT64_MAX;
}
// Claim ONE free granule, the highest there is.  Single granules (the granule behind a slab, one-granule blocks) are
// taken from the top of the pool and runs from the bottom, so that long-lived small blocks do not cut the space the
// runs need into pieces (unit_host frag: a third fewer failed large requests at 70-85 % occupancy).
uint64_t find_one_down(std::atomic<uint64_t> *bm, uint64_t N) {
    for (uint64_t w = (N + 63) / 64; w-- > 0;) {
        uint64_t word = bm[w].load(std::memory_order_relaxed);
        while (word != ~0ull) {
            const unsigned b = 63u - (unsigned)__builtin_clzll(~word);
            const uint64_t idx = (w << 6) + b;
            if (idx < N) {
                uint64_t conflict = 0;
                if (bitmap_claim(bm, idx, 1, &conflict)) return idx;
            }
            word = bm[w].load(std::memory_order_relaxed) | (idx >= N ? (1ull << b) : 0);
        }
    }
    return UINT64_MAX;
}
} // namespace

uint64_t default_granule(uint64_t pool_bytes, Tier tier) {
    uint64_t g = tier == TIER_GPU ? (64ull << 10) : (4ull << 10);
    while (pool_bytes / g > kMaxGranules) g <<= 1;
    return g;
}

// ------------------------------------------------------------------- shm --
bool shm_create(const std::string &name, size_t bytes, bool replace, ShmMapping *out, std::string *err) {
    if (replace) shm_unlink(name.c_str());
    int fd = shm_open(name.c_str(), O_RDWR | O_CREAT | O_EXCL, 0600);
    if (fd < 0) {
        if (err) *err = "shm_open(create " + name + "): " + strerror(errno);
        return false;
    }
    if (ftruncate(fd, (off_t)bytes) != 0) {
        if (err) *err = "ftruncate(" + name + "): " + strerror(errno);
        close(fd);
        shm_unlink(name.c_str());
        return false;
    }
    void *p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (p == MAP_FAILED) {
        if (err) *err = "mmap(" + name + "): " + strerror(errno);
        close(fd);
        shm_unlink(name.c_str());
        return false;
    }
    out->base = p; out->bytes = bytes; out->fd = fd; out->owner = true; out->name = name;
    return true;
}

bool shm_open_existing(const std::string &name, ShmMapping *out, std::string *err) {
    int fd = shm_open(name.c_str(), O_RDWR, 0600);
    if (fd < 0) {
        if (err) *err = "shm_open(" + name + "): " + strerror(errno);
        return false;
    }
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size == 0) {
        if (err) *err = "fstat(" + name + "): empty or " + strerror(errno);
        close(fd);
        return false;
    }
    void *p = mmap(nullptr, (size_t)st.st_size, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (p == MAP_FAILED) {
        if (err) *err = "mmap(" + name + "): " + strerror(errno);
        close(fd);
        return false;
    }
    out->base = p; out->bytes = (size_t)st.st_size; out->fd = fd; out->owner = false; out->name = name;
    return true;
}

void shm_close(ShmMapping *m, bool unlink_it) {
    if (m->base) munmap(m->base, m->bytes);
    if (m->fd >= 0) close(m->fd);
    if (unlink_it && !m->name.empty()) shm_unlink(m->name.c_str());
    m->base = nullptr; m->fd = -1; m->bytes = 0;
}

// ---------------------------------------------------------------- layout --
static inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
// slabs per tier: never more than the pool has granules, a whole number of classes
static inline uint32_t slab_count(uint64_t n_granules) {
    return (uint32_t)(std::min<uint64_t>(kMaxSlabs, n_granules) / kSlabClasses * kSlabClasses);
}
constexpr int32_t kSlabOwnerPid = -2;   // owner_pid of the granule block that backs a slab

size_t table_segment_bytes(uint64_t gpu_pool_bytes, uint64_t host_pool_bytes, uint64_t gg, uint64_t hg) {
    if (!gg) gg = default_granule(gpu_pool_bytes, TIER_GPU);
    if (!hg) hg = default_granule(host_pool_bytes, TIER_HOST);
    size_t n_gpu = gpu_pool_bytes / gg, n_host = host_pool_bytes / hg;
    size_t total = round_up(sizeof(TableHeader), 4096);
    for (size_t n : {n_gpu, n_host}) {
        total += round_up(n * sizeof(Region), 4096);
        total += round_up((n + 63) / 64 * 8, 4096);
        total += round_up((size_t)slab_count(n) * sizeof(Slab), 4096);
        total += round_up((size_t)slab_count(n) * kSlabSubs * sizeof(Region), 4096);
    }
    total += round_up((size_t)kSigExtRecords * sizeof(SigExt), 4096);
    return total;
}

void table_format(void *base, size_t seg_bytes, int rank, int n_ranks, int daemon_pid,
                  uint64_t gpu_pool_bytes, uint64_t host_pool_bytes, uint64_t gg, uint64_t hg) {
    if (!gg) gg = default_granule(gpu_pool_bytes, TIER_GPU);
    if (!hg) hg = default_granule(host_pool_bytes, TIER_HOST);
    auto *h = static_cast<TableHeader *>(base);
    h->magic = kTableMagic;
    h->version = kTableVersion;
    h->rank = rank;
    h->n_ranks = n_ranks;
    h->daemon_pid = daemon_pid;
    h->gpu_index = -1;
    h->gpu_share_mode = GPU_SHARE_NONE;
    h->segment_bytes = seg_bytes;
    h->next_id.store(1);
    h->scrub_enabled.store(1);
    uint64_t off = round_up(sizeof(TableHeader), 4096);
    const uint64_t sizes[TIER_COUNT] = {gpu_pool_bytes, host_pool_bytes};
    const uint64_t grans[TIER_COUNT] = {gg, hg};
    for (int t = 0; t < TIER_COUNT; ++t) {
        PoolDesc &p = h->pool[t];
        p.granule = grans[t];
        p.n_granules = (uint32_t)(sizes[t] / grans[t]);
        p.bytes = (uint64_t)p.n_granules * grans[t];
        p.tier = t;
        p.regions_offset = off;
        off += round_up((size_t)p.n_granules * sizeof(Region), 4096);
        p.bitmap_offset = off;
        p.bitmap_words = (p.n_granules + 63) / 64;
        off += round_up((size_t)p.bitmap_words * 8, 4096);
        // granules past the end of the pool are permanently taken
        if (p.n_granules & 63) {
            auto *bm = table_bitmap(h, t);
            bm[p.bitmap_words - 1].store(~((1ull << (p.n_granules & 63)) - 1), std::memory_order_relaxed);
        }
        p.n_slabs = slab_count(p.n_granules);
        p.slabs_offset = off;
        off += round_up((size_t)p.n_slabs * sizeof(Slab), 4096);
        p.subs_offset = off;
        off += round_up((size_t)p.n_slabs * kSlabSubs * sizeof(Region), 4096);
    }
    h->slabs_enabled.store(1);
    h->sigext_offset = off;
    h->n_sigext = kSigExtRecords;
    off += round_up((size_t)kSigExtRecords * sizeof(SigExt), 4096);
    OCM_BUG(off > seg_bytes);
}

bool table_validate(const TableHeader *h, size_t mapped) {
    if (mapped < sizeof(TableHeader)) return false;
    if (h->magic != kTableMagic || h->version != kTableVersion) return false;
    if (h->segment_bytes > mapped) return false;
    for (int t = 0; t < TIER_COUNT; ++t) {
        const PoolDesc &p = h->pool[t];
        if (!p.n_granules) continue;
        if (p.regions_offset + (uint64_t)p.n_granules * sizeof(Region) > mapped) return false;
        if (p.bitmap_offset + (uint64_t)p.bitmap_words * 8 > mapped) return false;
        if (p.granule & (p.granule - 1)) return false;
        if (p.n_slabs > kMaxSlabs || p.slabs_offset + (uint64_t)p.n_slabs * sizeof(Slab) > mapped) return false;
        if (p.subs_offset + (uint64_t)p.n_slabs * kSlabSubs * sizeof(Region) > mapped) return false;
    }
    if (h->sigext_offset + (uint64_t)h->n_sigext * sizeof(SigExt) > mapped) return false;
    return true;
}

// ------------------------------------------------------------- allocator --
// ---------------------------------------------------- signal extensions --
SigExt *sigext_claim(TableHeader *h, int tier, uint32_t region) {
    Region *rp = table_region_rec(h, tier, region);
    if (!rp) return nullptr;
    Region &r = *rp;
    SigExt *ext = table_sigext(h);
    uint32_t have = r.ext.load(std::memory_order_acquire);
    if (have) return &ext[have - 1];
    const uint64_t tag = (((uint64_t)tier << 32) | region) + 1;
    for (uint32_t i = 0; i < h->n_sigext; ++i) {
        uint64_t expect = 0;
        if (ext[i].owner.load(std::memory_order_relaxed) != 0 ||
            !ext[i].owner.compare_exchange_strong(expect, tag, std::memory_order_acq_rel))
            continue;
        ext[i].id = r.id;
        for (auto &s : ext[i].slot) s.store(0, std::memory_order_relaxed);
        uint32_t none = 0;
        if (r.ext.compare_exchange_strong(none, i + 1, std::memory_order_acq_rel)) return &ext[i];
        ext[i].owner.store(0, std::memory_order_release);   // another thread of the owner won
        return &ext[none - 1];
    }
    return nullptr;
}

// the block leaves the LIVE state: its extension record (if any) becomes free again
static void sigext_release(TableHeader *h, Region &r) {
    const uint32_t e = r.ext.exchange(0, std::memory_order_acq_rel);
    if (e && e <= h->n_sigext) table_sigext(h)[e - 1].owner.store(0, std::memory_order_release);
}

uint64_t region_signal_total(TableHeader *h, const Region &r) {
    uint64_t v = r.signal.load(std::memory_order_acquire) + r.signal_cpu.load(std::memory_order_acquire);
    const uint32_t e = r.ext.load(std::memory_order_acquire);
    if (e && e <= h->n_sigext) {
        const SigExt &x = table_sigext(h)[e - 1];
        for (const auto &s : x.slot) v += s.load(std::memory_order_acquire);
    }
    return v;
}

static void sub_release_clean(TableHeader *h, int tier, uint32_t slab, uint32_t sub);
static size_t slab_trim(TableHeader *h, int tier);

// A free sub-block of the class that fits `bytes`, from an active slab or from a slab made of a fresh granule.
// OCM_ENOMEM when the class has no slab header or the pool no granule left (the caller then takes a whole granule).
static int granule_alloc(TableHeader *h, int tier, uint64_t bytes, int pid, int rank, BlockRef *out, bool count);

static int slab_alloc(TableHeader *h, int tier, uint64_t bytes, int pid, int rank, BlockRef *out) {
    PoolDesc &p = h->pool[tier];
    // smallest sub-block that holds the request: class c has sub-blocks of granule >> (c + 1), at least 256 bytes
    int cls = -1;
    for (int c = (int)kSlabClasses - 1; c >= 0; --c) {
        const uint64_t sz = p.granule >> (c + 1);
        if (sz >= 256 && sz >= bytes) {
            cls = c;
            break;
        }
    }
    if (cls < 0 || p.n_slabs < kSlabClasses) return OCM_ENOMEM;
    const uint32_t n_sub = class_subs((uint32_t)cls);
    const uint64_t sub_size = p.granule >> (cls + 1);
    const uint32_t per_class = p.n_slabs / kSlabClasses;
    Slab *slabs = table_slabs(h, tier);
    uint32_t slab = 0, bit = 0;
    bool got = false;
    for (int attempt = 0; attempt < 3 && !got; ++attempt) {
        if (p.class_free[cls].load(std::memory_order_acquire) > 0) {
            const uint32_t start = p.class_hint[cls].load(std::memory_order_relaxed) % per_class;
            for (uint32_t k = 0; k < per_class && !got; ++k) {
                const uint32_t i = ((start + k) % per_class) * kSlabClasses + (uint32_t)cls;
                Slab &s = slabs[i];
                if (s.state.load(std::memory_order_acquire) != SLAB_ACTIVE) continue;
                uint64_t m = s.free_mask.load(std::memory_order_acquire);
                while (m) {
                    const uint32_t b = (uint32_t)__builtin_ctzll(m);
                    if (s.free_mask.compare_exchange_weak(m, m & ~(1ull << b), std::memory_order_acq_rel)) {
                        slab = i;
                        bit = b;
                        got = true;
                        p.class_free[cls].fetch_sub(1, std::memory_order_acq_rel);
                        p.class_hint[cls].store((start + k) % per_class, std::memory_order_relaxed);
                        break;
                    }
                }
            }
            if (got) break;
        }
        // no free sub-block anywhere: turn a fresh granule into a slab of this class
        uint32_t fresh = UINT32_MAX;
        for (uint32_t k = 0; k < per_class; ++k) {
            const uint32_t i = k * kSlabClasses + (uint32_t)cls;
            uint32_t expect = SLAB_FREE;
            if (slabs[i].state.load(std::memory_order_relaxed) == SLAB_FREE &&
                slabs[i].state.compare_exchange_strong(expect, SLAB_INIT, std::memory_order_acq_rel)) {
                fresh = i;
                break;
            }
        }
        if (fresh == UINT32_MAX) {
            if (p.class_free[cls].load(std::memory_order_acquire) > 0) continue;   // somebody freed one meanwhile
            return OCM_ENOMEM;
        }
        BlockRef g;
        int err = granule_alloc(h, tier, p.granule, kSlabOwnerPid, rank, &g, false);
        if (err) {
            slabs[fresh].state.store(SLAB_FREE, std::memory_order_release);
            return err;
        }
        Slab &s = slabs[fresh];
        s.region = g.region;
        s.backing_id = g.id;
        // sub-block 0 is ours, the others are published with the mask (release: region/backing_id first)
        const uint64_t all = n_sub == 64 ? ~0ull : ((1ull << n_sub) - 1);
        s.free_mask.store(all & ~1ull, std::memory_order_release);
        s.state.store(SLAB_ACTIVE, std::memory_order_release);
        p.class_free[cls].fetch_add((int64_t)n_sub - 1, std::memory_order_acq_rel);
        slab = fresh;
        bit = 0;
        got = true;
    }
    if (!got) return OCM_ENOMEM;
    Slab &s = slabs[slab];
    Region &r = table_subs(h, tier)[(size_t)slab * kSlabSubs + bit];
    r.n_granules = 0;
    r.owner_pid = pid;
    r.owner_rank = rank;
    r.bytes = bytes;
    r.id = h->next_id.fetch_add(1, std::memory_order_relaxed);
    r.generation += 1;
    r.attached.store(0, std::memory_order_relaxed);
    r.ext.store(0, std::memory_order_relaxed);
    r.signal.store(0, std::memory_order_relaxed);
    r.signal_cpu.store(0, std::memory_order_relaxed);
    r.state.store(REGION_LIVE, std::memory_order_release);
    p.sub_live.fetch_add(1, std::memory_order_relaxed);
    p.sub_bytes.fetch_add(sub_size, std::memory_order_relaxed);
    p.live.fetch_add(1, std::memory_order_relaxed);
    p.total_allocs.fetch_add(1, std::memory_order_relaxed);
    out->region = kSlabFlag | (slab << 6) | bit;
    out->generation = r.generation;
    out->offset = (uint64_t)s.region * p.granule + (uint64_t)bit * sub_size;
    out->block_bytes = sub_size;
    out->id = r.id;
    return OCM_OK;
}

// The sub-block is clean again: publish it; a slab that is completely free goes back to the pool when the class
// keeps at least another slab's worth of free sub-blocks (no create/retire ping-pong around one small block).
static void sub_release_clean(TableHeader *h, int tier, uint32_t slab, uint32_t sub) {
    PoolDesc &p = h->pool[tier];
    Slab &s = table_slabs(h, tier)[slab];
    const uint32_t cls = slab_class(slab), n_sub = class_subs(cls);
    const uint64_t all = n_sub == 64 ? ~0ull : ((1ull << n_sub) - 1);
    const uint64_t before = s.free_mask.fetch_or(1ull << sub, std::memory_order_acq_rel);
    const int64_t free_now = p.class_free[cls].fetch_add(1, std::memory_order_acq_rel) + 1;
    if ((before | (1ull << sub)) != all || free_now < 2 * (int64_t)n_sub) return;
    uint64_t expect = all;
    if (!s.free_mask.compare_exchange_strong(expect, 0, std::memory_order_acq_rel)) return;   // somebody took a block
    s.state.store(SLAB_RETIRING, std::memory_order_release);
    p.class_free[cls].fetch_sub((int64_t)n_sub, std::memory_order_acq_rel);
    const uint32_t region = s.region;
    const uint64_t id = s.backing_id;
    pool_free(h, tier, region, id);
    s.state.store(SLAB_FREE, std::memory_order_release);
}

// Retire every slab that has no live sub-block.  Returns how many granules went back.
static size_t slab_trim(TableHeader *h, int tier) {
    PoolDesc &p = h->pool[tier];
    Slab *slabs = table_slabs(h, tier);
    size_t n = 0;
    for (uint32_t i = 0; i < p.n_slabs; ++i) {
        Slab &s = slabs[i];
        if (s.state.load(std::memory_order_acquire) != SLAB_ACTIVE) continue;
        const uint32_t cls = slab_class(i), n_sub = class_subs(cls);
        uint64_t expect = n_sub == 64 ? ~0ull : ((1ull << n_sub) - 1);
        if (!s.free_mask.compare_exchange_strong(expect, 0, std::memory_order_acq_rel)) continue;
        s.state.store(SLAB_RETIRING, std::memory_order_release);
        p.class_free[cls].fetch_sub((int64_t)n_sub, std::memory_order_acq_rel);
        const uint32_t region = s.region;
        const uint64_t id = s.backing_id;
        pool_free(h, tier, region, id);
        s.state.store(SLAB_FREE, std::memory_order_release);
        ++n;
    }
    return n;
}

static int sub_free(TableHeader *h, int tier, uint32_t region, uint64_t id) {
    PoolDesc &p = h->pool[tier];
    Region *rp = table_region_rec(h, tier, region);
    if (!rp) return OCM_EINVAL;
    Region &r = *rp;
    if (r.state.load(std::memory_order_acquire) != REGION_LIVE || r.id != id) return OCM_ESTALE;
    const bool scrub = h->scrub_enabled.load(std::memory_order_relaxed) != 0;
    uint32_t expect = REGION_LIVE;
    if (!r.state.compare_exchange_strong(expect, scrub ? REGION_DIRTY : REGION_FREE, std::memory_order_acq_rel))
        return OCM_ESTALE;
    sigext_release(h, r);
    const uint32_t slab = sub_slab(region), sub = sub_index(region);
    const uint64_t sub_size = p.granule >> (slab_class(slab) + 1);
    p.sub_live.fetch_sub(1, std::memory_order_relaxed);
    p.sub_bytes.fetch_sub(sub_size, std::memory_order_relaxed);
    p.live.fetch_sub(1, std::memory_order_relaxed);
    p.total_frees.fetch_add(1, std::memory_order_relaxed);
    if (scrub) {
        p.dirty_bytes.fetch_add(sub_size, std::memory_order_relaxed);
        stack_push(p.sub_dirty_head, table_subs(h, tier), slab * kSlabSubs + sub);
    } else {
        sub_release_clean(h, tier, slab, sub);
    }
    return OCM_OK;
}

int pool_alloc(TableHeader *h, int tier, uint64_t bytes, int pid, int rank, BlockRef *out) {
    if (tier < 0 || tier >= TIER_COUNT || !out) return OCM_EINVAL;
    PoolDesc &p = h->pool[tier];
    if (p.n_granules == 0) return OCM_ENOMEM;
    if (bytes == 0) bytes = 1;
    if (bytes > p.bytes) return OCM_ENOMEM;   // before any arithmetic: a huge size must not wrap to a small block
    if (bytes <= p.granule / 2 && h->slabs_enabled.load(std::memory_order_relaxed) && slab_alloc(h, tier, bytes, pid, rank, out) == OCM_OK)
        return OCM_OK;
    int err = granule_alloc(h, tier, bytes, pid, rank, out, true);
    // out of granules while empty slabs are being kept around for the next small request: give them back and retry
    if (err == OCM_ENOMEM && slab_trim(h, tier)) err = granule_alloc(h, tier, bytes, pid, rank, out, true);
    return err;
}

static int granule_alloc(TableHeader *h, int tier, uint64_t bytes, int pid, int rank, BlockRef *out, bool count) {
    PoolDesc &p = h->pool[tier];
    const uint64_t n = (bytes + p.granule - 1) / p.granule;
    if (n == 0 || n > p.n_granules) return OCM_ENOMEM;
    std::atomic<uint64_t> *bm = table_bitmap(h, tier);
    const uint64_t N = p.n_granules;
    uint64_t start;
    if (n == 1) {
        start = find_one_down(bm, N);
        if (start == UINT64_MAX) return OCM_ENOMEM;
    } else if (n < kFirstFitMinGranules) {
        // short runs: next fit from a rotating start - cheapest search, and concurrent allocators do not meet on the
        // same bitmap words
        uint64_t hint = p.rot_hint.load(std::memory_order_relaxed);
        if (hint >= N) hint = 0;
        start = find_run(bm, hint, N, n);
        if (start == UINT64_MAX && hint) start = find_run(bm, 0, std::min<uint64_t>(N, hint + n), n);
        if (start == UINT64_MAX) return OCM_ENOMEM;
        p.rot_hint.store(start + n >= N ? 0 : start + n, std::memory_order_relaxed);
    } else {
        // long runs: first fit from the bottom, so that they pack towards low addresses and long runs stay available
        // (unit_host frag, 1 GiB pool 70-85 % full, random mix: 14 % of run requests refused with one rotating next
        // fit for everything, 1 % with single granules moved to the top, 0.03 % with this split; first fit for every
        // run: 0 %, but 3x the search time at that occupancy and all allocators meeting on the lowest free words).
        // p.hint is a lower bound of the lowest free granule, so a full prefix is skipped without reading it; frees
        // lower it (hint_lower), the search raises it.  A stale bound can only hide space, never hand out a taken
        // granule - hence the second search from 0 before giving up.
        uint64_t h0 = p.hint.load(std::memory_order_acquire);
        if (h0 >= N) h0 = 0;
        const uint64_t first = next_free(bm, h0, N);
        start = first + n <= N ? find_run(bm, first, N, n) : UINT64_MAX;
        if (first > h0) p.hint.compare_exchange_strong(h0, first, std::memory_order_acq_rel, std::memory_order_relaxed);
        if (start == UINT64_MAX && h0) start = find_run(bm, 0, N, n);
        if (start == UINT64_MAX) return OCM_ENOMEM;
    }

    Region *regs = table_regions(h, tier);
    Region &r = regs[start];
    const uint64_t block_bytes = n * p.granule;
    r.n_granules = (uint32_t)n;
    r.owner_pid = pid;
    r.owner_rank = rank;
    r.bytes = bytes;
    r.id = h->next_id.fetch_add(1, std::memory_order_relaxed);
    r.generation += 1;
    r.attached.store(0, std::memory_order_relaxed);
    r.ext.store(0, std::memory_order_relaxed);
    r.signal.store(0, std::memory_order_relaxed);
    r.signal_cpu.store(0, std::memory_order_relaxed);
    r.state.store(REGION_LIVE, std::memory_order_release);
    p.used_bytes.fetch_add(block_bytes, std::memory_order_relaxed);
    if (count) {
        p.live.fetch_add(1, std::memory_order_relaxed);
        p.total_allocs.fetch_add(1, std::memory_order_relaxed);
    }
    out->region = (uint32_t)start;
    out->generation = r.generation;
    out->offset = start * p.granule;
    out->block_bytes = block_bytes;
    out->id = r.id;
    return OCM_OK;
}

int pool_free(TableHeader *h, int tier, uint32_t region, uint64_t id) {
    if (tier < 0 || tier >= TIER_COUNT) return OCM_EINVAL;
    if (region_is_sub(region)) return sub_free(h, tier, region, id);
    PoolDesc &p = h->pool[tier];
    if (region >= p.n_granules) return OCM_EINVAL;
    Region *regs = table_regions(h, tier);
    Region &r = regs[region];
    if (r.state.load(std::memory_order_acquire) != REGION_LIVE || r.id != id) return OCM_ESTALE;
    const bool scrub = h->scrub_enabled.load(std::memory_order_relaxed) != 0;
    const uint32_t n = r.n_granules;
    uint32_t expect = REGION_LIVE;
    if (!r.state.compare_exchange_strong(expect, scrub ? REGION_DIRTY : REGION_FREE, std::memory_order_acq_rel))
        return OCM_ESTALE; // lost a race with another free / the reclaimer
    sigext_release(h, r);
    const uint64_t block_bytes = (uint64_t)n * p.granule;
    p.used_bytes.fetch_sub(block_bytes, std::memory_order_relaxed);
    if (r.owner_pid != kSlabOwnerPid) {   // the granule behind a slab was never counted as an allocation
        p.live.fetch_sub(1, std::memory_order_relaxed);
        p.total_frees.fetch_add(1, std::memory_order_relaxed);
    }
    if (scrub) {
        p.dirty_bytes.fetch_add(block_bytes, std::memory_order_relaxed);
        stack_push(p.dirty_head, regs, region);
    } else {
        bitmap_clear(table_bitmap(h, tier), region, n);
        hint_lower(p.hint, region);
    }
    return OCM_OK;
}

// Detach one dirty stack; returns its first element (index + 1) or 0.
static uint32_t detach_stack(std::atomic<uint64_t> &head) {
    uint64_t hd = head.load(std::memory_order_acquire);
    for (;;) {
        if (!head_idx(hd)) return 0;
        if (head.compare_exchange_weak(hd, pack_head(head_tag(hd) + 1, 0), std::memory_order_acq_rel)) return head_idx(hd);
    }
}

size_t pool_take_dirty(TableHeader *h, int tier, std::vector<DirtyBlock> *out) {
    PoolDesc &p = h->pool[tier];
    if (!p.n_granules) return 0;
    Region *regs = table_regions(h, tier);
    size_t n = 0;
    for (uint32_t ip1 = detach_stack(p.dirty_head); ip1;) {
        Region &r = regs[ip1 - 1];
        uint32_t nxt = r.next.load(std::memory_order_relaxed);
        r.state.store(REGION_SCRUBBING, std::memory_order_relaxed);
        out->push_back({ip1 - 1, (uint64_t)(ip1 - 1) * p.granule, (uint64_t)r.n_granules * p.granule});
        ++n;
        ip1 = nxt;
    }
    // released sub-blocks of slabs: same treatment, named by their sub-block reference
    if (p.n_slabs) {
        Region *subs = table_subs(h, tier);
        for (uint32_t ip1 = detach_stack(p.sub_dirty_head); ip1;) {
            Region &r = subs[ip1 - 1];
            uint32_t nxt = r.next.load(std::memory_order_relaxed);
            r.state.store(REGION_SCRUBBING, std::memory_order_relaxed);
            const uint32_t slab = (ip1 - 1) / kSlabSubs, sub = (ip1 - 1) % kSlabSubs;
            const uint32_t ref = kSlabFlag | (slab << 6) | sub;
            out->push_back({ref, table_region_offset(h, tier, ref), p.granule >> (slab_class(slab) + 1)});
            ++n;
            ip1 = nxt;
        }
    }
    return n;
}

void pool_release_scrubbed(TableHeader *h, int tier, uint32_t region) {
    PoolDesc &p = h->pool[tier];
    if (region_is_sub(region)) {
        Region *sr = table_region_rec(h, tier, region);
        if (!sr) return;
        p.dirty_bytes.fetch_sub(p.granule >> (slab_class(sub_slab(region)) + 1), std::memory_order_relaxed);
        sr->state.store(REGION_FREE, std::memory_order_release);
        sub_release_clean(h, tier, sub_slab(region), sub_index(region));
        return;
    }
    Region &r = table_regions(h, tier)[region];
    const uint32_t n = r.n_granules;
    p.dirty_bytes.fetch_sub((uint64_t)n * p.granule, std::memory_order_relaxed);
    r.state.store(REGION_FREE, std::memory_order_release);
    bitmap_clear(table_bitmap(h, tier), region, n);
    hint_lower(p.hint, region);
}

uint64_t pool_scrub_host_inline(TableHeader *h, void *host_pool_base) {
    if (!host_pool_base) return 0;
    std::vector<DirtyBlock> blocks;
    if (!pool_take_dirty(h, TIER_HOST, &blocks)) return 0;
    uint64_t total = 0;
    for (auto &b : blocks) {
        // only the part the owner could touch needs clearing
        const Region *rec = table_region_rec(h, TIER_HOST, b.region);
        uint64_t n = std::min<uint64_t>(b.bytes, ((rec ? rec->bytes : b.bytes) + 4095) & ~4095ull);
        memset(static_cast<char *>(host_pool_base) + b.offset, 0, n);
        total += b.bytes;
        pool_release_scrubbed(h, TIER_HOST, b.region);
    }
    return total;
}

// Visit the start granule of every block whose bits are set.  Blocks are found by
// walking the bitmap: a set bit whose Region is not FREE starts a block.
template <typename F> static void for_each_block(TableHeader *h, int tier, F &&fn) {
    PoolDesc &p = h->pool[tier];
    std::atomic<uint64_t> *bm = table_bitmap(h, tier);
    Region *regs = table_regions(h, tier);
    const uint64_t N = p.n_granules;
    for (uint64_t w = 0; w < p.bitmap_words; ++w) {
        uint64_t word = bm[w].load(std::memory_order_acquire);
        while (word) {
            unsigned b = (unsigned)__builtin_ctzll(word);
            word &= word - 1;
            uint64_t i = (w << 6) + b;
            if (i >= N) break;
            Region &r = regs[i];
            if (r.state.load(std::memory_order_acquire) != REGION_FREE) fn((uint32_t)i, r);
        }
    }
}

size_t pool_reclaim(TableHeader *h, int tier, bool (*pred)(int, int, void *), void *ctx, uint64_t *bytes_out,
                    std::vector<BlockRef> *reclaimed) {
    PoolDesc &p = h->pool[tier];
    if (!p.n_granules) return 0;
    size_t n = 0;
    // sub-blocks of active slabs first (their backing granules belong to nobody and stay)
    {
        Slab *slabs = table_slabs(h, tier);
        Region *subs = table_subs(h, tier);
        for (uint32_t s = 0; s < p.n_slabs; ++s) {
            if (slabs[s].state.load(std::memory_order_acquire) != SLAB_ACTIVE) continue;
            const uint32_t n_sub = class_subs(slab_class(s));
            const uint64_t sub_size = p.granule >> (slab_class(s) + 1);
            for (uint32_t k = 0; k < n_sub; ++k) {
                Region &r = subs[(size_t)s * kSlabSubs + k];
                if (r.state.load(std::memory_order_acquire) != REGION_LIVE || !pred(r.owner_pid, r.owner_rank, ctx)) continue;
                const uint32_t ref = kSlabFlag | (s << 6) | k;
                const uint64_t id = r.id;
                BlockRef b;
                b.region = ref; b.id = id; b.offset = table_region_offset(h, tier, ref); b.block_bytes = sub_size;
                b.generation = r.generation;
                if (pool_free(h, tier, ref, id) != OCM_OK) continue;
                ++n;
                if (bytes_out) *bytes_out += sub_size;
                if (reclaimed) reclaimed->push_back(b);
            }
        }
    }
    for_each_block(h, tier, [&](uint32_t i, Region &r) {
        if (r.state.load(std::memory_order_acquire) != REGION_LIVE || !pred(r.owner_pid, r.owner_rank, ctx)) return;
        const uint64_t id = r.id;
        const uint32_t g = r.n_granules;
        if (pool_free(h, tier, i, id) != OCM_OK) return;
        ++n;
        if (bytes_out) *bytes_out += (uint64_t)g * p.granule;
        if (reclaimed) {
            BlockRef b;
            b.region = i; b.id = id; b.offset = (uint64_t)i * p.granule;
            b.block_bytes = (uint64_t)g * p.granule; b.generation = r.generation;
            reclaimed->push_back(b);
        }
    });
    return n;
}

size_t pool_count_live(TableHeader *h, int tier) {
    if (!h->pool[tier].n_granules) return 0;
    size_t n = 0;
    for_each_block(h, tier, [&](uint32_t, Region &r) {
        if (r.state.load(std::memory_order_acquire) == REGION_LIVE && r.owner_pid != kSlabOwnerPid) ++n;
    });
    PoolDesc &p = h->pool[tier];
    Slab *slabs = table_slabs(h, tier);
    Region *subs = table_subs(h, tier);
    for (uint32_t s = 0; s < p.n_slabs; ++s) {
        if (slabs[s].state.load(std::memory_order_acquire) != SLAB_ACTIVE) continue;
        for (uint32_t k = 0; k < class_subs(slab_class(s)); ++k)
            if (subs[(size_t)s * kSlabSubs + k].state.load(std::memory_order_acquire) == REGION_LIVE) ++n;
    }
    return n;
}

uint64_t pool_payload_bytes(TableHeader *h, int tier) {
    PoolDesc &p = h->pool[tier];
    uint64_t slab_granules = 0;
    Slab *slabs = table_slabs(h, tier);
    for (uint32_t s = 0; s < p.n_slabs; ++s)
        if (slabs[s].state.load(std::memory_order_acquire) == SLAB_ACTIVE) ++slab_granules;
    const uint64_t used = p.used_bytes.load(std::memory_order_acquire);
    const uint64_t slab_bytes = slab_granules * p.granule;
    return (used > slab_bytes ? used - slab_bytes : 0) + p.sub_bytes.load(std::memory_order_acquire);
}

int pool_audit(TableHeader *h, int tier, TableAudit *out) {
    TableAudit a;
    memset(&a, 0, sizeof a);
    PoolDesc &p = h->pool[tier];
    if (p.n_granules) {
        std::atomic<uint64_t> *bm = table_bitmap(h, tier);
        uint64_t bits = 0;
        for (uint32_t w = 0; w < p.bitmap_words; ++w) bits += (uint64_t)__builtin_popcountll(bm[w].load(std::memory_order_acquire));
        bits -= (uint64_t)p.bitmap_words * 64 - p.n_granules;          // the permanently-taken tail
        a.bitmap_bytes = bits * p.granule;
        a.used_bytes = p.used_bytes.load(std::memory_order_acquire);
        a.dirty_bytes = p.dirty_bytes.load(std::memory_order_acquire);
        a.live_records = pool_count_live(h, tier);
        a.live_counter = p.live.load(std::memory_order_acquire);
        // every taken granule is either in use or waiting to be zeroed (dirty sub-blocks sit inside used slabs)
        if (a.bitmap_bytes < a.used_bytes || a.bitmap_bytes > a.used_bytes + a.dirty_bytes) ++a.problems;
        if (a.live_records != a.live_counter) ++a.problems;
        if (a.used_bytes > p.bytes) ++a.problems;
    }
    if (out) *out = a;
    return a.problems;
}

int64_t mc_range_alloc(ClusterTable *c, uint64_t bytes, int owner_pid, int owner_rank) {
    if (c->mc_state.load(std::memory_order_acquire) != MC_READY || !bytes || bytes > c->mc_bytes) return -1;
    const uint64_t n = (bytes + kMcGranule - 1) / kMcGranule;
    if (n > c->mc_n_granules) return -1;
    uint64_t start = find_run(c->mc_bitmap, 0, c->mc_n_granules, n);
    if (start == UINT64_MAX) return -1;
    c->mc_scrubbed[start].store(0, std::memory_order_relaxed);
    c->mc_owner[start].store(owner_pid > 0 ? mc_owner_pack(owner_pid, (uint32_t)n, owner_rank, MC_RANGE_OWNED) : 0,
                             std::memory_order_release);
    return (int64_t)(start * kMcGranule);
}

void mc_range_free(ClusterTable *c, uint64_t offset, uint64_t bytes) {
    const uint64_t n = (bytes + kMcGranule - 1) / kMcGranule;
    c->mc_owner[offset / kMcGranule].store(0, std::memory_order_release);
    bitmap_clear(c->mc_bitmap, offset / kMcGranule, n);
}

size_t mc_mark_dead(ClusterTable *c, int rank, bool (*pred)(int pid)) {
    size_t marked = 0;
    for (uint32_t g = 0; g < c->mc_n_granules && g < kMcMaxGranules; ++g) {
        uint64_t w = c->mc_owner[g].load(std::memory_order_acquire);
        if (mc_owner_state(w) != MC_RANGE_OWNED || mc_owner_rank(w) != rank || !pred(mc_owner_pid(w))) continue;
        const uint64_t dead = mc_owner_pack(mc_owner_pid(w), mc_owner_granules(w), rank, MC_RANGE_DEAD);
        if (c->mc_owner[g].compare_exchange_strong(w, dead, std::memory_order_acq_rel)) ++marked;
    }
    return marked;
}

size_t mc_dead_ranges(ClusterTable *c, int rank, std::vector<std::pair<uint32_t, uint32_t>> *out) {
    size_t found = 0;
    for (uint32_t g = 0; g < c->mc_n_granules && g < kMcMaxGranules; ++g) {
        const uint64_t w = c->mc_owner[g].load(std::memory_order_acquire);
        if (mc_owner_state(w) != MC_RANGE_DEAD) continue;
        if (c->mc_scrubbed[g].load(std::memory_order_acquire) & (1ull << rank)) continue;
        out->emplace_back(g, mc_owner_granules(w));
        ++found;
    }
    return found;
}

bool mc_range_scrubbed(ClusterTable *c, int rank, uint32_t start) {
    const uint64_t w = c->mc_owner[start].load(std::memory_order_acquire);
    if (mc_owner_state(w) != MC_RANGE_DEAD) return false;
    const uint64_t all = c->n_ranks >= 64 ? ~0ull : ((1ull << c->n_ranks) - 1);
    const uint64_t before = c->mc_scrubbed[start].fetch_or(1ull << rank, std::memory_order_acq_rel);
    if ((before | (1ull << rank)) != all || (before & (1ull << rank))) return false;
    // we completed the set: the range is zero on every GPU and goes back to the bitmap
    c->mc_owner[start].store(0, std::memory_order_release);
    bitmap_clear(c->mc_bitmap, start, mc_owner_granules(w));
    return true;
}

// ---------------------------------------------------------------- cluster --
void cluster_format(ClusterTable *c, int n_ranks) {
    c->magic = kTableMagic;
    c->version = kTableVersion;
    c->n_ranks = n_ranks;
}

bool cluster_reserve(ClusterTable *c, int rank, int tier, uint64_t bytes) {
    if (rank < 0 || rank >= c->n_ranks || rank >= kMaxRanks || tier < 0 || tier >= TIER_COUNT) return false;
    NodeSlot &n = c->node[rank];
    uint64_t cap = n.pool_bytes[tier];
    uint64_t u = n.used[tier].load(std::memory_order_relaxed);
    for (;;) {
        if (bytes > cap || u > cap - bytes) {   // written so that a huge request cannot wrap around
            c->rejected.fetch_add(1, std::memory_order_relaxed);
            return false;
        }
        if (n.used[tier].compare_exchange_weak(u, u + bytes, std::memory_order_acq_rel)) break;
    }
    n.live.fetch_add(1, std::memory_order_relaxed);
    c->total_allocs.fetch_add(1, std::memory_order_relaxed);
    return true;
}

void cluster_unreserve(ClusterTable *c, int rank, int tier, uint64_t bytes) {
    if (rank < 0 || rank >= c->n_ranks || rank >= kMaxRanks || tier < 0 || tier >= TIER_COUNT) return;
    NodeSlot &n = c->node[rank];
    uint64_t u = n.used[tier].load(std::memory_order_relaxed);
    for (;;) {
        uint64_t nv = u >= bytes ? u - bytes : 0;
        if (n.used[tier].compare_exchange_weak(u, nv, std::memory_order_acq_rel)) break;
    }
    n.live.fetch_sub(1, std::memory_order_relaxed);
    c->total_frees.fetch_add(1, std::memory_order_relaxed);
}

} // namespace ocm
