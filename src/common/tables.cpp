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

size_t table_segment_bytes(uint64_t gpu_pool_bytes, uint64_t host_pool_bytes, uint64_t gg, uint64_t hg) {
    if (!gg) gg = default_granule(gpu_pool_bytes, TIER_GPU);
    if (!hg) hg = default_granule(host_pool_bytes, TIER_HOST);
    size_t n_gpu = gpu_pool_bytes / gg, n_host = host_pool_bytes / hg;
    size_t total = round_up(sizeof(TableHeader), 4096);
    for (size_t n : {n_gpu, n_host}) {
        total += round_up(n * sizeof(Region), 4096);
        total += round_up((n + 63) / 64 * 8, 4096);
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
    }
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
    }
    if (h->sigext_offset + (uint64_t)h->n_sigext * sizeof(SigExt) > mapped) return false;
    return true;
}

// ------------------------------------------------------------- allocator --
// ---------------------------------------------------- signal extensions --
SigExt *sigext_claim(TableHeader *h, int tier, uint32_t region) {
    Region &r = table_regions(h, tier)[region];
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

int pool_alloc(TableHeader *h, int tier, uint64_t bytes, int pid, int rank, BlockRef *out) {
    if (tier < 0 || tier >= TIER_COUNT || !out) return OCM_EINVAL;
    PoolDesc &p = h->pool[tier];
    if (p.n_granules == 0) return OCM_ENOMEM;
    if (bytes == 0) bytes = 1;
    if (bytes > p.bytes) return OCM_ENOMEM;   // before any arithmetic: a huge size must not wrap to a small block
    const uint64_t n = (bytes + p.granule - 1) / p.granule;
    if (n == 0 || n > p.n_granules) return OCM_ENOMEM;
    std::atomic<uint64_t> *bm = table_bitmap(h, tier);
    const uint64_t N = p.n_granules;
    uint64_t hint = p.hint.load(std::memory_order_relaxed);
    if (hint >= N) hint = 0;
    uint64_t start = find_run(bm, hint, N, n);
    if (start == UINT64_MAX && hint) start = find_run(bm, 0, std::min<uint64_t>(N, hint + n), n);
    if (start == UINT64_MAX) return OCM_ENOMEM;
    p.hint.store(start + n >= N ? 0 : start + n, std::memory_order_relaxed);

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
    p.live.fetch_add(1, std::memory_order_relaxed);
    p.total_allocs.fetch_add(1, std::memory_order_relaxed);
    out->region = (uint32_t)start;
    out->generation = r.generation;
    out->offset = start * p.granule;
    out->block_bytes = block_bytes;
    out->id = r.id;
    return OCM_OK;
}

int pool_free(TableHeader *h, int tier, uint32_t region, uint64_t id) {
    if (tier < 0 || tier >= TIER_COUNT) return OCM_EINVAL;
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
    p.live.fetch_sub(1, std::memory_order_relaxed);
    p.total_frees.fetch_add(1, std::memory_order_relaxed);
    if (scrub) {
        p.dirty_bytes.fetch_add(block_bytes, std::memory_order_relaxed);
        stack_push(p.dirty_head, regs, region);
    } else {
        bitmap_clear(table_bitmap(h, tier), region, n);
    }
    return OCM_OK;
}

size_t pool_take_dirty(TableHeader *h, int tier, std::vector<DirtyBlock> *out) {
    PoolDesc &p = h->pool[tier];
    if (!p.n_granules) return 0;
    Region *regs = table_regions(h, tier);
    uint64_t hd = p.dirty_head.load(std::memory_order_acquire);
    for (;;) {
        if (!head_idx(hd)) return 0;
        if (p.dirty_head.compare_exchange_weak(hd, pack_head(head_tag(hd) + 1, 0), std::memory_order_acq_rel)) break;
    }
    size_t n = 0;
    for (uint32_t ip1 = head_idx(hd); ip1;) {
        Region &r = regs[ip1 - 1];
        uint32_t nxt = r.next.load(std::memory_order_relaxed);
        r.state.store(REGION_SCRUBBING, std::memory_order_relaxed);
        out->push_back({ip1 - 1, (uint64_t)(ip1 - 1) * p.granule, (uint64_t)r.n_granules * p.granule});
        ++n;
        ip1 = nxt;
    }
    return n;
}

void pool_release_scrubbed(TableHeader *h, int tier, uint32_t region) {
    PoolDesc &p = h->pool[tier];
    Region &r = table_regions(h, tier)[region];
    const uint32_t n = r.n_granules;
    p.dirty_bytes.fetch_sub((uint64_t)n * p.granule, std::memory_order_relaxed);
    r.state.store(REGION_FREE, std::memory_order_release);
    bitmap_clear(table_bitmap(h, tier), region, n);
}

uint64_t pool_scrub_host_inline(TableHeader *h, void *host_pool_base) {
    if (!host_pool_base) return 0;
    std::vector<DirtyBlock> blocks;
    if (!pool_take_dirty(h, TIER_HOST, &blocks)) return 0;
    Region *regs = table_regions(h, TIER_HOST);
    uint64_t total = 0;
    for (auto &b : blocks) {
        // only the part the owner could touch needs clearing
        uint64_t n = std::min<uint64_t>(b.bytes, (regs[b.region].bytes + 4095) & ~4095ull);
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
        if (r.state.load(std::memory_order_acquire) == REGION_LIVE) ++n;
    });
    return n;
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
