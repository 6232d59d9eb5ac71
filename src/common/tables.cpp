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

// ---------------------------------------------------------------- classes --
namespace {
struct ClassTable {
    uint32_t g[kNumClasses];
    int n;
    ClassTable() {
        n = 0;
        for (uint32_t s = 1; s <= 8; ++s) g[n++] = s;
        // four steps per doubling: 10 12 14 16 | 20 24 28 32 | ...
        for (uint64_t base = 8; n < kNumClasses;) {
            uint64_t step = base / 4;
            for (int i = 1; i <= 4 && n < kNumClasses; ++i) {
                uint64_t v = base + step * i;
                if (v > 0xFFFFFFFFull) { g[n++] = 0xFFFFFFFFu; continue; }
                g[n++] = (uint32_t)v;
            }
            base *= 2;
            if (base > (1ull << 33)) break;
        }
        for (; n < kNumClasses; ++n) g[n] = 0xFFFFFFFFu;
    }
};
const ClassTable &classes() {
    static ClassTable t;
    return t;
}
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
// returns idx+1 or 0
uint32_t stack_pop(std::atomic<uint64_t> &head, Region *regs) {
    uint64_t h = head.load(std::memory_order_acquire);
    for (;;) {
        uint32_t ip1 = head_idx(h);
        if (!ip1) return 0;
        uint32_t nxt = regs[ip1 - 1].next.load(std::memory_order_relaxed);
        uint64_t nh = pack_head(head_tag(h) + 1, nxt);
        if (head.compare_exchange_weak(h, nh, std::memory_order_acq_rel, std::memory_order_acquire)) return ip1;
    }
}
} // namespace

uint32_t class_granules(int cls) { return (cls >= 0 && cls < kNumClasses) ? classes().g[cls] : 0; }

int class_of(uint64_t granules) {
    if (granules == 0) granules = 1;
    const uint32_t *g = classes().g;
    const uint32_t *e = g + kNumClasses;
    if (granules > 0xFFFFFFFEull) return -1;
    const uint32_t *p = std::lower_bound(g, e, (uint32_t)granules);
    if (p == e || *p == 0xFFFFFFFFu) return -1;
    return (int)(p - g);
}

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
    return round_up(sizeof(TableHeader), 4096) + round_up(n_gpu * sizeof(Region), 4096) +
           round_up(n_host * sizeof(Region), 4096);
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
    }
    OCM_BUG(off > seg_bytes);
}

bool table_validate(const TableHeader *h, size_t mapped) {
    if (mapped < sizeof(TableHeader)) return false;
    if (h->magic != kTableMagic || h->version != kTableVersion) return false;
    if (h->segment_bytes > mapped) return false;
    for (int t = 0; t < TIER_COUNT; ++t) {
        const PoolDesc &p = h->pool[t];
        if (p.n_granules && (p.regions_offset + (uint64_t)p.n_granules * sizeof(Region) > mapped)) return false;
        if (p.n_granules && (p.granule & (p.granule - 1))) return false;
    }
    return true;
}

// ------------------------------------------------------------- allocator --
int pool_alloc(TableHeader *h, int tier, uint64_t bytes, int pid, int rank, BlockRef *out) {
    if (tier < 0 || tier >= TIER_COUNT || !out) return OCM_EINVAL;
    PoolDesc &p = h->pool[tier];
    if (p.n_granules == 0) return OCM_ENOMEM;
    if (bytes == 0) bytes = 1;
    uint64_t need = (bytes + p.granule - 1) / p.granule;
    int cls = class_of(need);
    if (cls < 0 || class_granules(cls) > p.n_granules) return OCM_ENOMEM;
    Region *regs = table_regions(h, tier);

    uint32_t idx_p1 = stack_pop(p.class_head[cls], regs);
    int got_cls = cls;
    if (!idx_p1) {
        // carve a fresh block from the frontier
        uint32_t g = class_granules(cls);
        uint64_t f = p.frontier.load(std::memory_order_relaxed);
        for (;;) {
            if (f + g > p.n_granules) break;
            if (p.frontier.compare_exchange_weak(f, f + g, std::memory_order_acq_rel)) {
                idx_p1 = (uint32_t)f + 1;
                regs[f].cls_plus1 = (uint32_t)cls + 1;
                break;
            }
        }
        if (!idx_p1) {
            // frontier exhausted: the tail may still fit an exact (unrounded) block
            f = p.frontier.load(std::memory_order_relaxed);
            // otherwise reuse a larger free block whole (no splitting)
            for (int c = cls + 1; c < kNumClasses && class_granules(c) <= p.n_granules; ++c) {
                idx_p1 = stack_pop(p.class_head[c], regs);
                if (idx_p1) { got_cls = c; break; }
            }
        }
    }
    if (!idx_p1) return OCM_ENOMEM;
    uint32_t idx = idx_p1 - 1;
    Region &r = regs[idx];
    uint64_t block_bytes = (uint64_t)class_granules(got_cls) * p.granule;
    r.owner_pid = pid;
    r.owner_rank = rank;
    r.bytes = bytes;
    r.id = h->next_id.fetch_add(1, std::memory_order_relaxed);
    r.generation += 1;
    r.flags = 0;
    r.signal.store(0, std::memory_order_relaxed);
    r.state.store(REGION_LIVE, std::memory_order_release);
    p.used_bytes.fetch_add(block_bytes, std::memory_order_relaxed);
    p.live.fetch_add(1, std::memory_order_relaxed);
    p.total_allocs.fetch_add(1, std::memory_order_relaxed);
    out->region = idx;
    out->generation = r.generation;
    out->offset = (uint64_t)idx * p.granule;
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
    if (r.cls_plus1 == 0) return OCM_ESTALE;
    if (r.state.load(std::memory_order_acquire) != REGION_LIVE || r.id != id) return OCM_ESTALE;
    bool scrub = h->scrub_enabled.load(std::memory_order_relaxed) != 0;
    uint32_t expect = REGION_LIVE;
    if (!r.state.compare_exchange_strong(expect, scrub ? REGION_DIRTY : REGION_FREE, std::memory_order_acq_rel))
        return OCM_ESTALE; // lost a race with another free / the reclaimer
    uint64_t block_bytes = (uint64_t)class_granules((int)r.cls_plus1 - 1) * p.granule;
    p.used_bytes.fetch_sub(block_bytes, std::memory_order_relaxed);
    p.live.fetch_sub(1, std::memory_order_relaxed);
    p.total_frees.fetch_add(1, std::memory_order_relaxed);
    if (scrub) {
        p.dirty_bytes.fetch_add(block_bytes, std::memory_order_relaxed);
        stack_push(p.dirty_head, regs, region);
    } else {
        stack_push(p.class_head[r.cls_plus1 - 1], regs, region);
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
        uint64_t bb = (uint64_t)class_granules((int)r.cls_plus1 - 1) * p.granule;
        out->push_back({ip1 - 1, (uint64_t)(ip1 - 1) * p.granule, bb});
        ++n;
        ip1 = nxt;
    }
    return n;
}

void pool_release_scrubbed(TableHeader *h, int tier, uint32_t region) {
    PoolDesc &p = h->pool[tier];
    Region *regs = table_regions(h, tier);
    Region &r = regs[region];
    uint64_t bb = (uint64_t)class_granules((int)r.cls_plus1 - 1) * p.granule;
    p.dirty_bytes.fetch_sub(bb, std::memory_order_relaxed);
    r.state.store(REGION_FREE, std::memory_order_release);
    stack_push(p.class_head[r.cls_plus1 - 1], regs, region);
}

size_t pool_reclaim(TableHeader *h, int tier, bool (*pred)(int, int, void *), void *ctx, uint64_t *bytes_out,
                    std::vector<BlockRef> *reclaimed) {
    PoolDesc &p = h->pool[tier];
    if (!p.n_granules) return 0;
    Region *regs = table_regions(h, tier);
    uint64_t frontier = p.frontier.load(std::memory_order_acquire);
    size_t n = 0;
    for (uint64_t i = 0; i < frontier;) {
        Region &r = regs[i];
        if (r.cls_plus1 == 0) break; // carve in progress; next sweep sees it
        uint32_t g = class_granules((int)r.cls_plus1 - 1);
        if (r.state.load(std::memory_order_acquire) == REGION_LIVE && pred(r.owner_pid, r.owner_rank, ctx)) {
            uint64_t id = r.id;
            if (pool_free(h, tier, (uint32_t)i, id) == OCM_OK) {
                ++n;
                if (bytes_out) *bytes_out += (uint64_t)g * p.granule;
                if (reclaimed) {
                    BlockRef b;
                    b.region = (uint32_t)i; b.id = id; b.offset = i * p.granule;
                    b.block_bytes = (uint64_t)g * p.granule; b.generation = r.generation;
                    reclaimed->push_back(b);
                }
            }
        }
        i += g;
    }
    return n;
}

size_t pool_count_live(TableHeader *h, int tier) {
    PoolDesc &p = h->pool[tier];
    if (!p.n_granules) return 0;
    Region *regs = table_regions(h, tier);
    uint64_t frontier = p.frontier.load(std::memory_order_acquire);
    size_t n = 0;
    for (uint64_t i = 0; i < frontier;) {
        Region &r = regs[i];
        if (r.cls_plus1 == 0) break;
        if (r.state.load(std::memory_order_acquire) == REGION_LIVE) ++n;
        i += class_granules((int)r.cls_plus1 - 1);
    }
    return n;
}

// ---------------------------------------------------------------- cluster --
void cluster_format(ClusterTable *c, int n_ranks) {
    c->magic = kTableMagic;
    c->version = kTableVersion;
    c->n_ranks = n_ranks;
}

bool cluster_reserve(ClusterTable *c, int rank, int tier, uint64_t bytes) {
    if (rank < 0 || rank >= c->n_ranks || rank >= kMaxRanks) return false;
    NodeSlot &n = c->node[rank];
    uint64_t cap = n.pool_bytes[tier];
    uint64_t u = n.used[tier].load(std::memory_order_relaxed);
    for (;;) {
        if (u + bytes > cap) {
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
    if (rank < 0 || rank >= c->n_ranks || rank >= kMaxRanks) return;
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
