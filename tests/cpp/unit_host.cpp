// Native unit tests of the host substrate (no daemon, no GPU needed):
//   unit_host <suite>|all     suites: msg nodefile ring tables tables_mt slabs governor topo pmsg sock oracle
// The reference has no unit layer at all; these are new.
#include "common/governor.hpp"
#include "common/msg.hpp"
#include "common/names.hpp"
#include "common/nodefile.hpp"
#include "common/ring.hpp"
#include "common/tables.hpp"
#include "common/topo.hpp"
#include <fstream>
#include <sys/stat.h>
#include "ipc/pmsg.hpp"
#include "ipc/sock.hpp"
#include "oracle/oracle.hpp"
#include "../../include/oncillamem.h"
#include <algorithm>
#include <chrono>
#include <cmath>
#include <random>
#include <cstdio>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <sys/socket.h>
#include <sys/wait.h>
#include <thread>
#include <unistd.h>
#include <vector>

static int g_fail = 0;
#define CHECK(cond)                                                                      \
    do {                                                                                 \
        if (!(cond)) {                                                                   \
            fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);              \
            ++g_fail;                                                                    \
        }                                                                                \
    } while (0)

using namespace ocm;

static void t_msg() {
    Message m;
    m.init(MSG_REQ_ALLOC);
    CHECK(m.valid());
    CHECK(sizeof(Message) == 192);
    CHECK(m.type == MSG_REQ_ALLOC && m.status == MSG_REQUEST);
    CHECK(MSG_CONNECT == 1 && MSG_RELEASE_APP == 9 && MSG_MAX == 11); // reference numbering, inc/msg.h:24-45
    CHECK(!strcmp(msg_type_str(MSG_DO_FREE), "MSG_DO_FREE"));
    CHECK(!strcmp(msg_status_str(MSG_RESPONSE), "MSG_RESPONSE"));
    m.magic = 0;
    CHECK(!m.valid());
    CHECK(OCM_LOCAL_HOST == 1 && OCM_REMOTE_RDMA == 5 && OCM_LOCAL_GPU == 6 && OCM_REMOTE_GPU == 7);
    CHECK(sizeof(struct ocm_params) == 48 && sizeof(struct ocm_alloc_params) == 24);
}

static void t_nodefile() {
    Nodefile nf;
    std::string err;
    const char *ref = "#rank dns ethernet_ip ocm_port rdmacm_port\n"
                      "0 shiva.cc.gt.atl.ga.us 143.215.131.65 12345 67890\n"
                      "1 ifrit.cc.gt.atl.ga.us 143.215.131.185 12345 67890\n";
    CHECK(parse_nodefile_text(ref, &nf, &err));
    CHECK(nf.size() == 2 && nf.nodes[1].dns == "ifrit.cc.gt.atl.ga.us" && nf.nodes[1].ocm_port == 12345);
    CHECK(nf.nodes[0].rdmacm_port == 67890);
    CHECK(resolve_rank(nf, "ifrit.cc.gt.atl.ga.us", nullptr) == 1);
    CHECK(resolve_rank(nf, "nobody", nullptr) == -1);
    CHECK(resolve_rank(nf, "nobody", "1") == 1);
    CHECK(resolve_rank(nf, "nobody", "7") == -1);
    Nodefile rma;
    CHECK(parse_nodefile_text("0 peac14 10.0.0.1 12345\n1 peac15 10.0.0.2 12345\n", &rma, &err)); // .rma variant
    Nodefile ext;
    CHECK(parse_nodefile_text("0 localhost 127.0.0.1 19000 0 gpu=0 gpu_bytes=2G host_bytes=256M\n"
                              "1 localhost 127.0.0.1 19001 gpu=1 gpu_bytes=512M\n",
                              &ext, &err));
    CHECK(ext.nodes[0].gpu == 0 && ext.nodes[0].gpu_bytes == (2ull << 30) && ext.nodes[0].host_bytes == (256ull << 20));
    CHECK(ext.nodes[1].gpu == 1 && ext.nodes[1].gpu_bytes == (512ull << 20));
    CHECK(resolve_rank(ext, "localhost", nullptr) == 1); // highest rank first, as the reference
    Nodefile one;
    CHECK(parse_nodefile_text("0 whatever 127.0.0.1 19000\n", &one, &err));
    CHECK(resolve_rank(one, "other", nullptr) == 0);
    Nodefile bad;
    CHECK(!parse_nodefile_text("0 a 1.2.3.4\n", &bad, &err));
    CHECK(!parse_nodefile_text("0 a 1.2.3.4 1\n0 b 1.2.3.4 2\n", &bad, &err));
    CHECK(!parse_nodefile_text("5 a 1.2.3.4 1\n", &bad, &err));
    CHECK(!parse_nodefile_text("# only comments\n", &bad, &err));
    CHECK(!parse_nodefile_text("0 a 1.2.3.4 1 bogus=3\n", &bad, &err));
    uint64_t v;
    CHECK(parse_size("4096", &v) && v == 4096);
    CHECK(parse_size("1.5G", &v) && v == (3ull << 29));
    CHECK(parse_size("64KiB", &v) && v == 65536);
    CHECK(!parse_size("abc", &v));
}

static void t_nodefile_numa() {
    Nodefile nf;
    std::string err;
    CHECK(parse_nodefile_text("0 a 127.0.0.1 9000 gpu=0 numa=1\n1 b 127.0.0.1 9001 numa=-1\n2 c 127.0.0.1 9002\n", &nf, &err));
    CHECK(nf.nodes[0].numa == 1 && nf.nodes[1].numa == -1 && nf.nodes[2].numa == -2);
    CHECK(!parse_nodefile_text("0 a 127.0.0.1 9000 numa=x\n", &nf, &err));
    CHECK(!parse_nodefile_text("0 a 127.0.0.1 9000 numa=-7\n", &nf, &err));
}

static void t_ring() {
    MpscRing<Message> q(8);
    Message m;
    CHECK(q.empty() && !q.pop(&m));
    for (int i = 0; i < 8; ++i) {
        m.init(MSG_PING);
        m.seq = i;
        CHECK(q.push(m));
    }
    CHECK(!q.push(m)); // full
    for (int i = 0; i < 8; ++i) {
        CHECK(q.pop(&m) && m.seq == (uint32_t)i);
    }
    CHECK(q.empty());
    // many producers, one consumer, nothing lost or duplicated
    MpscRing<uint64_t> r(256);
    const int P = 4, N = 20000;
    std::vector<std::thread> th;
    for (int p = 0; p < P; ++p)
        th.emplace_back([&, p] {
            for (int i = 0; i < N; ++i) {
                uint64_t v = ((uint64_t)p << 32) | (uint32_t)i;
                while (!r.push(v)) std::this_thread::yield();
            }
        });
    std::vector<int> next(P, 0);
    int got = 0;
    while (got < P * N) {
        uint64_t v;
        if (!r.pop(&v)) {
            std::this_thread::yield();
            continue;
        }
        int p = (int)(v >> 32), i = (int)(uint32_t)v;
        CHECK(p < P && i == next[p]); // per-producer FIFO order
        next[p] = i + 1;
        ++got;
    }
    for (auto &t : th) t.join();
    CHECK(r.empty());
}

static TableHeader *fresh_table(std::vector<char> &buf, uint64_t gpu, uint64_t host) {
    size_t seg = table_segment_bytes(gpu, host);
    buf.assign(seg + 4096, 0);
    // page-aligned like the shm segments the tables really live in (Region / SigExt are cache-line aligned types)
    char *base = reinterpret_cast<char *>((reinterpret_cast<uintptr_t>(buf.data()) + 4095) & ~uintptr_t(4095));
    table_format(base, seg, 0, 1, 1234, gpu, host);
    auto *h = reinterpret_cast<TableHeader *>(base);
    CHECK(table_validate(h, seg));
    return h;
}

static void t_tables() {
    std::vector<char> buf;
    TableHeader *h = fresh_table(buf, 64ull << 20, 1ull << 20); // 1024 GPU granules of 64 KiB, 256 host granules
    h->slabs_enabled.store(0);   // this suite is about whole granules; t_slabs covers the size classes below them
    CHECK(h->pool[TIER_GPU].granule == 65536 && h->pool[TIER_GPU].n_granules == 1024);
    CHECK(h->pool[TIER_HOST].granule == 4096 && h->pool[TIER_HOST].n_granules == 256);
    CHECK(default_granule(160ull << 30, TIER_GPU) == (1ull << 20)); // 160 GiB pool -> <= 2^18 granules
    BlockRef a, b, c;
    // single granules come from the top of the pool, runs from the bottom (anti-fragmentation, see t_frag)
    CHECK(pool_alloc(h, TIER_GPU, 100, 11, 0, &a) == OCM_OK && a.offset == 1023 * 65536ull && a.block_bytes == 65536);
    CHECK(pool_alloc(h, TIER_GPU, 65537, 11, 0, &b) == OCM_OK && b.offset == 0 && b.block_bytes == 131072);
    CHECK(a.id != b.id);
    CHECK(h->pool[TIER_GPU].used_bytes.load() == 3 * 65536 && h->pool[TIER_GPU].live.load() == 2);
    Region *regs = table_regions(h, TIER_GPU);
    CHECK(regs[a.region].state.load() == REGION_LIVE && regs[a.region].owner_pid == 11 && regs[a.region].bytes == 100);
    // free -> dirty -> scrub -> reuse
    CHECK(pool_free(h, TIER_GPU, a.region, a.id) == OCM_OK);
    CHECK(pool_free(h, TIER_GPU, a.region, a.id) == OCM_ESTALE);       // double free
    CHECK(pool_free(h, TIER_GPU, b.region, b.id + 77) == OCM_ESTALE);   // wrong id
    CHECK(regs[a.region].state.load() == REGION_DIRTY);
    CHECK(pool_alloc(h, TIER_GPU, 100, 12, 0, &c) == OCM_OK && c.offset != a.offset); // dirty block not reused yet
    std::vector<DirtyBlock> dirty;
    CHECK(pool_take_dirty(h, TIER_GPU, &dirty) == 1 && dirty[0].region == a.region && dirty[0].bytes == 65536);
    CHECK(regs[a.region].state.load() == REGION_SCRUBBING);
    pool_release_scrubbed(h, TIER_GPU, a.region);
    CHECK(regs[a.region].state.load() == REGION_FREE);
    // exhaustion: the rest of the pool in one-granule blocks, then the hole left by `a` is found too
    CHECK(pool_alloc(h, TIER_GPU, 128ull << 20, 1, 0, &c) == OCM_ENOMEM);
    std::vector<BlockRef> many;
    for (;;) {
        BlockRef x;
        if (pool_alloc(h, TIER_GPU, 65536, 14, 0, &x) != OCM_OK) break;
        many.push_back(x);
    }
    CHECK(many.size() == 1024 - 3);
    CHECK(regs[a.region].generation == a.generation + 1); // the scrubbed hole at the top was found too
    CHECK(pool_count_live(h, TIER_GPU) == many.size() + 2);
    CHECK(h->pool[TIER_GPU].used_bytes.load() == h->pool[TIER_GPU].bytes);
    // coalescing: free everything owned by pid 14, then one block spanning all of it fits
    h->scrub_enabled.store(0);
    uint64_t bytes = 0;
    auto pred = [](int pid, int, void *) { return pid == 14; };
    CHECK(pool_reclaim(h, TIER_GPU, pred, nullptr, &bytes) == many.size());
    CHECK(bytes == many.size() * 65536ull);
    CHECK(pool_count_live(h, TIER_GPU) == 2);
    BlockRef big;
    CHECK(pool_alloc(h, TIER_GPU, (1024 - 4) * 65536ull, 15, 0, &big) == OCM_OK && big.offset == 2 * 65536ull);
    CHECK(pool_alloc(h, TIER_GPU, 2 * 65536, 15, 0, &c) == OCM_ENOMEM);   // only the single-granule hole is left
    CHECK(pool_alloc(h, TIER_GPU, 65536, 15, 0, &c) == OCM_OK && c.offset == a.offset);
    CHECK(pool_free(h, TIER_GPU, big.region, big.id) == OCM_OK);
    // host tier independent
    BlockRef hb;
    CHECK(pool_alloc(h, TIER_HOST, 5000, 1, 0, &hb) == OCM_OK && hb.block_bytes == 8192);
    CHECK(pool_alloc(h, 7, 1, 1, 0, &hb) == OCM_EINVAL);
    // per-writer completion counters of a shared block: claimed once, summed, released with the block
    Region &hr = table_regions(h, TIER_HOST)[hb.region];
    CHECK(hr.ext.load() == 0 && region_signal_total(h, hr) == 0);
    SigExt *x = sigext_claim(h, TIER_HOST, hb.region);
    CHECK(x != nullptr && hr.ext.load() != 0 && x->id == hb.id);
    CHECK(sigext_claim(h, TIER_HOST, hb.region) == x);                     // idempotent
    hr.signal.fetch_add(2);
    hr.signal_cpu.fetch_add(3);
    x->slot[1].fetch_add(5);
    x->slot[kSigCpuSlot].fetch_add(7);
    CHECK(region_signal_total(h, hr) == 17);
    BlockRef hb2;
    CHECK(pool_alloc(h, TIER_HOST, 100, 1, 0, &hb2) == OCM_OK);
    SigExt *y = sigext_claim(h, TIER_HOST, hb2.region);
    CHECK(y != nullptr && y != x);
    CHECK(pool_free(h, TIER_HOST, hb.region, hb.id) == OCM_OK);
    CHECK(x->owner.load() == 0 && hr.ext.load() == 0 && y->owner.load() != 0);   // record is free again
    BlockRef hb3;
    CHECK(pool_alloc(h, TIER_HOST, 100, 1, 0, &hb3) == OCM_OK);
    CHECK(sigext_claim(h, TIER_HOST, hb3.region) == x && x->slot[1].load() == 0);   // reused, counters cleared
}

// Size classes below the granule (tables.hpp, "slabs"): small requests share a granule, every sub-block is an
// allocation of its own (record, id, signals, share), released blocks are zeroed before reuse, empty slabs go back.
static void t_slabs() {
    std::vector<char> buf;
    TableHeader *h = fresh_table(buf, 64ull << 20, 1ull << 20);
    PoolDesc &p = h->pool[TIER_GPU];
    CHECK(p.n_slabs > 0 && p.n_slabs % kSlabClasses == 0 && h->slabs_enabled.load() == 1);
    // 100 x 4 KiB: 16 per 64 KiB granule -> 7 granules instead of 100
    std::vector<BlockRef> small;
    std::set<uint64_t> offs;
    for (int i = 0; i < 100; ++i) {
        BlockRef b;
        CHECK(pool_alloc(h, TIER_GPU, 4096, 21, 0, &b) == OCM_OK);
        CHECK(region_is_sub(b.region) && b.block_bytes == 4096 && b.offset % 4096 == 0 && offs.insert(b.offset).second);
        Region *r = table_region_rec(h, TIER_GPU, b.region);
        CHECK(r && r->state.load() == REGION_LIVE && r->owner_pid == 21 && r->bytes == 4096 && r->id == b.id);
        CHECK(table_region_offset(h, TIER_GPU, b.region) == b.offset);
        small.push_back(b);
    }
    CHECK(p.used_bytes.load() == 7 * 65536ull && p.live.load() == 100 && pool_count_live(h, TIER_GPU) == 100);
    CHECK(pool_payload_bytes(h, TIER_GPU) == 100 * 4096ull);
    TableAudit au;
    CHECK(pool_audit(h, TIER_GPU, &au) == 0 && au.live_records == 100);
    // other classes: 1000 B -> 1 KiB blocks, 20 KiB -> 32 KiB blocks; more than half a granule takes granules
    BlockRef k1, k20, k33;
    CHECK(pool_alloc(h, TIER_GPU, 1000, 21, 0, &k1) == OCM_OK && region_is_sub(k1.region) && k1.block_bytes == 1024);
    CHECK(pool_alloc(h, TIER_GPU, 20 << 10, 21, 0, &k20) == OCM_OK && region_is_sub(k20.region) && k20.block_bytes == 32768);
    CHECK(pool_alloc(h, TIER_GPU, 33 << 10, 21, 0, &k33) == OCM_OK && !region_is_sub(k33.region) && k33.block_bytes == 65536);
    CHECK(slab_class(sub_slab(k1.region)) == 5 && slab_class(sub_slab(k20.region)) == 0);
    // release: double free and wrong id are refused; the block is zeroed (dirty -> scrubbing -> free) before reuse
    BlockRef v = small[37];
    CHECK(pool_free(h, TIER_GPU, v.region, v.id + 1) == OCM_ESTALE);
    CHECK(pool_free(h, TIER_GPU, v.region, v.id) == OCM_OK);
    CHECK(pool_free(h, TIER_GPU, v.region, v.id) == OCM_ESTALE);
    CHECK(table_region_rec(h, TIER_GPU, v.region)->state.load() == REGION_DIRTY && pool_has_dirty(h, TIER_GPU));
    BlockRef w;
    CHECK(pool_alloc(h, TIER_GPU, 4096, 22, 0, &w) == OCM_OK && w.offset != v.offset);   // not handed out while dirty
    std::vector<DirtyBlock> dirty;
    CHECK(pool_take_dirty(h, TIER_GPU, &dirty) == 1 && dirty[0].region == v.region && dirty[0].offset == v.offset && dirty[0].bytes == 4096);
    pool_release_scrubbed(h, TIER_GPU, v.region);
    CHECK(!pool_has_dirty(h, TIER_GPU) && table_region_rec(h, TIER_GPU, v.region)->state.load() == REGION_FREE);
    BlockRef again;
    bool reused = false;
    std::vector<BlockRef> extra;
    for (int i = 0; i < 20 && !reused; ++i) {
        CHECK(pool_alloc(h, TIER_GPU, 4096, 22, 0, &again) == OCM_OK);
        reused = again.offset == v.offset;
        extra.push_back(again);
    }
    CHECK(reused && again.generation == v.generation + 1 && again.id != v.id);
    // a sub-block is a full allocation: completion counters and the share extension work on it
    Region *wr = table_region_rec(h, TIER_GPU, w.region);
    SigExt *x = sigext_claim(h, TIER_GPU, w.region);
    CHECK(x && wr->ext.load() != 0 && x->id == w.id);
    wr->signal.fetch_add(3);
    x->slot[2].fetch_add(4);
    CHECK(region_signal_total(h, *wr) == 7);
    CHECK(pool_free(h, TIER_GPU, w.region, w.id) == OCM_OK && x->owner.load() == 0);
    // dead owners: sub-blocks are reclaimed like granule blocks; the granules behind slabs belong to nobody
    uint64_t bytes = 0;
    std::vector<BlockRef> got;
    auto pred = [](int pid, int, void *) { return pid == 21; };
    const size_t n21 = 99 + 3;       // the 4 KiB blocks still held by pid 21, k1, k20, k33
    CHECK(pool_reclaim(h, TIER_GPU, pred, nullptr, &bytes, &got) == n21 && got.size() == n21);
    CHECK(bytes == 99 * 4096ull + 1024 + 32768 + 65536);
    // everything released and scrubbed: the slabs retire (one per class may stay cached), the audit still agrees
    for (auto &b : extra) CHECK(pool_free(h, TIER_GPU, b.region, b.id) == OCM_OK);
    for (int round = 0; round < 4; ++round) {
        dirty.clear();
        pool_take_dirty(h, TIER_GPU, &dirty);
        for (auto &d : dirty) pool_release_scrubbed(h, TIER_GPU, d.region);
    }
    CHECK(pool_count_live(h, TIER_GPU) == 0 && p.live.load() == 0 && p.sub_live.load() == 0 && p.sub_bytes.load() == 0);
    CHECK(p.used_bytes.load() <= 3 * 65536ull && !pool_has_dirty(h, TIER_GPU));
    CHECK(pool_audit(h, TIER_GPU, &au) == 0);
    // a request for the whole pool makes the allocator give the cached empty slabs back (they are zeroed first)
    BlockRef whole;
    CHECK(pool_alloc(h, TIER_GPU, p.bytes, 30, 0, &whole) == OCM_ENOMEM && pool_has_dirty(h, TIER_GPU));
    dirty.clear();
    pool_take_dirty(h, TIER_GPU, &dirty);
    for (auto &d : dirty) pool_release_scrubbed(h, TIER_GPU, d.region);
    CHECK(p.used_bytes.load() == 0);
    CHECK(pool_alloc(h, TIER_GPU, p.bytes, 30, 0, &whole) == OCM_OK && pool_free(h, TIER_GPU, whole.region, whole.id) == OCM_OK);
    dirty.clear();
    pool_take_dirty(h, TIER_GPU, &dirty);
    for (auto &d : dirty) pool_release_scrubbed(h, TIER_GPU, d.region);
    p.used_bytes.fetch_add(5 * 65536ull);      // what a misbehaving client could do
    CHECK(pool_audit(h, TIER_GPU, &au) > 0);
    p.used_bytes.fetch_sub(5 * 65536ull);
    // slabs off: the same request takes a whole granule again
    h->slabs_enabled.store(0);
    CHECK(pool_alloc(h, TIER_GPU, 4096, 23, 0, &w) == OCM_OK && !region_is_sub(w.region) && w.block_bytes == 65536);
    h->slabs_enabled.store(1);
    // host tier: 4 KiB granules -> 256-byte .. 2 KiB classes
    BlockRef hb;
    CHECK(pool_alloc(h, TIER_HOST, 200, 1, 0, &hb) == OCM_OK && region_is_sub(hb.region) && hb.block_bytes == 256);
    CHECK(pool_alloc(h, TIER_HOST, 3000, 1, 0, &hb) == OCM_OK && !region_is_sub(hb.region) && hb.block_bytes == 4096);

    // concurrent small allocations and releases from 8 threads: no byte handed out twice, nothing leaks
    std::vector<char> buf2;
    TableHeader *g = fresh_table(buf2, 32ull << 20, 0);
    g->scrub_enabled.store(0);
    const int T = 8, ITER = 20000;
    std::vector<std::atomic<int>> owner((32ull << 20) / 1024);
    for (auto &o : owner) o.store(0);
    std::atomic<int> errors{0};
    std::vector<std::thread> th;
    for (int t = 0; t < T; ++t)
        th.emplace_back([&, t] {
            std::vector<BlockRef> mine;
            uint64_t rng = 0x9E3779B97F4A7C15ull * (t + 1);
            for (int i = 0; i < ITER; ++i) {
                rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17;
                if (mine.size() < 24 && (rng & 1)) {
                    static const uint64_t sizes[] = {700, 1024, 3000, 4096, 9000, 16384, 30000, 70000};
                    BlockRef b;
                    if (pool_alloc(g, TIER_GPU, sizes[(rng >> 8) % 8], 100 + t, 0, &b) == OCM_OK) {
                        for (uint64_t k = b.offset / 1024; k < (b.offset + b.block_bytes) / 1024; ++k)
                            if (owner[k].exchange(t + 1) != 0) errors++;
                        mine.push_back(b);
                    }
                } else if (!mine.empty()) {
                    const size_t pick = (rng >> 16) % mine.size();
                    BlockRef b = mine[pick];
                    mine.erase(mine.begin() + pick);
                    for (uint64_t k = b.offset / 1024; k < (b.offset + b.block_bytes) / 1024; ++k)
                        if (owner[k].exchange(0) != t + 1) errors++;
                    if (pool_free(g, TIER_GPU, b.region, b.id) != OCM_OK) errors++;
                }
            }
            for (auto &b : mine) {
                for (uint64_t k = b.offset / 1024; k < (b.offset + b.block_bytes) / 1024; ++k)
                    if (owner[k].exchange(0) != t + 1) errors++;
                if (pool_free(g, TIER_GPU, b.region, b.id) != OCM_OK) errors++;
            }
        });
    for (auto &t : th) t.join();
    CHECK(errors.load() == 0);
    CHECK(pool_count_live(g, TIER_GPU) == 0 && g->pool[TIER_GPU].live.load() == 0 && g->pool[TIER_GPU].sub_bytes.load() == 0);
    CHECK(pool_audit(g, TIER_GPU, &au) == 0);
    CHECK(g->pool[TIER_GPU].used_bytes.load() <= (uint64_t)kSlabClasses * 2 * 65536);   // at most the cached slabs remain
}

static void t_tables_mt() {
    // concurrent alloc/free from several threads: no block handed out twice
    std::vector<char> buf;
    TableHeader *h = fresh_table(buf, 256ull << 20, 0);
    h->scrub_enabled.store(0);
    const int T = 8, ITER = 20000;
    std::vector<std::atomic<int>> owner(h->pool[TIER_GPU].n_granules);
    for (auto &o : owner) o.store(0);
    std::atomic<int> errors{0};
    std::vector<std::thread> th;
    for (int t = 0; t < T; ++t)
        th.emplace_back([&, t] {
            std::vector<BlockRef> mine;
            uint64_t rng = 88172645463325252ull + t;
            for (int i = 0; i < ITER; ++i) {
                rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17;
                if (mine.size() < 16 && (rng & 1)) {
                    BlockRef b;
                    uint64_t sz = ((rng >> 8) % 6 + 1) * 65536;
                    if (pool_alloc(h, TIER_GPU, sz, 100 + t, 0, &b) == OCM_OK) {
                        uint32_t g = (uint32_t)(b.block_bytes / 65536);
                        for (uint32_t k = 0; k < g; ++k)
                            if (owner[b.region + k].exchange(t + 1) != 0) errors++;
                        mine.push_back(b);
                    }
                } else if (!mine.empty()) {
                    BlockRef b = mine.back();
                    mine.pop_back();
                    uint32_t g = (uint32_t)(b.block_bytes / 65536);
                    for (uint32_t k = 0; k < g; ++k)
                        if (owner[b.region + k].exchange(0) != t + 1) errors++;
                    if (pool_free(h, TIER_GPU, b.region, b.id) != OCM_OK) errors++;
                }
            }
            for (auto &b : mine) {
                uint32_t g = (uint32_t)(b.block_bytes / 65536);
                for (uint32_t k = 0; k < g; ++k) owner[b.region + k].store(0);
                if (pool_free(h, TIER_GPU, b.region, b.id) != OCM_OK) errors++;
            }
        });
    for (auto &t : th) t.join();
    CHECK(errors.load() == 0);
    CHECK(h->pool[TIER_GPU].live.load() == 0 && h->pool[TIER_GPU].used_bytes.load() == 0);
    CHECK(h->pool[TIER_GPU].total_allocs.load() == h->pool[TIER_GPU].total_frees.load());
    CHECK(pool_count_live(h, TIER_GPU) == 0);
}

static void t_governor() {
    std::vector<char> buf(sizeof(ClusterTable), 0);
    auto *c = reinterpret_cast<ClusterTable *>(buf.data());
    cluster_format(c, 4);
    NodeConfig cfg;
    memset(&cfg, 0, sizeof cfg);
    uint64_t pools[TIER_COUNT] = {1ull << 30, 1ull << 28};
    for (int r = 0; r < 3; ++r) { // rank 3 never joins
        cfg.gpu_index = r;
        CHECK(governor_add_node(c, r, cfg, pools) == OCM_OK);
    }
    CHECK(governor_joined(c) == 3);
    AllocRequest rq;
    memset(&rq, 0, sizeof rq);
    Allocation a;
    rq.orig_rank = 1; rq.bytes = 1 << 20; rq.type = ALLOC_MEM_REMOTE_GPU; rq.policy = OCM_POLICY_DEFAULT;
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.remote_rank == 2 && a.tier == TIER_GPU); // neighbour (r+1)%N
    rq.orig_rank = 2;
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.remote_rank == 0); // rank 3 absent -> next neighbour
    rq.type = ALLOC_MEM_HOST;
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.remote_rank == 2 && a.tier == TIER_HOST); // local kinds stay
    rq.type = ALLOC_MEM_RDMA; rq.orig_rank = 0;
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.remote_rank == 1 && a.tier == TIER_HOST && a.type == ALLOC_MEM_RDMA);
    rq.type = ALLOC_MEM_REMOTE_GPU; rq.policy = OCM_POLICY_EXPLICIT; rq.remote_rank = 2;
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.remote_rank == 2);
    rq.remote_rank = 3;
    CHECK(governor_place(c, rq, &a) == OCM_ENOMEM); // not joined
    rq.remote_rank = 9;
    CHECK(governor_place(c, rq, &a) == OCM_EINVAL);
    // capacity: fill rank 1's GPU pool, then the request spills to the host tier of rank 1, then moves on
    rq.policy = OCM_POLICY_EXPLICIT; rq.remote_rank = 1; rq.bytes = 1ull << 30;
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.tier == TIER_GPU);
    rq.bytes = 1ull << 27;
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.tier == TIER_HOST && a.type == ALLOC_MEM_REMOTE_HOST);
    governor_release(c, 1, TIER_GPU, 1ull << 30);
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.tier == TIER_GPU);
    // least loaded picks the emptiest pool
    rq.policy = OCM_POLICY_LEAST_LOADED; rq.orig_rank = 0; rq.bytes = 1 << 20;
    uint64_t u0 = c->node[0].used[TIER_GPU].load(), u1 = c->node[1].used[TIER_GPU].load(), u2 = c->node[2].used[TIER_GPU].load();
    CHECK(governor_place(c, rq, &a) == OCM_OK);
    uint64_t mn = std::min(u0, std::min(u1, u2));
    uint64_t chosen = a.remote_rank == 0 ? u0 : a.remote_rank == 1 ? u1 : u2;
    CHECK(chosen == mn);
    // round robin cycles over joined ranks
    rq.policy = OCM_POLICY_ROUND_ROBIN;
    std::set<int> seen;
    for (int i = 0; i < 6; ++i) {
        CHECK(governor_place(c, rq, &a) == OCM_OK);
        seen.insert(a.remote_rank);
    }
    CHECK(seen.size() == 3);
    // single-rank cluster keeps remote requests on rank 0 (the reference downgrades to host, src/alloc.c:82-83)
    std::vector<char> b1(sizeof(ClusterTable), 0);
    auto *c1 = reinterpret_cast<ClusterTable *>(b1.data());
    cluster_format(c1, 1);
    CHECK(governor_add_node(c1, 0, cfg, pools) == OCM_OK);
    rq.policy = OCM_POLICY_DEFAULT; rq.orig_rank = 0;
    CHECK(governor_place(c1, rq, &a) == OCM_OK && a.remote_rank == 0 && a.tier == TIER_GPU);
    // liveness: a rank whose heartbeat is stale is skipped; explicit requests for it report a dead daemon
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    const uint64_t now = (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec;
    CHECK(governor_rank_alive(c, 2) && !governor_rank_alive(c, 3));          // never beaten = alive; not joined = not
    c->node[2].heartbeat_ns.store(now - 5000000000ull);                       // 5 s ago
    c->node[0].heartbeat_ns.store(now);
    CHECK(!governor_rank_alive(c, 2) && governor_rank_alive(c, 0));
    rq.policy = OCM_POLICY_DEFAULT; rq.orig_rank = 1; rq.bytes = 1 << 20; rq.type = ALLOC_MEM_REMOTE_GPU;
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.remote_rank == 0);        // neighbour 2 is dead -> next one
    rq.policy = OCM_POLICY_EXPLICIT; rq.remote_rank = 2;
    CHECK(governor_place(c, rq, &a) == OCM_ENODAEMON);
    cfg.daemon_pid = 4242;                                                     // a new daemon process re-joins as rank 2
    c->node[2].used[TIER_GPU].store(12345);
    CHECK(governor_add_node(c, 2, cfg, pools) == OCM_OK && governor_rank_alive(c, 2));
    CHECK(c->node[2].used[TIER_GPU].load() == 0);                             // new incarnation: accounting restarts
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.remote_rank == 2);

    // multicast ranges: owner tracking and reclamation after the owner's death (every rank zeroes its slice first)
    c->mc_n_granules = 64;
    c->mc_bytes = 64 * kMcGranule;
    c->mc_state.store(MC_READY);
    const int64_t m0 = mc_range_alloc(c, 3 * kMcGranule, 777, 1);
    const int64_t m1 = mc_range_alloc(c, kMcGranule, 778, 0);
    CHECK(m0 == 0 && m1 == (int64_t)(3 * kMcGranule));
    CHECK(mc_owner_pid(c->mc_owner[0].load()) == 777 && mc_owner_granules(c->mc_owner[0].load()) == 3 &&
          mc_owner_rank(c->mc_owner[0].load()) == 1 && mc_owner_state(c->mc_owner[0].load()) == MC_RANGE_OWNED);
    auto dead777 = [](int pid) { return pid == 777; };
    CHECK(mc_mark_dead(c, 0, dead777) == 0);                                  // not rank 0's application
    CHECK(mc_mark_dead(c, 1, dead777) == 1 && mc_mark_dead(c, 1, dead777) == 0);
    std::vector<std::pair<uint32_t, uint32_t>> dr;
    int freed = 0;
    for (int r = 0; r < 4; ++r) {                                             // n_ranks = 4: all four must scrub
        dr.clear();
        CHECK(mc_dead_ranges(c, r, &dr) == 1 && dr[0].first == 0 && dr[0].second == 3);
        CHECK(mc_range_alloc(c, 60 * kMcGranule + 1, 1, 0) == -1 || r == 3);  // still taken while scrubbing
        freed += mc_range_scrubbed(c, r, 0) ? 1 : 0;
        dr.clear();
        CHECK(mc_dead_ranges(c, r, &dr) == 0);
    }
    CHECK(freed == 1 && c->mc_owner[0].load() == 0);
    CHECK(mc_range_alloc(c, 3 * kMcGranule, 779, 2) == 0);                    // the range is back
    mc_range_free(c, 0, 3 * kMcGranule);
    mc_range_free(c, (uint64_t)m1, kMcGranule);
    CHECK(c->mc_owner[3].load() == 0);
}

// Topology discovery (common/topo.hpp): list parsing, sysfs lookup on a fake tree, binding helpers, and the
// placement rule built on the peer rows the daemons report.
static void t_topo() {
    using namespace ocm::topo;
    CHECK((parse_cpulist("0-3,8,10-11\n") == std::vector<int>{0, 1, 2, 3, 8, 10, 11}));
    CHECK(parse_cpulist("").empty() && parse_cpulist("garbage").empty());
    CHECK((parse_cpulist("5") == std::vector<int>{5}) && (parse_cpulist("7-") == std::vector<int>{7}));
    std::vector<int> nodes = numa_nodes();
    CHECK(!nodes.empty());
    // a fake sysfs with two GPUs on different nodes; the CUDA spelling of the address (8-digit domain, upper case) resolves too
    char tmpl[] = "/tmp/ocm_topo_XXXXXX";
    std::string root = mkdtemp(tmpl);
    for (auto kv : {std::pair<const char *, const char *>{"0000:1b:00.0", "0"}, {"0000:9c:00.0", "1"}, {"0000:aa:00.0", "-1"}}) {
        std::string d = root + "/bus/pci/devices/" + kv.first;
        std::string cmd = "mkdir -p " + d;
        CHECK(system(cmd.c_str()) == 0);
        std::ofstream(d + "/numa_node") << kv.second << "\n";
    }
    CHECK(pci_numa_node("0000:1b:00.0", root) == 0);
    CHECK(pci_numa_node("00000000:9C:00.0", root) == 1);
    CHECK(pci_numa_node("0000:aa:00.0", root) == -1);
    CHECK(pci_numa_node("0000:ff:00.0", root) == -1);
    std::string rm = "rm -rf " + root;
    CHECK(system(rm.c_str()) == 0);
    // binding: unknown nodes are refused, the first real node works and leaves us with a non-empty CPU set
    CHECK(!bind_thread_to_node(-1) && !bind_thread_to_node(100000));
    std::thread([&] { CHECK(bind_thread_to_node(nodes[0]) || node_cpus(nodes[0]).empty()); }).join();
    std::vector<char> page(1 << 16, 7);
    touch_range(page.data(), page.size(), nodes[0], 3);
    CHECK(page[0] == 7 && page[page.size() - 1] == 7);   // placement touches never change the content
    prefer_node_for_range(page.data(), 0, 0);            // degenerate arguments are refused, not crashed on

    // placement: a far side in HBM only on GPUs the requester's GPU reaches; otherwise that rank's host tier
    std::vector<char> buf(sizeof(ClusterTable), 0);
    auto *c = reinterpret_cast<ClusterTable *>(buf.data());
    cluster_format(c, 3);
    uint64_t pools[TIER_COUNT] = {1ull << 30, 1ull << 28};
    NodeConfig cfg;
    for (int r = 0; r < 3; ++r) {
        memset(&cfg, 0, sizeof cfg);
        cfg.gpu_index = r;
        cfg.numa_node = (int16_t)(r / 2);
        cfg.peer_mask = r == 2 ? 0x4 : 0x3;      // GPUs 0 and 1 see each other; GPU 2 is an island
        cfg.caps = NODE_CAP_VMM_FD;
        snprintf(cfg.pci_bdf, sizeof cfg.pci_bdf, "0000:%02x:00.0", 0x1b + r);
        CHECK(governor_add_node(c, r, cfg, pools) == OCM_OK);
    }
    CHECK(c->node[1].numa_node == 0 && c->node[2].numa_node == 1 && c->node[2].peer_mask == 0x4 &&
          !strcmp(c->node[1].pci_bdf, "0000:1c:00.0"));
    CHECK(governor_gpu_reachable(c, 0, 1) && governor_gpu_reachable(c, 1, 0) && governor_gpu_reachable(c, 2, 2));
    CHECK(!governor_gpu_reachable(c, 0, 2) && !governor_gpu_reachable(c, 2, 0) && !governor_gpu_reachable(c, 0, 7));
    AllocRequest rq;
    memset(&rq, 0, sizeof rq);
    Allocation a;
    rq.bytes = 1 << 20; rq.type = ALLOC_MEM_REMOTE_GPU; rq.policy = OCM_POLICY_DEFAULT;
    rq.orig_rank = 0;
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.remote_rank == 1 && a.tier == TIER_GPU);
    rq.orig_rank = 1;    // the neighbour is the island: its pinned host pool is what rank 1 can use there
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.remote_rank == 2 && a.tier == TIER_HOST && a.type == ALLOC_MEM_REMOTE_HOST);
    rq.policy = OCM_POLICY_EXPLICIT; rq.remote_rank = 2; rq.orig_rank = 2;
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.tier == TIER_GPU);   // its own GPU is always reachable
    // daemons that report no peer row (peer_mask 0) keep the old behaviour: decided when the pool is mapped
    c->node[1].peer_mask = 0;
    rq.policy = OCM_POLICY_DEFAULT; rq.orig_rank = 1;
    CHECK(governor_place(c, rq, &a) == OCM_OK && a.remote_rank == 2 && a.tier == TIER_GPU);
}

static void t_pmsg() {
    // equivalent of the reference's pmsg_daemon / pmsg_client pair (test/pmsg_daemon.c, test/pmsg_client.c)
    setenv("OCM_NS", ("ut" + std::to_string(getpid())).c_str(), 1);
    struct Toy { int type; int pid; char text[256]; };
    Mailbox daemon;
    CHECK(daemon.open(mq_daemon_name(0), sizeof(Toy)) == 0);
    CHECK(daemon.pending() == 0);
    pid_t child = fork();
    if (child == 0) {
        Mailbox me;
        if (me.open(mq_app_name(getpid()), sizeof(Toy))) _exit(10);
        if (me.attach(mq_daemon_name(0))) _exit(11);
        Toy t = {1, (int)getpid(), "hello"};
        if (me.send(mq_daemon_name(0), &t)) _exit(12);
        Toy r;
        if (me.recv(&r, 5000) != 1) _exit(13);
        if (strcmp(r.text, "nice to meet you")) _exit(14);
        me.detach(mq_daemon_name(0));
        me.close();
        _exit(0);
    }
    Toy t;
    CHECK(daemon.recv(&t, 5000) == 1);
    CHECK(t.pid == child && !strcmp(t.text, "hello"));
    CHECK(daemon.attach(mq_app_name(t.pid)) == 0);
    Toy r = {2, (int)getpid(), "nice to meet you"};
    CHECK(daemon.send(mq_app_name(t.pid), &r) == 0);
    int st = 0;
    waitpid(child, &st, 0);
    CHECK(WIFEXITED(st) && WEXITSTATUS(st) == 0);
    CHECK(daemon.detach(mq_app_name(t.pid)) == 0);
    CHECK(daemon.recv(&t, 0) == 0);       // non-blocking poll, empty
    CHECK(daemon.recv(&t, 20) == 0);      // timed wait, empty
    CHECK(daemon.send("/ocm_mq_nobody", &r) == -ENOTCONN);
    // a second owner of the same name steals a stale queue only when asked to
    Mailbox dup;
    CHECK(dup.open(mq_daemon_name(0), sizeof(Toy), 8, false) == -EEXIST);
    // queue depth back-pressure: 8 messages fit, the 9th times out
    Mailbox self;
    CHECK(self.attach(mq_daemon_name(0)) == 0);
    for (int i = 0; i < 8; ++i) CHECK(self.send(mq_daemon_name(0), &r, 0) == 0);
    CHECK(self.send(mq_daemon_name(0), &r, 30) == -ETIMEDOUT);
    CHECK(daemon.pending() == 8);
    CHECK(daemon.close() == 0);
    // C wrappers with the reference's signatures
    setenv("OCM_RANK", "0", 1);
    CHECK(pmsg_init(sizeof(Toy)) == 0);
    CHECK(pmsg_open(PMSG_DAEMON_PID) == 0);
    CHECK(pmsg_attach(PMSG_DAEMON_PID) == 0);
    CHECK(pmsg_send(PMSG_DAEMON_PID, &r) == 0);
    CHECK(pmsg_pending() == 1);
    CHECK(pmsg_recv(&t, true) == 0 && t.type == 2);
    CHECK(pmsg_recv(&t, false) != 0);
    CHECK(pmsg_detach(PMSG_DAEMON_PID) == 0);
    CHECK(pmsg_close() == 0);
}

static void t_sock() {
    std::string name = "ocm_ut_" + std::to_string(getpid());
    int l = uds_listen(name);
    CHECK(l >= 0);
    int pipefd[2];
    CHECK(pipe(pipefd) == 0);
    pid_t child = fork();
    if (child == 0) {
        int s = uds_connect(name, 1000);
        if (s < 0) _exit(20);
        uint64_t payload = 0;
        int fd = -1;
        if (uds_recv_fd(s, &payload, sizeof payload, &fd, 2000) != 1) _exit(21);
        if (payload != 0xfeedfacecafebeefull || fd < 0) _exit(22);
        const char msg[] = "through the passed fd";
        if (write(fd, msg, sizeof msg) != (ssize_t)sizeof msg) _exit(23);
        _exit(0);
    }
    int s = accept(l, nullptr, nullptr);
    CHECK(s >= 0);
    uint64_t payload = 0xfeedfacecafebeefull;
    CHECK(uds_send_fd(s, &payload, sizeof payload, pipefd[1]) == 1);
    char buf[64] = {0};
    CHECK(read(pipefd[0], buf, sizeof buf) > 0 && !strcmp(buf, "through the passed fd"));
    int st = 0;
    waitpid(child, &st, 0);
    CHECK(WIFEXITED(st) && WEXITSTATUS(st) == 0);
    sock_close(s);
    sock_close(l);
    // TCP round trip of one wire message, 1/0/-errno tri-state
    int port = 20000 + (getpid() % 20000);
    int tl = tcp_listen("127.0.0.1", port);
    CHECK(tl >= 0);
    int cfd = tcp_connect("127.0.0.1", port, 1000);
    CHECK(cfd >= 0);
    int afd = tcp_accept(tl);
    CHECK(afd >= 0);
    Message m, r;
    m.init(MSG_PING);
    m.seq = 42;
    CHECK(sock_put(cfd, &m, sizeof m) == 1);
    CHECK(sock_get(afd, &r, sizeof r, 1000) == 1 && r.valid() && r.seq == 42);
    CHECK(sock_get(afd, &r, sizeof r, 20) == -ETIMEDOUT);
    sock_close(cfd);
    CHECK(sock_get(afd, &r, sizeof r, 1000) == 0); // peer closed
    sock_close(afd);
    sock_close(tl);
    CHECK(tcp_connect("127.0.0.1", port, 200) < 0);
}

static void t_oracle() {
    // e4m3: every finite code point round-trips; rounding is to nearest even; saturation at 448
    for (int v = 0; v < 256; ++v) {
        if ((v & 0x7f) == 0x7f) continue;
        float f = ocm_oracle_e4m3_to_f32((uint8_t)v);
        uint8_t back = ocm_oracle_f32_to_e4m3(f);
        CHECK(back == v || (f == 0.0f && (back & 0x7f) == 0));
    }
    CHECK(ocm_oracle_e4m3_to_f32(0x7e) == 448.0f && ocm_oracle_e4m3_to_f32(0x01) == ldexpf(1.0f, -9));
    CHECK(ocm_oracle_f32_to_e4m3(1e9f) == 0x7e && ocm_oracle_f32_to_e4m3(-1e9f) == 0xfe);
    CHECK(ocm_oracle_f32_to_e4m3(17.0f) == ocm_oracle_f32_to_e4m3(16.0f));   // tie -> even (16 | 18)
    CHECK(ocm_oracle_f32_to_e4m3(19.0f) == ocm_oracle_f32_to_e4m3(20.0f));   // tie -> even (18 | 20)
    CHECK(ocm_oracle_f32_to_e4m3(465.0f) == 0x7e);
    CHECK(ocm_oracle_f32_to_bf16(1.0f) == 0x3f80 && ocm_oracle_bf16_to_f32(0x3f80) == 1.0f);
    CHECK(ocm_oracle_f32_to_bf16(1.00390625f) == 0x3f80);                    // tie -> even
    CHECK(ocm_oracle_f32_to_bf16(1.01171875f) == 0x3f82);                    // tie -> even (odd rounds up)
    // fill / verify / checksum agree with each other
    std::vector<uint8_t> b(1003);
    ocm_oracle_fill(b.data(), b.size(), 0x0807060504030201ull);
    CHECK(b[0] == 1 && b[7] == 8 && b[8] == 1 && b[1002] == (1002 % 8) + 1);
    CHECK(ocm_oracle_verify(b.data(), b.size(), 0x0807060504030201ull) == 0);
    b[500] ^= 0xff;
    CHECK(ocm_oracle_verify(b.data(), b.size(), 0x0807060504030201ull) == 1);
    uint64_t w[2] = {5, 7};
    CHECK(ocm_oracle_checksum(w, 16) == 12 && ocm_oracle_checksum(w, 9) == 5 + 7);
    // strided quantise -> dequantise stays within the e4m3 step (2^-3 relative, half of it after RNE)
    const int R = 6, E = 96;
    std::vector<float> src(R * 128), deq(R * E);
    std::vector<uint8_t> q(R * E);
    std::vector<float> sc(R * 3);
    for (size_t i = 0; i < src.size(); ++i) src[i] = sinf((float)i * 0.37f) * (1.0f + (float)(i % 7));
    ocm_oracle_strided d;
    memset(&d, 0, sizeof d);
    d.src = src.data(); d.dst = q.data(); d.row_bytes = E * 4; d.n[0] = R; d.n[1] = d.n[2] = 1;
    d.src_stride[0] = 128 * 4; d.dst_stride[0] = E; d.src_dtype = 1; d.dst_dtype = 3; d.scale_block = 32; d.scales = sc.data();
    CHECK(ocm_oracle_strided_copy(&d) == 0);
    ocm_oracle_strided u = d;
    u.src = q.data(); u.dst = deq.data(); u.row_bytes = E; u.src_stride[0] = E; u.dst_stride[0] = E * 4;
    u.src_dtype = 3; u.dst_dtype = 1;
    CHECK(ocm_oracle_strided_copy(&u) == 0);
    for (int r = 0; r < R; ++r)
        for (int e = 0; e < E; ++e) {
            float x = src[r * 128 + e], y = deq[r * E + e], s = sc[r * 3 + e / 32];
            CHECK(fabsf(x - y) <= 0.0625f * fabsf(x) + s * ldexpf(1.0f, -10) + 1e-12f);
        }
    // GUPS is an involution: applying the same stream twice restores the table
    std::vector<uint64_t> tab(1 << 10, 3);
    ocm_oracle_gups(tab.data(), tab.size(), 5000, 99, 128);
    bool changed = false;
    for (auto x : tab) changed |= x != 3;
    CHECK(changed);
    ocm_oracle_gups(tab.data(), tab.size(), 5000, 99, 128);
    for (auto x : tab) CHECK(x == 3);
}

// Placement by run length (tables.cpp granule_alloc) in a pool whose granule count is not a multiple of 64: the bits
// past the end of the bitmap's last word are never handed out, singles descend from the top, short runs rotate, long
// runs pack from the bottom and the lowest-free bound follows frees.
static void t_placement() {
    std::vector<char> buf;
    TableHeader *h = fresh_table(buf, (1000ull << 16) + (29ull << 16), 1ull << 20);   // 1029 granules of 64 KiB
    h->slabs_enabled.store(0);
    PoolDesc &p = h->pool[TIER_GPU];
    CHECK(p.n_granules == 1029 && p.granule == 65536);
    BlockRef s1, s2, r3, l20, l30;
    CHECK(pool_alloc(h, TIER_GPU, 1, 5, 0, &s1) == OCM_OK && s1.offset == 1028 * 65536ull);
    CHECK(pool_alloc(h, TIER_GPU, 65536, 5, 0, &s2) == OCM_OK && s2.offset == 1027 * 65536ull);
    CHECK(pool_alloc(h, TIER_GPU, 3 * 65536, 5, 0, &r3) == OCM_OK && r3.offset == 0);                 // short run: rotating start
    CHECK(pool_alloc(h, TIER_GPU, 20 * 65536, 5, 0, &l20) == OCM_OK && l20.offset == 3 * 65536ull);   // long run: lowest free
    CHECK(pool_alloc(h, TIER_GPU, 30 * 65536, 5, 0, &l30) == OCM_OK && l30.offset == 23 * 65536ull);
    // free the first long run (no scrubbing in this suite: the bits clear at once): the next long run goes back there
    h->scrub_enabled.store(0);
    CHECK(pool_free(h, TIER_GPU, l20.region, l20.id) == OCM_OK);
    BlockRef l16, l17;
    CHECK(pool_alloc(h, TIER_GPU, 16 * 65536, 5, 0, &l16) == OCM_OK && l16.offset == 3 * 65536ull);
    CHECK(pool_alloc(h, TIER_GPU, 17 * 65536, 5, 0, &l17) == OCM_OK && l17.offset == 53 * 65536ull);  // hole of 4 is too small
    // everything that is left, one granule at a time: exactly the free granules, none past the end
    size_t singles = 0;
    uint64_t lowest = ~0ull;
    for (;;) {
        BlockRef x;
        if (pool_alloc(h, TIER_GPU, 65536, 6, 0, &x) != OCM_OK) break;
        CHECK(x.offset / 65536 < 1029);
        lowest = std::min(lowest, x.offset / 65536);
        ++singles;
    }
    CHECK(singles == 1029 - 2 - 3 - 30 - 16 - 17 && lowest == 19);
    CHECK(p.used_bytes.load() == p.bytes);
    TableAudit au;
    CHECK(pool_audit(h, TIER_GPU, &au) == 0);
}

// Fragmentation: a 1 GiB pool kept 70-85 % full by random allocations of 1 granule ... 64 granules and 1-32 KiB
// blocks, released in random order (zeroing simulated at once).  Checks that the accounting survives and that large
// requests still succeed while space exists in runs, and prints the latency of pool_alloc under that load
// (`unit_host frag`; numbers in profiles/README.md).
static void t_frag() {
    std::vector<char> buf;
    TableHeader *h = fresh_table(buf, 1ull << 30, 1ull << 20);
    PoolDesc &p = h->pool[TIER_GPU];
    std::mt19937_64 rng(getenv("FRAG_SEED") ? strtoull(getenv("FRAG_SEED"), nullptr, 0) : 7);
    std::vector<BlockRef> live;
    std::vector<double> ns_small, ns_large;
    uint64_t enomem_large = 0, tries_large = 0;
    auto scrub = [&] {
        std::vector<DirtyBlock> d;
        pool_take_dirty(h, TIER_GPU, &d);
        for (auto &b : d) pool_release_scrubbed(h, TIER_GPU, b.region);
    };
    for (int it = 0; it < 200000; ++it) {
        const double fill = (double)p.used_bytes.load() / (double)p.bytes;
        const bool do_alloc = fill < 0.70 || (fill < 0.85 && (rng() & 1));
        if (do_alloc) {
            const bool small = getenv("FRAG_NO_SMALL") ? false : (rng() % 4) != 0;
            const uint64_t bytes = small ? 1024 + rng() % (31 << 10) : p.granule * (1 + rng() % 64);
            BlockRef b;
            const auto t0 = std::chrono::steady_clock::now();
            int err = pool_alloc(h, TIER_GPU, bytes, 33, 0, &b);
            const double ns = std::chrono::duration<double, std::nano>(std::chrono::steady_clock::now() - t0).count();
            (small ? ns_small : ns_large).push_back(ns);
            if (!small) {
                ++tries_large;
                enomem_large += err == OCM_ENOMEM;
            }
            if (err == OCM_OK) live.push_back(b);
            else CHECK(err == OCM_ENOMEM);
        } else if (!live.empty()) {
            const size_t k = rng() % live.size();
            CHECK(pool_free(h, TIER_GPU, live[k].region, live[k].id) == OCM_OK);
            live[k] = live.back();
            live.pop_back();
            if ((it & 15) == 0) scrub();
        }
    }
    TableAudit au;
    CHECK(pool_audit(h, TIER_GPU, &au) == 0);
    CHECK(pool_count_live(h, TIER_GPU) == live.size());
    // below 85 % full a request of up to 64 granules (4 MiB of 1 GiB) must almost always find a run (< 0.5 % refused)
    CHECK(tries_large > 1000 && enomem_large * 200 < tries_large);
    for (auto &b : live) CHECK(pool_free(h, TIER_GPU, b.region, b.id) == OCM_OK);
    scrub();
    CHECK(p.live.load() == 0 && pool_payload_bytes(h, TIER_GPU) == 0);
    auto pct = [](std::vector<double> &v, double q) {
        std::sort(v.begin(), v.end());
        return v.empty() ? 0.0 : v[(size_t)(q * (v.size() - 1))];
    };
    printf("frag: %zu small allocs p50 %.0f ns p99 %.0f ns; %zu large allocs p50 %.0f ns p99 %.0f ns; large ENOMEM %llu of %llu\n",
           ns_small.size(), pct(ns_small, 0.5), pct(ns_small, 0.99), ns_large.size(), pct(ns_large, 0.5), pct(ns_large, 0.99),
           (unsigned long long)enomem_large, (unsigned long long)tries_large);
}

int main(int argc, char **argv) {
    std::string which = argc > 1 ? argv[1] : "all";
    std::map<std::string, void (*)()> suites = {
        {"msg", t_msg}, {"nodefile", t_nodefile}, {"nodefile_numa", t_nodefile_numa}, {"ring", t_ring}, {"tables", t_tables}, {"tables_mt", t_tables_mt}, {"slabs", t_slabs}, {"frag", t_frag}, {"placement", t_placement},
        {"governor", t_governor}, {"topo", t_topo}, {"pmsg", t_pmsg}, {"sock", t_sock}, {"oracle", t_oracle}};
    int ran = 0;
    for (auto &kv : suites) {
        if (which != "all" && which != kv.first) continue;
        int before = g_fail;
        kv.second();
        printf("%s: %s\n", kv.first.c_str(), g_fail == before ? "pass" : "FAIL");
        ++ran;
    }
    if (!ran) {
        fprintf(stderr, "unknown suite %s\n", which.c_str());
        return 2;
    }
    return g_fail ? 1 : 0;
}
