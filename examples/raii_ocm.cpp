// raii_ocm - the C++ convenience layer (include/oncillamem.hpp): sessions and allocations that clean up after
// themselves, errors as exceptions.  Host kinds only, so it runs on machines without a GPU.
//   build/bin/raii_ocm          (needs a running cluster)
#include <oncillamem.hpp>
#include <cstdio>
#include <cstring>
#include <vector>

int main() {
    try {
        ocm::Session session;
        std::printf("rank %d of %d\n", session.rank(), session.num_ranks());

        const size_t n = 1 << 20;
        ocm::Allocation pair(OCM_REMOTE_HOST, n, n);           // pinned window here, far side on the neighbour
        std::memset(pair.local_ptr(), 0x3C, n);
        pair.put(n);
        std::memset(pair.local_ptr(), 0, n);
        pair.get(n);
        const auto *w = static_cast<const unsigned char *>(pair.local_ptr());
        for (size_t i = 0; i < n; ++i)
            if (w[i] != 0x3C) throw std::runtime_error("data mismatch after put/get");
        if (pair.signal() != 1) throw std::runtime_error("the put did not bump the signal");

        // a second handle on the far side through a share token (what another application would do)
        ocm::Allocation peer = ocm::Allocation::attach(pair.share());
        if (peer.remote_ptr() != pair.remote_ptr() || peer.remote_bytes() != n) throw std::runtime_error("attach");
        ocm::Allocation src(OCM_LOCAL_HOST, n);
        std::memset(src.local_ptr(), 0x77, n);
        ocm::copy(peer, src, n);                                // lands in pair's far side, counted by its signal
        if (pair.signal() != 2) throw std::runtime_error("the peer's copy was not counted");

        // three puts as one batch (one kernel launch when a GPU side is involved), counted one by one
        struct ocm_params b[3] = {};
        for (int i = 0; i < 3; ++i) {
            b[i].src_offset = b[i].dest_offset = (uint64_t)i * 4096;
            b[i].bytes = 4096;
            b[i].op_flag = 1;
        }
        pair.batch(b, 3);
        ocm_quiet();
        if (pair.signal() != 5) throw std::runtime_error("the batch was not counted put by put");
        if (ocm::prefetch_views(false, true) < 1) throw std::runtime_error("prefetch_views");

        // ownership moves, errors are exceptions
        std::vector<ocm::Allocation> many;
        for (int i = 0; i < 8; ++i) many.emplace_back(OCM_REMOTE_HOST, 4096, 65536);
        ocm::Allocation moved = std::move(many[3]);
        if (many[3] || !moved) throw std::runtime_error("move semantics");
        bool threw = false;
        try {
            pair.put(n + 1);
        } catch (const ocm::Error &e) {
            threw = e.code() == OCM_ERANGE;
        }
        if (!threw) throw std::runtime_error("oversize put must throw OCM_ERANGE");
        peer.reset();
        struct ocm_stats st;
        ocm_get_stats(&st);
        std::printf("live allocations before scope exit: %llu\n", (unsigned long long)st.allocs_live);
    } catch (const std::exception &e) {
        std::printf("FAIL: %s\n", e.what());
        return 1;
    }
    // everything was freed and the session detached by the destructors
    std::printf("pass: raii_ocm\n");
    return 0;
}
