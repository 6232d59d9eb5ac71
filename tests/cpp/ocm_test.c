/*
 * ocm_test - end-to-end exercise of the public C API against running daemons.
 * Same four scenarios and CLI shape as the reference program (reference
 * test/ocm_test.c:18-30: 1=allocation 2=copy-onesided 3=copy-twosided 4=read/write BW)
 * but every scenario verifies data, releases what it allocated and the
 * bandwidth sweep actually times the transfers (the reference has no timer).
 * Plain C on purpose: it proves the header is a C header.
 *
 *   ocm_test <test> <alloc1 MB> <alloc2 MB> [kind] [iterations]
 *     kind: 1 host, 2 gpu, 3 remote-host (legacy RDMA name), 4 remote-host (legacy RMA name),
 *           7 remote-gpu, 8 remote-host
 */
#include <oncillamem.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

static int g_fail = 0;
#define EXPECT(cond, ...)                                    \
    do {                                                     \
        if (!(cond)) {                                       \
            printf("FAIL: " __VA_ARGS__);                    \
            printf(" (%s:%d)\n", __FILE__, __LINE__);        \
            g_fail++;                                        \
        }                                                    \
    } while (0)

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static enum ocm_kind kind_of(int sub) {
    switch (sub) {
    case 1: return OCM_LOCAL_HOST;
    case 2: return OCM_LOCAL_GPU;
    case 3: return OCM_REMOTE_RDMA;
    case 4: return OCM_REMOTE_RMA;
    case 7: return OCM_REMOTE_GPU;
    case 8: return OCM_REMOTE_HOST;
    default: return OCM_LOCAL_HOST;
    }
}

static void fill(unsigned char *p, size_t n, unsigned seed) {
    size_t i;
    for (i = 0; i < n; ++i) p[i] = (unsigned char)((i * 2654435761u + seed) >> 13);
}

static int alloc_test(int sub, uint64_t local_b, uint64_t rem_b) {
    struct ocm_alloc_params ap;
    ocm_alloc_t a[3];
    int i;
    memset(&ap, 0, sizeof ap);
    ap.kind = kind_of(sub);
    ap.local_alloc_bytes = local_b;
    ap.rem_alloc_bytes = rem_b;
    for (i = 0; i < 3; ++i) {
        void *buf = NULL;
        size_t len = 0, rlen = 0;
        a[i] = ocm_alloc(&ap);
        EXPECT(a[i] != NULL, "ocm_alloc kind %d: %s", (int)ap.kind, ocm_strerror(ocm_last_error()));
        if (!a[i]) return -1;
        EXPECT(ocm_localbuf(a[i], &buf, &len) == 0 && buf && len == local_b, "ocm_localbuf");
        printf("alloc %d: granted kind %d, local %zu B @ %p", i, (int)ocm_alloc_kind(a[i]), len, buf);
        if (ocm_is_remote(a[i])) {
            EXPECT(ocm_remote_sz(a[i], &rlen) == 0 && rlen == rem_b, "ocm_remote_sz");
            printf(", remote %zu B on rank %d\n", rlen, ocm_remote_rank(a[i]));
        } else {
            EXPECT(ocm_remote_sz(a[i], &rlen) == -1, "local kinds have no remote size");
            printf(", local\n");
        }
    }
    for (i = 0; i < 3; ++i) EXPECT(ocm_free(a[i]) == 0, "ocm_free");
    return 0;
}

static int onesided_test(int sub, uint64_t size_b) {
    struct ocm_alloc_params ap;
    struct ocm_params cp;
    ocm_alloc_t a;
    unsigned char *host = malloc(size_b), *back = malloc(size_b);
    uint64_t sig = 0;
    memset(&ap, 0, sizeof ap);
    ap.kind = kind_of(sub);
    ap.local_alloc_bytes = size_b + 1;      /* the reference also allocates size+1 (test/ocm_test.c:139) */
    ap.rem_alloc_bytes = size_b + 1;
    a = ocm_alloc(&ap);
    EXPECT(a != NULL, "ocm_alloc: %s", ocm_strerror(ocm_last_error()));
    if (!a) return -1;
    memset(&cp, 0, sizeof cp);
    cp.bytes = size_b;
    /* remote memory starts zeroed: a read must deliver zeros */
    cp.op_flag = 0;
    EXPECT(ocm_copy_onesided(a, &cp) == 0, "one-sided read");
    /* write a pattern through the window, clear the window, read it back */
    fill(host, size_b, 7);
    {
        /* stage the pattern into the window with a host->window copy via a local host allocation */
        struct ocm_alloc_params hp;
        struct ocm_params p2;
        ocm_alloc_t h;
        void *hb;
        size_t hl;
        memset(&hp, 0, sizeof hp);
        hp.kind = OCM_LOCAL_HOST;
        hp.local_alloc_bytes = size_b;
        h = ocm_alloc(&hp);
        EXPECT(h != NULL, "host alloc");
        ocm_localbuf(h, &hb, &hl);
        memcpy(hb, host, size_b);
        memset(&p2, 0, sizeof p2);
        p2.bytes = size_b;
        p2.op_flag = 1;
        EXPECT(ocm_copy(a, h, &p2) == 0, "host -> pair");       /* window + far side */
        EXPECT(ocm_signal_read(a, &sig) == 0 && sig >= 1, "completion signal after put (got %llu)", (unsigned long long)sig);
        memset(hb, 0, size_b);
        p2.op_flag = 0;                                         /* read: pair -> host */
        EXPECT(ocm_copy(a, h, &p2) == 0, "pair -> host");
        memcpy(back, hb, size_b);
        EXPECT(memcmp(back, host, size_b) == 0, "data mismatch after put/get");
        cp.op_flag = 1;
        EXPECT(ocm_copy_onesided(a, &cp) == 0, "one-sided write");
        cp.bytes = size_b + 2;
        EXPECT(ocm_copy_onesided(a, &cp) != 0, "oversized one-sided copy must be rejected");
        EXPECT(ocm_free(h) == 0, "free host");
    }
    EXPECT(ocm_free(a) == 0, "ocm_free");
    free(host);
    free(back);
    return 0;
}

static int twosided_test(uint64_t local_b, uint64_t rem_b, int have_gpu) {
    struct ocm_alloc_params ap;
    struct ocm_params cp;
    ocm_alloc_t h1, h2, g = NULL, r;
    void *b1, *b2;
    size_t l1, l2;
    memset(&ap, 0, sizeof ap);
    ap.local_alloc_bytes = local_b;
    ap.kind = OCM_LOCAL_HOST;
    h1 = ocm_alloc(&ap);
    h2 = ocm_alloc(&ap);
    EXPECT(h1 && h2, "host allocs");
    if (have_gpu) {
        ap.kind = OCM_LOCAL_GPU;
        g = ocm_alloc(&ap);
        EXPECT(g != NULL, "gpu alloc: %s", ocm_strerror(ocm_last_error()));
    }
    ap.kind = have_gpu ? OCM_REMOTE_GPU : OCM_REMOTE_RDMA;
    ap.rem_alloc_bytes = rem_b;
    r = ocm_alloc(&ap);
    EXPECT(r != NULL, "remote alloc: %s", ocm_strerror(ocm_last_error()));
    if (!h1 || !h2 || !r) return -1;
    ocm_localbuf(h1, &b1, &l1);
    ocm_localbuf(h2, &b2, &l2);
    fill(b1, local_b, 3);
    memset(&cp, 0, sizeof cp);
    cp.bytes = local_b < rem_b ? local_b : rem_b;
    cp.op_flag = 1;
    /* the reference's matrix (test/ocm_test.c:283-320), with verification */
    if (g) EXPECT(ocm_copy(g, h1, &cp) == 0, "host -> GPU");
    if (g) EXPECT(ocm_copy(r, g, &cp) == 0, "GPU -> remote");
    EXPECT(ocm_copy(h2, h1, &cp) == 0, "host -> host");
    EXPECT(memcmp(b1, b2, cp.bytes) == 0, "host -> host data");
    if (!g) EXPECT(ocm_copy(r, h1, &cp) == 0, "host -> remote");
    memset(b2, 0, local_b);
    EXPECT(ocm_copy(h2, r, &cp) == 0, "remote -> host");
    EXPECT(memcmp(b1, b2, cp.bytes) == 0, "remote -> host data");
    if (g) {
        memset(b2, 0, local_b);
        EXPECT(ocm_copy(g, r, &cp) == 0, "remote -> GPU");
        EXPECT(ocm_copy(h2, g, &cp) == 0, "GPU -> host");
        EXPECT(memcmp(b1, b2, cp.bytes) == 0, "remote -> GPU -> host data");
    }
    /* copy_in / copy_out are real here (stubs in the reference, src/lib.c:491-499) */
    {
        unsigned char *u = malloc(local_b), *v = malloc(local_b);
        fill(u, local_b, 11);
        EXPECT(ocm_copy_in(h1, u) == 0, "copy_in host");
        EXPECT(memcmp(b1, u, local_b) == 0, "copy_in data");
        EXPECT(ocm_copy_out(v, h1) == 0 && memcmp(u, v, local_b) == 0, "copy_out host");
        if (local_b == rem_b) {
            fill(u, local_b, 12);
            memset(v, 0, local_b);
            EXPECT(ocm_copy_in(r, u) == 0, "copy_in pair");
            EXPECT(ocm_copy_out(v, r) == 0 && memcmp(u, v, local_b) == 0, "copy_out pair");
        }
        free(u);
        free(v);
    }
    EXPECT(ocm_free(h1) == 0 && ocm_free(h2) == 0 && ocm_free(r) == 0, "frees");
    if (g) EXPECT(ocm_free(g) == 0, "free gpu");
    EXPECT(ocm_free(r) != 0, "double free must be rejected");
    return 0;
}

static int bw_test(int sub, uint64_t max_b, int iters) {
    struct ocm_alloc_params ap;
    struct ocm_params cp;
    ocm_alloc_t a;
    uint64_t sz;
    int dir;
    memset(&ap, 0, sizeof ap);
    ap.kind = kind_of(sub);
    ap.local_alloc_bytes = max_b + 1;
    ap.rem_alloc_bytes = max_b + 1;
    a = ocm_alloc(&ap);
    EXPECT(a != NULL, "ocm_alloc: %s", ocm_strerror(ocm_last_error()));
    if (!a) return -1;
    memset(&cp, 0, sizeof cp);
    printf("# dir bytes iters seconds GB/s (host wall clock; device-timed numbers come from bench.py)\n");
    for (dir = 0; dir < 2; ++dir) {
        cp.op_flag = dir;
        for (sz = 64; sz <= max_b; sz *= 2) {
            int i;
            double t0, t1;
            cp.bytes = sz;
            ocm_copy_onesided(a, &cp); /* warm-up */
            t0 = now_s();
            for (i = 0; i < iters; ++i)
                if (ocm_copy_onesided(a, &cp)) {
                    EXPECT(0, "copy failed at %llu", (unsigned long long)sz);
                    break;
                }
            t1 = now_s();
            printf("%s %llu %d %.6f %.3f\n", dir ? "write" : "read", (unsigned long long)sz, iters, t1 - t0,
                   (double)sz * iters / (t1 - t0) / 1e9);
        }
    }
    EXPECT(ocm_free(a) == 0, "free");
    return 0;
}

/* Test 5: open an allocation a second time through a share token, write through the attached handle,
 * watch the owner's signal, check the bounds and the stale-handle detection. */
static int share_test(int sub, uint64_t size_b) {
    struct ocm_alloc_params ap;
    struct ocm_params cp;
    struct ocm_share_token tok;
    ocm_alloc_t owner, src, att = NULL, att2 = NULL;
    uint64_t sig = 99;
    size_t rsz = 0;
    const int gpu = sub == 2 || sub == 7;
    memset(&ap, 0, sizeof ap);
    ap.kind = kind_of(sub);
    ap.local_alloc_bytes = size_b;
    ap.rem_alloc_bytes = size_b;
    owner = ocm_alloc(&ap);
    EXPECT(owner != NULL, "ocm_alloc owner: %s", ocm_strerror(ocm_last_error()));
    ap.kind = gpu ? OCM_LOCAL_GPU : OCM_LOCAL_HOST;
    src = ocm_alloc(&ap);
    EXPECT(src != NULL, "ocm_alloc src");
    if (!owner || !src) return -1;
    EXPECT(ocm_share(owner, &tok) == 0, "ocm_share");
    EXPECT(tok.bytes == size_b, "token bytes");
    EXPECT(ocm_attach(&tok, &att) == 0 && att, "ocm_attach");
    if (!att) return -1;
    EXPECT(ocm_attach_count(owner) == 1, "attach count %d", ocm_attach_count(owner));
    EXPECT(ocm_is_remote(att) && ocm_remote_sz(att, &rsz) == 0 && rsz == size_b, "attached handle is a pair of %zu", rsz);
    EXPECT(ocm_remote_rank(att) == tok.rank, "owner rank");
    memset(&cp, 0, sizeof cp);
    cp.op_flag = 1;
    cp.bytes = size_b;
    EXPECT(ocm_copy(att, src, &cp) == 0, "copy through the attached handle: %s", ocm_strerror(ocm_last_error()));
    EXPECT(ocm_signal_read(owner, &sig) == 0 && sig == 1, "owner signal %llu", (unsigned long long)sig);
    cp.bytes = size_b + 1;
    EXPECT(ocm_copy(att, src, &cp) != 0 && ocm_last_error() == OCM_ERANGE, "oversize copy must be rejected");
    cp.bytes = size_b;
    tok.generation += 1;
    EXPECT(ocm_attach(&tok, &att2) == -OCM_ESTALE && !att2, "forged token must be rejected");
    tok.generation -= 1;
    EXPECT(ocm_free(owner) == 0, "owner free");
    EXPECT(ocm_copy(att, src, &cp) != 0 && ocm_last_error() == OCM_ESTALE, "stale handle must be rejected");
    EXPECT(ocm_attach(&tok, &att2) == -OCM_ESTALE, "attach after free must be rejected");
    EXPECT(ocm_free(att) == 0, "free attached");
    EXPECT(ocm_free(src) == 0, "free src");
    return 0;
}

int main(int argc, char **argv) {
    int test, sub = 1, iters = 5;
    double mb1 = 1.0, mb2 = 1.0;
    uint64_t b1, b2;
    if (argc < 2) {
        fprintf(stderr, "Usage: %s <1=alloc|2=onesided|3=twosided|4=bw|5=share> [alloc1 MB] [alloc2 MB] [kind] [iters]\n", argv[0]);
        return 2;
    }
    test = atoi(argv[1]);
    if (argc > 2) mb1 = atof(argv[2]);
    if (argc > 3) mb2 = atof(argv[3]);
    if (argc > 4) sub = atoi(argv[4]);
    if (argc > 5) iters = atoi(argv[5]);
    b1 = (uint64_t)(mb1 * 1048576.0);
    b2 = (uint64_t)(mb2 * 1048576.0);
    if (ocm_init() != 0) {
        printf("FAIL: cannot connect to OCM: %s\n", ocm_strerror(ocm_last_error()));
        return 1;
    }
    printf("attached: rank %d of %d, device %d\n", ocm_rank(), ocm_num_ranks(), ocm_device());
    switch (test) {
    case 1: alloc_test(sub, b1, b2); break;
    case 2: onesided_test(sub < 3 ? 3 : sub, b1); break;
    case 3: twosided_test(b1, b2, ocm_device() >= 0 && sub != 1); break;
    case 4: bw_test(sub < 3 ? 3 : sub, b1, iters); break;
    case 5: share_test(sub, b1); break;
    default: printf("FAIL: unknown test\n"); g_fail++; break;
    }
    if (ocm_tini() != 0) {
        printf("FAIL: ocm_tini\n");
        g_fail++;
    }
    printf(g_fail ? "FAIL: %d check(s) failed\n" : "pass: OCM test completed successfully\n", g_fail);
    return g_fail ? 1 : 0;
}
