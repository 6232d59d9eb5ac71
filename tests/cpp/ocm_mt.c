/*
 * ocm_mt - several threads of ONE application use the library at the same time: allocate pairs on both allocation
 * paths, move data through them in both directions, share/attach among themselves, free.  Every byte is checked.
 * The reference keeps a global list without covering every access and documents no threading contract
 * (reference src/lib.c:81-93); here the contract is in INTEGRATION.md ("Threading") and this program holds it to it.
 *
 *   ocm_mt [threads=8] [iterations=200] [gpu]
 * With "gpu" the pairs are OCM_REMOTE_GPU: every blocking copy is then one kernel whose completion the caller
 * waits for on its own completion word, and the bytes are read back and checked right after each call - a thread
 * that returned early because ANOTHER thread's kernel finished would see stale data here.
 */
#include <oncillamem.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static int g_iters = 200;
static int g_gpu = 0;
static int g_fail = 0;
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
/* one shared block all threads put into: its signal must count every copy exactly once */
static ocm_alloc_t g_shared;
static struct ocm_share_token g_token;
static unsigned long g_expected_signals = 0;

#define FAIL(...)                          \
    do {                                   \
        pthread_mutex_lock(&g_mu);         \
        printf("FAIL: " __VA_ARGS__);      \
        printf("\n");                      \
        g_fail++;                          \
        pthread_mutex_unlock(&g_mu);       \
    } while (0)

static void *worker(void *arg) {
    const int id = (int)(long)arg;
    unsigned seed = 1234u + (unsigned)id * 7919u;
    unsigned long my_signals = 0;
    ocm_alloc_t att = NULL;
    int it;
    if (ocm_attach(&g_token, &att) != 0) {
        FAIL("thread %d: ocm_attach", id);
        return NULL;
    }
    for (it = 0; it < g_iters; ++it) {
        const size_t bytes = 64 + (rand_r(&seed) % (256 * 1024));
        struct ocm_alloc_params_ex ap;
        struct ocm_params cp;
        ocm_alloc_t pair, host;
        unsigned char *w, *h;
        void *p = NULL;
        size_t len = 0, i;
        const unsigned char tag = (unsigned char)(id * 16 + it);
        memset(&ap, 0, sizeof ap);
        ap.base.kind = g_gpu ? OCM_REMOTE_GPU : OCM_REMOTE_HOST;
        ap.base.local_alloc_bytes = bytes;
        ap.base.rem_alloc_bytes = bytes;
        ap.policy = (it % 3 == 0) ? OCM_POLICY_ROUND_ROBIN : OCM_POLICY_DEFAULT;
        ap.rank = -1;
        ap.flags = (it % 5 == 0) ? OCM_ALLOC_SLOW_PATH : 0;
        pair = ocm_alloc_ex(&ap);
        ap.base.kind = OCM_LOCAL_HOST;
        ap.flags = 0;
        host = ocm_alloc_ex(&ap);
        if (!pair || !host) {
            FAIL("thread %d it %d: alloc: %s", id, it, ocm_strerror(ocm_last_error()));
            if (pair) ocm_free(pair);
            if (host) ocm_free(host);
            continue;
        }
        ocm_localbuf(pair, &p, &len);
        w = p;
        ocm_localbuf(host, &p, &len);
        h = p;
        for (i = 0; i < bytes; ++i) {
            if (!g_gpu && w[i] != 0) {
                FAIL("thread %d it %d: window not zeroed at %zu", id, it, i);
                break;
            }
            h[i] = (unsigned char)(tag + i);
        }
        memset(&cp, 0, sizeof cp);
        cp.bytes = bytes;
        cp.op_flag = 1;
        if (ocm_copy(pair, host, &cp) != 0) FAIL("thread %d it %d: copy host->pair", id, it);   /* host -> window -> far */
        if (!g_gpu) memset(w, 0, bytes);
        memset(h, 0, bytes);
        if (g_gpu) {
            /* far -> window -> host allocation, one kernel; the host bytes are checked as soon as the call returns */
            if (ocm_copy(host, pair, &cp) != 0) FAIL("thread %d it %d: copy pair->host", id, it);
            w = h;
        } else {
            cp.op_flag = 0;
            if (ocm_copy_onesided(pair, &cp) != 0) FAIL("thread %d it %d: get", id, it);       /* far -> window */
        }
        for (i = 0; i < bytes; ++i)
            if (w[i] != (unsigned char)(tag + i)) {
                FAIL("thread %d it %d: byte %zu is %u, want %u", id, it, i, w[i], (unsigned char)(tag + i));
                break;
            }
        /* everybody also writes 64 bytes of its own slot of the shared block */
        cp.op_flag = 1;
        cp.bytes = 64;
        cp.src_offset = 0;
        cp.dest_offset = (uint64_t)id * 64;
        if (ocm_copy(att, pair, &cp) == 0) ++my_signals;   /* far(pair) -> far(shared): counted by the shared block's signal */
        else FAIL("thread %d it %d: copy into the shared block: %s", id, it, ocm_strerror(ocm_last_error()));
        if (ocm_free(pair) != 0 || ocm_free(host) != 0) FAIL("thread %d it %d: free", id, it);
    }
    ocm_free(att);
    pthread_mutex_lock(&g_mu);
    g_expected_signals += my_signals;
    pthread_mutex_unlock(&g_mu);
    return NULL;
}

int main(int argc, char **argv) {
    const int n_threads = argc > 1 ? atoi(argv[1]) : 8;
    pthread_t th[64];
    struct ocm_alloc_params ap;
    struct ocm_stats st;
    uint64_t sig = 0;
    int i;
    if (argc > 2) g_iters = atoi(argv[2]);
    if (argc > 3 && !strcmp(argv[3], "gpu")) g_gpu = 1;
    if (n_threads < 1 || n_threads > 64) return 2;
    if (ocm_init() != 0) {
        printf("FAIL: cannot connect to OCM: %s\n", ocm_strerror(ocm_last_error()));
        return 1;
    }
    memset(&ap, 0, sizeof ap);
    ap.kind = OCM_REMOTE_HOST;
    ap.local_alloc_bytes = 4096;
    ap.rem_alloc_bytes = 64 * 64;
    g_shared = ocm_alloc(&ap);
    if (!g_shared || ocm_share(g_shared, &g_token) != 0) {
        printf("FAIL: shared block\n");
        return 1;
    }
    for (i = 0; i < n_threads; ++i) pthread_create(&th[i], NULL, worker, (void *)(long)i);
    for (i = 0; i < n_threads; ++i) pthread_join(th[i], NULL);
    ocm_signal_read(g_shared, &sig);
    if (sig != g_expected_signals) {
        printf("FAIL: shared signal %llu, expected %lu\n", (unsigned long long)sig, g_expected_signals);
        g_fail++;
    }
    if (ocm_attach_count(g_shared) != 0) {
        printf("FAIL: attach count %d after all threads detached\n", ocm_attach_count(g_shared));
        g_fail++;
    }
    ocm_free(g_shared);
    ocm_get_stats(&st);
    if (st.allocs_live != 0) {
        printf("FAIL: %llu allocations still live\n", (unsigned long long)st.allocs_live);
        g_fail++;
    }
    ocm_tini();
    printf(g_fail ? "FAIL: %d check(s) failed\n" : "pass: ocm_mt %d threads x %d iterations, %lu shared signals\n", g_fail ? g_fail : n_threads,
           g_iters, g_expected_signals);
    return g_fail ? 1 : 0;
}
