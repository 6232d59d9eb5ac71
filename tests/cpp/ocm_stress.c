/*
 * ocm_stress - one client of the many-client allocation stress (BASELINE.json
 * config 5: 64 apps x 10k ocm_alloc/ocm_free).  Nothing like it exists in the
 * reference ("Nothing tests ... concurrent apps").
 *
 *   ocm_stress <iterations> <max_bytes> [kind=8 remote-host|7 remote-gpu|1 host|6 gpu] [hold=8]
 *
 * Keeps up to <hold> allocations live, frees them in random order, writes and
 * checks a tag in every far side it can address from the CPU, and prints the
 * alloc/free latency distribution (host clock).
 */
#include <oncillamem.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

static double now_us(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}
static int cmp_d(const void *a, const void *b) {
    double x = *(const double *)a, y = *(const double *)b;
    return x < y ? -1 : x > y;
}

int main(int argc, char **argv) {
    int iters = argc > 1 ? atoi(argv[1]) : 1000;
    uint64_t max_bytes = argc > 2 ? strtoull(argv[2], NULL, 0) : 65536;
    int kind = argc > 3 ? atoi(argv[3]) : OCM_REMOTE_HOST;
    int hold = argc > 4 ? atoi(argv[4]) : 8;
    ocm_alloc_t *live = calloc(hold, sizeof *live);
    uint64_t *tag = calloc(hold, sizeof *tag);
    double *ta = malloc(sizeof(double) * iters), *tf = malloc(sizeof(double) * iters);
    int n_live = 0, i, fails = 0, nf = 0;
    uint64_t rng = 0x9E3779B97F4A7C15ull ^ (uint64_t)getpid();
    const int cpu_visible = kind == OCM_REMOTE_HOST || kind == OCM_LOCAL_HOST || kind == OCM_REMOTE_RDMA;
    if (ocm_init() != 0) {
        printf("FAIL: ocm_init: %s\n", ocm_strerror(ocm_last_error()));
        return 1;
    }
    for (i = 0; i < iters; ++i) {
        struct ocm_alloc_params ap;
        double t0, t1;
        rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17;
        if (n_live == hold) {
            int v = (int)(rng % (uint64_t)n_live);
            void *p = ocm_remote_ptr(live[v]);
            if (cpu_visible && p && *(volatile uint64_t *)p != tag[v]) {
                printf("FAIL: far side of allocation %d was overwritten\n", v);
                fails++;
            }
            t0 = now_us();
            if (ocm_free(live[v]) != 0) fails++;
            t1 = now_us();
            tf[nf++] = t1 - t0;
            live[v] = live[n_live - 1];
            tag[v] = tag[n_live - 1];
            n_live--;
        }
        memset(&ap, 0, sizeof ap);
        ap.kind = (enum ocm_kind)kind;
        ap.local_alloc_bytes = 4096;
        ap.rem_alloc_bytes = 4096 + (rng >> 20) % max_bytes;
        t0 = now_us();
        live[n_live] = ocm_alloc(&ap);
        t1 = now_us();
        ta[i] = t1 - t0;
        if (!live[n_live]) {
            printf("FAIL: ocm_alloc(%llu): %s\n", (unsigned long long)ap.rem_alloc_bytes, ocm_strerror(ocm_last_error()));
            fails++;
            continue;
        }
        if (cpu_visible) {
            volatile uint64_t *p = (volatile uint64_t *)ocm_remote_ptr(live[n_live]);
            if (p) {
                if (*p != 0) {
                    printf("FAIL: fresh far side not zeroed (%llx)\n", (unsigned long long)*p);
                    fails++;
                }
                tag[n_live] = rng | 1;
                *p = tag[n_live];
            }
        }
        n_live++;
    }
    while (n_live > 0) {
        if (ocm_free(live[--n_live]) != 0) fails++;
    }
    qsort(ta, iters, sizeof(double), cmp_d);
    qsort(tf, nf, sizeof(double), cmp_d);
    printf("alloc_us p50=%.2f p99=%.2f max=%.2f  free_us p50=%.2f  iters=%d\n", ta[iters / 2], ta[(int)(iters * 0.99)],
           ta[iters - 1], nf ? tf[nf / 2] : 0.0, iters);
    if (ocm_tini() != 0) fails++;
    printf(fails ? "FAIL: %d error(s)\n" : "pass\n", fails);
    return fails ? 1 : 0;
}
