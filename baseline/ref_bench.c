/*
 * ref_bench.c - timing harness around the reference's PUBLIC API (inc/oncillamem.h),
 * linked against the reference's own libocm.so built from unmodified sources.
 * The reference's bandwidth programs contain no timer, so the
 * clock lives here, outside the reference code.  Same metric and shapes as
 * bench.py's own arm:
 *   copy   : K x ocm_copy(dest = "remote" pair, src = GPU allocation, SIZE bytes)
 *   e2e    : K x { ocm_copy(dest = GPU allocation, src = host allocation)   -- cudaMemcpy H2D, reference src/lib.c:550-558
 *                  ocm_copy(dest = "remote" pair, src = GPU allocation) }  -- cudaMemcpy D2H + ib_write, src/lib.c:643-649
 *            (host -> local GPU -> remote: the same end-to-end path as bench.py's own arm)
 *   h2r    : K x ocm_copy(dest = "remote" pair, src = host allocation, SIZE bytes), all on the CPU side
 *   put/get: K x ocm_copy_onesided (the reference's own BW test shape, test/ocm_test.c:323-425)
 *   alloc  : p50 of ocm_alloc / ocm_free for the remote kind
 * Prints one JSON object.   Usage: ref_bench <size_bytes> <steps> <warmup> <alloc_iters>
 *
 * With several ranks (REF_SYNC_DIR / REF_RANK / REF_WORLD set by ref_harness.py) the
 * ALLOCATION phases of the ranks are serialised and only the timed copy loops run
 * concurrently: the reference's rank-0 governor links a stack-allocated record into
 * a global list (reference src/mem.c:211-214 -> src/alloc.c:133-137)
 * and its daemons crash when two applications allocate at the same time.
 */
#include <oncillamem.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}
static int cmp_d(const void *a, const void *b) {
    double x = *(const double *)a, y = *(const double *)b;
    return x < y ? -1 : x > y;
}

static const char *g_dir;
static int g_rank, g_world = 1;

static void touch(const char *what, int r) {
    char p[512];
    FILE *f;
    if (!g_dir) return;
    snprintf(p, sizeof p, "%s/%s_%d", g_dir, what, r);
    f = fopen(p, "w");
    if (f) fclose(f);
}
static int wait_for(const char *what, int r, double timeout_s) {
    char p[512];
    double t0 = now_s();
    if (!g_dir) return 0;
    snprintf(p, sizeof p, "%s/%s_%d", g_dir, what, r);
    while (access(p, F_OK) != 0) {
        if (now_s() - t0 > timeout_s) return -1;
        usleep(2000);
    }
    return 0;
}

int main(int argc, char **argv) {
    uint64_t size = argc > 1 ? strtoull(argv[1], NULL, 0) : (256ull << 20);
    int steps = argc > 2 ? atoi(argv[2]) : 10, warm = argc > 3 ? atoi(argv[3]) : 3;
    int alloc_iters = argc > 4 ? atoi(argv[4]) : 50;
    struct ocm_alloc_params ap;
    struct ocm_params cp;
    ocm_alloc_t gpu, host, rem;
    double t0, t1, copy_s, e2e_s, h2r_s, put_s = 0, get_s = 0, *lat;
    int i, have_pair;
    void *hb;
    size_t hl;

    g_dir = getenv("REF_SYNC_DIR");
    if (getenv("REF_RANK")) g_rank = atoi(getenv("REF_RANK"));
    if (getenv("REF_WORLD")) g_world = atoi(getenv("REF_WORLD"));

    /* ---- allocation phase, one rank at a time -------------------------------------- */
    if (g_rank > 0 && wait_for("alloc_done", g_rank - 1, 300.0)) {
        printf("{\"error\": \"timed out waiting for rank %d\"}\n", g_rank - 1);
        return 1;
    }
    if (ocm_init() != 0) {
        printf("{\"error\": \"ocm_init failed\"}\n");
        touch("alloc_done", g_rank);
        return 1;
    }
    memset(&ap, 0, sizeof ap);
    ap.local_alloc_bytes = size;
    ap.kind = OCM_LOCAL_GPU;
    gpu = ocm_alloc(&ap);
    ap.kind = OCM_LOCAL_HOST;
    host = ocm_alloc(&ap);
    ap.kind = OCM_REMOTE_RDMA;
    ap.rem_alloc_bytes = size;
    rem = ocm_alloc(&ap);
    if (!gpu || !host || !rem) {
        printf("{\"error\": \"ocm_alloc failed (gpu=%p host=%p remote=%p)\"}\n", (void *)gpu, (void *)host, (void *)rem);
        touch("alloc_done", g_rank);
        return 1;
    }
    lat = malloc(sizeof(double) * (alloc_iters > 0 ? alloc_iters : 1));
    ap.local_alloc_bytes = 4096;
    ap.rem_alloc_bytes = 1 << 20;
    ap.kind = OCM_REMOTE_RDMA;
    for (i = 0; i < alloc_iters; ++i) {
        ocm_alloc_t a;
        t0 = now_s();
        a = ocm_alloc(&ap);
        t1 = now_s();
        lat[i] = (t1 - t0) * 1e6;
        if (!a) { alloc_iters = i; break; }
        ocm_free(a);
    }
    qsort(lat, alloc_iters, sizeof(double), cmp_d);
    touch("alloc_done", g_rank);
    if (wait_for("alloc_done", g_world - 1, 600.0)) {
        printf("{\"error\": \"timed out waiting for the last rank\"}\n");
        return 1;
    }

    /* ---- timed phase, all ranks together ------------------------------------------------ */
    ocm_localbuf(host, &hb, &hl);
    memset(hb, 0x5a, hl);
    /* a one-rank cluster is downgraded to OCM_LOCAL_HOST by the reference (src/alloc.c:82-83) */
    have_pair = ocm_alloc_kind(rem) == OCM_REMOTE_RDMA;
    memset(&cp, 0, sizeof cp);
    cp.bytes = size;
    cp.op_flag = 1;
    for (i = 0; i < warm; ++i) ocm_copy(rem, gpu, &cp);
    t0 = now_s();
    for (i = 0; i < steps; ++i)
        if (ocm_copy(rem, gpu, &cp)) { printf("{\"error\": \"ocm_copy gpu->remote failed\"}\n"); return 1; }
    t1 = now_s();
    copy_s = t1 - t0;

    for (i = 0; i < 2; ++i) ocm_copy(rem, host, &cp);
    t0 = now_s();
    for (i = 0; i < steps; ++i)
        if (ocm_copy(rem, host, &cp)) { printf("{\"error\": \"ocm_copy host->remote failed\"}\n"); return 1; }
    t1 = now_s();
    h2r_s = t1 - t0;

    for (i = 0; i < 2; ++i) { ocm_copy(gpu, host, &cp); ocm_copy(rem, gpu, &cp); }
    t0 = now_s();
    for (i = 0; i < steps; ++i)
        if (ocm_copy(gpu, host, &cp) || ocm_copy(rem, gpu, &cp)) {
            printf("{\"error\": \"ocm_copy host->gpu->remote failed\"}\n");
            return 1;
        }
    t1 = now_s();
    e2e_s = t1 - t0;

    if (have_pair) {
        ocm_copy_onesided(rem, &cp);
        t0 = now_s();
        for (i = 0; i < steps; ++i) ocm_copy_onesided(rem, &cp);
        t1 = now_s();
        put_s = t1 - t0;
        cp.op_flag = 0;
        ocm_copy_onesided(rem, &cp);
        t0 = now_s();
        for (i = 0; i < steps; ++i) ocm_copy_onesided(rem, &cp);
        t1 = now_s();
        get_s = t1 - t0;
    }

    printf("{\"size\": %llu, \"steps\": %d, \"granted_kind\": %d, \"copy_s\": %.9f, \"e2e_s\": %.9f, \"h2r_s\": %.9f, \"put_s\": %.9f, "
           "\"get_s\": %.9f, \"alloc_p50_us\": %.2f, \"alloc_iters\": %d}\n",
           (unsigned long long)size, steps, (int)ocm_alloc_kind(rem), copy_s, e2e_s, h2r_s, put_s, get_s,
           alloc_iters ? lat[alloc_iters / 2] : -1.0, alloc_iters);
    fflush(stdout);
    /* ---- teardown, again one rank at a time ------------------------------------------------ */
    touch("timed_done", g_rank);
    for (i = 0; i < g_world; ++i) wait_for("timed_done", i, 600.0);
    if (g_rank > 0) wait_for("free_done", g_rank - 1, 120.0);
    ocm_free(rem);
    ocm_free(gpu);
    ocm_free(host);
    ocm_tini();
    touch("free_done", g_rank);
    /* our daemon is stopped when this program exits, but it may still own another rank's far side:
     * stay until every rank has released its allocations */
    for (i = 0; i < g_world; ++i) wait_for("free_done", i, 120.0);
    return 0;
}
