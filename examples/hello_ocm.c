/*
 * hello_ocm - the smallest complete OCM program: what an application written against the reference's
 * inc/oncillamem.h looks like here (same calls, same structs; compare reference test/ocm_test.c:60-130).
 *
 *   OCM_RANK=0 build/bin/oncillamem nodefile &          # one daemon per rank, rank 0 first
 *   build/bin/hello_ocm [kind]                          # kind: 8 = remote host (default), 7 = remote GPU
 *
 * Three allocations: two local host buffers and a pair (local window + far side on the neighbour rank).
 * Data goes  host buffer -> window -> far side  (ocm_copy, two hops in one call), is pulled back into the
 * window with a one-sided get, and returns  far side -> window -> second host buffer.  Links against libocm only.
 */
#include <oncillamem.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static ocm_alloc_t alloc_or_die(enum ocm_kind kind, size_t local_bytes, size_t far_bytes) {
    struct ocm_alloc_params ap;
    ocm_alloc_t a;
    memset(&ap, 0, sizeof ap);
    ap.kind = kind;
    ap.local_alloc_bytes = local_bytes;
    ap.rem_alloc_bytes = far_bytes;
    a = ocm_alloc(&ap);
    if (!a) {
        fprintf(stderr, "ocm_alloc(kind %d): %s\n", (int)kind, ocm_strerror(ocm_last_error()));
        exit(1);
    }
    return a;
}

int main(int argc, char **argv) {
    const enum ocm_kind kind = argc > 1 ? (enum ocm_kind)atoi(argv[1]) : OCM_REMOTE_HOST;
    const size_t bytes = 1 << 20;
    struct ocm_params cp;
    ocm_alloc_t src, dst, pair;
    void *sp = NULL, *dp = NULL;
    size_t len = 0, far_len = 0, i;

    if (ocm_init() != 0) {
        fprintf(stderr, "no oncillamem daemon for this rank: %s\n", ocm_strerror(ocm_last_error()));
        return 1;
    }
    src = alloc_or_die(OCM_LOCAL_HOST, bytes, 0);
    dst = alloc_or_die(OCM_LOCAL_HOST, bytes, 0);
    pair = alloc_or_die(kind, bytes, bytes);
    ocm_localbuf(src, &sp, &len);
    ocm_localbuf(dst, &dp, &len);
    ocm_remote_sz(pair, &far_len);
    printf("rank %d of %d: pair of kind %d, far side %zu bytes on rank %d\n", ocm_rank(), ocm_num_ranks(),
           (int)ocm_alloc_kind(pair), far_len, ocm_remote_rank(pair));
    for (i = 0; i < bytes; ++i) ((unsigned char *)sp)[i] = (unsigned char)(i * 7 + 3);

    memset(&cp, 0, sizeof cp);
    cp.bytes = bytes;
    cp.op_flag = 1;                                  /* "write": data flows src -> dst argument */
    if (ocm_copy(pair, src, &cp) != 0) return 2;     /* host buffer -> window -> far side */
    cp.op_flag = 0;
    if (ocm_copy_onesided(pair, &cp) != 0) return 2; /* get: far side -> window */
    cp.op_flag = 1;
    if (ocm_copy(dst, pair, &cp) != 0) return 2;     /* far side -> window -> second host buffer */

    if (memcmp(sp, dp, bytes) != 0) {
        printf("FAIL: data mismatch\n");
        return 3;
    }
    ocm_free(pair);
    ocm_free(dst);
    ocm_free(src);
    ocm_tini();
    printf("pass: hello_ocm\n");
    return 0;
}
