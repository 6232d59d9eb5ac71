/*
 * ib_shim.c - out-of-tree loop-back provider of the nine ib_* symbols declared in
 * the reference's inc/io/rdma.h, so that the UNMODIFIED reference daemon and
 * library (src/main.c mem.c alloc.c nodefile.c pmsg.c queue.c sock.c lib.c) link
 * and run on a box without InfiniBand hardware, libibverbs or librdmacm.
 *
 * This file is NOT part of the product and is never linked into it; it exists only
 * so that `bench.py --impl reference` can measure the reference's own code path
 * ("reference, remote (socket/loop-back ib_* shim)").
 *
 * Semantics kept from the reference transport (reference src/rdma*.c):
 *   - one connection per allocation; the server side blocks in ib_connect until
 *     the client connects (src/rdma_server.c:72-82);
 *   - the server publishes {address, length} of its buffer to the client at
 *     connect time (there: RDMA-CM private data, src/rdma_server.c:141-158);
 *   - ib_write/ib_read move bytes between the client's buffer at src_offset and
 *     the server's buffer at dest_offset, bounds-checked against the remote
 *     length (src/rdma.c:241-263); ib_poll waits for completion.
 * Data path: one-sided process_vm_writev/readv into the daemon's address space
 * (the closest CPU analogue of RDMA: single copy, no server involvement); if the
 * kernel refuses that, the bytes are streamed over the TCP connection instead.
 */
#define _GNU_SOURCE
#include <arpa/inet.h>
#include <errno.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <pthread.h>
#include <stdbool.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/uio.h>
#include <unistd.h>

#include <io/rdma.h>

struct ib_alloc {
    struct ib_params params;
    int      sock, lsock;
    bool     is_server, use_vm, pending;
    pid_t    peer_pid;
    uint64_t remote_va, remote_len;
    pthread_t thr;
    bool     thr_running;
};

struct hello { int32_t pid; int32_t pad; uint64_t va, len; };
struct req   { uint32_t op; uint32_t pad; uint64_t off, len; };   /* op 1 write, 2 read, 0 bye */

static int put_all(int fd, const void *b, size_t n) {
    const char *p = b;
    while (n) {
        ssize_t k = send(fd, p, n, MSG_NOSIGNAL);
        if (k < 0 && errno == EINTR) continue;
        if (k <= 0) return -1;
        p += k; n -= (size_t)k;
    }
    return 0;
}
static int get_all(int fd, void *b, size_t n) {
    char *p = b;
    while (n) {
        ssize_t k = recv(fd, p, n, 0);
        if (k < 0 && errno == EINTR) continue;
        if (k <= 0) return -1;
        p += k; n -= (size_t)k;
    }
    return 0;
}

/* Several reference daemons share one network namespace here and every one of them numbers its
 * per-allocation ports from the same start (reference src/mem.c:38,355-356), so each daemon gets its own
 * loop-back address, derived from the trailing digits of its hostname: "ocmref3" -> 127.0.1.4. */
static void shim_ip(char *out, size_t len) {
    char host[256] = "";
    size_t n;
    unsigned v = 0, mul = 1;
    int digits = 0;
    gethostname(host, sizeof host - 1);
    n = strlen(host);
    while (n > 0 && host[n - 1] >= '0' && host[n - 1] <= '9' && digits < 3) {
        v += (unsigned)(host[n - 1] - '0') * mul;
        mul *= 10;
        --n;
        ++digits;
    }
    if (digits) snprintf(out, len, "127.0.1.%u", (v % 250) + 1);
    else snprintf(out, len, "127.0.0.1");
}

int ib_init(void) { return 0; }

ib_t ib_new(struct ib_params *p) {
    struct ib_alloc *ib;
    if (!p) return NULL;
    ib = calloc(1, sizeof *ib);
    if (!ib) return NULL;
    ib->params = *p;
    if (p->addr) ib->params.addr = strdup(p->addr);
    ib->sock = ib->lsock = -1;
    return ib;
}

int ib_free(ib_t ib) {
    if (!ib) return -1;
    if (ib->params.buf) free(ib->params.buf);   /* as the reference does (src/rdma.c:159-176) */
    if (ib->params.addr) free(ib->params.addr);
    free(ib);
    return 0;
}

/* fallback service loop: only used when process_vm_* is not permitted */
static void *serve(void *arg) {
    struct ib_alloc *ib = arg;
    struct req r;
    char ack = 1;
    while (get_all(ib->sock, &r, sizeof r) == 0 && r.op) {
        if (r.off + r.len > ib->params.buf_len) break;
        if (r.op == 1) {
            if (get_all(ib->sock, (char *)ib->params.buf + r.off, r.len)) break;
            if (put_all(ib->sock, &ack, 1)) break;
        } else if (r.op == 2) {
            if (put_all(ib->sock, (char *)ib->params.buf + r.off, r.len)) break;
        }
    }
    return NULL;
}

static int server_connect(struct ib_alloc *ib) {
    struct sockaddr_in a;
    struct hello h;
    int one = 1;
    ib->is_server = true;
    ib->lsock = socket(AF_INET, SOCK_STREAM, 0);
    if (ib->lsock < 0) return -1;
    setsockopt(ib->lsock, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    memset(&a, 0, sizeof a);
    a.sin_family = AF_INET;
    {
        char ip[64];
        shim_ip(ip, sizeof ip);
        if (inet_pton(AF_INET, ip, &a.sin_addr) != 1) a.sin_addr.s_addr = htonl(INADDR_ANY);
    }
    a.sin_port = htons((uint16_t)ib->params.port);   /* the reference truncates its 67980+ ports the same way */
    if (bind(ib->lsock, (struct sockaddr *)&a, sizeof a) || listen(ib->lsock, 1)) return -1;
    ib->sock = accept(ib->lsock, NULL, NULL);        /* blocks until the application connects */
    if (ib->sock < 0) return -1;
    setsockopt(ib->sock, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    h.pid = getpid(); h.pad = 0;
    h.va = (uint64_t)(uintptr_t)ib->params.buf;
    h.len = ib->params.buf_len;
    if (put_all(ib->sock, &h, sizeof h)) return -1;
    ib->thr_running = pthread_create(&ib->thr, NULL, serve, ib) == 0;
    return 0;
}

static int client_connect(struct ib_alloc *ib) {
    struct sockaddr_in a;
    struct hello h;
    int one = 1, tries;
    char probe = 0;
    struct iovec l, r;
    memset(&a, 0, sizeof a);
    a.sin_family = AF_INET;
    a.sin_port = htons((uint16_t)ib->params.port);
    if (inet_pton(AF_INET, ib->params.addr ? ib->params.addr : "127.0.0.1", &a.sin_addr) != 1)
        a.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
    /* the daemon replies before it listens ("XXX possible race", reference src/mem.c:351-354): retry */
    for (tries = 0; tries < 2000; ++tries) {
        ib->sock = socket(AF_INET, SOCK_STREAM, 0);
        if (ib->sock < 0) return -1;
        if (connect(ib->sock, (struct sockaddr *)&a, sizeof a) == 0) break;
        close(ib->sock);
        ib->sock = -1;
        usleep(1000);
    }
    if (ib->sock < 0) return -1;
    setsockopt(ib->sock, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    if (get_all(ib->sock, &h, sizeof h)) return -1;
    ib->peer_pid = h.pid;
    ib->remote_va = h.va;
    ib->remote_len = h.len;
    /* can we write the daemon's memory directly? (zero-length style probe with one byte read) */
    l.iov_base = &probe; l.iov_len = 1;
    r.iov_base = (void *)(uintptr_t)h.va; r.iov_len = 1;
    ib->use_vm = h.len > 0 && process_vm_readv(h.pid, &l, 1, &r, 1, 0) == 1 && !getenv("OCM_REF_SHIM_TCP");
    return 0;
}

int ib_connect(ib_t ib, bool is_server) {
    if (!ib) return -1;
    return is_server ? server_connect(ib) : client_connect(ib);
}

int ib_disconnect(ib_t ib, bool is_server) {
    (void)is_server;
    if (!ib) return -1;
    if (!ib->is_server && ib->sock >= 0) {
        struct req r = {0, 0, 0, 0};
        put_all(ib->sock, &r, sizeof r);
    }
    if (ib->sock >= 0) { shutdown(ib->sock, SHUT_RDWR); close(ib->sock); ib->sock = -1; }
    if (ib->lsock >= 0) { close(ib->lsock); ib->lsock = -1; }
    if (ib->thr_running) { pthread_join(ib->thr, NULL); ib->thr_running = false; }
    return 0;
}

static int vm_xfer(struct ib_alloc *ib, bool write, size_t local_off, size_t remote_off, size_t len) {
    size_t done = 0;
    while (done < len) {
        struct iovec l, r;
        ssize_t k;
        size_t chunk = len - done > (1u << 30) ? (1u << 30) : len - done;
        l.iov_base = (char *)ib->params.buf + local_off + done; l.iov_len = chunk;
        r.iov_base = (void *)(uintptr_t)(ib->remote_va + remote_off + done); r.iov_len = chunk;
        k = write ? process_vm_writev(ib->peer_pid, &l, 1, &r, 1, 0) : process_vm_readv(ib->peer_pid, &l, 1, &r, 1, 0);
        if (k <= 0) return -1;
        done += (size_t)k;
    }
    return 0;
}

int ib_write(ib_t ib, size_t src_offset, size_t dest_offset, size_t len) {
    struct req r;
    if (!ib || dest_offset + len > ib->remote_len || src_offset + len > ib->params.buf_len) return -1;
    if (ib->use_vm) return vm_xfer(ib, true, src_offset, dest_offset, len);
    r.op = 1; r.pad = 0; r.off = dest_offset; r.len = len;
    if (put_all(ib->sock, &r, sizeof r) || put_all(ib->sock, (char *)ib->params.buf + src_offset, len)) return -1;
    ib->pending = true;
    return 0;
}

int ib_read(ib_t ib, size_t src_offset, size_t dest_offset, size_t len) {
    struct req r;
    if (!ib || dest_offset + len > ib->remote_len || src_offset + len > ib->params.buf_len) return -1;
    if (ib->use_vm) return vm_xfer(ib, false, src_offset, dest_offset, len);
    r.op = 2; r.pad = 0; r.off = dest_offset; r.len = len;
    if (put_all(ib->sock, &r, sizeof r)) return -1;
    return get_all(ib->sock, (char *)ib->params.buf + src_offset, len);
}

int ib_poll(ib_t ib) {
    char ack;
    if (!ib) return -1;
    if (ib->pending) {
        ib->pending = false;
        return get_all(ib->sock, &ack, 1);
    }
    return 0;
}

int ib_nic_ip(int idx, char *ip_str, size_t len) {
    (void)idx;
    shim_ip(ip_str, len);
    return 0;
}
