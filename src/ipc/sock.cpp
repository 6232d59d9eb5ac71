#include "sock.hpp"
#include "../common/log.hpp"
#include <arpa/inet.h>
#include <cerrno>
#include <cstring>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

namespace ocm {

int set_nonblocking(int fd, bool on) {
    int fl = fcntl(fd, F_GETFL, 0);
    if (fl < 0) return -errno;
    fl = on ? (fl | O_NONBLOCK) : (fl & ~O_NONBLOCK);
    return fcntl(fd, F_SETFL, fl) < 0 ? -errno : 0;
}

void sock_close(int fd) {
    if (fd >= 0) close(fd);
}

static int wait_fd(int fd, short ev, int timeout_ms) {
    struct pollfd p = {fd, ev, 0};
    for (;;) {
        int r = poll(&p, 1, timeout_ms);
        if (r < 0 && errno == EINTR) continue;
        if (r < 0) return -errno;
        return r; // 0 timeout, 1 ready
    }
}

static int finish_connect(int fd, int timeout_ms) {
    int r = wait_fd(fd, POLLOUT, timeout_ms);
    if (r <= 0) return r == 0 ? -ETIMEDOUT : r;
    int err = 0;
    socklen_t l = sizeof err;
    if (getsockopt(fd, SOL_SOCKET, SO_ERROR, &err, &l) < 0) return -errno;
    return err ? -err : 0;
}

int tcp_listen(const std::string &ip, int port, int backlog) {
    int fd = socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (fd < 0) return -errno;
    int one = 1;
    setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    struct sockaddr_in a;
    memset(&a, 0, sizeof a);
    a.sin_family = AF_INET;
    a.sin_port = htons((uint16_t)port);
    if (ip.empty() || ip == "*" || inet_pton(AF_INET, ip.c_str(), &a.sin_addr) != 1) a.sin_addr.s_addr = htonl(INADDR_ANY);
    if (bind(fd, (struct sockaddr *)&a, sizeof a) < 0 || listen(fd, backlog) < 0) {
        int e = errno;
        close(fd);
        return -e;
    }
    return fd;
}

int tcp_connect(const std::string &host, int port, int timeout_ms) {
    struct addrinfo hints, *res = nullptr;
    memset(&hints, 0, sizeof hints);
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    char ps[16];
    snprintf(ps, sizeof ps, "%d", port);
    if (getaddrinfo(host.c_str(), ps, &hints, &res) != 0 || !res) return -EHOSTUNREACH;
    int rc = -ECONNREFUSED;
    for (struct addrinfo *ai = res; ai; ai = ai->ai_next) {
        int fd = socket(ai->ai_family, ai->ai_socktype | SOCK_CLOEXEC, ai->ai_protocol);
        if (fd < 0) { rc = -errno; continue; }
        set_nonblocking(fd, true);
        int r = connect(fd, ai->ai_addr, ai->ai_addrlen);
        if (r < 0 && errno == EINPROGRESS) r = finish_connect(fd, timeout_ms);
        else if (r < 0) r = -errno;
        if (r == 0) {
            set_nonblocking(fd, false);
            int one = 1;
            setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
            freeaddrinfo(res);
            return fd;
        }
        rc = r;
        close(fd);
    }
    freeaddrinfo(res);
    return rc;
}

int tcp_accept(int lfd, std::string *peer) {
    struct sockaddr_in a;
    socklen_t l = sizeof a;
    int fd = accept4(lfd, (struct sockaddr *)&a, &l, SOCK_CLOEXEC);
    if (fd < 0) return -errno;
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    if (peer) {
        char b[64];
        inet_ntop(AF_INET, &a.sin_addr, b, sizeof b);
        *peer = std::string(b) + ":" + std::to_string(ntohs(a.sin_port));
    }
    return fd;
}

static socklen_t fill_abstract(struct sockaddr_un *a, const std::string &name) {
    memset(a, 0, sizeof *a);
    a->sun_family = AF_UNIX;
    size_t n = std::min(name.size(), sizeof(a->sun_path) - 2);
    memcpy(a->sun_path + 1, name.data(), n); // sun_path[0] == 0: abstract namespace
    return (socklen_t)(offsetof(struct sockaddr_un, sun_path) + 1 + n);
}

int uds_listen(const std::string &name, int backlog) {
    int fd = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (fd < 0) return -errno;
    struct sockaddr_un a;
    socklen_t l = fill_abstract(&a, name);
    if (bind(fd, (struct sockaddr *)&a, l) < 0 || listen(fd, backlog) < 0) {
        int e = errno;
        close(fd);
        return -e;
    }
    return fd;
}

int uds_connect(const std::string &name, int timeout_ms) {
    int fd = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (fd < 0) return -errno;
    struct sockaddr_un a;
    socklen_t l = fill_abstract(&a, name);
    // abstract unix connects complete immediately or fail; retry briefly for start-up races
    int waited = 0;
    for (;;) {
        if (connect(fd, (struct sockaddr *)&a, l) == 0) return fd;
        if ((errno != ECONNREFUSED && errno != ENOENT && errno != EAGAIN) || waited >= timeout_ms) {
            int e = errno;
            close(fd);
            return -e;
        }
        usleep(2000);
        waited += 2;
    }
}

int sock_put(int fd, const void *buf, size_t len) {
    const char *p = static_cast<const char *>(buf);
    while (len) {
        ssize_t n = send(fd, p, len, MSG_NOSIGNAL);
        if (n < 0) {
            if (errno == EINTR) continue;
            if (errno == EAGAIN || errno == EWOULDBLOCK) {
                wait_fd(fd, POLLOUT, 1000);
                continue;
            }
            if (errno == EPIPE || errno == ECONNRESET) return 0;
            return -errno;
        }
        p += n;
        len -= (size_t)n;
    }
    return 1;
}

int sock_get(int fd, void *buf, size_t len, int timeout_ms) {
    char *p = static_cast<char *>(buf);
    while (len) {
        if (timeout_ms >= 0) {
            int r = wait_fd(fd, POLLIN, timeout_ms);
            if (r == 0) return -ETIMEDOUT;
            if (r < 0) return r;
        }
        ssize_t n = recv(fd, p, len, 0);
        if (n == 0) return 0;
        if (n < 0) {
            if (errno == EINTR) continue;
            if (errno == EAGAIN || errno == EWOULDBLOCK) {
                wait_fd(fd, POLLIN, timeout_ms < 0 ? 1000 : timeout_ms);
                continue;
            }
            if (errno == ECONNRESET) return 0;
            return -errno;
        }
        p += n;
        len -= (size_t)n;
    }
    return 1;
}

int uds_send_fd(int sock, const void *payload, size_t len, int fd_to_pass) {
    struct msghdr mh;
    memset(&mh, 0, sizeof mh);
    struct iovec iov = {const_cast<void *>(payload), len};
    mh.msg_iov = &iov;
    mh.msg_iovlen = 1;
    alignas(struct cmsghdr) char ctl[CMSG_SPACE(sizeof(int))];
    if (fd_to_pass >= 0) {
        memset(ctl, 0, sizeof ctl);
        mh.msg_control = ctl;
        mh.msg_controllen = sizeof ctl;
        struct cmsghdr *c = CMSG_FIRSTHDR(&mh);
        c->cmsg_level = SOL_SOCKET;
        c->cmsg_type = SCM_RIGHTS;
        c->cmsg_len = CMSG_LEN(sizeof(int));
        memcpy(CMSG_DATA(c), &fd_to_pass, sizeof(int));
    }
    for (;;) {
        ssize_t n = sendmsg(sock, &mh, MSG_NOSIGNAL);
        if (n < 0 && errno == EINTR) continue;
        if (n < 0) return -errno;
        if ((size_t)n != len) return -EIO; // small fixed payloads always go out whole
        return 1;
    }
}

int uds_recv_fd(int sock, void *payload, size_t len, int *fd_out, int timeout_ms) {
    if (fd_out) *fd_out = -1;
    if (timeout_ms >= 0) {
        int r = wait_fd(sock, POLLIN, timeout_ms);
        if (r == 0) return -ETIMEDOUT;
        if (r < 0) return r;
    }
    struct msghdr mh;
    memset(&mh, 0, sizeof mh);
    struct iovec iov = {payload, len};
    mh.msg_iov = &iov;
    mh.msg_iovlen = 1;
    alignas(struct cmsghdr) char ctl[CMSG_SPACE(sizeof(int))];
    mh.msg_control = ctl;
    mh.msg_controllen = sizeof ctl;
    ssize_t n;
    do n = recvmsg(sock, &mh, MSG_CMSG_CLOEXEC | MSG_WAITALL);
    while (n < 0 && errno == EINTR);
    if (n == 0) return 0;
    if (n < 0) return -errno;
    if ((size_t)n != len) return -EIO;
    for (struct cmsghdr *c = CMSG_FIRSTHDR(&mh); c; c = CMSG_NXTHDR(&mh, c))
        if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS && fd_out)
            memcpy(fd_out, CMSG_DATA(c), sizeof(int));
    return 1;
}

} // namespace ocm
