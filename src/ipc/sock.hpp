// Socket helpers: TCP for daemon<->daemon control, abstract AF_UNIX sockets for
// handing CUDA VMM file descriptors to applications (SCM_RIGHTS) - POSIX message
// queues cannot carry descriptors.
// Counterpart of the reference's conn_* wrappers (reference src/sock.c:18-261,
// inc/sock.h:30-47) but connections are persistent and non-leaking; the
// reference opens and abandons one TCP connection per message
// (src/mem.c:62-111, 420-427).
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>

namespace ocm {

// All return a descriptor >= 0 or -errno.
int tcp_listen(const std::string &ip, int port, int backlog = 64);
int tcp_connect(const std::string &host, int port, int timeout_ms = 2000);
int tcp_accept(int listen_fd, std::string *peer = nullptr);

int uds_listen(const std::string &abstract_name, int backlog = 64);
int uds_connect(const std::string &abstract_name, int timeout_ms = 2000);

int set_nonblocking(int fd, bool on);

// Blocking transfer of exactly len bytes.  1 ok, 0 peer closed, -errno error
// (same tri-state as the reference's conn_put/conn_get, src/sock.c:215-253).
int sock_put(int fd, const void *buf, size_t len);
int sock_get(int fd, void *buf, size_t len, int timeout_ms = -1);

// One datagram-like exchange on a stream socket: fixed payload plus up to one
// descriptor.  send: fd_to_pass < 0 sends none.  recv: *fd_out = -1 if none came.
int uds_send_fd(int sock, const void *payload, size_t len, int fd_to_pass);
int uds_recv_fd(int sock, void *payload, size_t len, int *fd_out, int timeout_ms = 5000);

void sock_close(int fd);

} // namespace ocm
