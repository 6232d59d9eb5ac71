// Leveled logger with a cached environment check.
// Replaces the reference's printd/BUG/ABORT macros (reference inc/debug.h:22-65),
// which call getenv() on every print and mix stderr/stdout.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <unistd.h>
#include <sys/syscall.h>

namespace ocm {

enum LogLevel { LOG_ERROR = 0, LOG_WARN = 1, LOG_INFO = 2, LOG_DEBUG = 3 };

inline int log_level() {
    static int lvl = [] {
        if (const char *e = getenv("OCM_LOG")) {
            if (!strcasecmp(e, "debug")) return (int)LOG_DEBUG;
            if (!strcasecmp(e, "info")) return (int)LOG_INFO;
            if (!strcasecmp(e, "warn")) return (int)LOG_WARN;
            if (!strcasecmp(e, "error")) return (int)LOG_ERROR;
            return atoi(e);
        }
        // OCM_VERBOSE (existence only) is the reference's switch: README:45-49.
        return getenv("OCM_VERBOSE") ? (int)LOG_DEBUG : (int)LOG_WARN;
    }();
    return lvl;
}

inline const char *log_tag() {
    static char tag[48] = "";
    return tag;
}
inline void set_log_tag(const char *t) {
    char *dst = const_cast<char *>(log_tag());
    snprintf(dst, 48, "%s", t);
}

__attribute__((format(printf, 5, 6)))
inline void log_write(int lvl, const char *file, const char *func, int line, const char *fmt, ...) {
    static const char *names[] = {"E", "W", "I", "D"};
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    struct timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    const char *base = strrchr(file, '/');
    fprintf(stderr, "[ocm %s %ld.%03ld %s(%d:%ld) %s::%s:%d] %s\n", names[lvl & 3],
            (long)(ts.tv_sec % 100000), ts.tv_nsec / 1000000, log_tag(), (int)getpid(),
            (long)syscall(SYS_gettid), base ? base + 1 : file, func, line, buf);
}

} // namespace ocm

#define OCM_LOG(lvl, ...)                                                            \
    do {                                                                             \
        if ((lvl) <= ::ocm::log_level())                                             \
            ::ocm::log_write((lvl), __FILE__, __func__, __LINE__, __VA_ARGS__);      \
    } while (0)
#define LOGE(...) OCM_LOG(::ocm::LOG_ERROR, __VA_ARGS__)
#define LOGW(...) OCM_LOG(::ocm::LOG_WARN, __VA_ARGS__)
#define LOGI(...) OCM_LOG(::ocm::LOG_INFO, __VA_ARGS__)
#define LOGD(...) OCM_LOG(::ocm::LOG_DEBUG, __VA_ARGS__)

// Invariant violation in our own code: print and abort (never compiled out).
#define OCM_BUG(expr)                                                                \
    do {                                                                             \
        if (__builtin_expect(!!(expr), 0)) {                                         \
            ::ocm::log_write(::ocm::LOG_ERROR, __FILE__, __func__, __LINE__,         \
                             "BUG: %s", #expr);                                      \
            abort();                                                                 \
        }                                                                            \
    } while (0)
