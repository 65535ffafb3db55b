/*
 * common.h - small helpers shared by the host code.
 *
 * Provides the names the reference's include/common.h:19-38 provides (min,
 * divup, roundup, unlikely, unused, NS_IN_SEC, get_time_ns) because lib/salz.h
 * includes this header and consumers of the reference inherit them.
 */
#ifndef SALZ_COMMON_H
#define SALZ_COMMON_H

#include <errno.h>
#include <stdint.h>
#include <time.h>

#ifndef min
#define min(a, b) (((a) < (b)) ? (a) : (b))
#endif
#define divup(a, b) (((a) + (b) - 1) / (b))
#define roundup(a, b) (divup(a, b) * b)

#define unlikely(x) __builtin_expect((x), 0)
#define unused(x) ((void)(x))

#define NS_IN_SEC (1 * 1000 * 1000 * 1000)

/* Monotonic clock in nanoseconds; returns 0 or errno. */
static inline int get_time_ns(uint64_t *res)
{
    struct timespec now;
    if (clock_gettime(CLOCK_MONOTONIC, &now) != 0)
        return errno;
    *res = (uint64_t)now.tv_sec * NS_IN_SEC + (uint64_t)now.tv_nsec;
    return 0;
}

#endif /* SALZ_COMMON_H */
