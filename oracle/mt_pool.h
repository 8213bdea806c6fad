/*
 * mt_pool.h — a minimal pthread work queue for the CPU checker libraries. TEST INFRASTRUCTURE ONLY
 * (see me_oracle.h): used by the multi-threaded CPU baseline of bench.py.
 *
 * The reference encoder runs its ME stage on a pool of ME threads that pull (picture, segment)
 * tasks from one FIFO (Source/Lib/Codec/me_process.c:97-124, thread counts
 * Source/Lib/Globals/enc_handle.c:768-805, segments :443-461). mt_run is that scheme reduced to
 * its core: n_items tasks, `threads` workers, each worker takes the next task index from a shared
 * atomic counter until none are left.
 */
#ifndef MT_POOL_H
#define MT_POOL_H
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>

typedef void (*MtItemFn)(void *ctx, uint32_t item, uint32_t thread);

typedef struct MtShared {
    MtItemFn fn;
    void    *ctx;
    uint32_t n_items;
    uint32_t next; /* atomic */
} MtShared;

typedef struct MtWorker {
    MtShared *sh;
    uint32_t  index;
    pthread_t tid;
} MtWorker;

static void *mt_worker_main(void *arg) {
    MtWorker *w = (MtWorker *)arg;
    for (;;) {
        const uint32_t i = __atomic_fetch_add(&w->sh->next, 1u, __ATOMIC_RELAXED);
        if (i >= w->sh->n_items) break;
        w->sh->fn(w->sh->ctx, i, w->index);
    }
    return NULL;
}

/* Runs fn(ctx, item, thread) for item in [0, n_items) on `threads` workers (the caller is worker 0).
 * Returns 0, or -1 if a thread could not be created (the work is still completed by the others). */
static int mt_run(uint32_t n_items, uint32_t threads, MtItemFn fn, void *ctx) {
    if (threads < 1) threads = 1;
    if (threads > n_items) threads = n_items ? n_items : 1;
    MtShared  sh = {fn, ctx, n_items, 0};
    MtWorker *w  = (MtWorker *)calloc(threads, sizeof(*w));
    if (!w) return -1;
    int rc = 0;
    uint32_t started = 0;
    for (uint32_t t = 1; t < threads; ++t) {
        w[t].sh    = &sh;
        w[t].index = t;
        if (pthread_create(&w[t].tid, NULL, mt_worker_main, &w[t]) != 0) {
            rc = -1;
            break;
        }
        started = t;
    }
    w[0].sh    = &sh;
    w[0].index = 0;
    mt_worker_main(&w[0]);
    for (uint32_t t = 1; t <= started; ++t) pthread_join(w[t].tid, NULL);
    free(w);
    return rc;
}
#endif
