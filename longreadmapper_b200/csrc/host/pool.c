/*
 * pool.c -- a small persistent thread pool for the ingest and emit stages (host tooling, plain C + pthreads).
 * Workers sleep on a condition variable between jobs: the stages of accaln_b200 run side by side with the threads that
 * drive the GPUs, and a pool that spins while idle (OpenMP's default) takes the cores those threads need to launch kernels
 * and to wake up on their events.
 */
#include "accseq_host.h"

#include <pthread.h>
#include <stdlib.h>

struct as_pool {
    int n;
    pthread_t *th;
    pthread_mutex_t mu;
    pthread_cond_t go, done;
    void (*fn)(int, int, void *);
    void *arg;
    unsigned long gen;          /* job number; a worker runs each generation once */
    int pending;                /* workers that have not finished the current job */
    int stop;
};

typedef struct {
    struct as_pool *p;
    int id;
} worker_arg;

static void *worker(void *v) {
    worker_arg *wa = v;
    struct as_pool *p = wa->p;
    const int id = wa->id;
    free(wa);
    unsigned long seen = 0;
    pthread_mutex_lock(&p->mu);
    for (;;) {
        while (!p->stop && p->gen == seen) pthread_cond_wait(&p->go, &p->mu);
        if (p->stop) break;
        seen = p->gen;
        void (*fn)(int, int, void *) = p->fn;
        void *arg = p->arg;
        pthread_mutex_unlock(&p->mu);
        fn(id, p->n, arg);
        pthread_mutex_lock(&p->mu);
        if (--p->pending == 0) pthread_cond_signal(&p->done);
    }
    pthread_mutex_unlock(&p->mu);
    return NULL;
}

as_pool *as_pool_new(int threads) {
    if (threads < 1) threads = 1;
    as_pool *p = calloc(1, sizeof *p);
    if (!p) return NULL;
    p->n = threads;
    pthread_mutex_init(&p->mu, NULL);
    pthread_cond_init(&p->go, NULL);
    pthread_cond_init(&p->done, NULL);
    p->th = calloc((size_t)threads, sizeof *p->th);
    if (!p->th) {
        free(p);
        return NULL;
    }
    for (int i = 1; i < threads; i++) {            /* worker 0 is the caller of as_pool_run */
        worker_arg *wa = malloc(sizeof *wa);
        wa->p = p;
        wa->id = i;
        if (pthread_create(&p->th[i], NULL, worker, wa) != 0) {
            free(wa);
            p->n = i;
            break;
        }
    }
    return p;
}

int as_pool_threads(const as_pool *p) { return p ? p->n : 1; }

/* fn(t, n, arg) runs once for every t in [0, n); returns when all of them have */
void as_pool_run(as_pool *p, void (*fn)(int, int, void *), void *arg) {
    if (!p || p->n == 1) {
        fn(0, 1, arg);
        return;
    }
    pthread_mutex_lock(&p->mu);
    p->fn = fn;
    p->arg = arg;
    p->pending = p->n - 1;
    p->gen++;
    pthread_cond_broadcast(&p->go);
    pthread_mutex_unlock(&p->mu);
    fn(0, p->n, arg);
    pthread_mutex_lock(&p->mu);
    while (p->pending) pthread_cond_wait(&p->done, &p->mu);
    pthread_mutex_unlock(&p->mu);
}

void as_pool_free(as_pool *p) {
    if (!p) return;
    pthread_mutex_lock(&p->mu);
    p->stop = 1;
    pthread_cond_broadcast(&p->go);
    pthread_mutex_unlock(&p->mu);
    for (int i = 1; i < p->n; i++) pthread_join(p->th[i], NULL);
    free(p->th);
    free(p);
}
