#include "ssn_pool.h"

#include <pthread.h>
#include <stdlib.h>

typedef struct group {
    ssn_pool_fn fn;
    void *arg;
    int n, next, done;          /* next index to hand out; finished count */
    struct group *link;
} group;

struct ssn_pool {
    int n_threads, stop;
    pthread_t *th;
    pthread_mutex_t mu;
    pthread_cond_t work, fin;
    group *head;                /* groups that still have indices to hand out */
};

static void unlink_group(ssn_pool *p, group *g)
{
    group **pp = &p->head;
    while (*pp && *pp != g) pp = &(*pp)->link;
    if (*pp) *pp = g->link;
}

/* mu held on entry and exit */
static void run_one(ssn_pool *p, group *g)
{
    const int i = g->next++;
    if (g->next >= g->n) unlink_group(p, g);
    pthread_mutex_unlock(&p->mu);
    g->fn(g->arg, i);
    pthread_mutex_lock(&p->mu);
    if (++g->done == g->n) pthread_cond_broadcast(&p->fin);
}

static void *worker(void *v)
{
    ssn_pool *p = (ssn_pool*)v;
    pthread_mutex_lock(&p->mu);
    for (;;) {
        while (!p->stop && !p->head) pthread_cond_wait(&p->work, &p->mu);
        if (p->stop) break;
        run_one(p, p->head);
    }
    pthread_mutex_unlock(&p->mu);
    return 0;
}

ssn_pool *ssn_pool_create(int n_threads)
{
    ssn_pool *p = (ssn_pool*)calloc(1, sizeof(*p));
    int i;
    if (!p) return 0;
    pthread_mutex_init(&p->mu, 0);
    pthread_cond_init(&p->work, 0);
    pthread_cond_init(&p->fin, 0);
    if (n_threads < 0) n_threads = 0;
    p->th = (pthread_t*)calloc((size_t)n_threads + 1, sizeof(pthread_t));
    for (i = 0; i < n_threads; ++i) {
        if (pthread_create(&p->th[p->n_threads], 0, worker, p) != 0) break;
        ++p->n_threads;
    }
    return p;
}

void ssn_pool_destroy(ssn_pool *p)
{
    int i;
    if (!p) return;
    pthread_mutex_lock(&p->mu);
    p->stop = 1;
    pthread_cond_broadcast(&p->work);
    pthread_mutex_unlock(&p->mu);
    for (i = 0; i < p->n_threads; ++i) pthread_join(p->th[i], 0);
    pthread_mutex_destroy(&p->mu);
    pthread_cond_destroy(&p->work);
    pthread_cond_destroy(&p->fin);
    free(p->th);
    free(p);
}

int ssn_pool_threads(const ssn_pool *p) { return p ? p->n_threads : 0; }

void ssn_pool_for(ssn_pool *p, int n, ssn_pool_fn fn, void *arg)
{
    group g;
    int i;
    if (n <= 0) return;
    if (!p || p->n_threads == 0 || n == 1) {
        for (i = 0; i < n; ++i) fn(arg, i);
        return;
    }
    g.fn = fn; g.arg = arg; g.n = n; g.next = 0; g.done = 0; g.link = 0;
    pthread_mutex_lock(&p->mu);
    {   /* append so that older groups drain first */
        group **pp = &p->head;
        while (*pp) pp = &(*pp)->link;
        *pp = &g;
    }
    pthread_cond_broadcast(&p->work);
    while (g.next < g.n) run_one(p, &g);        /* the caller works on its own group */
    while (g.done < g.n) pthread_cond_wait(&p->fin, &p->mu);
    pthread_mutex_unlock(&p->mu);
}
