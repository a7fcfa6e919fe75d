/* oracle/par.h — tiny pthread parallel-for for the oracle's multi-threaded CPU baselines (test infrastructure). */
#ifndef KVO_PAR_H
#define KVO_PAR_H
#include <pthread.h>
#include <stdlib.h>
#include <unistd.h>

typedef void (*kvo_body_fn)(long begin, long end, void* ctx);
struct kvo_par_arg {
  kvo_body_fn fn;
  void* ctx;
  long begin, end;
};
static void* kvo_par_tramp(void* a_) {
  struct kvo_par_arg* a = (struct kvo_par_arg*)a_;
  a->fn(a->begin, a->end, a->ctx);
  return NULL;
}
/* static contiguous partition of [0,total) over `threads` workers (threads<=1: inline) */
static inline void kvo_parallel_for(long total, int threads, kvo_body_fn fn, void* ctx) {
  if (threads <= 1 || total <= 1) {
    fn(0, total, ctx);
    return;
  }
  if (threads > total) threads = (int)total;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
  struct kvo_par_arg* args = (struct kvo_par_arg*)malloc(sizeof(struct kvo_par_arg) * (size_t)threads);
  for (int t = 0; t < threads; ++t) {
    args[t].fn = fn;
    args[t].ctx = ctx;
    args[t].begin = total * t / threads;
    args[t].end = total * (t + 1) / threads;
    pthread_create(&th[t], NULL, kvo_par_tramp, &args[t]);
  }
  for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
  free(th);
  free(args);
}
static inline int kvo_hw_threads(void) {
  long n = sysconf(_SC_NPROCESSORS_ONLN);
  return n > 0 ? (int)n : 1;
}
#endif
