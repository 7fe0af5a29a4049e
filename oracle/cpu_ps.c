/*
 * oracle/cpu_ps.c — CPU restatement of the reference's parameter-server step.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT. Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / `--impl reference` legs may load this file's
 * library. libaccel_ps.so never links, loads or calls it.
 *
 * PARITY UNPINNED: the arithmetic of this path lives in TensorFlow v1.4.0, a
 * third-party dependency pinned only by the image tag in the reference
 * (examples/tfjob/dist.yml:13,32; examples/tfjob/local.yml:13) and absent from
 * /root/reference; TensorFlow is not importable in this image and the reference
 * holds no test, golden vector or fixture for the path (its only test is
 * pkg/checker/checker_test.go:10-38). What follows restates the published
 * TF-1.4 algorithms:
 *   tensorflow/core/kernels/training_ops.cc   ApplyGradientDescent<CPU>,
 *                                             ApplyMomentum<CPU>, ApplyAdamNonCuda
 *   tensorflow/core/kernels/conditional_accumulator.h   (sum in arrival order;
 *                                             take_grad divides by the count)
 *   tensorflow/python/ops/clip_ops.py         clip_by_global_norm, global_norm
 *   tensorflow/python/training/adam.py        _prepare / _finish (beta powers)
 * anchored on the reference's call sites: examples/workdir/mnist_replica.py:170
 * (AdamOptimizer), :172-182 (SyncReplicasOptimizer), :184 (minimize), :256
 * (the step); examples/workdir/mnist_softmax.py:59 (GradientDescentOptimizer).
 * Known-answer vectors: SURVEY.md §8(c), held in tests/golden/kat_survey.json.
 *
 * Shape of the computation, kept on purpose: ONE MEMORY PASS PER STATEMENT, the
 * way Eigen evaluates each `x.device(d) = expr` of those functors (Adam = 3
 * passes, Momentum = 2, aggregation = one pass per arriving gradient), fp32,
 * each operation rounded on its own (build with -ffp-contract=off: the stock
 * TF-1.4 wheel has no FMA). Passes are split over `threads` host threads by
 * contiguous ranges, as Eigen's ThreadPoolDevice does, on a persistent pool.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_F32 0
#define ORACLE_BF16 1

typedef struct oracle_hparams { /* same layout as accel_hparams */
  float lr, momentum, beta1, beta2, epsilon, beta1_power, beta2_power, clip_norm;
  int32_t use_nesterov, aggregate;
} oracle_hparams;

/* ---- a pass = one statement over [0, count), split across threads -------- */
typedef void (*pass_fn)(void *ctx, uint64_t lo, uint64_t hi, int tid);
/* A persistent pool, as TF's intra-op thread pool is: workers are created once and park on a
 * condition variable between passes (creating and joining threads for each of the 9 passes of a
 * step would handicap this baseline). The calling thread works too (range 0). One pass at a
 * time: callers are serialised by g_call. */
#define POOL_MAX 256
static struct {
  pthread_mutex_t mu;
  pthread_cond_t start, done;
  pthread_t th[POOL_MAX];
  int n_workers;          /* threads created so far (worker t handles range t + 1) */
  unsigned long gen;      /* bumped once per pass */
  int pending;            /* workers that have not finished the current pass */
  pass_fn fn;
  void *ctx;
  uint64_t count, per;
  int ranges;             /* ranges of the current pass, range 0 is the caller's */
} g_pool = {PTHREAD_MUTEX_INITIALIZER, PTHREAD_COND_INITIALIZER, PTHREAD_COND_INITIALIZER,
            {0}, 0, 0, 0, NULL, NULL, 0, 0, 0};
static pthread_mutex_t g_call = PTHREAD_MUTEX_INITIALIZER;

static void run_range(pass_fn fn, void *ctx, uint64_t count, uint64_t per, int r) {
  uint64_t lo = per * (uint64_t)r, hi = lo + per;
  if (lo >= count) return;
  if (hi > count) hi = count;
  fn(ctx, lo, hi, r);
}

static void *pool_worker(void *p) {
  const int r = (int)(intptr_t)p;   /* this worker's range index, >= 1 */
  unsigned long seen = 0;
  pthread_mutex_lock(&g_pool.mu);
  for (;;) {
    while (g_pool.gen == seen) pthread_cond_wait(&g_pool.start, &g_pool.mu);
    seen = g_pool.gen;
    const pass_fn fn = g_pool.fn;
    void *ctx = g_pool.ctx;
    const uint64_t count = g_pool.count, per = g_pool.per;
    const int mine = r < g_pool.ranges;
    pthread_mutex_unlock(&g_pool.mu);
    if (mine) run_range(fn, ctx, count, per, r);
    pthread_mutex_lock(&g_pool.mu);
    if (--g_pool.pending == 0) pthread_cond_signal(&g_pool.done);
  }
  return NULL;
}

static void run_pass(pass_fn fn, void *ctx, uint64_t count, int threads) {
  if (threads < 1) threads = 1;
  if (threads > POOL_MAX) threads = POOL_MAX;
  if (threads == 1 || count < 4096) {
    fn(ctx, 0, count, 0);
    return;
  }
  /* ranges on 16-element boundaries so vector loops stay aligned */
  const uint64_t per = ((count + threads - 1) / threads + 15) & ~(uint64_t)15;
  pthread_mutex_lock(&g_call);
  pthread_mutex_lock(&g_pool.mu);
  while (g_pool.n_workers < threads - 1) {   /* grow the pool on first use */
    const int r = g_pool.n_workers + 1;
    if (pthread_create(&g_pool.th[g_pool.n_workers], NULL, pool_worker, (void *)(intptr_t)r))
      break;
    pthread_detach(g_pool.th[g_pool.n_workers]);
    ++g_pool.n_workers;
  }
  int ranges = g_pool.n_workers + 1;
  if (ranges > threads) ranges = threads;
  /* if thread creation fell short, the ranges that have no worker are run by the caller */
  const uint64_t per_eff = ranges == threads
                               ? per
                               : ((count + ranges - 1) / ranges + 15) & ~(uint64_t)15;
  g_pool.fn = fn;
  g_pool.ctx = ctx;
  g_pool.count = count;
  g_pool.per = per_eff;
  g_pool.ranges = ranges;
  g_pool.pending = g_pool.n_workers;
  ++g_pool.gen;
  pthread_cond_broadcast(&g_pool.start);
  pthread_mutex_unlock(&g_pool.mu);
  run_range(fn, ctx, count, per_eff, 0);
  pthread_mutex_lock(&g_pool.mu);
  while (g_pool.pending != 0) pthread_cond_wait(&g_pool.done, &g_pool.mu);
  pthread_mutex_unlock(&g_pool.mu);
  pthread_mutex_unlock(&g_call);
}

/* ---- aggregation: ConditionalAccumulator ---------------------------------- */
typedef struct {
  float *acc;
  const void *g;
  int dtype;
  float div;
} agg_ctx;

static inline float bf16_to_f32(uint16_t b) {
  uint32_t u = (uint32_t)b << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

static void pass_acc_set(void *c, uint64_t lo, uint64_t hi, int tid) { /* acc = g0 */
  (void)tid;
  agg_ctx *a = (agg_ctx *)c;
  if (a->dtype == ORACLE_BF16) {
    const uint16_t *g = (const uint16_t *)a->g;
    for (uint64_t i = lo; i < hi; ++i) a->acc[i] = bf16_to_f32(g[i]);
  } else {
    const float *g = (const float *)a->g;
    for (uint64_t i = lo; i < hi; ++i) a->acc[i] = g[i];
  }
}
static void pass_acc_add(void *c, uint64_t lo, uint64_t hi, int tid) { /* acc += g_w */
  (void)tid;
  agg_ctx *a = (agg_ctx *)c;
  if (a->dtype == ORACLE_BF16) {
    const uint16_t *g = (const uint16_t *)a->g;
    for (uint64_t i = lo; i < hi; ++i) a->acc[i] = a->acc[i] + bf16_to_f32(g[i]);
  } else {
    const float *g = (const float *)a->g;
    for (uint64_t i = lo; i < hi; ++i) a->acc[i] = a->acc[i] + g[i];
  }
}
static void pass_acc_div(void *c, uint64_t lo, uint64_t hi, int tid) { /* take_grad: acc / N */
  (void)tid;
  agg_ctx *a = (agg_ctx *)c;
  for (uint64_t i = lo; i < hi; ++i) a->acc[i] = a->acc[i] / a->div;
}

/* acc[count] <- aggregate of n gradients, worker order, fp32. mean != 0: / n. */
void oracle_aggregate(float *acc, const void *const *grads, int n, int dtype, uint64_t count,
                      int mean, int threads) {
  agg_ctx a = {acc, grads[0], dtype, (float)n};
  run_pass(pass_acc_set, &a, count, threads);
  for (int w = 1; w < n; ++w) {
    a.g = grads[w];
    run_pass(pass_acc_add, &a, count, threads);
  }
  if (mean) run_pass(pass_acc_div, &a, count, threads);
}

/* ---- clip_by_global_norm --------------------------------------------------- */
typedef struct {
  float *g;
  double partial[256];
  float scale;
} clip_ctx;

static void pass_sumsq(void *c, uint64_t lo, uint64_t hi, int tid) {
  clip_ctx *k = (clip_ctx *)c;
  /* fp32 products accumulated in 16 fp32 lanes (Eigen's packet reduction is a
   * lane-wise fp32 tree too); the order of a tree reduction is unspecified in
   * TF, so results are compared with the path's tolerance, not bit-exactly. */
  float lane[16] = {0};
  uint64_t i = lo;
  for (; i + 16 <= hi; i += 16)
    for (int l = 0; l < 16; ++l) lane[l] = lane[l] + k->g[i + l] * k->g[i + l];
  double s = 0.0;
  for (int l = 0; l < 16; ++l) s += lane[l];
  for (; i < hi; ++i) s += (double)(k->g[i] * k->g[i]);
  k->partial[tid] = s;
}
static void pass_scale(void *c, uint64_t lo, uint64_t hi, int tid) {
  (void)tid;
  clip_ctx *k = (clip_ctx *)c;
  for (uint64_t i = lo; i < hi; ++i) k->g[i] = k->g[i] * k->scale;
}

float oracle_sumsq(const float *g, uint64_t count, int threads) {
  clip_ctx k;
  memset(&k, 0, sizeof(k));
  k.g = (float *)g;
  run_pass(pass_sumsq, &k, count, threads);
  double s = 0.0;
  for (int t = 0; t < 256; ++t) s += k.partial[t];
  return (float)s;
}

/* scale = clip_norm * min(1/global_norm, 1/clip_norm); g *= scale */
float oracle_clip_scale(float sumsq, float clip_norm) {
  float gn = sqrtf(sumsq);
  float a = 1.0f / gn, b = 1.0f / clip_norm;
  return clip_norm * (a < b ? a : b);
}

void oracle_clip(float *g, uint64_t count, float clip_norm, float sumsq, int threads) {
  clip_ctx k;
  memset(&k, 0, sizeof(k));
  k.g = g;
  k.scale = oracle_clip_scale(sumsq, clip_norm);
  run_pass(pass_scale, &k, count, threads);
}

/* ---- optimizers -------------------------------------------------------------- */
typedef struct {
  float *var, *s0, *s1;
  const float *g;
  float lr, momentum, beta1, omb1, omb2, alpha, eps;
} opt_ctx;

/* ApplyGradientDescent: var -= grad * lr */
static void pass_sgd(void *c, uint64_t lo, uint64_t hi, int tid) {
  (void)tid;
  opt_ctx *o = (opt_ctx *)c;
  for (uint64_t i = lo; i < hi; ++i) o->var[i] = o->var[i] - o->g[i] * o->lr;
}
/* ApplyMomentum statement 1: accum = accum * momentum + grad */
static void pass_mom_accum(void *c, uint64_t lo, uint64_t hi, int tid) {
  (void)tid;
  opt_ctx *o = (opt_ctx *)c;
  for (uint64_t i = lo; i < hi; ++i) o->s0[i] = o->s0[i] * o->momentum + o->g[i];
}
/* statement 2: var -= accum * lr */
static void pass_mom_var(void *c, uint64_t lo, uint64_t hi, int tid) {
  (void)tid;
  opt_ctx *o = (opt_ctx *)c;
  for (uint64_t i = lo; i < hi; ++i) o->var[i] = o->var[i] - o->s0[i] * o->lr;
}
/* statement 2, nesterov: var -= grad * lr + accum * momentum * lr */
static void pass_mom_var_nesterov(void *c, uint64_t lo, uint64_t hi, int tid) {
  (void)tid;
  opt_ctx *o = (opt_ctx *)c;
  for (uint64_t i = lo; i < hi; ++i) {
    float t = o->g[i] * o->lr + (o->s0[i] * o->momentum) * o->lr;
    o->var[i] = o->var[i] - t;
  }
}
/* ApplyAdam statement 1: m += (grad - m) * (1 - beta1) */
static void pass_adam_m(void *c, uint64_t lo, uint64_t hi, int tid) {
  (void)tid;
  opt_ctx *o = (opt_ctx *)c;
  for (uint64_t i = lo; i < hi; ++i) o->s0[i] = o->s0[i] + (o->g[i] - o->s0[i]) * o->omb1;
}
/* statement 2: v += (grad^2 - v) * (1 - beta2) */
static void pass_adam_v(void *c, uint64_t lo, uint64_t hi, int tid) {
  (void)tid;
  opt_ctx *o = (opt_ctx *)c;
  for (uint64_t i = lo; i < hi; ++i)
    o->s1[i] = o->s1[i] + (o->g[i] * o->g[i] - o->s1[i]) * o->omb2;
}
/* statement 3: var -= (m * alpha) / (sqrt(v) + epsilon) */
static void pass_adam_var(void *c, uint64_t lo, uint64_t hi, int tid) {
  (void)tid;
  opt_ctx *o = (opt_ctx *)c;
  for (uint64_t i = lo; i < hi; ++i)
    o->var[i] = o->var[i] - (o->s0[i] * o->alpha) / (sqrtf(o->s1[i]) + o->eps);
}
/* statement 3, nesterov: var -= ((grad*(1-beta1) + beta1*m) * alpha) / (sqrt(v) + epsilon) */
static void pass_adam_var_nesterov(void *c, uint64_t lo, uint64_t hi, int tid) {
  (void)tid;
  opt_ctx *o = (opt_ctx *)c;
  for (uint64_t i = lo; i < hi; ++i) {
    float num = (o->g[i] * o->omb1 + o->beta1 * o->s0[i]) * o->alpha;
    o->var[i] = o->var[i] - num / (sqrtf(o->s1[i]) + o->eps);
  }
}

float oracle_adam_alpha(float lr, float beta1_power, float beta2_power) {
  float num = lr * sqrtf(1.0f - beta2_power);
  return num / (1.0f - beta1_power);
}

void oracle_apply_sgd(float *var, const float *g, uint64_t count, float lr, int threads) {
  opt_ctx o = {var, NULL, NULL, g, lr, 0, 0, 0, 0, 0, 0};
  run_pass(pass_sgd, &o, count, threads);
}

void oracle_apply_momentum(float *var, float *accum, const float *g, uint64_t count, float lr,
                           float momentum, int nesterov, int threads) {
  opt_ctx o = {var, accum, NULL, g, lr, momentum, 0, 0, 0, 0, 0};
  run_pass(pass_mom_accum, &o, count, threads);
  run_pass(nesterov ? pass_mom_var_nesterov : pass_mom_var, &o, count, threads);
}

void oracle_apply_adam(float *var, float *m, float *v, const float *g, uint64_t count, float lr,
                       float beta1, float beta2, float epsilon, float beta1_power,
                       float beta2_power, int nesterov, int threads) {
  opt_ctx o = {var, m, v, g, lr, 0, beta1, 1.0f - beta1, 1.0f - beta2,
               oracle_adam_alpha(lr, beta1_power, beta2_power), epsilon};
  run_pass(pass_adam_m, &o, count, threads);
  run_pass(pass_adam_v, &o, count, threads);
  run_pass(nesterov ? pass_adam_var_nesterov : pass_adam_var, &o, count, threads);
}

/* One PS step: aggregate -> (clip) -> apply. optimizer: 0 SGD, 1 Momentum, 2 Adam.
 * acc: caller's scratch of `count` floats (the accumulator tensor). Returns the
 * sum of squares used for clipping (0 when clipping is off). */
float oracle_ps_step(int optimizer, float *var, float *slot0, float *slot1,
                     const void *const *grads, int n_workers, int dtype, uint64_t count,
                     const oracle_hparams *hp, float *acc, int threads) {
  float ss = 0.0f;
  oracle_aggregate(acc, grads, n_workers, dtype, count, hp->aggregate == 1, threads);
  if (hp->clip_norm > 0.0f) {
    ss = oracle_sumsq(acc, count, threads);
    oracle_clip(acc, count, hp->clip_norm, ss, threads);
  }
  if (optimizer == 0)
    oracle_apply_sgd(var, acc, count, hp->lr, threads);
  else if (optimizer == 1)
    oracle_apply_momentum(var, slot0, acc, count, hp->lr, hp->momentum, hp->use_nesterov,
                          threads);
  else
    oracle_apply_adam(var, slot0, slot1, acc, count, hp->lr, hp->beta1, hp->beta2, hp->epsilon,
                      hp->beta1_power, hp->beta2_power, hp->use_nesterov, threads);
  return ss;
}
