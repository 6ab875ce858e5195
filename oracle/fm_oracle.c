/*
 * fm_oracle.c -- CPU restatement of the fast_tffm sparse-FM hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under fast_tffm_b200/ (the product) may
 * import, link or call this file.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs use it, and only as the
 * checker or as the timed CPU baseline -- never as the thing shipped.
 *
 * The reference's own ops cannot be built here (they need TensorFlow 1.x and
 * Eigen's unsupported Tensor headers; neither is in this image), so this file
 * restates the algorithm function by function, citing the reference
 * file:line each function follows (paths relative to /root/reference).
 *
 * Parity status: the parser / scorer / grad restatements are pinned by the
 * reference's own test vectors (test/fm_parser_op_test.py:7-15,
 * test/fm_scorer_op_test.py:10-63; see tests/test_oracle_golden.py).
 * Hash64 and the stock-TF glue (unique, cumsum, embedding_lookup,
 * sigmoid_cross_entropy, SparseApplyAdagrad) live in TensorFlow, which is
 * absent and unpinned (peteris.yaml:13 installs "tensorflow" without a
 * version): for those pieces parity is UNPINNED -- the published algorithm is
 * restated and anchored on the reference's call sites.
 *
 * Two flavours of the float kernels:
 *   *_det       double accumulation, fixed order: the parity checker.
 *   *_faithful  float arithmetic per output coefficient exactly as the Eigen
 *               generators do it (including powf), parallel-for over
 *               coefficients on a pthread pool: the timed CPU baseline.
 */
#define _GNU_SOURCE
#include <errno.h>
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define FMO_API __attribute__((visibility("default")))

/* error kinds, in the order the reference can raise them (fm_parser_op.cc) */
enum {
  FMO_OK = 0,
  FMO_ERR_LABEL = 1,   /* :70  "Label could not be read in example: " */
  FMO_ERR_FORMAT = 2,  /* :77, :88  "Invalid format in example: " */
  FMO_ERR_RANGE = 3,   /* :89-93 "Invalid feature id. Should be in range [0, vocabulary_size)." */
  FMO_ERR_VALUE = 4,   /* :100 "Invalid feature value. " */
  FMO_ERR_CAPACITY = 5 /* not a reference error: caller's output buffer too small */
};

/* ------------------------------------------------------------------------
 * tensorflow::Hash64 (tensorflow/core/lib/hash/hash.cc; third party, absent
 * from /root/reference, version unpinned).  Called at fm_parser_op.cc:84.
 * Murmur64A-style; default seed 0xDECAFCAFFE.
 * ---------------------------------------------------------------------- */
FMO_API uint64_t fmo_hash64_seed(const uint8_t* data, uint64_t n, uint64_t seed) {
  const uint64_t m = 0xc6a4a7935bd1e995ULL;
  const int r = 47;
  uint64_t h = seed ^ (n * m);
  while (n >= 8) {
    uint64_t k;
    memcpy(&k, data, 8); /* little-endian DecodeFixed64 */
    data += 8;
    n -= 8;
    k *= m;
    k ^= k >> r;
    k *= m;
    h ^= k;
    h *= m;
  }
  switch (n) {
    case 7: h ^= (uint64_t)data[6] << 48; /* fallthrough */
    case 6: h ^= (uint64_t)data[5] << 40; /* fallthrough */
    case 5: h ^= (uint64_t)data[4] << 32; /* fallthrough */
    case 4: h ^= (uint64_t)data[3] << 24; /* fallthrough */
    case 3: h ^= (uint64_t)data[2] << 16; /* fallthrough */
    case 2: h ^= (uint64_t)data[1] << 8;  /* fallthrough */
    case 1: h ^= (uint64_t)data[0]; h *= m;
  }
  h ^= h >> r;
  h *= m;
  h ^= h >> r;
  return h;
}

FMO_API uint64_t fmo_hash64(const uint8_t* data, uint64_t n) {
  return fmo_hash64_seed(data, n, 0xDECAFCAFFEULL);
}

/* ------------------------------------------------------------------------
 * FmParserOp::ParseLine, fm_parser_op.cc:58-109, control flow kept as is.
 * strtof / strtoll are glibc's: the very functions the reference calls.
 * `line` must be NUL-terminated (std::string::c_str()).
 * Returns an error kind; *cnt_out = number of features pushed.
 * ---------------------------------------------------------------------- */
static int parse_line(const char* line, int hash, int64_t vocab, float* label,
                      int64_t* ids, float* vals, int64_t cap, int32_t* cnt_out) {
  const char* p = line;
  char* next;
  float fv = strtof(p, &next);                       /* :69 */
  if (p == next) return FMO_ERR_LABEL;               /* :70 */
  *label = fv;                                       /* :72 */
  p = next;
  int32_t cnt;
  for (cnt = 0; *p != '\0'; ++cnt) {                 /* :75 */
    if (*p != ' ') return FMO_ERR_FORMAT;            /* :76 */
    ++p;
    if (*p == '\0') break;                           /* :79-81 trailing single space */
    int64_t id;
    if (hash) {                                      /* :82-84 */
      const char* q = p;
      while (*q != ' ' && *q != ':' && *q != '\0') ++q;
      id = (int64_t)(fmo_hash64((const uint8_t*)p, (uint64_t)(q - p)) % (uint64_t)vocab);
      next = (char*)q;
    } else {                                         /* :86-93 */
      id = strtoll(p, &next, 10);
      if (p == next) return FMO_ERR_FORMAT;
      if (!(id >= 0 && id < vocab)) return FMO_ERR_RANGE;
    }
    p = next;
    if (*p == ':') {                                 /* :95-101 */
      p += 1;
      fv = strtof(p, &next);
      if (p == next) return FMO_ERR_VALUE;
      p = next;
    } else {
      fv = 1;                                        /* :103 */
    }
    if (cnt >= cap) return FMO_ERR_CAPACITY;
    ids[cnt] = id;                                   /* :105-106 */
    vals[cnt] = fv;
  }
  *cnt_out = cnt;                                    /* :108 */
  return FMO_OK;
}

/*
 * FmParserOp::Compute, fm_parser_op.cc:29-51: serial loop over the batch.
 * Line i is text[off[i] .. off[i+1]-1): every line is followed by exactly one
 * terminator byte (the '\n' of a file, or padding), which is not part of it.
 * The first failing line wins (OpKernelContext keeps the first status).
 * Returns the error kind, *bad_line the failing line (or -1), *nnz the number
 * of features written.
 */
FMO_API int fmo_parse(const uint8_t* text, const int64_t* off, int64_t n_lines,
                      int64_t vocab, int hash, float* labels, int32_t* sizes,
                      int64_t* ids, float* vals, int64_t cap, int64_t* nnz,
                      int64_t* bad_line) {
  int64_t n = 0;
  size_t bufcap = 1024;
  char* buf = (char*)malloc(bufcap);
  *bad_line = -1;
  for (int64_t i = 0; i < n_lines; ++i) {
    size_t len = (size_t)(off[i + 1] - 1 - off[i]);
    if (len + 1 > bufcap) {
      bufcap = 2 * (len + 1);
      buf = (char*)realloc(buf, bufcap);
    }
    memcpy(buf, text + off[i], len);
    buf[len] = '\0';
    int32_t cnt = 0;
    int rc = parse_line(buf, hash, vocab, &labels[i], ids + n, vals + n, cap - n, &cnt);
    if (rc != FMO_OK) {
      free(buf);
      *bad_line = i;
      *nnz = n;
      return rc;
    }
    sizes[i] = cnt;
    n += cnt;
  }
  free(buf);
  *nnz = n;
  return FMO_OK;
}

/* Upper bound on the features of a text block: each feature needs a ' ' */
FMO_API int64_t fmo_count_spaces(const uint8_t* text, int64_t nbytes) {
  int64_t c = 0;
  for (int64_t i = 0; i < nbytes; ++i) c += (text[i] == ' ');
  return c;
}

/* glibc strtof on one token: used to pin the GPU number scanner bit for bit */
FMO_API int fmo_strtof(const char* s, float* out) {
  char* e;
  *out = strtof(s, &e);
  return (int)(e - s);
}
FMO_API int fmo_strtoll(const char* s, int64_t* out) {
  char* e;
  *out = strtoll(s, &e, 10);
  return (int)(e - s);
}

/* ------------------------------------------------------------------------
 * tf.unique (tffm/fm_model.py:72): distinct values in order of first
 * occurrence, plus for every input its index into that list.
 * ---------------------------------------------------------------------- */
FMO_API int64_t fmo_unique(const int64_t* ids, int64_t n, int64_t* ori_ids, int32_t* local) {
  uint64_t capm = 16;
  while (capm < (uint64_t)n * 2 + 1) capm <<= 1;
  int64_t* keys = (int64_t*)malloc(capm * sizeof(int64_t));
  int32_t* slot = (int32_t*)malloc(capm * sizeof(int32_t));
  for (uint64_t i = 0; i < capm; ++i) slot[i] = -1;
  int64_t u = 0;
  for (int64_t j = 0; j < n; ++j) {
    uint64_t h = (uint64_t)ids[j] * 0x9E3779B97F4A7C15ULL;
    uint64_t s = (h >> 20) & (capm - 1);
    while (slot[s] >= 0 && keys[s] != ids[j]) s = (s + 1) & (capm - 1);
    if (slot[s] < 0) {
      keys[s] = ids[j];
      slot[s] = (int32_t)u;
      ori_ids[u++] = ids[j];
    }
    local[j] = slot[s];
  }
  free(keys);
  free(slot);
  return u;
}

/* feature_poses = concat([0], cumsum(sizes))  (tffm/fm_model.py:73) */
FMO_API void fmo_poses(const int32_t* sizes, int64_t b, int32_t* poses) {
  int32_t acc = 0;
  poses[0] = 0;
  for (int64_t i = 0; i < b; ++i) {
    acc += sizes[i];
    poses[i + 1] = acc;
  }
}

/* ------------------------------------------------------------------------
 * tf.nn.embedding_lookup(vocab_blocks, ori_ids), default "mod" strategy
 * (tffm/fm_model.py:269-284, :291): nb blocks of (vocab / nb + 1) rows of
 * (K+1) floats; id -> block id % nb, row id / nb.  `blocks` holds the nb
 * blocks back to back.
 * ---------------------------------------------------------------------- */
FMO_API void fmo_lookup_mod(const float* blocks, int64_t vocab, int64_t nb, int64_t k,
                            const int64_t* ori_ids, int64_t u, float* out) {
  const int64_t rows = vocab / nb + 1, w = k + 1;
  for (int64_t i = 0; i < u; ++i) {
    const int64_t id = ori_ids[i];
    const float* src = blocks + ((id % nb) * rows + id / nb) * w;
    memcpy(out + i * w, src, (size_t)w * sizeof(float));
  }
}

/* ------------------------------------------------------------------------
 * Parity flavour of FmScorer: fm_scorer_op.h:22-36 (rank sums), :60-86 (bias,
 * bias^2, factor^2 sums), :132-138 (combination), accumulated in double.
 * params is [rows, ld] with ld >= K+1 (the reference has ld == K+1).
 * ---------------------------------------------------------------------- */
FMO_API void fmo_score_det(const int32_t* fids, const float* params, int64_t ld,
                           const float* vals, const int32_t* poses, int64_t b, int64_t k,
                           float lam_f, float lam_b, float* pred, float* reg) {
  double* s = (double*)malloc((size_t)(k > 0 ? k : 1) * sizeof(double));
  double bias_sq = 0.0, fac_sq = 0.0;
  for (int64_t i = 0; i < b; ++i) {
    double bias = 0.0, q = 0.0;
    for (int64_t c = 0; c < k; ++c) s[c] = 0.0;
    for (int64_t j = poses[i]; j < poses[i + 1]; ++j) {
      const float* row = params + (int64_t)fids[j] * ld;
      const double x = vals[j];
      bias += x * row[0];
      bias_sq += (double)row[0] * row[0];
      for (int64_t c = 0; c < k; ++c) {
        const double t = row[c + 1];
        s[c] += x * t;
        q += (x * t) * (x * t);
        fac_sq += t * t;
      }
    }
    double r1 = 0.0;
    for (int64_t c = 0; c < k; ++c) r1 += s[c] * s[c];
    pred[i] = (float)(bias - 0.5 * q + 0.5 * r1);
  }
  *reg = (float)(0.5 * lam_f * fac_sq + 0.5 * lam_b * bias_sq);
  free(s);
}

/* Parity flavour of FmGrad: fm_grad_op.h:37-49 (factor sums), :80-106
 * (per-occurrence terms), with a plain (correct) sum instead of :3-16. */
FMO_API void fmo_grad_det(const int32_t* fids, const float* params, int64_t ld,
                          const float* vals, const int32_t* poses, int64_t b, int64_t k,
                          float lam_f, float lam_b, const float* pred_grad, float reg_grad,
                          int64_t rows, float* grad_out /* [rows, ld] */) {
  const int64_t w = k + 1;
  double* g = (double*)calloc((size_t)(rows * w), sizeof(double));
  double* s = (double*)malloc((size_t)(k > 0 ? k : 1) * sizeof(double));
  for (int64_t i = 0; i < b; ++i) {
    for (int64_t c = 0; c < k; ++c) s[c] = 0.0;
    for (int64_t j = poses[i]; j < poses[i + 1]; ++j) {
      const float* row = params + (int64_t)fids[j] * ld;
      for (int64_t c = 0; c < k; ++c) s[c] += (double)vals[j] * row[c + 1];
    }
    const double pg = pred_grad[i];
    for (int64_t j = poses[i]; j < poses[i + 1]; ++j) {
      const float* row = params + (int64_t)fids[j] * ld;
      double* go = g + (int64_t)fids[j] * w;
      const double x = vals[j];
      go[0] += pg * x + (double)reg_grad * lam_b * row[0];
      for (int64_t c = 0; c < k; ++c) {
        const double t = row[c + 1];
        go[c + 1] += (double)reg_grad * lam_f * t - pg * x * x * t + pg * x * s[c];
      }
    }
  }
  for (int64_t r = 0; r < rows; ++r) {
    for (int64_t c = 0; c < w; ++c) grad_out[r * ld + c] = (float)g[r * w + c];
    for (int64_t c = w; c < ld; ++c) grad_out[r * ld + c] = 0.f;
  }
  free(g);
  free(s);
}

/* ------------------------------------------------------------------------
 * Loss (tffm/fm_model.py:311-316 logistic, :324-326 mse) and its gradient
 * w.r.t. pred.  loss = mean_i w_i * l(pred_i, y_i); pred_grad = dloss/dpred.
 * logistic: l = max(x,0) - x*y + log1p(exp(-|x|))  (TF's
 * sigmoid_cross_entropy_with_logits), dl/dx = sigmoid(x) - y.
 * kind: 0 = logistic, 1 = mse.
 * ---------------------------------------------------------------------- */
FMO_API double fmo_loss(const float* pred, const float* labels, const float* weights,
                        int64_t b, int kind, float* pred_grad) {
  double acc = 0.0;
  for (int64_t i = 0; i < b; ++i) {
    const double x = pred[i], y = labels[i], w = weights ? weights[i] : 1.0;
    double l, d;
    if (kind == 0) {
      l = (x > 0 ? x : 0) - x * y + log1p(exp(-fabs(x)));
      d = 1.0 / (1.0 + exp(-x)) - y;
    } else {
      l = (x - y) * (x - y);
      d = 2.0 * (x - y);
    }
    acc += w * l;
    if (pred_grad) pred_grad[i] = (float)(w * d / (double)b);
  }
  return b > 0 ? acc / (double)b : 0.0;
}

/* ------------------------------------------------------------------------
 * TF1 SparseApplyAdagrad on the mod-partitioned table (the update
 * tf.train.AdagradOptimizer.minimize emits for tffm/fm_model.py:341-348):
 * per touched row  acc += g*g ; w -= lr * g / sqrt(acc).  No epsilon.
 * ---------------------------------------------------------------------- */
FMO_API void fmo_adagrad_mod(float* blocks, float* acc_blocks, int64_t vocab, int64_t nb,
                             int64_t k, const int64_t* ori_ids, int64_t u,
                             const float* grad /* [u, k+1] */, float lr) {
  const int64_t rows = vocab / nb + 1, w = k + 1;
  for (int64_t i = 0; i < u; ++i) {
    const int64_t id = ori_ids[i];
    const int64_t base = ((id % nb) * rows + id / nb) * w;
    for (int64_t c = 0; c < w; ++c) {
      const double g = grad[i * w + c];
      const double a = (double)acc_blocks[base + c] + g * g;
      acc_blocks[base + c] = (float)a;
      blocks[base + c] = (float)((double)blocks[base + c] - (double)lr * g / sqrt((double)(float)a));
    }
  }
}

/* ========================================================================
 * Faithful-threaded flavour: the timed CPU baseline.
 * One work item per OUTPUT COEFFICIENT, row-major, each coefficient running
 * the reference generator's own inner loop in float (Eigen ThreadPoolDevice
 * evaluates generate() by splitting the flat coefficient range over its
 * threads).
 * ====================================================================== */
typedef struct {
  void (*fn)(int64_t lo, int64_t hi, void* ctx);
  void* ctx;
  int64_t lo, hi;
} fmo_job;

static void* fmo_thread_main(void* a) {
  fmo_job* j = (fmo_job*)a;
  j->fn(j->lo, j->hi, j->ctx);
  return NULL;
}

static void parallel_for(int64_t n, int nthreads, void (*fn)(int64_t, int64_t, void*), void* ctx) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads == 1 || n < 2 * nthreads) {
    fn(0, n, ctx);
    return;
  }
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
  fmo_job* jobs = (fmo_job*)malloc(sizeof(fmo_job) * (size_t)nthreads);
  for (int t = 0; t < nthreads; ++t) {
    jobs[t].fn = fn;
    jobs[t].ctx = ctx;
    jobs[t].lo = n * t / nthreads;
    jobs[t].hi = n * (t + 1) / nthreads;
    pthread_create(&th[t], NULL, fmo_thread_main, &jobs[t]);
  }
  for (int t = 0; t < nthreads; ++t) pthread_join(th[t], NULL);
  free(th);
  free(jobs);
}

typedef struct {
  const int32_t* fids;
  const float* params;
  const float* vals;
  const int32_t* poses;
  int64_t b, k;
  float* ranks;  /* [2, b, k] */
  float* biases; /* [3, b] */
  /* grad */
  const float* pred_grad;
  float reg_grad, lam_f, lam_b;
  float* factor_sum; /* [b, k] */
  float* grad;       /* [rows, k+1] */
} fmo_ctx;

/* generator::BiasGenerator, fm_scorer_op.h:60-86; coefficient (p, i) */
static void bias_gen(int64_t lo, int64_t hi, void* c_) {
  fmo_ctx* c = (fmo_ctx*)c_;
  for (int64_t idx = lo; idx < hi; ++idx) {
    const int64_t p = idx / c->b, i = idx % c->b;
    float r = 0;
    for (int64_t j = c->poses[i]; j < c->poses[i + 1]; ++j) {
      const int64_t off = (int64_t)c->fids[j] * (c->k + 1);
      const float fval = c->vals[j], bias = c->params[off];
      if (p == 0) r += fval * bias;
      else if (p == 1) r += bias * bias;
      else
        for (int64_t q = 0; q < c->k; ++q) {
          const float t = c->params[off + q + 1];
          r += t * t;
        }
    }
    c->biases[idx] = r;
  }
}

/* generator::FeatureRankGenerator, fm_scorer_op.h:22-36; coefficient (p,i,k) */
static void rank_gen(int64_t lo, int64_t hi, void* c_) {
  fmo_ctx* c = (fmo_ctx*)c_;
  for (int64_t idx = lo; idx < hi; ++idx) {
    const int64_t p = idx / (c->b * c->k), rem = idx % (c->b * c->k);
    const int64_t i = rem / c->k, q = rem % c->k;
    float r = 0;
    for (int64_t j = c->poses[i]; j < c->poses[i + 1]; ++j) {
      const int64_t off = (int64_t)c->fids[j] * (c->k + 1);
      const float t = c->params[off + q + 1];
      r += powf(c->vals[j] * t, (float)(p + 1));
    }
    c->ranks[idx] = r;
  }
}

/* functor::FmScorer, fm_scorer_op.h:101-140 (params has ld == K+1) */
FMO_API void fmo_score_faithful(const int32_t* fids, const float* params, const float* vals,
                                const int32_t* poses, int64_t b, int64_t k, float lam_f,
                                float lam_b, float* pred, float* reg, int nthreads) {
  fmo_ctx c;
  memset(&c, 0, sizeof(c));
  c.fids = fids; c.params = params; c.vals = vals; c.poses = poses; c.b = b; c.k = k;
  c.ranks = (float*)malloc(sizeof(float) * (size_t)(2 * b * k + 1));
  c.biases = (float*)malloc(sizeof(float) * (size_t)(3 * b + 1));
  parallel_for(3 * b, nthreads, bias_gen, &c);
  parallel_for(2 * b * k, nthreads, rank_gen, &c);
  float s1 = 0, s2 = 0;
  for (int64_t i = 0; i < b; ++i) {
    float r2 = 0, r1 = 0;
    for (int64_t q = 0; q < k; ++q) {
      r2 += c.ranks[b * k + i * k + q];
      const float a = c.ranks[i * k + q];
      r1 += 0.5f * a * a;
    }
    pred[i] = c.biases[i] - 0.5f * r2 + r1;           /* :134-136 */
    s1 += c.biases[b + i];
    s2 += c.biases[2 * b + i];
  }
  *reg = 0.5f * lam_f * s2 + 0.5f * lam_b * s1;       /* :138 */
  free(c.ranks);
  free(c.biases);
}

/* a CORRECT CAS float add (the reference's fm_grad_op.h:3-16 can double-add
 * under contention; not replicated) */
static inline void atomic_addf(float* target, float inc) {
  uint32_t* t = (uint32_t*)target;
  uint32_t old = __atomic_load_n(t, __ATOMIC_RELAXED), nw;
  float f;
  do {
    memcpy(&f, &old, 4);
    f += inc;
    memcpy(&nw, &f, 4);
  } while (!__atomic_compare_exchange_n(t, &old, nw, 0, __ATOMIC_RELAXED, __ATOMIC_RELAXED));
}

/* generator::FactorSumGenerator, fm_grad_op.h:37-49; coefficient (i,k) */
static void factor_sum_gen(int64_t lo, int64_t hi, void* c_) {
  fmo_ctx* c = (fmo_ctx*)c_;
  for (int64_t idx = lo; idx < hi; ++idx) {
    const int64_t i = idx / c->k, q = idx % c->k;
    float r = 0;
    for (int64_t j = c->poses[i]; j < c->poses[i + 1]; ++j)
      r += c->vals[j] * c->params[(int64_t)c->fids[j] * (c->k + 1) + q + 1];
    c->factor_sum[idx] = r;
  }
}

/* generator::ParamGradGenerator, fm_grad_op.h:80-106; coefficient (i, k in [0,K]) */
static void param_grad_gen(int64_t lo, int64_t hi, void* c_) {
  fmo_ctx* c = (fmo_ctx*)c_;
  const int64_t w = c->k + 1;
  for (int64_t idx = lo; idx < hi; ++idx) {
    const int64_t i = idx / w, q = idx % w;
    for (int64_t j = c->poses[i]; j < c->poses[i + 1]; ++j) {
      const int64_t fid = c->fids[j];
      const float fval = c->vals[j];
      const float t = c->params[fid * w + q];
      float v;
      if (q == 0)
        v = c->pred_grad[i] * fval + c->reg_grad * c->lam_b * t;
      else
        v = c->reg_grad * c->lam_f * t - c->pred_grad[i] * fval * fval * t +
            c->pred_grad[i] * fval * c->factor_sum[i * c->k + q - 1];
      atomic_addf(&c->grad[fid * w + q], v);
    }
  }
}

/* functor::FmGrad, fm_grad_op.h:126-163 */
FMO_API void fmo_grad_faithful(const int32_t* fids, const float* params, const float* vals,
                               const int32_t* poses, int64_t b, int64_t k, float lam_f,
                               float lam_b, const float* pred_grad, float reg_grad,
                               int64_t rows, float* grad_out, int nthreads) {
  fmo_ctx c;
  memset(&c, 0, sizeof(c));
  c.fids = fids; c.params = params; c.vals = vals; c.poses = poses; c.b = b; c.k = k;
  c.pred_grad = pred_grad; c.reg_grad = reg_grad; c.lam_f = lam_f; c.lam_b = lam_b;
  c.factor_sum = (float*)malloc(sizeof(float) * (size_t)(b * k + 1));
  c.grad = grad_out;
  parallel_for(b * k, nthreads, factor_sum_gen, &c);            /* :148 */
  memset(grad_out, 0, sizeof(float) * (size_t)(rows * (k + 1))); /* :155 */
  parallel_for(b * (k + 1), nthreads, param_grad_gen, &c);      /* :157-161 */
  free(c.factor_sum);
}

/* float flavours of the stock-TF glue for the timed baseline */
FMO_API float fmo_loss_f32(const float* pred, const float* labels, const float* weights,
                           int64_t b, int kind, float* pred_grad) {
  float acc = 0;
  for (int64_t i = 0; i < b; ++i) {
    const float x = pred[i], y = labels[i], w = weights ? weights[i] : 1.f;
    float l, d;
    if (kind == 0) {
      l = (x > 0 ? x : 0) - x * y + log1pf(expf(-fabsf(x)));
      d = 1.f / (1.f + expf(-x)) - y;
    } else {
      l = (x - y) * (x - y);
      d = 2.f * (x - y);
    }
    acc += w * l;
    pred_grad[i] = w * d / (float)b;
  }
  return b > 0 ? acc / (float)b : 0.f;
}

FMO_API void fmo_adagrad_mod_f32(float* blocks, float* acc_blocks, int64_t vocab, int64_t nb,
                                 int64_t k, const int64_t* ori_ids, int64_t u,
                                 const float* grad, float lr) {
  const int64_t rows = vocab / nb + 1, w = k + 1;
  for (int64_t i = 0; i < u; ++i) {
    const int64_t id = ori_ids[i];
    const int64_t base = ((id % nb) * rows + id / nb) * w;
    for (int64_t c = 0; c < w; ++c) {
      const float g = grad[i * w + c];
      const float a = acc_blocks[base + c] + g * g;
      acc_blocks[base + c] = a;
      blocks[base + c] -= lr * g / sqrtf(a);
    }
  }
}

/*
 * The optimizer's row loop the way the TF runtime schedules it: AdagradOptimizer.minimize over nb block variables is
 * one SparseApplyAdagrad per block (tffm/fm_model.py:269-291, :341-348), and independent ops run side by side on the
 * inter-op pool.  One task per block, each serial over its rows like the op kernel; the result equals the serial loop
 * (distinct rows, disjoint blocks).  The lookup stays the serial gather: as per-block tasks writing one output it was
 * SLOWER on the 16-core box (119 against 85 ms: the threads contend on the first-touch page faults of the 161 MB
 * result), and TF's form adds a dynamic_stitch copy on top, so the serial loop is the kinder estimate.
 */
typedef struct {
  float* blocks; float* acc; int64_t vocab, nb, k;
  const int64_t* ori; const float* grad; float lr;
  const int64_t* start;      /* [nb + 1]: block b owns order[start[b] .. start[b + 1]) */
  const int64_t* order;      /* positions 0..u-1 grouped by block (tf.dynamic_partition's output, serial like the op) */
} fmo_block_ctx;
static void partition_by_block(const int64_t* ori, int64_t u, int64_t nb, int64_t* start, int64_t* order) {
  for (int64_t b = 0; b <= nb; ++b) start[b] = 0;
  for (int64_t i = 0; i < u; ++i) ++start[ori[i] % nb + 1];
  for (int64_t b = 0; b < nb; ++b) start[b + 1] += start[b];
  int64_t* at = (int64_t*)malloc(sizeof(int64_t) * (size_t)nb);
  for (int64_t b = 0; b < nb; ++b) at[b] = start[b];
  for (int64_t i = 0; i < u; ++i) order[at[ori[i] % nb]++] = i;
  free(at);
}
static void adagrad_block_task(int64_t lo, int64_t hi, void* p) {
  const fmo_block_ctx* c = (const fmo_block_ctx*)p;
  const int64_t rows = c->vocab / c->nb + 1, w = c->k + 1;
  for (int64_t blk = lo; blk < hi; ++blk)
    for (int64_t j = c->start[blk]; j < c->start[blk + 1]; ++j) {
      const int64_t i = c->order[j];
      const int64_t base = (blk * rows + c->ori[i] / c->nb) * w;
      for (int64_t col = 0; col < w; ++col) {
        const float g = c->grad[i * w + col];
        const float a = c->acc[base + col] + g * g;
        c->acc[base + col] = a;
        c->blocks[base + col] -= c->lr * g / sqrtf(a);
      }
    }
}
/* one thread per task (parallel_for keeps small n serial: right for its per-element jobs, wrong for ten heavy ones) */
static void parallel_tasks(int64_t n, int nthreads, void (*fn)(int64_t, int64_t, void*), void* ctx) {
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
  fmo_job* jobs = (fmo_job*)malloc(sizeof(fmo_job) * (size_t)nthreads);
  for (int t = 0; t < nthreads; ++t) {
    jobs[t].fn = fn;
    jobs[t].ctx = ctx;
    jobs[t].lo = n * t / nthreads;
    jobs[t].hi = n * (t + 1) / nthreads;
    pthread_create(&th[t], NULL, fmo_thread_main, &jobs[t]);
  }
  for (int t = 0; t < nthreads; ++t) pthread_join(th[t], NULL);
  free(th);
  free(jobs);
}
static void by_block(fmo_block_ctx* c, int64_t u, int nthreads, void (*task)(int64_t, int64_t, void*)) {
  int64_t* start = (int64_t*)malloc(sizeof(int64_t) * (size_t)(c->nb + 1));
  int64_t* order = (int64_t*)malloc(sizeof(int64_t) * (size_t)(u + 1));
  partition_by_block(c->ori, u, c->nb, start, order);
  c->start = start;
  c->order = order;
  const int tasks = (int)(c->nb < nthreads ? c->nb : nthreads);
  if (tasks <= 1) task(0, c->nb, c); else parallel_tasks(c->nb, tasks, task, c);
  free(order);
  free(start);
}
static void adagrad_by_block(float* blocks, float* acc, int64_t vocab, int64_t nb, int64_t k, const int64_t* ori,
                             int64_t u, const float* grad, float lr, int nthreads) {
  fmo_block_ctx c = {blocks, acc, vocab, nb, k, ori, grad, lr, NULL, NULL};
  by_block(&c, u, nthreads, adagrad_block_task);
}

/*
 * One whole reference train step on the CPU, timed per stage (seconds in
 * t[0..6]: parse, unique+poses, lookup, score, loss, grad, adagrad), the way
 * tffm/fm_model.py strings the ops together (:69-73, :291-293, :311-348).
 * Returns the loss; <0 on a parse error.
 */
#include <time.h>
static double now_s(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

FMO_API double fmo_train_step_faithful(const uint8_t* text, const int64_t* off, int64_t n_lines,
                                       int64_t vocab, int hash, int64_t nb, int64_t k,
                                       float* blocks, float* acc_blocks, const float* weights,
                                       int kind, float lam_f, float lam_b, float lr,
                                       int64_t cfg_batch, int64_t cap, int nthreads,
                                       int do_update, double* t, float* pred_out) {
  float* labels = (float*)malloc(sizeof(float) * (size_t)n_lines);
  int32_t* sizes = (int32_t*)malloc(sizeof(int32_t) * (size_t)n_lines);
  int32_t* poses = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n_lines + 1));
  int64_t* ids = (int64_t*)malloc(sizeof(int64_t) * (size_t)cap);
  float* vals = (float*)malloc(sizeof(float) * (size_t)cap);
  int64_t nnz = 0, bad = -1;
  double loss = -1.0;
  double t0 = now_s();
  int rc = fmo_parse(text, off, n_lines, vocab, hash, labels, sizes, ids, vals, cap, &nnz, &bad);
  double t1 = now_s();
  t[0] = t1 - t0;
  if (rc == FMO_OK) {
    int64_t* ori = (int64_t*)malloc(sizeof(int64_t) * (size_t)(nnz + 1));
    int32_t* local = (int32_t*)malloc(sizeof(int32_t) * (size_t)(nnz + 1));
    int64_t u = fmo_unique(ids, nnz, ori, local);
    fmo_poses(sizes, n_lines, poses);
    double t2 = now_s();
    t[1] = t2 - t1;
    float* lp = (float*)malloc(sizeof(float) * (size_t)((u + 1) * (k + 1)));
    fmo_lookup_mod(blocks, vocab, nb, k, ori, u, lp);
    double t3 = now_s();
    t[2] = t3 - t2;
    float* pred = (float*)malloc(sizeof(float) * (size_t)n_lines);
    float reg = 0;
    fmo_score_faithful(local, lp, vals, poses, n_lines, k, lam_f, lam_b, pred, &reg, nthreads);
    double t4 = now_s();
    t[3] = t4 - t3;
    float* pg = (float*)malloc(sizeof(float) * (size_t)n_lines);
    loss = fmo_loss_f32(pred, labels, weights, n_lines, kind, pg);
    double t5 = now_s();
    t[4] = t5 - t4;
    t[5] = t[6] = 0;
    if (do_update) {
      float* g = (float*)malloc(sizeof(float) * (size_t)((u + 1) * (k + 1)));
      fmo_grad_faithful(local, lp, vals, poses, n_lines, k, lam_f, lam_b, pg,
                        1.0f / (float)cfg_batch, u, g, nthreads);
      double t6 = now_s();
      t[5] = t6 - t5;
      adagrad_by_block(blocks, acc_blocks, vocab, nb, k, ori, u, g, lr, nthreads);
      t[6] = now_s() - t6;
      free(g);
    }
    if (pred_out) memcpy(pred_out, pred, sizeof(float) * (size_t)n_lines);
    free(pg); free(pred); free(lp); free(local); free(ori);
  }
  free(vals); free(ids); free(poses); free(sizes); free(labels);
  return loss;
}
