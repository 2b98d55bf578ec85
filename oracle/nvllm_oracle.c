/*
 * nvllm_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C, fp32, single thread) of the reference's forward
 * pass so that the CUDA path can be checked against it.  Nothing in the
 * product path may link, import or call this file: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 *
 * PARITY UNPINNED: the reference (hashneo/nano-vllm-go) is Go and there is no
 * Go toolchain in this image, so the reference itself cannot be run here, and
 * none of its own tests executes a forward pass.  What IS
 * pinned: the Falcon QKV de-interleave KAT (purego/tensor/falcon_split_test.go)
 * and the block-manager / sequence tests (restated in
 * nano-vllm-go_b200/engine.py, checked in tests/).  The numerics below are a
 * loop-for-loop restatement (same accumulation order, transcendental calls in
 * double exactly where the Go code goes through float64), self-checked by
 * cached == uncached forward and cross-checked against HuggingFace
 * transformers on tiny random-init models (tests/golden/).
 *
 * Build: gcc -O3 -ffp-contract=off -fno-fast-math  (no FMA contraction: Go on
 * amd64 never fuses a*b+c; with that flag the auto-vectoriser keeps every
 * per-element accumulation order, so results equal the scalar Go loops).
 *
 * Every function cites the reference file:line it follows
 * (paths relative to /root/reference).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/nvllm.h"

#define MAX_LAYERS 128

typedef struct {
  const float *attn_norm_w, *attn_norm_b, *ffn_norm_w, *ffn_norm_b;
  const float *q, *k, *v, *kv, *o;
  const float *q_b, *k_b, *v_b, *o_b;
  const float *w1, *b1, *w2, *b2;
  const float *router, *moe_in, *moe_out;
  /* Mamba2Layer (mamba2.go:9-28), un-transposed as the loader leaves them */
  const float *m_in_proj, *m_conv_w, *m_conv_b, *m_dt_bias, *m_a_log, *m_d, *m_norm, *m_out_proj;
} layer_w;

typedef struct oracle {
  nvllm_config cfg;
  const float *wte, *wpe, *lnf_w, *lnf_b, *lm_head; /* lm_head [H,V] (post-load) or NULL => tied */
  float* lm_head_owned;
  layer_w L[MAX_LAYERS];
  float *rope_cos, *rope_sin; /* [max_seq_len, hd]  rope.go:18-50 */
  int is_mamba[MAX_LAYERS];   /* ModelConfig.HybridLayers (config.go:113) */
  char err[256];
} oracle;

/* purego/tensor/kv_cache.go:4-39 — per-layer K,V [n_kv, len, hd]; the
 * reference grows them with Concatenate (tensor.go:283-321); we keep capacity
 * max_seq_len and a length, the values are identical. */
typedef struct okv {
  int n_layers, n_kv, hd, cap, len;
  float* k; /* [L][n_kv][cap][hd] */
  float* v;
  /* Mamba2Layer.SSMState (mamba2.go:27) of every layer, [L][heads][head_dim][state]; the
   * reference keeps it in the layer object and resets it on a new prompt
   * (generic_model.go:285-292) — per sequence here, zero at creation / reset */
  float* ssm;
  size_t ssm_layer;
} okv;

/* ---------------------------------------------------------------- helpers */

/* tensor.go:62-88 MatMul: i-p-j, result zero-initialised, fp32 accumulate. */
static void matmul(const float* a, const float* b, float* c, int m, int k, int n) {
  memset(c, 0, sizeof(float) * (size_t)m * n);
  for (int i = 0; i < m; i++) {
    const float* arow = a + (size_t)i * k;
    float* crow = c + (size_t)i * n;
    for (int p = 0; p < k; p++) {
      const float av = arow[p];
      const float* brow = b + (size_t)p * n;
      for (int j = 0; j < n; j++) crow[j] += av * brow[j];
    }
  }
}

static void add_bias(float* x, const float* b, int rows, int n) {
  if (!b) return;
  for (int i = 0; i < rows; i++)
    for (int j = 0; j < n; j++) x[(size_t)i * n + j] += b[j];
}

/* tensor.go:128-155 Softmax (2-D branch): max starts at element 0, exp in f64. */
static void softmax_rows(float* x, int rows, int cols) {
  for (int i = 0; i < rows; i++) {
    float* r = x + (size_t)i * cols;
    float mx = r[0];
    for (int j = 1; j < cols; j++)
      if (r[j] > mx) mx = r[j];
    float sum = 0.f;
    for (int j = 0; j < cols; j++) {
      float v = (float)exp((double)(r[j] - mx));
      r[j] = v;
      sum += v;
    }
    for (int j = 0; j < cols; j++) r[j] /= sum;
  }
}

/* tensor.go:193-250 LayerNorm; RMSNorm when bias == NULL. */
static void layer_norm(const float* x, const float* w, const float* b, float eps, float* y,
                       int rows, int hs) {
  for (int i = 0; i < rows; i++) {
    const float* xr = x + (size_t)i * hs;
    float* yr = y + (size_t)i * hs;
    if (!b) {
      float rms = 0.f;
      for (int j = 0; j < hs; j++) {
        float v = xr[j];
        rms += v * v;
      }
      rms = (float)sqrt((double)(rms / (float)hs + eps));
      for (int j = 0; j < hs; j++) {
        float nrm = xr[j] / rms;
        yr[j] = nrm * w[j];
      }
    } else {
      float mean = 0.f;
      for (int j = 0; j < hs; j++) mean += xr[j];
      mean /= (float)hs;
      float var = 0.f;
      for (int j = 0; j < hs; j++) {
        float d = xr[j] - mean;
        var += d * d;
      }
      var /= (float)hs;
      float sd = (float)sqrt((double)(var + eps));
      for (int j = 0; j < hs; j++) {
        float nrm = (xr[j] - mean) / sd;
        yr[j] = nrm * w[j] + b[j];
      }
    }
  }
}

/* tensor.go:181-190 GELU (tanh approximation, used for every model incl. Falcon). */
static inline float gelu1(float x) {
  float x3 = x * x * x;
  float t = x + 0.044715f * x3;
  double inner = sqrt(2.0 / M_PI) * (double)t;
  float th = (float)tanh(inner);
  return 0.5f * x * (1.0f + th);
}

/* mamba2.go:360-367 SiLU: sigmoid in f64. */
static inline float silu1(float x) {
  double s = 1.0 / (1.0 + exp(-(double)x));
  return x * (float)s;
}

/* rope.go:18-50 NewRoPECache. */
static void build_rope(oracle* o, double base) {
  int hd = o->cfg.head_dim, half = hd / 2, n = o->cfg.max_seq_len;
  o->rope_cos = (float*)malloc(sizeof(float) * (size_t)n * hd);
  o->rope_sin = (float*)malloc(sizeof(float) * (size_t)n * hd);
  for (int pos = 0; pos < n; pos++)
    for (int i = 0; i < half; i++) {
      double freq = 1.0 / pow(base, (double)(2 * i) / (double)hd);
      double ang = (double)pos * freq;
      float c = (float)cos(ang), s = (float)sin(ang);
      o->rope_cos[(size_t)pos * hd + i] = c;
      o->rope_cos[(size_t)pos * hd + half + i] = c;
      o->rope_sin[(size_t)pos * hd + i] = s;
      o->rope_sin[(size_t)pos * hd + half + i] = s;
    }
}

/* rope.go:153-205 ApplyRoPESingleTensor (== the Q / K loops of ApplyRoPE :55-148)
 * t: [heads, seq, hd]. Returns -1 when pos >= max_seq_len (the Go code panics). */
static int apply_rope(const oracle* o, float* t, int heads, int seq, int start) {
  int hd = o->cfg.head_dim, half = hd / 2;
  float orig[1024];
  for (int h = 0; h < heads; h++)
    for (int s = 0; s < seq; s++) {
      int pos = start + s;
      if (pos >= o->cfg.max_seq_len) return -1;
      float* p = t + ((size_t)h * seq + s) * hd;
      const float* cs = o->rope_cos + (size_t)pos * hd;
      const float* sn = o->rope_sin + (size_t)pos * hd;
      memcpy(orig, p, sizeof(float) * hd);
      for (int i = 0; i < hd; i++) {
        float rot = (i < half) ? -orig[i + half] : orig[i - half];
        p[i] = orig[i] * cs[i] + rot * sn[i];
      }
    }
  return 0;
}

/* [seq, heads*hd] -> [heads, seq, hd]  (attention.go:88-106,300-331; mqa.go:142-158) */
static void split_heads(const float* x, float* y, int seq, int heads, int hd) {
  for (int s = 0; s < seq; s++)
    for (int h = 0; h < heads; h++)
      memcpy(y + ((size_t)h * seq + s) * hd, x + ((size_t)s * heads + h) * hd, sizeof(float) * hd);
}
/* [heads, seq, hd] -> [seq, heads*hd]  (attention.go:108-125,423-437; mqa.go:247-267) */
static void merge_heads(const float* x, float* y, int seq, int heads, int hd) {
  for (int h = 0; h < heads; h++)
    for (int s = 0; s < seq; s++)
      memcpy(y + ((size_t)s * heads + h) * hd, x + ((size_t)h * seq + s) * hd, sizeof(float) * hd);
}

/* Append new K/V [n_kv, seq, hd] to the cache of `layer` (Concatenate dim 2). */
static void kv_append(okv* c, int layer, const float* k, const float* v, int seq) {
  for (int h = 0; h < c->n_kv; h++) {
    size_t dst = (((size_t)layer * c->n_kv + h) * c->cap + c->len) * c->hd;
    memcpy(c->k + dst, k + (size_t)h * seq * c->hd, sizeof(float) * (size_t)seq * c->hd);
    memcpy(c->v + dst, v + (size_t)h * seq * c->hd, sizeof(float) * (size_t)seq * c->hd);
  }
}

/*
 * Scaled dot-product attention for one query head against kv head `kvh`.
 * style 0: MHA (attention.go:127-191) and MQA (mqa.go:184-243):
 *          scores = (q.k) * scale, THEN masked entries overwritten with -1e10,
 *          Softmax with max initialised to element 0 (tensor.go:128-155).
 * style 1: GQA (attention.go:354-397,439-470,399-421): masked entries are
 *          -1e10 without computing the dot product, softmax max initialised
 *          to -1e10.
 * q: [qs, hd]; K,V: cache rows [kvs, hd]; out: [qs, hd]
 */
static void sdpa_head(const float* q, const float* K, const float* V, float* out, float* scores,
                      int qs, int kvs, int hd, float scale, int style) {
  for (int i = 0; i < qs; i++) {
    float* sr = scores + (size_t)i * kvs;
    int maxpos = kvs - qs + i;
    for (int j = 0; j < kvs; j++) {
      if (style == 1 && j > maxpos) {
        sr[j] = -1e10f;
        continue;
      }
      float sum = 0.f;
      const float* qr = q + (size_t)i * hd;
      const float* kr = K + (size_t)j * hd;
      for (int d = 0; d < hd; d++) sum += qr[d] * kr[d];
      sr[j] = sum * scale;
    }
    if (style == 0)
      for (int j = maxpos + 1; j < kvs; j++) sr[j] = -1e10f;
  }
  if (style == 0) {
    softmax_rows(scores, qs, kvs);
  } else {
    for (int i = 0; i < qs; i++) {
      float* sr = scores + (size_t)i * kvs;
      float mx = -1e10f;
      for (int j = 0; j < kvs; j++)
        if (sr[j] > mx) mx = sr[j];
      float sum = 0.f;
      for (int j = 0; j < kvs; j++) {
        sr[j] = (float)exp((double)(sr[j] - mx));
        sum += sr[j];
      }
      for (int j = 0; j < kvs; j++) sr[j] /= sum;
    }
  }
  for (int i = 0; i < qs; i++)
    for (int d = 0; d < hd; d++) {
      float sum = 0.f;
      const float* sr = scores + (size_t)i * kvs;
      for (int j = 0; j < kvs; j++) sum += sr[j] * V[(size_t)j * hd + d];
      out[(size_t)i * hd + d] = sum;
    }
}

/* transformer.go:40-96 FeedForward.Forward. x: [T,H] -> y: [T,H] */
static void ffn_forward(const oracle* o, const layer_w* w, const float* x, float* y, int T) {
  int H = o->cfg.hidden, F = o->cfg.ffn_dim;
  if (o->cfg.activation_type == NVLLM_ACT_SWIGLU) {
    float* h = (float*)malloc(sizeof(float) * (size_t)T * 2 * F);
    float* g = (float*)malloc(sizeof(float) * (size_t)T * F);
    matmul(x, w->w1, h, T, H, 2 * F);
    for (int i = 0; i < T; i++)
      for (int j = 0; j < F; j++) {
        float gate = silu1(h[(size_t)i * 2 * F + j]);
        g[(size_t)i * F + j] = gate * h[(size_t)i * 2 * F + F + j];
      }
    matmul(g, w->w2, y, T, F, H);
    free(h);
    free(g);
  } else {
    float* h = (float*)malloc(sizeof(float) * (size_t)T * F);
    matmul(x, w->w1, h, T, H, F);
    add_bias(h, w->b1, T, F);
    for (size_t i = 0; i < (size_t)T * F; i++) h[i] = gelu1(h[i]);
    matmul(h, w->w2, y, T, F, H);
    free(h);
  }
  add_bias(y, w->b2, T, H);
}

typedef struct {
  int index;
  float score;
} escore;

/* moe.go:43-128 MoELayer.Forward + :167-226 forwardExpertSeparate.
 * sort.Slice is not stable; ties between router probabilities have measure
 * zero with random-init weights — we use a stable insertion sort (descending). */
static void moe_forward(const oracle* o, const layer_w* w, const float* x, float* y, int T) {
  int H = o->cfg.hidden, E = o->cfg.num_experts, K = o->cfg.num_experts_per_tok;
  int I = o->cfg.ffn_dim;
  int in_out = 2 * I, in_in = H, out_out = H, out_in = I;
  float* probs = (float*)malloc(sizeof(float) * (size_t)T * E);
  matmul(x, w->router, probs, T, H, E);
  softmax_rows(probs, T, E);
  memset(y, 0, sizeof(float) * (size_t)T * H);
  escore* es = (escore*)malloc(sizeof(escore) * E);
  float* proj1 = (float*)malloc(sizeof(float) * in_out);
  float* inter = (float*)malloc(sizeof(float) * out_in);
  float* eout = (float*)malloc(sizeof(float) * out_out);
  for (int t = 0; t < T; t++) {
    for (int j = 0; j < E; j++) {
      es[j].index = j;
      es[j].score = probs[(size_t)t * E + j];
    }
    for (int a = 1; a < E; a++) { /* stable, descending */
      escore key = es[a];
      int b = a - 1;
      while (b >= 0 && es[b].score < key.score) {
        es[b + 1] = es[b];
        b--;
      }
      es[b + 1] = key;
    }
    float sum_scores = 0.f;
    for (int s = 0; s < K; s++) sum_scores += es[s].score;
    const float* xin = x + (size_t)t * H;
    for (int s = 0; s < K; s++) {
      int e = es[s].index;
      float weight = es[s].score / sum_scores;
      const float* win = w->moe_in + (size_t)e * in_out * in_in;
      const float* wout = w->moe_out + (size_t)e * out_out * out_in;
      for (int i = 0; i < in_out; i++) {
        float sum = 0.f;
        for (int j = 0; j < in_in; j++) sum += xin[j] * win[(size_t)i * in_in + j];
        proj1[i] = sum;
      }
      for (int i = 0; i < out_in; i++) {
        float gate = proj1[i], up = proj1[i + out_in];
        float sig = 1.0f / (1.0f + (float)exp((double)(-gate))); /* moe.go:208: fp32 divide */
        float ga = gate * sig;
        inter[i] = ga * up;
      }
      for (int i = 0; i < out_out; i++) {
        float sum = 0.f;
        for (int j = 0; j < out_in; j++) sum += inter[j] * wout[(size_t)i * out_in + j];
        eout[i] = sum;
      }
      for (int j = 0; j < H; j++) y[(size_t)t * H + j] += weight * eout[j];
    }
  }
  free(probs);
  free(es);
  free(proj1);
  free(inter);
  free(eout);
}

/* Mamba2Layer.Forward (mamba2.go:72-207) for one sequence: x [T, H] -> out [T, H]; `state` is
 * the layer's SSMState [heads][head_dim][state_size] (updated in place).  Every loop keeps the
 * reference's accumulation order; exp / log / sqrt go through double where the Go code does.
 * Quirk kept on purpose: the causal convolution is zero-padded at the start of EVERY call
 * (causalConv1d :209-270 never reads ConvCache), so a decode step convolves its single token
 * with nothing but the last tap. */
static void mamba_forward(const oracle* o, const layer_w* w, const float* x, float* out, int T,
                          float* state) {
  const nvllm_config* g = &o->cfg;
  const int H = g->hidden, nh = g->mamba_num_heads, hd = g->mamba_head_dim, N = g->mamba_state_size;
  const int G = g->mamba_n_groups, K = g->mamba_conv_kernel;
  const int E = nh * hd, conv = E + 2 * G * N, P = E + conv + nh;
  float* proj = (float*)malloc(sizeof(float) * (size_t)T * P);
  float* xbc = (float*)malloc(sizeof(float) * (size_t)T * conv);
  float* y = (float*)malloc(sizeof(float) * (size_t)T * E);
  /* :86-87 projected = MatMul(xFlat, Transpose(InProj)), InProj [P, H] */
  for (int t = 0; t < T; t++)
    for (int n = 0; n < P; n++) {
      float acc = 0.f;
      const float* wr = w->m_in_proj + (size_t)n * H;
      for (int p = 0; p < H; p++) acc += x[(size_t)t * H + p] * wr[p];
      proj[(size_t)t * P + n] = acc;
    }
  /* :95-98 split [gate | xBC | dt]; :105 causal conv on xBC (:209-270); :108 SiLU */
  for (int t = 0; t < T; t++)
    for (int c = 0; c < conv; c++) {
      float sum = 0.f;
      for (int k = 0; k < K; k++) {
        const int src = t + k - (K - 1); /* padded position t + k; the first K - 1 are zeros */
        const float v = src >= 0 ? proj[(size_t)src * P + E + c] : 0.f;
        sum += v * w->m_conv_w[(size_t)c * K + k];
      }
      if (w->m_conv_b) sum += w->m_conv_b[c];
      xbc[(size_t)t * conv + c] = silu1(sum);
    }
  /* :121-131 delta = softplus(dt + bias); :134 selectiveScan (:273-351) */
  for (int t = 0; t < T; t++) {
    for (int h = 0; h < nh; h++) {
      float dt = proj[(size_t)t * P + E + conv + h];
      if (w->m_dt_bias) dt += w->m_dt_bias[h];
      dt = (float)log(1.0 + exp((double)dt));
      float abar = 1.0f;
      if (w->m_a_log) {
        const float A = -(float)exp((double)w->m_a_log[h]);
        abar = (float)exp((double)(A * dt));
      }
      int grp = h * G / nh;
      if (grp >= G) grp = G - 1;
      const float* u = xbc + (size_t)t * conv + (size_t)h * hd;
      const float* Bt = xbc + (size_t)t * conv + E + (size_t)grp * N;
      const float* Ct = xbc + (size_t)t * conv + E + (size_t)G * N + (size_t)grp * N;
      float* st = state + (size_t)h * hd * N;
      for (int d = 0; d < hd; d++)
        for (int s2 = 0; s2 < N; s2++) {
          const float old = st[(size_t)d * N + s2];
          st[(size_t)d * N + s2] = abar * old + dt * Bt[s2] * u[d];
        }
      for (int d = 0; d < hd; d++) {
        float sum = 0.f;
        for (int s2 = 0; s2 < N; s2++) sum += Ct[s2] * st[(size_t)d * N + s2];
        if (w->m_d) sum += w->m_d[h] * u[d];
        y[(size_t)t * E + (size_t)h * hd + d] = sum;
      }
    }
  }
  /* :139-173 RMSNormGated: y *= silu(gate); rms over E with eps 1e-5; * Norm */
  for (int t = 0; t < T; t++) {
    float* yr = y + (size_t)t * E;
    for (int d = 0; d < E; d++) yr[d] *= silu1(proj[(size_t)t * P + d]);
    float var = 0.f;
    for (int d = 0; d < E; d++) var += yr[d] * yr[d];
    var /= (float)E;
    const float rms = 1.0f / (float)sqrt((double)(var + 1e-5f));
    for (int d = 0; d < E; d++) yr[d] *= rms;
    if (w->m_norm)
      for (int d = 0; d < E; d++) yr[d] *= w->m_norm[d];
  }
  /* :177-179 output = MatMul(yFlat, Transpose(OutProj)), OutProj [H, E] */
  for (int t = 0; t < T; t++)
    for (int n = 0; n < H; n++) {
      float acc = 0.f;
      const float* wr = w->m_out_proj + (size_t)n * E;
      for (int p = 0; p < E; p++) acc += y[(size_t)t * E + p] * wr[p];
      out[(size_t)t * H + n] = acc;
    }
  free(proj);
  free(xbc);
  free(y);
}

/* residual combine used by every sequential branch (generic_model.go:320-326 etc.) */
static void residual_add(const oracle* o, float* x, const float* residual, size_t n) {
  float rm = o->cfg.residual_multiplier;
  if (rm != 0.f) {
    for (size_t j = 0; j < n; j++) x[j] = residual[j] + rm * x[j];
  } else {
    for (size_t j = 0; j < n; j++) x[j] = x[j] + residual[j]; /* Add(x, residual) tensor.go:91-100 */
  }
}

/*
 * Attention for one layer, all three flavours.
 * xin: [T,H] normed input.  out: [T,H].  Appends to cache (c->len is advanced
 * by the caller after the last layer).
 */
static int attention_forward(const oracle* o, const layer_w* w, int layer, const float* xin,
                             float* out, int T, okv* c, int pos_offset) {
  const nvllm_config* g = &o->cfg;
  int H = g->hidden, nh = g->num_heads, hd = g->head_dim;
  int nkv = (g->attention_type == NVLLM_ATTN_MQA) ? 1
            : (g->attention_type == NVLLM_ATTN_MHA ? nh : g->num_kv_heads);
  int qdim = nh * hd, kvdim = nkv * hd;
  float* Q = (float*)malloc(sizeof(float) * (size_t)T * qdim);
  float* Kf = (float*)malloc(sizeof(float) * (size_t)T * kvdim);
  float* Vf = (float*)malloc(sizeof(float) * (size_t)T * kvdim);
  float* Qh = (float*)malloc(sizeof(float) * (size_t)T * qdim);
  float* Kh = (float*)malloc(sizeof(float) * (size_t)T * kvdim);
  float* Vh = (float*)malloc(sizeof(float) * (size_t)T * kvdim);
  int rc = 0;

  if (g->attention_type == NVLLM_ATTN_MQA) {
    /* mqa.go:101-140 projectQ / projectKV (fused [K|V]) */
    matmul(xin, w->q, Q, T, H, qdim);
    add_bias(Q, w->q_b, T, qdim);
    float* KV = (float*)malloc(sizeof(float) * (size_t)T * 2 * hd);
    matmul(xin, w->kv, KV, T, H, 2 * hd);
    for (int i = 0; i < T; i++)
      for (int j = 0; j < hd; j++) {
        Kf[(size_t)i * hd + j] = KV[(size_t)i * 2 * hd + j];
        Vf[(size_t)i * hd + j] = KV[(size_t)i * 2 * hd + hd + j];
      }
    free(KV);
  } else {
    matmul(xin, w->q, Q, T, H, qdim);
    matmul(xin, w->k, Kf, T, H, kvdim);
    matmul(xin, w->v, Vf, T, H, kvdim);
    if (g->attention_type == NVLLM_ATTN_MHA) { /* attention.go:66-86 project with bias */
      add_bias(Q, w->q_b, T, qdim);
      add_bias(Kf, w->k_b, T, kvdim);
      add_bias(Vf, w->v_b, T, kvdim);
    }
  }
  split_heads(Q, Qh, T, nh, hd);
  split_heads(Kf, Kh, T, nkv, hd);
  split_heads(Vf, Vh, T, nkv, hd);

  /* RoPE: always for MQA (mqa.go:65-67, base hard-wired 10000 mqa.go:35);
   * for GQA only when PositionRoPE (generic_model.go:109-111, attention.go:234-237);
   * never for MHA. */
  int use_rope = (g->attention_type == NVLLM_ATTN_MQA) ||
                 (g->attention_type == NVLLM_ATTN_GQA && g->position_type == NVLLM_POS_ROPE);
  if (use_rope) {
    if (apply_rope(o, Qh, nh, T, pos_offset) || apply_rope(o, Kh, nkv, T, pos_offset)) {
      rc = -1;
      goto done;
    }
  }
  if (c->len + T > c->cap) {
    rc = -1;
    goto done;
  }
  kv_append(c, layer, Kh, Vh, T);
  {
    int kvs = c->len + T;
    float scale;
    int style;
    if (g->attention_type == NVLLM_ATTN_GQA) {
      style = 1;
      scale = g->attention_multiplier; /* attention.go:361-364 */
      if (scale == 0.f) scale = 1.0f / (float)sqrt((double)hd);
    } else {
      style = 0;
      scale = (float)(1.0 / sqrt((double)hd)); /* attention.go:136, mqa.go:185 */
    }
    float* scores = (float*)malloc(sizeof(float) * (size_t)T * kvs);
    float* Oh = (float*)malloc(sizeof(float) * (size_t)T * qdim);
    int rep = nh / nkv;
    for (int h = 0; h < nh; h++) {
      int kvh = h / rep; /* attention.go:337-339; MQA: 0; MHA: h */
      const float* Kc = c->k + (((size_t)layer * c->n_kv + kvh) * c->cap) * hd;
      const float* Vc = c->v + (((size_t)layer * c->n_kv + kvh) * c->cap) * hd;
      sdpa_head(Qh + (size_t)h * T * hd, Kc, Vc, Oh + (size_t)h * T * hd, scores, T, kvs, hd,
                scale, style);
    }
    float* merged = (float*)malloc(sizeof(float) * (size_t)T * qdim);
    merge_heads(Oh, merged, T, nh, hd);
    matmul(merged, w->o, out, T, qdim, H);
    if (g->attention_type != NVLLM_ATTN_GQA) add_bias(out, w->o_b, T, H);
    free(scores);
    free(Oh);
    free(merged);
  }
done:
  free(Q);
  free(Kf);
  free(Vf);
  free(Qh);
  free(Kh);
  free(Vh);
  return rc;
}

/* ------------------------------------------------------------ public API */

oracle* oracle_create(const nvllm_config* cfg) {
  if (cfg->num_layers > MAX_LAYERS || cfg->head_dim > 1024) return NULL;
  oracle* o = (oracle*)calloc(1, sizeof(oracle));
  o->cfg = *cfg;
  if (cfg->attention_type == NVLLM_ATTN_MQA)
    build_rope(o, 10000.0); /* mqa.go:35 / generic_model.go:92 */
  else if (cfg->attention_type == NVLLM_ATTN_GQA && cfg->position_type == NVLLM_POS_ROPE)
    build_rope(o, cfg->rope_base);
  return o;
}

void oracle_destroy(oracle* o) {
  if (!o) return;
  free(o->rope_cos);
  free(o->rope_sin);
  free(o->lm_head_owned);
  free(o);
}

/* Keeps the pointer (the python side keeps the numpy arrays alive). */
int oracle_set_weight(oracle* o, int id, int layer, const float* p) {
  if (id < 16) {
    switch (id) {
      case NVLLM_W_TOKEN_EMBEDDING: o->wte = p; break;
      case NVLLM_W_POS_EMBEDDING: o->wpe = p; break;
      case NVLLM_W_FINAL_NORM_W: o->lnf_w = p; break;
      case NVLLM_W_FINAL_NORM_B: o->lnf_b = p; break;
      case NVLLM_W_LM_HEAD: o->lm_head = p; break;
      default: return -1;
    }
    return 0;
  }
  if (layer < 0 || layer >= o->cfg.num_layers) return -1;
  layer_w* w = &o->L[layer];
  switch (id) {
    case NVLLM_W_ATTN_NORM_W: w->attn_norm_w = p; break;
    case NVLLM_W_ATTN_NORM_B: w->attn_norm_b = p; break;
    case NVLLM_W_FFN_NORM_W: w->ffn_norm_w = p; break;
    case NVLLM_W_FFN_NORM_B: w->ffn_norm_b = p; break;
    case NVLLM_W_Q: w->q = p; break;
    case NVLLM_W_K: w->k = p; break;
    case NVLLM_W_V: w->v = p; break;
    case NVLLM_W_KV: w->kv = p; break;
    case NVLLM_W_O: w->o = p; break;
    case NVLLM_W_Q_BIAS: w->q_b = p; break;
    case NVLLM_W_K_BIAS: w->k_b = p; break;
    case NVLLM_W_V_BIAS: w->v_b = p; break;
    case NVLLM_W_O_BIAS: w->o_b = p; break;
    case NVLLM_W_FFN_W1: w->w1 = p; break;
    case NVLLM_W_FFN_B1: w->b1 = p; break;
    case NVLLM_W_FFN_W2: w->w2 = p; break;
    case NVLLM_W_FFN_B2: w->b2 = p; break;
    case NVLLM_W_MOE_ROUTER: w->router = p; break;
    case NVLLM_W_MOE_INPUT: w->moe_in = p; break;
    case NVLLM_W_MOE_OUTPUT: w->moe_out = p; break;
    case NVLLM_W_MAMBA_IN_PROJ: w->m_in_proj = p; break;
    case NVLLM_W_MAMBA_CONV_W: w->m_conv_w = p; break;
    case NVLLM_W_MAMBA_CONV_B: w->m_conv_b = p; break;
    case NVLLM_W_MAMBA_DT_BIAS: w->m_dt_bias = p; break;
    case NVLLM_W_MAMBA_A_LOG: w->m_a_log = p; break;
    case NVLLM_W_MAMBA_D: w->m_d = p; break;
    case NVLLM_W_MAMBA_NORM: w->m_norm = p; break;
    case NVLLM_W_MAMBA_OUT_PROJ: w->m_out_proj = p; break;
    default: return -1;
  }
  return 0;
}

okv* oracle_kv_create(const oracle* o) {
  const nvllm_config* g = &o->cfg;
  okv* c = (okv*)calloc(1, sizeof(okv));
  c->n_layers = g->num_layers;
  c->n_kv = (g->attention_type == NVLLM_ATTN_MQA) ? 1
            : (g->attention_type == NVLLM_ATTN_MHA ? g->num_heads : g->num_kv_heads);
  c->hd = g->head_dim;
  c->cap = g->max_seq_len;
  size_t n = (size_t)c->n_layers * c->n_kv * c->cap * c->hd;
  c->k = (float*)malloc(sizeof(float) * n);
  c->v = (float*)malloc(sizeof(float) * n);
  c->ssm_layer = (size_t)g->mamba_num_heads * g->mamba_head_dim * g->mamba_state_size;
  if (c->ssm_layer) c->ssm = (float*)calloc((size_t)c->n_layers * c->ssm_layer, sizeof(float));
  return c;
}
void oracle_kv_destroy(okv* c) {
  if (!c) return;
  free(c->k);
  free(c->v);
  free(c->ssm);
  free(c);
}
void oracle_kv_reset(okv* c) {
  c->len = 0;
  if (c->ssm) memset(c->ssm, 0, sizeof(float) * (size_t)c->n_layers * c->ssm_layer);
}
int oracle_set_layer_types(oracle* o, const int32_t* types, int n) {
  if (n != o->cfg.num_layers) return -1;
  for (int i = 0; i < n; i++) o->is_mamba[i] = types[i] != 0;
  return 0;
}
int oracle_kv_len(const okv* c) { return c->len; }
/* copy out K or V of (layer, kv head): [len, hd] — lets tests compare the device KV pool */
void oracle_kv_read(const okv* c, int layer, int head, int which, float* out) {
  const float* src = (which ? c->v : c->k) + (((size_t)layer * c->n_kv + head) * c->cap) * c->hd;
  memcpy(out, src, sizeof(float) * (size_t)c->len * c->hd);
}

/*
 * generic_model.go:276-480 ForwardWithCache (batch = 1).
 * tokens[T] are processed at positions pos_offset .. pos_offset+T-1; the
 * reference always has pos_offset == cache length.
 * logits_out: if all_rows, [T,V]; else the last row only [V]
 * (GetLogitsForLastToken generic_model.go:595-604 — the reference computes all
 * rows (:467-470) and discards them; row T-1 is bit-identical either way
 * because MatMul rows are independent).
 * hidden_out (optional): final-normed hidden of the last token [H].
 */
int oracle_forward(oracle* o, const int32_t* tokens, int T, okv* c, int pos_offset,
                   float* logits_out, int all_rows, float* hidden_out) {
  const nvllm_config* g = &o->cfg;
  int H = g->hidden, V = g->vocab_size;
  if (T <= 0) return -1;
  size_t n = (size_t)T * H;
  float* x = (float*)malloc(sizeof(float) * n);
  float* xn = (float*)malloc(sizeof(float) * n);
  float* a = (float*)malloc(sizeof(float) * n);
  float* f = (float*)malloc(sizeof(float) * n);
  int rc = 0;

  /* generic_model.go:567-592 embedWithOffset */
  for (int i = 0; i < T; i++) {
    int tok = tokens[i];
    if (tok < 0 || tok >= V) {
      rc = -1;
      goto done;
    }
    memcpy(x + (size_t)i * H, o->wte + (size_t)tok * H, sizeof(float) * H);
    if (g->position_type == NVLLM_POS_LEARNED && o->wpe) {
      int ap = pos_offset + i;
      if (ap < g->max_seq_len)
        for (int j = 0; j < H; j++) x[(size_t)i * H + j] += o->wpe[(size_t)ap * H + j];
    }
  }
  if (g->embedding_multiplier != 0.f) /* :298-302 */
    for (size_t i = 0; i < n; i++) x[i] *= g->embedding_multiplier;

  /* generic_model.go:285-292: a brand-new prompt resets every Mamba2 state */
  if (pos_offset == 0 && T > 1 && c->ssm)
    memset(c->ssm, 0, sizeof(float) * (size_t)c->n_layers * c->ssm_layer);

  for (int l = 0; l < g->num_layers; l++) {
    const layer_w* w = &o->L[l];
    if (o->is_mamba[l]) {
      /* generic_model.go:457-459 -> GenericBlock.Forward -> forwardMamba2 (:162-202):
       * norm + Mamba2 + residual, norm + FFN (shared MLP, never MoE) + residual */
      layer_norm(x, w->attn_norm_w, w->attn_norm_b, g->norm_eps, xn, T, H);
      mamba_forward(o, w, xn, a, T, c->ssm + (size_t)l * c->ssm_layer);
      residual_add(o, a, x, n);
      memcpy(x, a, sizeof(float) * n);
      layer_norm(x, w->ffn_norm_w, w->ffn_norm_b, g->norm_eps, xn, T, H);
      ffn_forward(o, w, xn, f, T);
      residual_add(o, f, x, n);
      memcpy(x, f, sizeof(float) * n);
    } else if (g->attention_type == NVLLM_ATTN_MQA && g->block_style == NVLLM_BLOCK_PARALLEL) {
      /* :395-418 parallel block */
      layer_norm(x, w->attn_norm_w, w->attn_norm_b, g->norm_eps, xn, T, H);
      if (attention_forward(o, w, l, xn, a, T, c, pos_offset)) {
        rc = -1;
        goto done;
      }
      ffn_forward(o, w, xn, f, T);
      if (g->attention_multiplier != 0.f && g->residual_multiplier != 0.f) {
        for (size_t j = 0; j < n; j++)
          x[j] = x[j] + g->attention_multiplier * a[j] + g->residual_multiplier * f[j];
      } else {
        for (size_t j = 0; j < n; j++) x[j] = x[j] + a[j] + f[j];
      }
    } else {
      /* sequential: :308-346 (MHA), :347-389 (GQA), :420-455 (MQA sequential) */
      layer_norm(x, w->attn_norm_w, w->attn_norm_b, g->norm_eps, xn, T, H);
      if (attention_forward(o, w, l, xn, a, T, c, pos_offset)) {
        rc = -1;
        goto done;
      }
      residual_add(o, a, x, n);
      memcpy(x, a, sizeof(float) * n);
      layer_norm(x, w->ffn_norm_w, w->ffn_norm_b, g->norm_eps, xn, T, H);
      if (g->use_moe)
        moe_forward(o, w, xn, f, T);
      else
        ffn_forward(o, w, xn, f, T);
      residual_add(o, f, x, n);
      memcpy(x, f, sizeof(float) * n);
    }
  }
  c->len += T;

  /* :464 final norm */
  layer_norm(x, o->lnf_w, o->lnf_b, g->norm_eps, xn, T, H);
  if (hidden_out) memcpy(hidden_out, xn + (size_t)(T - 1) * H, sizeof(float) * H);

  /* :467-470 LM head [H,V] (pre-transposed at load; tied => Transpose(wte), generic_loader.go:255-259) */
  if (!o->lm_head) {
    if (!o->lm_head_owned) {
      o->lm_head_owned = (float*)malloc(sizeof(float) * (size_t)H * V);
      for (int i = 0; i < V; i++)
        for (int j = 0; j < H; j++) o->lm_head_owned[(size_t)j * V + i] = o->wte[(size_t)i * H + j];
    }
  }
  {
    const float* lm = o->lm_head ? o->lm_head : o->lm_head_owned;
    int rows = all_rows ? T : 1;
    const float* src = all_rows ? xn : xn + (size_t)(T - 1) * H;
    matmul(src, lm, logits_out, rows, H, V);
    if (g->logits_scaling != 0.f) /* :473-477 */
      for (size_t i = 0; i < (size_t)rows * V; i++) logits_out[i] /= g->logits_scaling;
  }
done:
  free(x);
  free(xn);
  free(a);
  free(f);
  return rc;
}

/* cmd/ask/main.go:389-402 argmax — first index of the maximum (strict >). */
int oracle_argmax(const float* data, int n) {
  if (n == 0) return 0;
  int mi = 0;
  float mv = data[0];
  for (int i = 0; i < n; i++)
    if (data[i] > mv) {
      mv = data[i];
      mi = i;
    }
  return mi;
}

typedef struct {
  int idx;
  float p;
} iprob;
static int cmp_desc(const void* a, const void* b) {
  const iprob* x = (const iprob*)a;
  const iprob* y = (const iprob*)b;
  if (x->p > y->p) return -1;
  if (x->p < y->p) return 1;
  return (x->idx > y->idx) - (x->idx < y->idx); /* make qsort deterministic on ties */
}

/*
 * sampling.go:33-102 SampleWithHistory, with rand.Float32() (:205) replaced by
 * the caller-supplied uniform u in [0,1) so the draw is reproducible.
 * rep_penalty == 1 disables (:43); top_k == 0 disables (:82); top_p >= 1 disables (:87).
 */
int oracle_sample(const float* logits, int V, const int32_t* history, int n_hist, float temperature,
                  float top_p, int top_k, float rep_penalty, float u) {
  float* lc = (float*)malloc(sizeof(float) * V);
  memcpy(lc, logits, sizeof(float) * V);
  if (rep_penalty != 1.0f && n_hist > 0) { /* :43-67 */
    int* counts = (int*)calloc(V, sizeof(int));
    for (int i = 0; i < n_hist; i++) {
      int w = (i >= n_hist - 10) ? 3 : 1;
      if (history[i] >= 0 && history[i] < V) counts[history[i]] += w;
    }
    for (int t = 0; t < V; t++)
      if (counts[t]) {
        float pen = rep_penalty * (float)counts[t];
        if (lc[t] > 0)
          lc[t] /= pen;
        else
          lc[t] *= pen;
      }
    free(counts);
  }
  if (temperature > 0 && temperature != 1.0f) /* :70-74 */
    for (int i = 0; i < V; i++) lc[i] /= temperature;
  /* :105-128 softmax */
  float mx = lc[0];
  for (int i = 1; i < V; i++)
    if (lc[i] > mx) mx = lc[i];
  float sum = 0.f;
  for (int i = 0; i < V; i++) {
    lc[i] = (float)exp((double)(lc[i] - mx));
    sum += lc[i];
  }
  for (int i = 0; i < V; i++) lc[i] /= sum;
  if ((top_k > 0 && top_k < V) || top_p < 1.0f) {
    iprob* ip = (iprob*)malloc(sizeof(iprob) * V);
    if (top_k > 0 && top_k < V) { /* :131-155 */
      for (int i = 0; i < V; i++) {
        ip[i].idx = i;
        ip[i].p = lc[i];
      }
      qsort(ip, V, sizeof(iprob), cmp_desc);
      memset(lc, 0, sizeof(float) * V);
      for (int i = 0; i < top_k; i++) lc[ip[i].idx] = ip[i].p;
    }
    if (top_p < 1.0f) { /* :158-193 */
      for (int i = 0; i < V; i++) {
        ip[i].idx = i;
        ip[i].p = lc[i];
      }
      qsort(ip, V, sizeof(iprob), cmp_desc);
      float cum = 0.f;
      int cutoff = V;
      for (int i = 0; i < V; i++) {
        cum += ip[i].p;
        if (cum >= top_p) {
          cutoff = i + 1;
          break;
        }
      }
      memset(lc, 0, sizeof(float) * V);
      for (int i = 0; i < cutoff; i++) lc[ip[i].idx] = ip[i].p;
    }
    free(ip);
  }
  sum = 0.f; /* :90-98 renormalise */
  for (int i = 0; i < V; i++) sum += lc[i];
  if (sum > 0)
    for (int i = 0; i < V; i++) lc[i] /= sum;
  /* :196-217 sampleMultinomial: fp32 running sum, first i with cum[i] >= r */
  float cum = 0.f, total = 0.f;
  for (int i = 0; i < V; i++) total += lc[i];
  float r = u * total;
  int idx = V - 1;
  for (int i = 0; i < V; i++) {
    cum += lc[i];
    if (cum >= r) {
      idx = i;
      break;
    }
  }
  free(lc);
  return idx;
}

/* generic_loader.go:705-748 splitFalconQKV + :750-765 combineMQAKV.
 * qkv: [hidden, (nh+2)*hd] (already transposed to [in,out]).  Pinned by the
 * KAT in purego/tensor/falcon_split_test.go:7-158. */
void oracle_split_falcon_qkv(const float* qkv, int hidden, int nh, int hd, float* Q, float* KV) {
  for (int row = 0; row < hidden; row++) {
    const float* src = qkv + (size_t)row * (nh + 2) * hd;
    memcpy(Q + (size_t)row * nh * hd, src, sizeof(float) * (size_t)nh * hd);
    memcpy(KV + (size_t)row * 2 * hd, src + (size_t)nh * hd, sizeof(float) * hd);
    memcpy(KV + (size_t)row * 2 * hd + hd, src + (size_t)(nh + 1) * hd, sizeof(float) * hd);
  }
}

/* generic_loader.go:674-702 splitGPT2QKV: c_attn [H, 3H] -> Q,K,V [H,H] by columns. */
void oracle_split_gpt2_qkv(const float* qkv, int hidden, float* Q, float* K, float* V) {
  for (int row = 0; row < hidden; row++) {
    const float* src = qkv + (size_t)row * 3 * hidden;
    memcpy(Q + (size_t)row * hidden, src, sizeof(float) * hidden);
    memcpy(K + (size_t)row * hidden, src + hidden, sizeof(float) * hidden);
    memcpy(V + (size_t)row * hidden, src + 2 * hidden, sizeof(float) * hidden);
  }
}

/* stand-alone ops for unit parity of single kernels */
/* one Mamba2 layer on its own (tests: checked against an independent numpy restatement) */
void oracle_mamba_layer(oracle* o, int layer, const float* x, float* out, int T, float* state) {
  mamba_forward(o, &o->L[layer], x, out, T, state);
}
void oracle_matmul(const float* a, const float* b, float* c, int m, int k, int n) {
  matmul(a, b, c, m, k, n);
}
void oracle_layer_norm(const float* x, const float* w, const float* b, float eps, float* y, int rows,
                       int hs) {
  layer_norm(x, w, b, eps, y, rows, hs);
}
void oracle_gelu(float* x, size_t n) {
  for (size_t i = 0; i < n; i++) x[i] = gelu1(x[i]);
}
void oracle_silu(float* x, size_t n) {
  for (size_t i = 0; i < n; i++) x[i] = silu1(x[i]);
}
