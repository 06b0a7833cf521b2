// Native self-test / micro-benchmark of libwm_b200 through its C ABI (no torch, fast start-up).
// Each invocation runs ONE case so that a trap or time-out in a kernel cannot take the others down:
//     selftest <case> [args...]
// CPU references here are plain loops over the same bf16-rounded inputs (fp32/double accumulate); they are
// test infrastructure only.  tests/native/run_selftest.sh drives the whole matrix on the GPU box.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/wm_b200.h"

#define CK(x)                                                                   \
  do {                                                                          \
    cudaError_t e_ = (x);                                                       \
    if (e_ != cudaSuccess) {                                                    \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(2);                                                                  \
    }                                                                           \
  } while (0)
#define WM(x)                                                      \
  do {                                                             \
    wm_status s_ = (x);                                            \
    if (s_ != WM_OK) {                                             \
      printf("wm error %d: %s (%s)\n", s_, wm_last_error(), #x);   \
      exit(3);                                                     \
    }                                                              \
  } while (0)

static uint32_t rng_state = 12345u;
static float frand() {  // uniform (-1, 1)
  rng_state = rng_state * 1664525u + 1013904223u;
  return ((rng_state >> 8) & 0xFFFFFF) / 8388608.0f - 1.0f;
}
static uint16_t f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  u += 0x7FFFu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static float bf2f(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
struct BF {  // host bf16 matrix with float mirror of the rounded values
  std::vector<uint16_t> h;
  std::vector<float> f;
  void fill(size_t n, float scale) {
    h.resize(n), f.resize(n);
    for (size_t i = 0; i < n; ++i) {
      h[i] = f2bf(frand() * scale);
      f[i] = bf2f(h[i]);
    }
  }
};
template <typename T>
static T* dev(const std::vector<T>& v) {
  T* p;
  CK(cudaMalloc(&p, v.size() * sizeof(T) + 16));
  CK(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  return p;
}
template <typename T>
static T* dev_alloc(size_t n, int fill_byte = 0xCD) {
  T* p;
  CK(cudaMalloc(&p, n * sizeof(T) + 16));
  CK(cudaMemset(p, fill_byte, n * sizeof(T)));
  return p;
}
template <typename T>
static std::vector<T> host(const T* p, size_t n) {
  std::vector<T> v(n);
  CK(cudaMemcpy(v.data(), p, n * sizeof(T), cudaMemcpyDeviceToHost));
  return v;
}
static std::vector<float> randf(size_t n, float scale, float mean = 0.f) {
  std::vector<float> v(n);
  for (auto& x : v) x = mean + frand() * scale;
  return v;
}

struct Err {
  double max_abs = 0, max_ref = 0;
  long long bad = 0, n = 0;
  void add(double got, double ref, double atol, double rtol) {
    const double d = fabs(got - ref);
    if (!(d <= atol + rtol * fabs(ref))) ++bad;  // also catches NaN
    if (d > max_abs || d != d) max_abs = d != d ? 1e30 : d;
    if (fabs(ref) > max_ref) max_ref = fabs(ref);
    ++n;
  }
  int report(const char* name) const {
    printf("%-44s n=%-9lld max_abs_err=%-11.4g max_ref=%-9.4g bad=%-8lld %s\n", name, n, max_abs, max_ref, bad,
           bad == 0 ? "PASS" : "FAIL");
    fflush(stdout);
    return bad == 0 ? 0 : 1;
  }
};

static double gelu_ref(double x) { return 0.5 * x * (1.0 + erf(x / sqrt(2.0))); }

// ------------------------------------------------------------------------------------------------ probe
static int case_probe(int mode, unsigned lbo, unsigned sbo) {
  BF P, V;
  P.fill(128 * 128, 1.0f);
  V.fill(128 * 64, 1.0f);
  uint16_t *dp = dev(P.h), *dv = dev(V.h);
  float* dd = dev_alloc<float>(128 * 64);
  WM(wm_umma_probe(0, dp, dv, dd, mode, lbo, sbo));
  CK(cudaDeviceSynchronize());
  auto d = host(dd, 128 * 64);
  Err e;
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < 64; ++n) {
      double acc = 0;
      for (int k = 0; k < 128; ++k) acc += (double)P.f[m * 128 + k] * V.f[k * 64 + n];
      e.add(d[m * 64 + n], acc, 1e-3, 1e-3);
    }
  char name[96];
  snprintf(name, sizeof name, "probe mode=%d (P in %s) v_lbo=%u v_sbo=%u", mode, mode ? "TMEM" : "smem", lbo, sbo);
  return e.report(name);
}

// ------------------------------------------------------------------------------------------------ layernorm
static int case_layernorm(int rows) {
  auto x = randf((size_t)rows * 1024, 2.0f, 0.3f), w = randf(1024, 0.2f, 1.0f), b = randf(1024, 0.1f);
  float *dx = dev(x), *dw = dev(w), *db = dev(b);
  uint16_t* dy = dev_alloc<uint16_t>((size_t)rows * 1024);
  float* dyf = dev_alloc<float>((size_t)rows * 1024);
  WM(wm_layernorm(0, dx, dw, db, 1e-5f, rows, dy, 0));
  WM(wm_layernorm(0, dx, dw, db, 1e-5f, rows, dyf, 1));
  CK(cudaDeviceSynchronize());
  auto y = host(dy, (size_t)rows * 1024);
  auto yf = host(dyf, (size_t)rows * 1024);
  Err e, ef;
  for (int r = 0; r < rows; ++r) {
    double mu = 0, var = 0;
    for (int c = 0; c < 1024; ++c) mu += x[(size_t)r * 1024 + c];
    mu /= 1024;
    for (int c = 0; c < 1024; ++c) var += (x[(size_t)r * 1024 + c] - mu) * (x[(size_t)r * 1024 + c] - mu);
    var /= 1024;
    const double rstd = 1.0 / sqrt(var + 1e-5);
    for (int c = 0; c < 1024; ++c) {
      const double ref = (x[(size_t)r * 1024 + c] - mu) * rstd * w[c] + b[c];
      e.add(bf2f(y[(size_t)r * 1024 + c]), ref, 1e-3, 8e-3);
      ef.add(yf[(size_t)r * 1024 + c], ref, 2e-5, 1e-5);
    }
  }
  return e.report("layernorm bf16 out") | ef.report("layernorm fp32 out");
}

// ------------------------------------------------------------------------------------------------ GEMMs
// ref[m][n] = sum_k A[m][k] W[n][k] for the sampled rows
static void gemm_ref_row(const BF& A, int lda, const BF& W, int ldw, int m, int N, int K, std::vector<double>& out) {
  out.assign(N, 0.0);
  for (int n = 0; n < N; ++n) {
    double acc = 0;
    const float* a = &A.f[(size_t)m * lda];
    const float* w = &W.f[(size_t)n * ldw];
    for (int k = 0; k < K; ++k) acc += (double)a[k] * w[k];
    out[n] = acc;
  }
}

static int case_gemm_bias(int M, int N, int K, int gelu, int row_step) {
  BF A, W;
  A.fill((size_t)M * K, 1.0f);
  W.fill((size_t)N * K, 1.0f / sqrtf((float)K) * 1.7f);
  auto bias = randf(N, 0.5f);
  uint16_t *da = dev(A.h), *dw = dev(W.h);
  float* db = dev(bias);
  uint16_t* dout = dev_alloc<uint16_t>((size_t)M * N);
  WM(wm_gemm_bias_act(0, da, K, dw, K, db, M, N, K, dout, N, gelu));
  CK(cudaDeviceSynchronize());
  auto out = host(dout, (size_t)M * N);
  Err e;
  std::vector<double> ref;
  for (int m = 0; m < M; m += row_step) {
    gemm_ref_row(A, K, W, K, m, N, K, ref);
    for (int n = 0; n < N; ++n) {
      double r = ref[n] + bias[n];
      if (gelu) r = gelu_ref(r);
      e.add(bf2f(out[(size_t)m * N + n]), r, 4e-3, 8e-3);
    }
  }
  char name[96];
  snprintf(name, sizeof name, "gemm_bias_act M=%d N=%d K=%d gelu=%d", M, N, K, gelu);
  return e.report(name);
}

static int case_gemm_resid(int M, int N, int K, int with_gamma, int with_out2, int row_step, int split = 0) {
  BF A, W;
  A.fill((size_t)M * K, 1.0f);
  W.fill((size_t)N * K, 1.0f / sqrtf((float)K) * 1.7f);
  auto bias = randf(N, 0.5f), gamma = randf(N, 0.3f, 0.5f), x0 = randf((size_t)M * N, 1.0f);
  uint16_t *da = dev(A.h), *dw = dev(W.h);
  float *db = dev(bias), *dg = dev(gamma), *dx = dev(x0);
  const int ld2 = 2 * N, col2 = N;
  float* d2 = dev_alloc<float>((size_t)M * ld2, 0);
  if (split) {  // tail tiles cut along K (gemm.cu): run twice on the zeroed workspace, the flags must come back to zero
    const size_t wsb = wm_gemm_workspace_bytes();
    uint8_t* ws = dev_alloc<uint8_t>(wsb, 0);
    float* dx_b = dev(x0);
    WM(wm_gemm_ls_residual_ws(0, da, K, dw, K, db, with_gamma ? dg : nullptr, M, N, K, dx, N, ws, wsb));
    WM(wm_gemm_ls_residual_ws(0, da, K, dw, K, db, with_gamma ? dg : nullptr, M, N, K, dx_b, N, ws, wsb));
    CK(cudaDeviceSynchronize());
    auto xa = host(dx, (size_t)M * N), xb = host(dx_b, (size_t)M * N);
    if (memcmp(xa.data(), xb.data(), xa.size() * 4) != 0) { printf("FAIL split GEMM: two runs differ bitwise\n"); return 1; }
  } else {
    WM(wm_gemm_ls_residual(0, da, K, dw, K, db, with_gamma ? dg : nullptr, M, N, K, dx, N));
  }
  if (with_out2) {
    if (N != 1024) { printf("out2 copy needs N=1024\n"); return 1; }
    WM(wm_copy_cols(0, dx, M, d2, ld2, col2));
  }
  CK(cudaDeviceSynchronize());
  auto x = host(dx, (size_t)M * N);
  auto o2 = host(d2, (size_t)M * ld2);
  Err e, e2;
  std::vector<double> ref;
  for (int m = 0; m < M; m += row_step) {
    gemm_ref_row(A, K, W, K, m, N, K, ref);
    for (int n = 0; n < N; ++n) {
      const double r = with_gamma ? x0[(size_t)m * N + n] + gamma[n] * (ref[n] + bias[n]) : ref[n] + bias[n];
      e.add(x[(size_t)m * N + n], r, 2e-4, 1e-4);
      if (with_out2) {
        e2.add(o2[(size_t)m * ld2 + col2 + n], r, 2e-4, 1e-4);
        e2.add(o2[(size_t)m * ld2 + n], 0.0, 0, 0);  // the other half must stay untouched
      }
    }
  }
  char name[96];
  snprintf(name, sizeof name, "gemm_ls_residual M=%d N=%d K=%d gamma=%d split=%d", M, N, K, with_gamma, split);
  int rc = e.report(name);
  if (with_out2) rc |= e2.report("  second (strided) output");
  return rc;
}

static int case_gemm_table(int M, int N, int K, int ld, int period) {
  BF A, W;
  A.fill((size_t)M * ld, 1.0f);
  W.fill((size_t)N * ld, 0.05f);
  auto table = randf((size_t)period * N, 1.0f);
  uint16_t *da = dev(A.h), *dw = dev(W.h);
  float* dt = dev(table);
  float* dout = dev_alloc<float>((size_t)M * N);
  WM(wm_gemm_table(0, da, ld, dw, ld, dt, period, M, N, K, dout, N));
  CK(cudaDeviceSynchronize());
  auto out = host(dout, (size_t)M * N);
  Err e;
  std::vector<double> ref;
  for (int m = 0; m < M; m += 3) {
    gemm_ref_row(A, ld, W, ld, m, N, K, ref);
    for (int n = 0; n < N; ++n) e.add(out[(size_t)m * N + n], ref[n] + table[(size_t)(m % period) * N + n], 2e-4, 1e-4);
  }
  char name[96];
  snprintf(name, sizeof name, "gemm_table M=%d N=%d K=%d ld=%d period=%d", M, N, K, ld, period);
  return e.report(name);
}

static int case_gemm_qkv(int M, int with_norm, int with_rope, int row_step) {
  const int N = 3072, K = 1024, P = 37, ps = 7, gw = 6;  // tokens per view 37 = 7 + 5x6 grid
  BF A, W;
  A.fill((size_t)M * K, 1.0f);
  W.fill((size_t)N * K, 1.0f / 32.0f * 1.7f);
  auto bias = randf(N, 0.3f), qw = randf(64, 0.2f, 1.0f), qb = randf(64, 0.1f), kw = randf(64, 0.2f, 1.0f), kb = randf(64, 0.1f);
  const int max_pos = 16;
  std::vector<float> cs((max_pos + 1) * 16), sn((max_pos + 1) * 16);
  wm_rope_tables(max_pos, cs.data(), sn.data());
  uint16_t *da = dev(A.h), *dw = dev(W.h);
  float *db = dev(bias), *dqw = dev(qw), *dqb = dev(qb), *dkw = dev(kw), *dkb = dev(kb), *dcs = dev(cs), *dsn = dev(sn);
  uint16_t* dq = dev_alloc<uint16_t>((size_t)M * 1024);
  uint16_t* dkv = dev_alloc<uint16_t>((size_t)M * 2048);
  WM(wm_gemm_qkv(0, da, dw, db, with_norm ? dqw : nullptr, with_norm ? dqb : nullptr, with_norm ? dkw : nullptr,
                 with_norm ? dkb : nullptr, 1e-5f, with_rope ? dcs : nullptr, with_rope ? dsn : nullptr, P, ps, gw, 1.0f, M, dq, dkv));
  CK(cudaDeviceSynchronize());
  auto q = host(dq, (size_t)M * 1024);
  auto kv = host(dkv, (size_t)M * 2048);
  Err e;
  std::vector<double> ref;
  for (int m = 0; m < M; m += row_step) {
    gemm_ref_row(A, K, W, K, m, N, K, ref);
    const int p = m % P;
    int py = 0, px = 0;
    if (p >= ps) py = (p - ps) / gw + 1, px = (p - ps) % gw + 1;
    for (int hd = 0; hd < 48; ++hd) {
      double x[64];
      for (int i = 0; i < 64; ++i) x[i] = ref[hd * 64 + i] + bias[hd * 64 + i];
      const int which = hd / 16;
      if (which < 2 && with_norm) {
        double mu = 0, var = 0;
        for (int i = 0; i < 64; ++i) mu += x[i];
        mu /= 64;
        for (int i = 0; i < 64; ++i) var += (x[i] - mu) * (x[i] - mu);
        var /= 64;
        const double rstd = 1.0 / sqrt(var + 1e-5);
        for (int i = 0; i < 64; ++i) x[i] = (x[i] - mu) * rstd * (which ? kw[i] : qw[i]) + (which ? kb[i] : qb[i]);
      }
      if (which < 2 && with_rope) {
        for (int half = 0; half < 2; ++half) {
          const int pos = half ? px : py;
          for (int i = 0; i < 16; ++i) {
            const double c = cs[pos * 16 + i], s = sn[pos * 16 + i];
            const double a = x[half * 32 + i], b = x[half * 32 + 16 + i];
            x[half * 32 + i] = a * c - b * s;
            x[half * 32 + 16 + i] = b * c + a * s;
          }
        }
      }
      for (int i = 0; i < 64; ++i) {
        const float got = which == 0 ? bf2f(q[(size_t)m * 1024 + hd * 64 + i]) : bf2f(kv[(size_t)m * 2048 + (hd - 16) * 64 + i]);
        e.add(got, x[i], 6e-3, 8e-3);
      }
    }
  }
  char name[96];
  snprintf(name, sizeof name, "gemm_qkv M=%d qk_norm=%d rope=%d", M, with_norm, with_rope);
  return e.report(name);
}

// ------------------------------------------------------------------------------------------------ attention
// q [batch*Lq, heads*64], kv [batch*Lkv, heads*128] (k | v); reference for every `row_step`-th query row.
static int case_attn(int batch, int Lq, int Lkv, int heads, float qscale, int variant, unsigned lbo, unsigned sbo,
                     int row_step) {
  const int ldq = heads * 64, ldkv = heads * 128;
  BF Q, KV;
  // variants >= 8 (bounded-logit mode): q arrives pre-multiplied by scale * log2(e) and P = 2^(q . k)
  const bool bounded = (variant >= 8 && variant <= 19) || variant >= 24;  // 20..23: general softmax of the pipelined kernel
  Q.fill((size_t)batch * Lq * ldq, bounded ? qscale * 0.125f * 1.4426950408889634f : qscale);
  KV.fill((size_t)batch * Lkv * ldkv, 1.0f);
  uint16_t *dq = dev(Q.h), *dkv = dev(KV.h);
  uint16_t* dout = dev_alloc<uint16_t>((size_t)batch * Lq * ldq);
  // lbo != 0: no workspace (whole items per CTA); default: balanced steps + partial merge
  void* ws = lbo ? nullptr : (void*)dev_alloc<uint8_t>(wm_attn_workspace_bytes());
  (void)sbo;
  WM(wm_flash_attn_ws(0, dq, ldq, dkv, dkv + heads * 64, ldkv, dout, ldq, batch, Lq, Lkv, heads, 0.125f, variant, ws,
                      ws ? wm_attn_workspace_bytes() : 0));
  CK(cudaDeviceSynchronize());
  auto out = host(dout, (size_t)batch * Lq * ldq);
  Err e;
  std::vector<double> s(Lkv);
  for (int b = 0; b < batch; ++b)
    for (int h = 0; h < heads; ++h)
      for (int i = (b * 7 + h * 3) % row_step; i < Lq; i += row_step) {
        const float* qr = &Q.f[((size_t)b * Lq + i) * ldq + h * 64];
        double mx = -1e300;
        for (int j = 0; j < Lkv; ++j) {
          const float* kr = &KV.f[((size_t)b * Lkv + j) * ldkv + h * 64];
          double acc = 0;
          for (int d = 0; d < 64; ++d) acc += (double)qr[d] * kr[d];
          s[j] = bounded ? acc * 0.6931471805599453 : acc * 0.125;
          if (s[j] > mx) mx = s[j];
        }
        double l = 0, o[64] = {0};
        for (int j = 0; j < Lkv; ++j) {
          const double p = exp(s[j] - mx);
          l += p;
          const float* vr = &KV.f[((size_t)b * Lkv + j) * ldkv + heads * 64 + h * 64];
          for (int d = 0; d < 64; ++d) o[d] += p * vr[d];
        }
        for (int d = 0; d < 64; ++d) e.add(bf2f(out[((size_t)b * Lq + i) * ldq + h * 64 + d]), o[d] / l, 6e-3, 2e-2);
      }
  char name[128];
  snprintf(name, sizeof name, "attn b=%d Lq=%d Lkv=%d h=%d qs=%.0f var=%d lbo=%u sbo=%u", batch, Lq, Lkv, heads, qscale,
           variant, lbo, sbo);
  return e.report(name);
}

// ------------------------------------------------------------------------------------------------ timing
static float time_ms(int iters, void (*fn)(void*), void* ctx) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  for (int i = 0; i < 3; ++i) fn(ctx);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  for (int i = 0; i < iters; ++i) fn(ctx);
  CK(cudaEventRecord(b));
  CK(cudaEventSynchronize(b));
  float ms;
  CK(cudaEventElapsedTime(&ms, a, b));
  return ms / iters;
}

struct GemmCtx {
  int kind, M, N, K;
  void* ws;
  uint16_t *a, *w, *out, *q, *kv;
  float *bias, *gamma, *x, *qw, *cs;
};
static void gemm_run(void* p) {
  GemmCtx* c = (GemmCtx*)p;
  if (c->kind == 0) WM(wm_gemm_bias_act(0, c->a, c->K, c->w, c->K, c->bias, c->M, c->N, c->K, c->out, c->N, 0));
  if (c->kind == 1) WM(wm_gemm_bias_act(0, c->a, c->K, c->w, c->K, c->bias, c->M, c->N, c->K, c->out, c->N, 1));
  if (c->kind == 2)
    WM(wm_gemm_qkv(0, c->a, c->w, c->bias, c->qw, c->qw, c->qw, c->qw, 1e-5f, c->cs, c->cs, 1376, 7, 37, 1.0f, c->M, c->q, c->kv));
  if (c->kind == 3) WM(wm_gemm_ls_residual(0, c->a, c->K, c->w, c->K, c->bias, c->gamma, c->M, c->N, c->K, c->x, c->N));
  if (c->kind == 5) WM(wm_gemm_ls_residual(0, c->a, c->K, c->w, c->K, c->bias, nullptr, c->M, c->N, c->K, c->x, c->N));  // plain store
  if (c->kind == 4)
    WM(wm_gemm_ls_residual_ws(0, c->a, c->K, c->w, c->K, c->bias, c->gamma, c->M, c->N, c->K, c->x, c->N, c->ws,
                              wm_gemm_workspace_bytes()));
}
static int bench_gemm(int kind, int M, int N, int K) {
  GemmCtx c{};
  c.kind = kind, c.M = M, c.N = N, c.K = K;
  BF A, W;
  A.fill((size_t)M * K, 1.0f);
  W.fill((size_t)N * K, 0.03f);
  c.a = dev(A.h), c.w = dev(W.h);
  c.out = dev_alloc<uint16_t>((size_t)M * N);
  c.q = dev_alloc<uint16_t>((size_t)M * 1024), c.kv = dev_alloc<uint16_t>((size_t)M * 2048);
  c.bias = dev(randf(N, 0.1f)), c.gamma = dev(randf(N, 0.01f)), c.x = dev(randf((size_t)M * N, 1.0f));
  c.qw = dev(randf(64, 0.1f, 1.0f));
  std::vector<float> cs(64 * 16), sn(64 * 16);
  wm_rope_tables(63, cs.data(), sn.data());
  c.cs = dev(cs);
  if (kind == 4) c.ws = dev_alloc<uint8_t>(wm_gemm_workspace_bytes(), 0);
  const float ms = time_ms(20, gemm_run, &c);
  const char* names[] = {"bias", "gelu", "qkv", "ls_residual", "ls_resid_split", "ls_store"};
  printf("BENCH gemm_%-12s M=%d N=%d K=%d  %.3f ms  %.1f TFLOP/s\n", names[kind], M, N, K, ms, 2.0 * M * N * K / ms * 1e-9);
  return 0;
}

struct AttnCtx {
  uint16_t *q, *kv, *o;
  int batch, Lq, Lkv, heads, variant;
  void* ws;
};
static void attn_run(void* p) {
  AttnCtx* c = (AttnCtx*)p;
  WM(wm_flash_attn_ws(0, c->q, c->heads * 64, c->kv, c->kv + c->heads * 64, c->heads * 128, c->o, c->heads * 64, c->batch,
                      c->Lq, c->Lkv, c->heads, 0.125f, c->variant, c->ws, c->ws ? wm_attn_workspace_bytes() : 0));
}
static int bench_attn(int batch, int Lq, int Lkv, int heads, int variant, int no_ws) {
  AttnCtx c{};
  c.batch = batch, c.Lq = Lq, c.Lkv = Lkv, c.heads = heads, c.variant = variant;
  c.ws = no_ws ? nullptr : (void*)dev_alloc<uint8_t>(wm_attn_workspace_bytes());
  BF Q, KV;
  Q.fill((size_t)batch * Lq * heads * 64, 1.0f);
  KV.fill((size_t)batch * Lkv * heads * 128, 1.0f);
  c.q = dev(Q.h), c.kv = dev(KV.h), c.o = dev_alloc<uint16_t>((size_t)batch * Lq * heads * 64);
  const float ms = time_ms(Lkv > 50000 ? 3 : 10, attn_run, &c);
  printf("BENCH attn batch=%d Lq=%d Lkv=%d heads=%d variant=%d balanced=%d  %.3f ms  %.1f TFLOP/s\n", batch, Lq, Lkv, heads, variant, !no_ws, ms,
         4.0 * batch * (double)Lq * Lkv * heads * 64 / ms * 1e-9);
  return 0;
}

static int bench_ln(int rows) {
  struct C {
    float *x, *w;
    uint16_t* y;
    int rows;
  } c;
  c.rows = rows;
  c.x = dev(randf((size_t)rows * 1024, 1.0f)), c.w = dev(randf(1024, 0.1f, 1.0f));
  c.y = dev_alloc<uint16_t>((size_t)rows * 1024);
  const float ms = time_ms(20, [](void* p) { C* c = (C*)p; WM(wm_layernorm(0, c->x, c->w, c->w, 1e-5f, c->rows, c->y, 0)); }, &c);
  printf("BENCH layernorm rows=%d  %.4f ms  %.0f GB/s\n", rows, ms, rows * 6144.0 / ms * 1e-6);
  return 0;
}

static int bench_copy(int rows) {  // a17: one half of cat([frame, global], -1): read 4096 B + write 4096 B per token
  struct C {
    float *x, *out;
    int rows;
  } c;
  c.rows = rows;
  c.x = dev(randf((size_t)rows * 1024, 1.0f));
  c.out = dev_alloc<float>((size_t)rows * 2048);
  const float ms = time_ms(20, [](void* p) { C* c = (C*)p; WM(wm_copy_cols(0, c->x, c->rows, c->out, 2048, 1024)); }, &c);
  printf("BENCH copy_cols rows=%d  %.4f ms  %.0f GB/s\n", rows, ms, rows * 8192.0 / ms * 1e-6);
  return 0;
}

static int bench_gather(int views, int H, int W) {  // a1 + a2: normalise + im2col(14 x 14): read 4 B, write 2 B per pixel-channel
  struct C {
    float* img;
    uint16_t* out;
    int n, H, W, rows;
  } c;
  c.n = views, c.H = H, c.W = W, c.rows = 5 + (H / 14) * (W / 14);
  c.img = dev(randf((size_t)views * 3 * H * W, 0.5f, 0.5f));
  c.out = dev_alloc<uint16_t>((size_t)views * c.rows * 640, 0);
  const float ms = time_ms(20, [](void* p) { C* c = (C*)p; WM(wm_patch_gather(0, c->img, c->n, 3, c->H, c->W, c->rows, 5, 640, 1, c->out)); }, &c);
  printf("BENCH patch_gather views=%d %dx%d  %.4f ms  %.0f GB/s\n", views, H, W, ms, (double)views * 3 * H * W * 6.0 / ms * 1e-6);
  return 0;
}

// MUFU.EX2 issue rate per data type: does the fp16 / bf16 form of ex2 run faster than the fp32 one on sm_100a?
// (cuobjdump: ex2.approx.f16x2 / .bf16x2 each become TWO MUFU.EX2.F16 / .BF16 + PRMT, so only a higher unit rate could
// make the packed forms pay.)  One CTA per SM, 16 warps, a long dependent-free stream of ex2 per thread.
template <int MODE>
__global__ void mufu_rate_kernel(float* out, int iters) {
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 0.1f, a2 = a0 + 0.2f, a3 = a0 + 0.3f;
  unsigned h0 = 0x3c003800u + threadIdx.x, h1 = h0 + 7, h2 = h0 + 13, h3 = h0 + 29;
  for (int i = 0; i < iters; ++i) {
    if (MODE == 0) {
      asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a0));
      asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a1));
      asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a2));
      asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a3));
    } else if (MODE == 1) {
      asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h0));
      asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h1));
      asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h2));
      asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h3));
    } else {
      asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h0));
      asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h1));
      asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h2));
      asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h3));
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + __uint_as_float(h0 ^ h1 ^ h2 ^ h3);
}
static int bench_mufu() {
  float* out = dev_alloc<float>(148 * 512);
  const int iters = 20000;
  for (int mode = 0; mode < 3; ++mode) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    for (int rep = 0; rep < 2; ++rep) {
      CK(cudaEventRecord(a));
      if (mode == 0) mufu_rate_kernel<0><<<148, 512>>>(out, iters);
      if (mode == 1) mufu_rate_kernel<1><<<148, 512>>>(out, iters);
      if (mode == 2) mufu_rate_kernel<2><<<148, 512>>>(out, iters);
      CK(cudaEventRecord(b));
      CK(cudaEventSynchronize(b));
    }
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    const double results = 148.0 * 512 * iters * 4 * (mode == 0 ? 1 : 2);
    printf("BENCH mufu ex2 %-7s %.3f ms  %.1f G results/s  (%.2f results/clk/SM at 1.9 GHz)\n",
           mode == 0 ? "f32" : mode == 1 ? "f16x2" : "bf16x2", ms, results / ms * 1e-6, results / ms * 1e-6 / 148 / 1.9);
  }
  return 0;
}

// ------------------------------------------------------------------------------------------------ softmax inner-loop model
// How much of the attention kernel's exponential throughput is lost to too few warps?  A stand-in for the softmax warps'
// inner loop without the tensor core: every thread reads CHUNK * 2 "scores" per step from shared memory (standing in for the
// TMEM load), turns them into P = 2^s with NPOLY of every 32 on the FMA pipe (the kernel's packed polynomial) and the rest on
// MUFU, adds them up, packs bf16 pairs and writes them back.  16 warps x 64 scores per step at <= 128 registers (the kernel's
// shape) against 32 warps x 32 scores per step at <= 64 registers (4 threads per query row): same elements per SM.
__device__ __forceinline__ float2 st_ffma2(float2 a, float2 b, float2 c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)),
      "l"(*reinterpret_cast<unsigned long long*>(&b)), "l"(*reinterpret_cast<unsigned long long*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 st_fadd2(float2 a, float2 b) {
  unsigned long long d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 st_exp2_poly2(float2 x) {
  const float2 magic = make_float2(12582912.0f, 12582912.0f);
  const float2 xr = st_fadd2(x, magic);
  const float2 n = st_fadd2(xr, make_float2(-12582912.0f, -12582912.0f));
  const float2 f = st_ffma2(n, make_float2(-1.0f, -1.0f), x);
  float2 p = st_ffma2(make_float2(0.05520551f, 0.05520551f), f, make_float2(0.24261397f, 0.24261397f));
  p = st_ffma2(p, f, make_float2(0.69325477f, 0.69325477f));
  p = st_ffma2(p, f, make_float2(0.9999277f, 0.9999277f));
  float2 r;
  r.x = __uint_as_float(__float_as_uint(p.x) + (__float_as_uint(xr.x) << 23));
  r.y = __uint_as_float(__float_as_uint(p.y) + (__float_as_uint(xr.y) << 23));
  return r;
}
template <int THREADS, int CHUNK, int NPOLY>
__global__ void __launch_bounds__(THREADS, 1) softmax_rate_kernel(float* out, int steps) {
  extern __shared__ float4 st_smem[];
  // CHUNK scores per half-step and thread, laid out [CHUNK / 4][THREADS] float4 so that a warp reads 512 contiguous bytes
  float4* src = st_smem;
  uint4* dst = reinterpret_cast<uint4*>(st_smem + (CHUNK / 4) * THREADS);
  for (int i = threadIdx.x; i < (CHUNK / 4) * THREADS; i += THREADS)
    src[i] = make_float4(-0.1f * (i & 7), -0.3f, -1.7f + 0.01f * (i & 3), -2.2f);
  __syncthreads();
  float2 sum2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
  for (int s = 0; s < steps; ++s) {
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      float v[CHUNK];
#pragma unroll
      for (int q = 0; q < CHUNK / 4; ++q) {
        const float4 t = src[q * THREADS + (threadIdx.x + 32 * s + 64 * half) % THREADS];  // (moves with s and half: nothing is invariant or shared)
        v[4 * q] = t.x, v[4 * q + 1] = t.y, v[4 * q + 2] = t.z, v[4 * q + 3] = t.w;
      }
      uint32_t pk[CHUNK / 2];
#pragma unroll
      for (int c = 0; c < CHUNK; c += 2) {
        const float2 xs = make_float2(v[c], v[c + 1]);
        float2 p;
        if (c >= CHUNK - NPOLY * CHUNK / 32) {
          p = st_exp2_poly2(xs);
        } else {
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p.x) : "f"(xs.x));
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p.y) : "f"(xs.y));
        }
        sum2[(c >> 1) & 1] = st_fadd2(sum2[(c >> 1) & 1], p);
        __nv_bfloat162 b2 = __floats2bfloat162_rn(p.x, p.y);
        pk[c >> 1] = *reinterpret_cast<uint32_t*>(&b2);
      }
#pragma unroll
      for (int q = 0; q < CHUNK / 8; ++q)
        dst[(q + half * (CHUNK / 8)) * THREADS + threadIdx.x] = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
    }
  }
  out[blockIdx.x * THREADS + threadIdx.x] = sum2[0].x + sum2[0].y + sum2[1].x + sum2[1].y;
}
template <int THREADS, int CHUNK, int NPOLY>
static void bench_softmax_one(float* out, const char* what) {
  const size_t smem = (size_t)(CHUNK / 4) * THREADS * 16 + (size_t)2 * (CHUNK / 8) * THREADS * 16;
  CK(cudaFuncSetAttribute(softmax_rate_kernel<THREADS, CHUNK, NPOLY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaFuncAttributes fa;
  CK(cudaFuncGetAttributes(&fa, softmax_rate_kernel<THREADS, CHUNK, NPOLY>));
  // same number of elements per SM for every configuration: THREADS * 2 * CHUNK per step
  const int steps = (int)(512ll * 64 * 2 * 4000 / ((long long)THREADS * 2 * CHUNK));
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  float ms = 0;
  for (int rep = 0; rep < 2; ++rep) {
    CK(cudaEventRecord(a));
    softmax_rate_kernel<THREADS, CHUNK, NPOLY><<<148, THREADS, smem>>>(out, steps);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    CK(cudaEventElapsedTime(&ms, a, b));
  }
  const double elems = 148.0 * THREADS * 2.0 * CHUNK * steps;
  printf("BENCH softmax-loop %-44s regs %3d  %.3f ms  %.2f elements/clk/SM at 1.9 GHz\n", what, fa.numRegs, ms,
         elems / ms * 1e-6 / 148 / 1.9);
}
static int bench_softmax() {
  float* out = dev_alloc<float>(148 * 1024);
  bench_softmax_one<512, 32, 12>(out, "16 warps x 64 scores/step, 12/32 polynomial");
  bench_softmax_one<1024, 16, 12>(out, "32 warps x 32 scores/step, 12/32 polynomial");
  bench_softmax_one<512, 32, 16>(out, "16 warps x 64 scores/step, 16/32 polynomial");
  bench_softmax_one<1024, 16, 16>(out, "32 warps x 32 scores/step, 16/32 polynomial");
  bench_softmax_one<512, 32, 0>(out, "16 warps x 64 scores/step, MUFU only");
  bench_softmax_one<1024, 16, 0>(out, "32 warps x 32 scores/step, MUFU only");
  bench_softmax_one<768, 16, 12>(out, "24 warps x 32 scores/step, 12/32 polynomial");
  bench_softmax_one<256, 32, 12>(out, " 8 warps x 64 scores/step, 12/32 polynomial");
  printf("(the attention kernel: 256 x 128 scores per 2170 clk = 15.1 elements/clk/SM; MUFU bound at 12/32: 25.6)\n");
  return 0;
}

int main(int argc, char** argv) {
  if (argc < 2) {
    printf("usage: selftest <case> [args]\n");
    return 64;
  }
  const std::string c = argv[1];
  auto I = [&](int i, int def) { return argc > i ? atoi(argv[i]) : def; };
  if (c == "probe") return case_probe(I(2, 0), I(3, 1024), I(4, 1024));
  if (c == "layernorm") return case_layernorm(I(2, 333));
  if (c == "gemm_bias") return case_gemm_bias(I(2, 300), I(3, 512), I(4, 192), I(5, 0), I(6, 1));
  if (c == "gemm_resid") return case_gemm_resid(I(2, 300), I(3, 512), I(4, 192), I(5, 1), I(6, 1), I(7, 1), I(8, 0));
  if (c == "gemm_table") return case_gemm_table(I(2, 300), I(3, 1024), I(4, 588), I(5, 640), I(6, 77));
  if (c == "gemm_qkv") return case_gemm_qkv(I(2, 300), I(3, 1), I(4, 1), I(5, 3));
  if (c == "attn") return case_attn(I(2, 2), I(3, 331), I(4, 331), I(5, 2), (float)I(6, 1), I(7, 0), I(8, 0), I(9, 0), I(10, 5));
  if (c == "bench_gemm") return bench_gemm(I(2, 0), I(3, 11008), I(4, 1024), I(5, 1024));
  if (c == "bench_attn") return bench_attn(I(2, 1), I(3, 11008), I(4, 11008), I(5, 16), I(6, 0), I(7, 0));
  if (c == "bench_ln") return bench_ln(I(2, 11008));
  if (c == "bench_copy") return bench_copy(I(2, 11008));
  if (c == "bench_gather") return bench_gather(I(2, 8), I(3, 518), I(4, 518));
  if (c == "bench_mufu") return bench_mufu();
  if (c == "bench_softmax") return bench_softmax();
  printf("unknown case %s\n", c.c_str());
  return 64;
}
