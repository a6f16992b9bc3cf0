// Shared helpers for the fedb200 native library (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <type_traits>

#include "../include/fedb200.h"

namespace fedb {

extern thread_local char g_err[512];

inline int set_err(int code, const char* fmt, const char* a = "", const char* b = "") {
  snprintf(g_err, sizeof(g_err), fmt, a, b);
  return code;
}

#define FEDB_CUDA_OK(expr)                                                                      \
  do {                                                                                          \
    cudaError_t e_ = (expr);                                                                    \
    if (e_ != cudaSuccess) {                                                                    \
      snprintf(fedb::g_err, sizeof(fedb::g_err), "%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, \
               cudaGetErrorString(e_));                                                         \
      return FEDB_E_CUDA;                                                                       \
    }                                                                                           \
  } while (0)

#define FEDB_LAUNCH_OK()                                                                        \
  do {                                                                                          \
    cudaError_t e_ = cudaGetLastError();                                                        \
    if (e_ != cudaSuccess) {                                                                    \
      snprintf(fedb::g_err, sizeof(fedb::g_err), "%s:%d: kernel launch failed: %s", __FILE__, __LINE__, \
               cudaGetErrorString(e_));                                                         \
      return FEDB_E_CUDA;                                                                       \
    }                                                                                           \
  } while (0)

#define FEDB_REQUIRE(cond, msg)                                                                 \
  do {                                                                                          \
    if (!(cond)) {                                                                              \
      snprintf(fedb::g_err, sizeof(fedb::g_err), "%s:%d: %s", __FILE__, __LINE__, msg);          \
      return FEDB_E_ARG;                                                                        \
    }                                                                                           \
  } while (0)

__host__ __device__ inline int ru4(int x) { return (x + 3) & ~3; }

// Offsets (floats) of layer l's weight / bias inside a slab; mirrors MLPSpec._layout().
struct SlabLayout {
  int w_off[FEDB_MAX_LAYERS];
  int b_off[FEDB_MAX_LAYERS];
  int ld[FEDB_MAX_LAYERS];
  int end;    // first float past the last bias (padding up to `total` is never touched)
  int total;
};

__host__ __device__ inline SlabLayout slab_layout(const int32_t* dims, int n_layers) {
  SlabLayout s;
  int off = 0;
  for (int l = 0; l < FEDB_MAX_LAYERS; ++l) {
    if (l < n_layers) {
      s.ld[l] = ru4(dims[l]);
      s.w_off[l] = off;
      off += dims[l + 1] * s.ld[l];
      s.b_off[l] = off;
      off += ru4(dims[l + 1]);
    } else {
      s.ld[l] = 0;
      s.w_off[l] = s.b_off[l] = off;
    }
  }
  s.end = off;
  s.total = (off + 31) & ~31;
  return s;
}

// Philox4x32-10 (Salmon et al.), bit-identical to models/init.py:philox4x32.
__host__ __device__ inline void philox4x32(uint32_t c[4], uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)M0 * c[0];
    uint64_t p1 = (uint64_t)M1 * c[2];
    uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
    uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
    uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += W0; k1 += W1;
  }
}

// keyed Feistel bijection on [0, n) by cycle walking: the per-epoch reshuffle of a shard (k_shuffle and the persistent fit kernels)
__device__ __forceinline__ uint32_t feistel_perm(uint32_t i, uint32_t n, int half_bits, uint32_t ka, uint32_t kb) {
  const uint32_t mask = (1u << half_bits) - 1u;
  uint32_t x = i;
  do {
    uint32_t l = x >> half_bits, r = x & mask;
#pragma unroll
    for (int round = 0; round < 4; ++round) {
      uint32_t c[4] = {r, (uint32_t)round, ka, kb};
      philox4x32(c, 0xA511E9B3u, 0x5851F42Du);
      const uint32_t nl = r, nr = l ^ (c[0] & mask);
      l = nl; r = nr;
    }
    x = (l << half_bits) | r;
  } while (x >= n);
  return x;
}

__device__ __forceinline__ float act_fwd(float z, int act) {
  switch (act) {
    case FEDB_ACT_RELU: return fmaxf(z, 0.f);
    case FEDB_ACT_LOGISTIC: return 1.f / (1.f + __expf(-z));
    case FEDB_ACT_TANH: return tanhf(z);
    default: return z;
  }
}
// derivative expressed through the activation value a = act(z) (SKB:142-157)
__device__ __forceinline__ float act_bwd(float a, int act) {
  switch (act) {
    case FEDB_ACT_RELU: return a > 0.f ? 1.f : 0.f;
    case FEDB_ACT_LOGISTIC: return a * (1.f - a);
    case FEDB_ACT_TANH: return 1.f - a * a;
    default: return 1.f;
  }
}

// ---- optimizer arithmetic shared by the SIMT and tensor-core kernels ----------------------
struct StepScalars {
  float lr_eff;   // torch Adam: lr / (1 - b1^t) ; sklearn Adam: lr*sqrt(1-b2^t)/(1-b1^t) ; SGD: lr
  float inv_sqrt_bc2;  // torch Adam: 1/sqrt(1 - b2^t)
};

__device__ inline StepScalars step_scalars(const FedbGroup& G, int t) {
  StepScalars s;
  double lr = (double)G.lr;
  if (G.lr_step_size > 0) lr *= pow((double)G.lr_gamma, (double)((t - 1) / G.lr_step_size));
  if (G.opt == FEDB_OPT_ADAM_TORCH) {
    const double bc1 = 1.0 - pow((double)G.beta1, (double)t), bc2 = 1.0 - pow((double)G.beta2, (double)t);
    s.lr_eff = (float)(lr / bc1);
    s.inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  } else if (G.opt == FEDB_OPT_ADAM_SKLEARN) {
    const double bc1 = 1.0 - pow((double)G.beta1, (double)t), bc2 = 1.0 - pow((double)G.beta2, (double)t);
    s.lr_eff = (float)(lr * sqrt(bc2) / bc1);
    s.inv_sqrt_bc2 = 1.f;
  } else {
    s.lr_eff = (float)lr;
    s.inv_sqrt_bc2 = 1.f;
  }
  return s;
}

// step_scalars() in single precision (exp2f/log2f instead of double pow): evaluated by one thread
// on the critical path of every step, where the double version costs ~3K cycles.  |rel err| < 1e-5.
__device__ inline StepScalars step_scalars_fast(const FedbGroup& G, int t) {
  StepScalars s;
  float lr = G.lr;
  if (G.lr_step_size > 0) lr *= exp2f((float)((t - 1) / G.lr_step_size) * log2f(G.lr_gamma));
  if (G.opt == FEDB_OPT_SGD) {
    s.lr_eff = lr; s.inv_sqrt_bc2 = 1.f;
    return s;
  }
  const float bc1 = 1.f - exp2f((float)t * log2f(G.beta1)), bc2 = 1.f - exp2f((float)t * log2f(G.beta2));
  if (G.opt == FEDB_OPT_ADAM_TORCH) { s.lr_eff = lr / bc1; s.inv_sqrt_bc2 = rsqrtf(bc2); }
  else { s.lr_eff = lr * sqrtf(bc2) / bc1; s.inv_sqrt_bc2 = 1.f; }
  return s;
}

__device__ __forceinline__ float opt_update(const FedbGroup& G, const StepScalars& sc, float w, float g, float& m,
                                            float& v) {
  if (G.opt == FEDB_OPT_ADAM_TORCH) {
    m = fmaf(G.beta1, m, (1.f - G.beta1) * g);
    v = fmaf(G.beta2, v, (1.f - G.beta2) * g * g);
    const float denom = sqrtf(v) * sc.inv_sqrt_bc2 + G.eps;
    return w - sc.lr_eff * (m / denom);
  } else if (G.opt == FEDB_OPT_ADAM_SKLEARN) {
    m = fmaf(G.beta1, m, (1.f - G.beta1) * g);
    v = fmaf(G.beta2, v, (1.f - G.beta2) * g * g);
    return w - sc.lr_eff * m / (sqrtf(v) + G.eps);
  } else {
    // SKO SGDOptimizer: velocity = momentum*velocity - lr*g ; nesterov looks ahead
    m = G.momentum * m - sc.lr_eff * g;
    const float upd = G.nesterov ? (G.momentum * m - sc.lr_eff * g) : m;
    return w + upd;
  }
}

// opt_update with the optimizer kind as a compile-time constant: callers that update many independent parameters per thread
// branch ONCE around straight-line code, so the square-root / division chains overlap.  Same expressions, same results.
template <int OPT>
__device__ __forceinline__ float opt_update_t(const FedbGroup& G, const StepScalars& sc, float w, float g, float& m, float& v) {
  if (OPT == FEDB_OPT_ADAM_TORCH) {
    m = fmaf(G.beta1, m, (1.f - G.beta1) * g);
    v = fmaf(G.beta2, v, (1.f - G.beta2) * g * g);
    const float denom = sqrtf(v) * sc.inv_sqrt_bc2 + G.eps;
    return w - sc.lr_eff * (m / denom);
  } else if (OPT == FEDB_OPT_ADAM_SKLEARN) {
    m = fmaf(G.beta1, m, (1.f - G.beta1) * g);
    v = fmaf(G.beta2, v, (1.f - G.beta2) * g * g);
    return w - sc.lr_eff * m / (sqrtf(v) + G.eps);
  } else {
    m = G.momentum * m - sc.lr_eff * g;
    const float upd = G.nesterov ? (G.momentum * m - sc.lr_eff * g) : m;
    return w + upd;
  }
}

// Register-resident form of the optimizer for streaming epilogues (the wgrad GEMMs): the group's constants are
// read once per thread, the two Adam flavours share one expression (sklearn's has inv_sqrt_bc2 = 1), and the square
// root / quotient use the SFU approximations (2 ulp) — the update feeds bf16 tensor-core operands, and a lone CTA's
// warps must keep up with their SM's share of HBM while the neighbouring CTA runs its main loop.
struct OptRegs {
  float b1, omb1, b2, omb2, eps, lr_eff, isb2, mom, l2;
  int adam, nesterov;
};
__device__ __forceinline__ OptRegs opt_regs(const FedbGroup& G, const StepScalars& sc, float l2) {
  OptRegs o;
  o.b1 = G.beta1; o.omb1 = 1.f - G.beta1; o.b2 = G.beta2; o.omb2 = 1.f - G.beta2; o.eps = G.eps;
  o.lr_eff = sc.lr_eff; o.isb2 = sc.inv_sqrt_bc2; o.mom = G.momentum; o.l2 = l2;
  o.adam = G.opt != FEDB_OPT_SGD; o.nesterov = G.nesterov;
  return o;
}
__device__ __forceinline__ float opt_update_fast(const OptRegs& o, float w, float g, float& m, float& v) {
  g = fmaf(o.l2, w, g);
  if (o.adam) {
    m = fmaf(o.b1, m, o.omb1 * g);
    v = fmaf(o.b2, v, o.omb2 * g * g);
    float sq;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(v));
    return fmaf(-o.lr_eff, __fdividef(m, fmaf(sq, o.isb2, o.eps)), w);
  }
  m = fmaf(o.mom, m, -o.lr_eff * g);
  return w + (o.nesterov ? fmaf(o.mom, m, -o.lr_eff * g) : m);
}


// ---- metrics from confusion counts + early stop (K3/K5), shared by the aggregation kernels ---------
// a / b to ~1e-14 relative: single-precision quotient plus one double-precision residual correction.
// IEEE double division is a ~100-instruction dependent chain; a round's metrics need a dozen of them
// on ONE warp on the critical path, which measured at ~4 us.
__device__ __forceinline__ double fast_ddiv(double a, double b) {
  const float rb = 1.0f / (float)b;
  const double q0 = (double)((float)a * rb);
  return fma(fma(-q0, b, a), (double)rb, q0);
}

// Binary problems (every model of the reference) take a short straight-line path; the general case
// is a compact rolled loop.  Code size matters here: this runs once per round on one warp, and a
// fully unrolled 8x8 version spent more time missing the instruction cache than computing.
__device__ inline void metrics_from_conf(const int32_t* cm, int C, double out[4]) {
  out[0] = out[1] = out[2] = out[3] = 0.0;
  if (C == 2) {
    const double c00 = (double)cm[0], c01 = (double)cm[1], c10 = (double)cm[FEDB_MAX_CLASSES], c11 = (double)cm[FEDB_MAX_CLASSES + 1];
    const double s0 = c00 + c01, s1 = c10 + c11, p0 = c00 + c10, p1 = c01 + c11, total = s0 + s1;
    if (total == 0) return;
    const double it = fast_ddiv(1.0, total), w0 = s0 * it, w1 = s1 * it;
    out[0] = (c00 + c11) * it;
    out[1] = (p0 > 0 ? fast_ddiv(c00, p0) : 0.0) * w0 + (p1 > 0 ? fast_ddiv(c11, p1) : 0.0) * w1;
    out[2] = (s0 > 0 ? fast_ddiv(c00, s0) : 0.0) * w0 + (s1 > 0 ? fast_ddiv(c11, s1) : 0.0) * w1;
    out[3] = (s0 + p0 > 0 ? fast_ddiv(2.0 * c00, s0 + p0) : 0.0) * w0 + (s1 + p1 > 0 ? fast_ddiv(2.0 * c11, s1 + p1) : 0.0) * w1;
    return;
  }
  double total = 0, tr = 0;
  for (int i = 0; i < C; ++i)
    for (int j = 0; j < C; ++j) {
      const double x = (double)cm[i * FEDB_MAX_CLASSES + j];
      total += x;
      if (i == j) tr += x;
    }
  if (total == 0) return;
  out[0] = fast_ddiv(tr, total);
#pragma unroll 1
  for (int c = 0; c < C; ++c) {
    double sup = 0, prd = 0;
    for (int j = 0; j < C; ++j) { sup += (double)cm[c * FEDB_MAX_CLASSES + j]; prd += (double)cm[j * FEDB_MAX_CLASSES + c]; }
    const double tp = (double)cm[c * FEDB_MAX_CLASSES + c];
    const double w = fast_ddiv(sup, total);
    const double p = prd > 0 ? fast_ddiv(tp, prd) : 0.0;
    const double r = sup > 0 ? fast_ddiv(tp, sup) : 0.0;
    const double f = sup + prd > 0 ? fast_ddiv(2.0 * tp, sup + prd) : 0.0;
    out[1] += p * w; out[2] += r * w; out[3] += f * w;
  }
}

// C:181-192: anchor / patience early-stop rule on the four global metrics.
__device__ inline void early_stop_update(const FedbFed& F, FedbFedState& S, const double glob[4]) {
  for (int i = 0; i < 4; ++i) S.last[i] = glob[i];
  if (F.patience <= 0) return;
  bool close = S.has_anchor != 0;
  if (close)
    for (int i = 0; i < 4; ++i)
      close = close && (fabs(glob[i] - S.anchor[i]) <= (double)F.atol + (double)F.rtol * fabs(S.anchor[i]));
  if (close) {
    S.count -= 1;
    if (S.count == 0) S.stop_pending = 1;
  } else {
    for (int i = 0; i < 4; ++i) S.anchor[i] = glob[i];
    S.has_anchor = 1;
    S.count = F.patience;
  }
}

// What fedb_round_begin does for one federation (C:130-142), callable from an aggregation kernel's tail.
__device__ inline void begin_next_round(FedbFedState& S) {
  if (S.stop_pending) S.stop = 1;
  if (!S.stop) S.round += 1;
}

// One WARP turns a federation's confusion counts into the round's metrics: lane k handles clients
// k, k+32, ... (per-client metrics -> hist_local), a fixed xor-tree adds them up (mean of local
// metrics, C:169) or the pooled counts are scored (S:130-134); lane 0 then records the global
// metrics and runs the early-stop rule.  `pool` = 64 ints of shared memory.  All 32 lanes must call.
// (F is taken BY VALUE: through a pointer every field access may be re-fetched from L2 after each store)
__device__ inline void round_metrics_warp(const FedbFed F, FedbFedState& S, int f, int n_groups, const int32_t* conf,
                                          float* hist_local, float* hist_global, int* pool, bool begin_next = false,
                                          FedbFedState* host_mirror = nullptr, void* host_result = nullptr, long long* stamps = nullptr) {
  const int lane = threadIdx.x & 31;
#define FEDB_MSTAMP(i) do { if (stamps && lane == 0) { long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); stamps[i] = t_; } } while (0)
  FEDB_MSTAMP(0);
  const int C = F.n_classes;
  const int r = S.round - 1;  // 0-based index of the round in progress
  const bool pooled = F.metric_mode == FEDB_METRIC_POOLED;
  for (int i = lane; i < FEDB_MAX_CLASSES * FEDB_MAX_CLASSES; i += 32) pool[i] = 0;
  __syncwarp();
  double acc[4] = {0, 0, 0, 0};
  for (int k = lane; k < F.n_clients; k += 32) {
    const int g = F.first_group + k;
    const int32_t* cm = conf + (size_t)g * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES;
    double mloc[4];
    metrics_from_conf(cm, C, mloc);
    if (hist_local && r >= 0 && r < F.max_rounds)
      *reinterpret_cast<float4*>(hist_local + ((size_t)r * n_groups + g) * 4) =
          make_float4((float)mloc[0], (float)mloc[1], (float)mloc[2], (float)mloc[3]);
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] += mloc[i];
    if (pooled)
      for (int i = 0; i < C; ++i)
        for (int j = 0; j < C; ++j) atomicAdd(&pool[i * FEDB_MAX_CLASSES + j], cm[i * FEDB_MAX_CLASSES + j]);
  }
  FEDB_MSTAMP(1);
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
  __syncwarp();
  FEDB_MSTAMP(2);
  if (lane == 0) {
    double glob[4];
    if (pooled) metrics_from_conf(pool, C, glob);
    else
      for (int i = 0; i < 4; ++i) glob[i] = fast_ddiv(acc[i], (double)F.n_clients);
    if (hist_global && r >= 0 && r < F.max_rounds)
      *reinterpret_cast<float4*>(hist_global + ((size_t)f * F.max_rounds + r) * 4) =
          make_float4((float)glob[0], (float)glob[1], (float)glob[2], (float)glob[3]);
    // work on a private copy: S lives in global memory and every dependent field access would be an
    // L2 round trip on the round's critical path
    FedbFedState st = S;
    early_stop_update(F, st, glob);
    st.done = st.round;
    if (begin_next) begin_next_round(st);
    S = st;
    if (host_mirror) {
      // mapped pinned host memory: the round's result reaches the host without a memcpy node (complete once the stream has
      // synchronised).  A host that does not want to wait for the kernel to retire polls `host_result` instead: four 16-byte
      // chunks {done, stop, metric i as a double}, each ONE store and so one write on the host link — every chunk carries the
      // round counter, which makes it valid on its own; no system-scope fence (a ~2 us round trip) is needed to order them.
      *host_mirror = st;
      if (host_result) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          reinterpret_cast<int4*>(host_result)[i] = make_int4(st.done, st.stop, __double2loint(st.last[i]), __double2hiint(st.last[i]));
      }
    }
  }
  FEDB_MSTAMP(3);
#undef FEDB_MSTAMP
  __syncwarp();
}

}  // namespace fedb
