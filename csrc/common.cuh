// Shared helpers for the fedb200 native library (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

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


}  // namespace fedb
