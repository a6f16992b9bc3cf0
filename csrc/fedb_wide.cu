// Wide-MLP path (BASELINE config 5), the pieces around the grouped tensor-core GEMM: operand
// preparation, the narrow classification head (a tcgen05 N of 2..8 would waste the instruction, so
// it runs on CUDA cores), bias gradients, and the bf16 shadow of the fp32 master weights.
//   k_wide_gather_x   rows of the resident dataset -> bf16 [G][B][ld0] activations of layer 0
//   k_wide_shadow     fp32 slab arena -> bf16 shadow arena (after init / aggregation)
//   k_wide_head_fwd   logits, softmax-CE or sigmoid/log-loss, dz of the head; or predictions -> counts
//   k_wide_head_bwd   head weight gradient + optimizer, and dZ of the last hidden layer
//   k_wide_bias_opt   bias gradient (column sums of dZ) + optimizer for a hidden layer
// Reference behaviour: C:63-73 / SK:301-393 for an MLP too wide for the whole-network kernels.
#include "common.cuh"
#include <cuda_bf16.h>

namespace fedb {

__global__ void k_wide_gather_x(const FedbGroup* __restrict__ groups, const float* __restrict__ X, int ldx,
                                const int32_t* __restrict__ idx, __nv_bfloat16* __restrict__ A0, int B, int ld0) {
  const int g = blockIdx.y;
  const FedbGroup& G = groups[g];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B * ld0; i += gridDim.x * blockDim.x) {
    const int r = i / ld0, c = i - r * ld0;
    float v = 0.f;
    if (r < G.n_rows && c < G.dims[0]) v = X[(size_t)idx[G.idx_off + r] * ldx + c];
    A0[((size_t)g * B + r) * ld0 + c] = __float2bfloat16_rn(v);
  }
}

__global__ void k_wide_shadow(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t n4) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(src)[i];
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    reinterpret_cast<uint2*>(dst)[i] = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
  }
}

// One warp per data row: logits = A[r] . W[c] + b[c] (fp32 master head weights), then loss / dz (train) or
// the prediction's confusion count (eval).  grid = (ceil(B / 8), G), 256 threads.
__global__ void __launch_bounds__(256)
k_wide_head_fwd(const FedbGroup* __restrict__ groups, const float* __restrict__ params, const __nv_bfloat16* __restrict__ A, int B,
                int d, const int32_t* __restrict__ y, const int32_t* __restrict__ idx, float* __restrict__ dz,
                float* __restrict__ loss_rows, int32_t* __restrict__ conf, int32_t* __restrict__ step,
                const FedbFedState* __restrict__ fed_state, int eval_mode) {
  const int g = blockIdx.y;
  const FedbGroup& G = groups[g];
  if (fed_state && (fed_state[G.fed].stop)) return;
  const int L = G.n_layers, C = G.dims[L];
  const SlabLayout lay = slab_layout(G.dims, L);
  const float* W = params + G.slab_off + lay.w_off[L - 1];
  const float* bias = params + G.slab_off + lay.b_off[L - 1];
  const int ldw = lay.ld[L - 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * 8 + warp;
  if (!eval_mode && blockIdx.x == 0 && threadIdx.x == 0) step[g] += 1;
  if (r >= B) return;
  const bool valid = r < G.n_rows;
  const __nv_bfloat16* a = A + ((size_t)g * B + r) * d;
  float z[FEDB_MAX_CLASSES];
#pragma unroll
  for (int c = 0; c < FEDB_MAX_CLASSES; ++c) z[c] = 0.f;
  for (int k = lane * 8; k < d; k += 256) {
    const uint4 q = *reinterpret_cast<const uint4*>(a + k);
    const __nv_bfloat16* ab = reinterpret_cast<const __nv_bfloat16*>(&q);
    float av[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) av[i] = __bfloat162float(ab[i]);
#pragma unroll
    for (int c = 0; c < FEDB_MAX_CLASSES; ++c)
      if (c < C) {
        const float4 w0 = __ldg(reinterpret_cast<const float4*>(W + (size_t)c * ldw + k));
        const float4 w1 = __ldg(reinterpret_cast<const float4*>(W + (size_t)c * ldw + k + 4));
        z[c] += av[0] * w0.x + av[1] * w0.y + av[2] * w0.z + av[3] * w0.w + av[4] * w1.x + av[5] * w1.y + av[6] * w1.z + av[7] * w1.w;
      }
  }
#pragma unroll
  for (int c = 0; c < FEDB_MAX_CLASSES; ++c)
    if (c < C) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) z[c] += __shfl_xor_sync(0xffffffffu, z[c], o);
      z[c] += __ldg(bias + c);
    }
  if (lane != 0) return;
  const int label = valid ? y[idx[G.idx_off + r]] : 0;
  if (eval_mode) {
    int pred = 0;
    if (G.head == FEDB_HEAD_SOFTMAX_CE) {
      float best = z[0];
      for (int c = 1; c < C; ++c)
        if (z[c] > best) { best = z[c]; pred = c; }
    } else {
      pred = z[0] > 0.f ? 1 : 0;
    }
    if (valid && label >= 0 && label < FEDB_MAX_CLASSES)
      atomicAdd(conf + (size_t)g * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES + label * FEDB_MAX_CLASSES + pred, 1);
    return;
  }
  const float inv_n = 1.f / (float)G.n_rows;
  float loss = 0.f, dzv[FEDB_MAX_CLASSES];
#pragma unroll
  for (int c = 0; c < FEDB_MAX_CLASSES; ++c) dzv[c] = 0.f;
  if (valid) {
    if (G.head == FEDB_HEAD_SOFTMAX_CE) {
      float mx = -INFINITY;
      for (int c = 0; c < C; ++c) mx = fmaxf(mx, z[c]);
      float se = 0.f;
      for (int c = 0; c < C; ++c) se += expf(z[c] - mx);
      const float lse = mx + logf(se);
      for (int c = 0; c < C; ++c) {
        if (c == label) loss = lse - z[c];
        dzv[c] = (expf(z[c] - lse) - (c == label ? 1.f : 0.f)) * inv_n;
      }
    } else {
      const float p = 1.f / (1.f + expf(-z[0]));
      const float eps = 1.1920929e-07f, pc = fminf(fmaxf(p, eps), 1.f - eps), yt = (float)label;
      loss = -(yt * logf(pc) + (1.f - yt) * logf(1.f - pc));
      dzv[0] = (p - yt) * inv_n;
    }
  }
  float* dzr = dz + ((size_t)g * B + r) * FEDB_MAX_CLASSES;
  *reinterpret_cast<float4*>(dzr) = make_float4(dzv[0], dzv[1], dzv[2], dzv[3]);
  *reinterpret_cast<float4*>(dzr + 4) = make_float4(dzv[4], dzv[5], dzv[6], dzv[7]);
  loss_rows[(size_t)g * B + r] = loss;
}

// Head weight gradient + optimizer and dZ of the last hidden layer.  A block owns 128 columns j of that layer (a lane
// owns 4 consecutive ones, so a warp moves a 256-byte row segment per access) and its 8 warps split the rows:
//   dZ[r][j] = (A[r][j] > 0) * sum_c dz[r][c] W_head[c][j]      (weights as they were before the update; row-independent)
//   dW_head[c][j] = sum_r dz[r][c] A[r][j]                       (per-warp partial sums, combined in warp order)
// Block (0, g) additionally updates the head bias and reduces the loss.  grid = (ceil(d / 128), G).
__global__ void __launch_bounds__(256)
k_wide_head_bwd(const FedbGroup* __restrict__ groups, float* __restrict__ params, float* __restrict__ m, float* __restrict__ v,
                __nv_bfloat16* __restrict__ shadow, const __nv_bfloat16* __restrict__ A, int B, int d, const float* __restrict__ dz,
                const float* __restrict__ loss_rows, __nv_bfloat16* __restrict__ dZ, const int32_t* __restrict__ step,
                const FedbFedState* __restrict__ fed_state, float* __restrict__ loss_out) {
  const int g = blockIdx.y;
  const FedbGroup& G = groups[g];
  if (fed_state && fed_state[G.fed].stop) return;
  const int L = G.n_layers, C = G.dims[L];
  const SlabLayout lay = slab_layout(G.dims, L);
  const size_t wbase = (size_t)G.slab_off + lay.w_off[L - 1], bbase = (size_t)G.slab_off + lay.b_off[L - 1];
  const int ldw = lay.ld[L - 1];
  constexpr int RCH = 256;                                   // rows of dz staged at a time
  __shared__ float dz_s[RCH][FEDB_MAX_CLASSES];
  __shared__ float red[8][FEDB_MAX_CLASSES][128];            // per-warp partial head gradients
  __shared__ StepScalars sc_s;
  if (threadIdx.x == 0) sc_s = step_scalars_fast(G, step[g]);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j0 = blockIdx.x * 128 + lane * 4;                // d is a multiple of 4 (padded layer width)
  const bool col_ok = j0 < d;
  float wj[FEDB_MAX_CLASSES][4], gw[FEDB_MAX_CLASSES][4];
#pragma unroll
  for (int c = 0; c < FEDB_MAX_CLASSES; ++c) {
    const float4 q = (c < C && col_ok) ? *reinterpret_cast<const float4*>(params + wbase + (size_t)c * ldw + j0) : make_float4(0.f, 0.f, 0.f, 0.f);
    wj[c][0] = q.x; wj[c][1] = q.y; wj[c][2] = q.z; wj[c][3] = q.w;
    gw[c][0] = gw[c][1] = gw[c][2] = gw[c][3] = 0.f;
  }
  const __nv_bfloat16* Ag = A + (size_t)g * B * d;
  __nv_bfloat16* dZg = dZ + (size_t)g * B * d;
  for (int r0 = 0; r0 < B; r0 += RCH) {
    if (r0) __syncthreads();
    for (int i = threadIdx.x; i < RCH * FEDB_MAX_CLASSES; i += 256)
      dz_s[i / FEDB_MAX_CLASSES][i % FEDB_MAX_CLASSES] = (r0 + i / FEDB_MAX_CLASSES < B) ? dz[((size_t)g * B + r0) * FEDB_MAX_CLASSES + i] : 0.f;
    __syncthreads();
    if (col_ok) {
      const int rend = min(RCH, B - r0);
      for (int rr = warp; rr < rend; rr += 32) {             // 4 rows of this warp in flight
        uint2 q[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (rr + 8 * u < rend) q[u] = __ldcs(reinterpret_cast<const uint2*>(Ag + (size_t)(r0 + rr + 8 * u) * d + j0));
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (rr + 8 * u < rend) {
            const __nv_bfloat16* ab = reinterpret_cast<const __nv_bfloat16*>(&q[u]);
            float a[4], da[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = __bfloat162float(ab[i]);
#pragma unroll
            for (int c = 0; c < FEDB_MAX_CLASSES; ++c)
              if (c < C) {
                const float z = dz_s[rr + 8 * u][c];
#pragma unroll
                for (int i = 0; i < 4; ++i) { gw[c][i] = fmaf(z, a[i], gw[c][i]); da[i] = fmaf(z, wj[c][i], da[i]); }
              }
            const __nv_bfloat162 lo = __floats2bfloat162_rn(a[0] > 0.f ? da[0] : 0.f, a[1] > 0.f ? da[1] : 0.f);
            const __nv_bfloat162 hi = __floats2bfloat162_rn(a[2] > 0.f ? da[2] : 0.f, a[3] > 0.f ? da[3] : 0.f);
            *reinterpret_cast<uint2*>(dZg + (size_t)(r0 + rr + 8 * u) * d + j0) =
                make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
          }
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < FEDB_MAX_CLASSES; ++c)
    if (c < C) *reinterpret_cast<float4*>(&red[warp][c][lane * 4]) = make_float4(gw[c][0], gw[c][1], gw[c][2], gw[c][3]);
  __syncthreads();
  const StepScalars sc = sc_s;
  const bool use_v = v && G.opt != FEDB_OPT_SGD;
  for (int t = threadIdx.x; t < C * 128; t += 256) {
    const int c = t >> 7, col = t & 127, j = blockIdx.x * 128 + col;
    if (j < d) {
      float gsum = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) gsum += red[w][c][col];
      const float l2 = G.alpha / (float)G.n_rows;
      const size_t o = wbase + (size_t)c * ldw + j;
      const float w0 = params[o];
      float mm = m ? m[o] : 0.f, vv = use_v ? v[o] : 0.f;
      const float w = opt_update(G, sc, w0, fmaf(l2, w0, gsum), mm, vv);
      params[o] = w;
      if (m) m[o] = mm;
      if (use_v) v[o] = vv;
      if (shadow) shadow[o] = __float2bfloat16_rn(w);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < 32) {   // head bias + loss: one warp, fixed order
    const int lane = threadIdx.x;
    float gb[FEDB_MAX_CLASSES], ls = 0.f;
#pragma unroll
    for (int c = 0; c < FEDB_MAX_CLASSES; ++c) gb[c] = 0.f;
    for (int r = lane; r < B; r += 32) {
      ls += loss_rows[(size_t)g * B + r];
#pragma unroll
      for (int c = 0; c < FEDB_MAX_CLASSES; ++c) gb[c] += dz[((size_t)g * B + r) * FEDB_MAX_CLASSES + c];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      ls += __shfl_xor_sync(0xffffffffu, ls, o);
#pragma unroll
      for (int c = 0; c < FEDB_MAX_CLASSES; ++c) gb[c] += __shfl_xor_sync(0xffffffffu, gb[c], o);
    }
    if (lane < C) {
      const size_t o = bbase + lane;
      float mm = m ? m[o] : 0.f, vv = use_v ? v[o] : 0.f;
      float gsel = 0.f;
#pragma unroll
      for (int c = 0; c < FEDB_MAX_CLASSES; ++c) gsel = (c == lane) ? gb[c] : gsel;
      params[o] = opt_update(G, sc, params[o], gsel, mm, vv);
      if (m) m[o] = mm;
      if (use_v) v[o] = vv;
    }
    if (lane == 0 && loss_out) loss_out[g] = ls / (float)G.n_rows;
  }
}

// db_l[j] = sum_r dZ[r][j]; bias update.  grid = (ceil(d / 256), G).
__global__ void __launch_bounds__(256)
k_wide_bias_opt(const FedbGroup* __restrict__ groups, float* __restrict__ params, float* __restrict__ m, float* __restrict__ v,
                const __nv_bfloat16* __restrict__ dZ, int B, int d, int layer, const int32_t* __restrict__ step,
                const FedbFedState* __restrict__ fed_state) {
  const int g = blockIdx.y;
  const FedbGroup& G = groups[g];
  if (fed_state && fed_state[G.fed].stop) return;
  const SlabLayout lay = slab_layout(G.dims, G.n_layers);
  __shared__ StepScalars sc_s;
  if (threadIdx.x == 0) sc_s = step_scalars_fast(G, step[g]);
  __syncthreads();
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= d) return;
  const __nv_bfloat16* z = dZ + (size_t)g * B * d + j;
  float s = 0.f;
#pragma unroll 8
  for (int r = 0; r < B; ++r) s += __bfloat162float(z[(size_t)r * d]);
  const size_t o = (size_t)G.slab_off + lay.b_off[layer] + j;
  const bool use_v = v && G.opt != FEDB_OPT_SGD;
  float mm = m ? m[o] : 0.f, vv = use_v ? v[o] : 0.f;
  params[o] = opt_update(G, sc_s, params[o], s, mm, vv);
  if (m) m[o] = mm;
  if (use_v) v[o] = vv;
}

}  // namespace fedb

using namespace fedb;

extern "C" int fedb_wide_gather_x(const FedbGroup* groups, int32_t n_groups, const float* X, int32_t ldx, const int32_t* idx,
                                  void* A0, int32_t B, int32_t ld0, void* stream) {
  dim3 grid((B * ld0 + 255) / 256, n_groups);
  k_wide_gather_x<<<grid, 256, 0, (cudaStream_t)stream>>>(groups, X, ldx, idx, (__nv_bfloat16*)A0, B, ld0);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}
extern "C" int fedb_wide_shadow(const float* src, void* dst, int64_t n_floats, void* stream) {
  FEDB_REQUIRE(n_floats % 4 == 0, "arena size must be a multiple of 4 floats");
  const size_t n4 = (size_t)n_floats / 4;
  const int blocks = (int)((n4 + 255) / 256 < 148 * 16 ? (n4 + 255) / 256 : 148 * 16);
  k_wide_shadow<<<blocks, 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, n4);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}
extern "C" int fedb_wide_head_fwd(const FedbGroup* groups, int32_t n_groups, const float* params, const void* A, int32_t B, int32_t d,
                                  const int32_t* y, const int32_t* idx, float* dz, float* loss_rows, int32_t* conf, int32_t* step,
                                  const FedbFedState* fed_state, int32_t eval_mode, void* stream) {
  FEDB_REQUIRE(d % 8 == 0, "hidden width must be a multiple of 8");
  dim3 grid((B + 7) / 8, n_groups);
  k_wide_head_fwd<<<grid, 256, 0, (cudaStream_t)stream>>>(groups, params, (const __nv_bfloat16*)A, B, d, y, idx, dz, loss_rows, conf,
                                                          step, fed_state, eval_mode);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}
extern "C" int fedb_wide_head_bwd(const FedbGroup* groups, int32_t n_groups, float* params, float* m, float* v, void* shadow,
                                  const void* A, int32_t B, int32_t d, const float* dz, const float* loss_rows, void* dZ,
                                  const int32_t* step, const FedbFedState* fed_state, float* loss_out, void* stream) {
  dim3 grid((d + 127) / 128, n_groups);
  k_wide_head_bwd<<<grid, 256, 0, (cudaStream_t)stream>>>(groups, params, m, v, (__nv_bfloat16*)shadow, (const __nv_bfloat16*)A, B, d, dz,
                                                          loss_rows, (__nv_bfloat16*)dZ, step, fed_state, loss_out);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}
extern "C" int fedb_wide_bias_opt(const FedbGroup* groups, int32_t n_groups, float* params, float* m, float* v, const void* dZ,
                                  int32_t B, int32_t d, int32_t layer, const int32_t* step, const FedbFedState* fed_state,
                                  void* stream) {
  dim3 grid((d + 255) / 256, n_groups);
  k_wide_bias_opt<<<grid, 256, 0, (cudaStream_t)stream>>>(groups, params, m, v, (const __nv_bfloat16*)dZ, B, d, layer, step, fed_state);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}
