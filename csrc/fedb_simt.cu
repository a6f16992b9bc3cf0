// fedb200 SIMT fp32 kernels: the shape-generic, bit-reproducible implementation of every hot
// step of a federated round.  One lane per data row (32-row tiles), activations feature-major
// in shared memory, weights read warp-uniformly through L1 (they are small and shared by all
// tiles of a client).  These kernels are (a) the numerics oracle for the tcgen05 kernels,
// (b) the path for shapes the tensor-core kernels do not cover, and (c) what the sklearn-mode
// trainer (minibatch 200) and the grid search run on.
//
// Reference behaviour being provided (see include/fedb200.h for the citation tags):
//   k_fwd_bwd     C:68-71 (forward, CrossEntropyLoss, backward); SK:301-393 (_backprop)
//   k_reduce_opt  C:72-73 (Adam.step, StepLR.step); SKO Adam/SGD updates; SK:235 L2
//   k_eval        C:75-83, S:104 (predict) ; counts replace the sklearn metric calls C:85-90
//   k_fedavg      C:101-120, S:108-122, H:24-46 (aggregation) + C:165-192 (global metrics,
//                 early stop)
#include <cooperative_groups.h>

#include <mutex>

#include "common.cuh"
#include <cuda_bf16.h>

namespace fedb {

thread_local char g_err[512] = "";

constexpr int LDS = FEDB_TILE_ROWS + 1;  // 33: feature-major rows, conflict-free both ways
constexpr int NTHREADS = 256;
// The tile kernels (forward/backward, evaluation) are bound by shared-memory latency: 16 warps hide it better than 8.
// (Results do not depend on the block size: every output element is computed by one thread in a fixed order.)
constexpr int TILE_THREADS = 512;

struct SmemPlan {
  int a_off[FEDB_MAX_LAYERS + 1];  // float offsets of activation buffers a[0..L]
  int dz_off[2];
  int misc_off;
  int total_floats;
};

__host__ __device__ inline SmemPlan smem_plan(const int32_t* dims, int L) {
  SmemPlan p;
  int off = 0;
  for (int l = 0; l <= FEDB_MAX_LAYERS; ++l) {
    p.a_off[l] = off;
    if (l <= L) off += ru4(dims[l]) * LDS;
  }
  // The backward pass ping-pongs between two dz buffers: buffer 0 holds dz of layers L, L-2, ..., buffer 1 that of layers
  // L-1, L-3, ...  Each is sized for the widths it actually holds (for 14-50-400-1 that is 52 + 400 rows instead of 2 x 400,
  // which is what lets that model's 86 KB of weights be staged next to the activations).
  int w0 = 4, w1 = 4;
  for (int l = L, par = 0; l >= 1; --l, par ^= 1) {
    const int w = ru4(dims[l]);
    if (par == 0) { if (w > w0) w0 = w; } else { if (w > w1) w1 = w; }
  }
  p.dz_off[0] = off; off += w0 * LDS;
  p.dz_off[1] = off; off += w1 * LDS;
  p.misc_off = off; off += 64;
  p.total_floats = off;
  return p;
}

// ---------------------------------------------------------------------------------------
// device building blocks
// ---------------------------------------------------------------------------------------

// z[j][r] = act(b[j] + sum_k a_in[k][r] * W[j][k]);  W row-major [d_out][ld], pad cols zero.
// Rows j in [d_out, ru4(d_out)) of a_out are written as 0 (they are the next layer's K pad).
// WS = where the weights are read from: 1 = staged in shared memory (plain loads), 0 = global memory through the
// read-only path (__ldg: the weights do not change during the kernel), 2 = global memory, L2-coherent (__ldcg: the
// persistent fit kernel, whose weights are rewritten by other CTAs between mini-batches).
template <int WS>
__device__ __forceinline__ float4 ldw4(const float* p) {
  if constexpr (WS == 1) return *reinterpret_cast<const float4*>(p);
  else if constexpr (WS == 2) return __ldcg(reinterpret_cast<const float4*>(p));
  else return __ldg(reinterpret_cast<const float4*>(p));
}
template <int WS>
__device__ __forceinline__ float ldw1(const float* p) {
  if constexpr (WS == 1) return *p;
  else if constexpr (WS == 2) return __ldcg(p);
  else return __ldg(p);
}

template <int WS>
__device__ void layer_forward(const float* __restrict__ W, const float* __restrict__ b, int ld, int d_out,
                              const float* a_in, float* a_out, int act, bool apply_act) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int njb = (d_out + 3) >> 2;
  for (int jb = warp; jb < njb; jb += nwarps) {
    const int j0 = jb * 4;
    const float* w[4];
    float acc[4];
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      int j = min(j0 + jj, d_out - 1);
      w[jj] = W + (size_t)j * ld;
      acc[jj] = ldw1<WS>(b + j);
    }
#pragma unroll 2
    for (int k = 0; k < ld; k += 4) {
      const float a0 = a_in[(k + 0) * LDS + lane], a1 = a_in[(k + 1) * LDS + lane];
      const float a2 = a_in[(k + 2) * LDS + lane], a3 = a_in[(k + 3) * LDS + lane];
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const float4 q = ldw4<WS>(w[jj] + k);
        acc[jj] = fmaf(a0, q.x, acc[jj]);
        acc[jj] = fmaf(a1, q.y, acc[jj]);
        acc[jj] = fmaf(a2, q.z, acc[jj]);
        acc[jj] = fmaf(a3, q.w, acc[jj]);
      }
    }
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const int j = j0 + jj;
      float v = 0.f;
      if (j < d_out) v = apply_act ? act_fwd(acc[jj], act) : acc[jj];
      a_out[j * LDS + lane] = v;
    }
  }
}

// dW[j][k] = sum_r dz[j][r] a[k][r]  (tile partial), db[j] = sum_r dz[j][r]
// Wide inputs (ld >= 160; measured: it loses on ld = 100, where a layer has 50 such tasks for 16 warps): a WARP takes 8
// outputs x 128 inputs, lane l the inputs l, l + 32, l + 64, l + 96 of the group, so the
// activation loads of a warp hit 32 different banks (the 4-consecutive-inputs-per-thread mapping below puts them on 8: a
// 4-way conflict on the loads that feed two thirds of the FMAs) and the dz loads are broadcasts; 12 shared-memory wavefronts
// per 32 FMA instructions instead of 20 per 16.  Each dW element still adds its rows in order: bit-identical.
__device__ void layer_wgrad_wide(const float* dz, const float* a, int d_in_ld, int d_out, float* __restrict__ gW) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int njb = (d_out + 7) >> 3, nkg = (d_in_ld + 127) >> 7;
  for (int task = warp; task < njb * nkg; task += nwarps) {
    const int jb = task / nkg, kg = task - jb * nkg;
    const int j0 = jb * 8, kbase = kg * 128 + lane;
    const float* dp[8];
    const float* xp[4];
#pragma unroll
    for (int i = 0; i < 8; ++i) dp[i] = dz + min(j0 + i, d_out - 1) * LDS;
#pragma unroll
    for (int c = 0; c < 4; ++c) xp[c] = a + min(kbase + 32 * c, d_in_ld - 1) * LDS;
    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
#pragma unroll 2
    for (int r = 0; r < FEDB_TILE_ROWS; ++r) {
      float x[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) x[c] = xp[c][r];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float d = dp[i][r];
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[i][c] = fmaf(d, x[c], acc[i][c]);
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (j0 + i < d_out) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          if (kbase + 32 * c < d_in_ld) gW[(size_t)(j0 + i) * d_in_ld + kbase + 32 * c] = acc[i][c];
      }
  }
}

__device__ void layer_wgrad(const float* dz, const float* a, int d_in_ld, int d_out, float* __restrict__ gW,
                            float* __restrict__ gb) {
  const int njb = (d_out + 3) >> 2, nkb = d_in_ld >> 2;
  if (d_in_ld >= 160) layer_wgrad_wide(dz, a, d_in_ld, d_out, gW);
  else
  for (int blk = threadIdx.x; blk < njb * nkb; blk += blockDim.x) {
    const int jb = blk / nkb, kb = blk - jb * nkb;
    const int j0 = jb * 4, k0 = kb * 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 4
    for (int r = 0; r < FEDB_TILE_ROWS; ++r) {
      float d[4], x[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        d[i] = dz[(j0 + i) * LDS + r];
        x[i] = a[(k0 + i) * LDS + r];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(d[i], x[j], acc[i][j]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (j0 + i < d_out)
        *reinterpret_cast<float4*>(gW + (size_t)(j0 + i) * d_in_ld + k0) =
            make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  }
  for (int j = threadIdx.x; j < ru4(d_out); j += blockDim.x) {
    float s = 0.f;
    if (j < d_out) {
#pragma unroll 8
      for (int r = 0; r < FEDB_TILE_ROWS; ++r) s += dz[j * LDS + r];
    }
    gb[j] = s;
  }
}

// dz_prev[k][r] = act'(a_l[k][r]) * sum_j dz[j][r] W[j][k]
template <int WS>
__device__ void layer_dgrad(const float* __restrict__ W, int ld, int d_in, int d_out, const float* dz,
                            const float* a_l, float* dz_prev, int act) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int nkb = ld >> 2;
  for (int kb = warp; kb < nkb; kb += nwarps) {
    const int k0 = kb * 4;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
    for (int j = 0; j < d_out; ++j) {
      const float d = dz[j * LDS + lane];
      const float4 q = ldw4<WS>(W + (size_t)j * ld + k0);
      acc[0] = fmaf(d, q.x, acc[0]);
      acc[1] = fmaf(d, q.y, acc[1]);
      acc[2] = fmaf(d, q.z, acc[2]);
      acc[3] = fmaf(d, q.w, acc[3]);
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const int k = k0 + kk;
      const float av = a_l[k * LDS + lane];
      dz_prev[k * LDS + lane] = (k < d_in) ? acc[kk] * act_bwd(av, act) : 0.f;
    }
  }
}

// ---- models whose weights do not fit next to the activations --------------------------------------------------------------
// Reading such a layer's weights straight from global memory makes every warp wait an L2 round trip per float4 (the
// activations' carve-out leaves L1 too small to hold a 160-320 KB matrix): the (400, 200) / (200, 400) / (100, 400) corners of
// the grid ran 5-10x slower per FLOP than the shapes that fit.  Instead the CTA streams the matrix through whatever shared
// memory is left (launches ask for the full 227 KB): a block of consecutive rows is copied with every thread's loads in flight
// (coalesced, one L2 round trip per block), the layer loops read it at shared-memory latency, next block.  The order in which
// each output accumulates its terms is the one of the plain loops, so the results are bit-identical.
// Asynchronous 16-byte copies global -> shared (through L2, so also right for weights other CTAs rewrote earlier in the
// kernel): block i + 1 of a matrix is in flight while the warps work on block i.
__device__ __forceinline__ void cp_async16(float* dst, const float* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_but_one() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// rows [0, n_rows) x columns [k0, k0 + kc) of a row-major matrix -> dst[row * kc + col] (kc a multiple of 4)
__device__ __forceinline__ void stage_cols_async(const float* __restrict__ W, int ld, int n_rows, int k0, int kc, float* dst) {
  const int q = kc >> 2;
  for (int e = threadIdx.x; e < n_rows * q; e += blockDim.x) {
    const int row = e / q, c4 = e - row * q;
    cp_async16(dst + row * kc + 4 * c4, W + (size_t)row * ld + k0 + 4 * c4);
  }
  cp_async_commit();
}
// n_floats consecutive floats (whole rows) -> dst
__device__ __forceinline__ void stage_rows_async(const float* __restrict__ W, int n_floats, float* dst) {
  for (int i = threadIdx.x * 4; i < n_floats; i += blockDim.x * 4) cp_async16(dst + i, W + i);
  cp_async_commit();
}

// Forward with streamed weights: the matrix goes through shared memory in blocks of COLUMNS (all outputs x kc inputs), each
// warp keeps the accumulators of its NT tasks of 8 outputs across the blocks (every warp has work in every block; blocks of
// rows left 11 of 16 warps idle on a 200 x 400 layer), inputs in ascending order as in layer_forward: bit-identical.
constexpr int FWD_TMAX = 4;   // tasks of 8 outputs per warp: d_out <= 8 * 4 * (warps per CTA)
template <int NT>
__device__ __forceinline__ void fwd_cols_nt(float (&acc)[FWD_TMAX][8], const float* a_in, const float* wb, int kc, int k0, int lane,
                                            int task0, int tstride) {
  // (the block holds ru8(d_out) rows: a last, partial task reads unstaged rows whose accumulators are never stored)
  const float* a = a_in + k0 * LDS + lane;
  const float* w0 = wb + task0 * 8 * kc;
  const int ts = tstride * 8 * kc;
#pragma unroll 2
  for (int k = 0; k < kc; k += 4) {
    const float a0 = a[(k + 0) * LDS], a1 = a[(k + 1) * LDS], a2 = a[(k + 2) * LDS], a3 = a[(k + 3) * LDS];
#pragma unroll
    for (int t = 0; t < NT; ++t)
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) {
        const float4 q = *reinterpret_cast<const float4*>(w0 + t * ts + jj * kc + k);
        acc[t][jj] = fmaf(a0, q.x, acc[t][jj]);
        acc[t][jj] = fmaf(a1, q.y, acc[t][jj]);
        acc[t][jj] = fmaf(a2, q.z, acc[t][jj]);
        acc[t][jj] = fmaf(a3, q.w, acc[t][jj]);
      }
  }
}

template <int WS>
__device__ void layer_forward_chunked(const float* __restrict__ W, const float* __restrict__ b, int ld, int d_out, const float* a_in,
                                      float* a_out, int act, bool apply_act, float* wbuf, int kc_max) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int njb = (d_out + 7) >> 3;
  const int nt = njb > warp ? (njb - warp + nwarps - 1) / nwarps : 0;   // warp-uniform, <= FWD_TMAX
  float acc[FWD_TMAX][8];
#pragma unroll
  for (int t = 0; t < FWD_TMAX; ++t)
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) acc[t][jj] = ldw1<WS>(b + min((warp + t * nwarps) * 8 + jj, d_out - 1));
  const int nblk = (ld + kc_max - 1) / kc_max;
  float* const buf0 = wbuf;
  float* const buf1 = wbuf + ((d_out + 7) & ~7) * kc_max;
  stage_cols_async(W, ld, d_out, 0, min(kc_max, ld), buf0);
  for (int c = 0; c < nblk; ++c) {
    const int k0 = c * kc_max, kc = min(kc_max, ld - k0);
    if (c + 1 < nblk) {
      stage_cols_async(W, ld, d_out, k0 + kc_max, min(kc_max, ld - k0 - kc_max), (c & 1) ? buf0 : buf1);
      cp_async_wait_but_one();
    } else {
      cp_async_wait_all();
    }
    __syncthreads();
    const float* wb = (c & 1) ? buf1 : buf0;
    switch (nt) {
      case 0: break;
      case 1: fwd_cols_nt<1>(acc, a_in, wb, kc, k0, lane, warp, nwarps); break;
      case 2: fwd_cols_nt<2>(acc, a_in, wb, kc, k0, lane, warp, nwarps); break;
      case 3: fwd_cols_nt<3>(acc, a_in, wb, kc, k0, lane, warp, nwarps); break;
      default: fwd_cols_nt<4>(acc, a_in, wb, kc, k0, lane, warp, nwarps); break;
    }
    __syncthreads();
  }
#pragma unroll
  for (int t = 0; t < FWD_TMAX; ++t)
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) {
      const int j = (warp + t * nwarps) * 8 + jj;
      if (j < ru4(d_out)) {                      // (also the zero rows up to ru4(d_out): the next layer's K pad)
        float v = 0.f;
        if (j < d_out) v = apply_act ? act_fwd(acc[t][jj], act) : acc[t][jj];
        a_out[j * LDS + lane] = v;
      }
    }
}

constexpr int DGRAD_TMAX = 8;   // column blocks of 4 a warp keeps accumulators for: ld <= 4 * 8 * (warps per CTA)
// NT = how many column blocks THIS warp owns (kb = warp + t * nwarps < nkb), a compile-time constant so that the inner loop is
// straight-line code: one dz load, NT broadcast float4 loads, 4 NT FMAs per weight row (with the test inside the loop the
// kernel executed three instructions per FMA).
template <int NT>
__device__ __forceinline__ void dgrad_rows_nt(float (&acc)[DGRAD_TMAX][4], const float* dp, const float* wr, int nr, int ld, int tstride) {
#pragma unroll 4
  for (int j = 0; j < nr; ++j) {
    const float d = dp[j * LDS];
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const float4 q = *reinterpret_cast<const float4*>(wr + j * ld + t * tstride);
      acc[t][0] = fmaf(d, q.x, acc[t][0]);
      acc[t][1] = fmaf(d, q.y, acc[t][1]);
      acc[t][2] = fmaf(d, q.z, acc[t][2]);
      acc[t][3] = fmaf(d, q.w, acc[t][3]);
    }
  }
}

template <int WS>
__device__ void layer_dgrad_chunked(const float* __restrict__ W, int ld, int d_in, int d_out, const float* dz, const float* a_l,
                                    float* dz_prev, int act, float* wbuf, int rpc) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int nkb = ld >> 2;
  const int nt = nkb > warp ? (nkb - warp + nwarps - 1) / nwarps : 0;   // warp-uniform, <= DGRAD_TMAX
  float acc[DGRAD_TMAX][4];
#pragma unroll
  for (int t = 0; t < DGRAD_TMAX; ++t) acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f;
  const int nblk = (d_out + rpc - 1) / rpc;     // blocks of rpc rows, two buffers: block c + 1 loads while block c is used
  float* const buf0 = wbuf;
  float* const buf1 = wbuf + rpc * ld;
  stage_rows_async(W, min(rpc, d_out) * ld, buf0);
  for (int c = 0; c < nblk; ++c) {
    const int j0 = c * rpc, nr = min(rpc, d_out - j0);
    if (c + 1 < nblk) {
      stage_rows_async(W + (size_t)(j0 + rpc) * ld, min(rpc, d_out - j0 - rpc) * ld, (c & 1) ? buf0 : buf1);
      cp_async_wait_but_one();
    } else {
      cp_async_wait_all();
    }
    __syncthreads();
    const float* dp = dz + j0 * LDS + lane;
    const float* wr = ((c & 1) ? buf1 : buf0) + 4 * warp;
    switch (nt) {
      case 0: break;
      case 1: dgrad_rows_nt<1>(acc, dp, wr, nr, ld, 4 * nwarps); break;
      case 2: dgrad_rows_nt<2>(acc, dp, wr, nr, ld, 4 * nwarps); break;
      case 3: dgrad_rows_nt<3>(acc, dp, wr, nr, ld, 4 * nwarps); break;
      case 4: dgrad_rows_nt<4>(acc, dp, wr, nr, ld, 4 * nwarps); break;
      case 5: dgrad_rows_nt<5>(acc, dp, wr, nr, ld, 4 * nwarps); break;
      case 6: dgrad_rows_nt<6>(acc, dp, wr, nr, ld, 4 * nwarps); break;
      case 7: dgrad_rows_nt<7>(acc, dp, wr, nr, ld, 4 * nwarps); break;
      default: dgrad_rows_nt<8>(acc, dp, wr, nr, ld, 4 * nwarps); break;
    }
    __syncthreads();
  }
#pragma unroll
  for (int t = 0; t < DGRAD_TMAX; ++t) {
    const int kb = warp + t * nwarps;
    if (kb < nkb) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const int k = kb * 4 + kk;
        const float av = a_l[k * LDS + lane];
        dz_prev[k * LDS + lane] = (k < d_in) ? acc[t][kk] * act_bwd(av, act) : 0.f;
      }
    }
  }
}

// Loads the tile's rows, feature-major, into a0; labels into ylab. Rows >= nvalid replicate
// the last valid row (finite values; their dz is forced to 0 later).
template <bool COH = false>   // COH: the row-index array is rewritten during the kernel (per-epoch shuffle of the fit kernel)
__device__ void load_tile(const float* __restrict__ X, int ldx, const int32_t* __restrict__ y,
                          const int32_t* idx, int64_t base, int nvalid, int ld0, float* a0,
                          int* ylab) {
  const int q4 = ld0 >> 2;
  for (int e = threadIdx.x; e < FEDB_TILE_ROWS * q4; e += blockDim.x) {
    const int r = e / q4, q = e - r * q4;
    const int row = COH ? __ldcg(idx + base + min(r, nvalid - 1)) : __ldg(idx + base + min(r, nvalid - 1));
    const float4 v = __ldg(reinterpret_cast<const float4*>(X + (size_t)row * ldx) + q);
    a0[(4 * q + 0) * LDS + r] = v.x;
    a0[(4 * q + 1) * LDS + r] = v.y;
    a0[(4 * q + 2) * LDS + r] = v.z;
    a0[(4 * q + 3) * LDS + r] = v.w;
  }
  if (threadIdx.x < FEDB_TILE_ROWS) {
    const int row = COH ? __ldcg(idx + base + min((int)threadIdx.x, nvalid - 1)) : __ldg(idx + base + min((int)threadIdx.x, nvalid - 1));
    ylab[threadIdx.x] = __ldg(y + row);
  }
}

// Cooperative, fully coalesced copy of a client's slab into shared memory (one L2 round trip
// with every thread's loads in flight) so that the layer loops read weights at smem latency.
template <bool COH = false>
__device__ void stage_weights(const float* src, float* dst, int n_floats) {
  const float4* s4 = reinterpret_cast<const float4*>(src);
  float4* d4 = reinterpret_cast<float4*>(dst);
  for (int i = threadIdx.x; i < (n_floats >> 2); i += blockDim.x) d4[i] = COH ? __ldcg(s4 + i) : __ldg(s4 + i);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---------------------------------------------------------------------------------------
// K4: init
// ---------------------------------------------------------------------------------------
__global__ void k_init(const FedbGroup* __restrict__ groups, float* __restrict__ params,
                       const int32_t* __restrict__ streams, int kind, uint32_t k0, uint32_t k1) {
  const FedbGroup& G = groups[blockIdx.y];
  const int sid = streams[blockIdx.y];
  if (sid < 0) return;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q * 4 >= G.slab_size) return;
  const SlabLayout lay = slab_layout(G.dims, G.n_layers);
  uint32_t c[4] = {(uint32_t)q, (uint32_t)sid, 0u, 0u};
  philox4x32(c, k0, k1);
  float out[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int e = q * 4 + i;
    float bound = 0.f;
    for (int l = 0; l < G.n_layers; ++l) {
      const int fi = G.dims[l], fo = G.dims[l + 1];
      float s;
      if (kind == FEDB_INIT_TORCH_DEFAULT) s = (float)(1.0 / sqrt((double)fi));
      else if (kind == FEDB_INIT_SKLEARN_GLOROT)
        s = (float)sqrt((G.act == FEDB_ACT_LOGISTIC ? 2.0 : 6.0) / (double)(fi + fo));
      else s = 0.f;
      if (e >= lay.w_off[l] && e < lay.b_off[l]) {
        const int col = (e - lay.w_off[l]) % lay.ld[l];
        if (col < fi) bound = s;
      } else if (e >= lay.b_off[l] && e < lay.b_off[l] + fo) {
        bound = s;
      }
    }
    const float u = (float)(c[i] >> 8) * 5.9604644775390625e-08f;  // 2^-24
    out[i] = __fmul_rn(__fsub_rn(__fmul_rn(2.0f, u), 1.0f), bound);
  }
  *reinterpret_cast<float4*>(params + G.slab_off + (size_t)q * 4) = make_float4(out[0], out[1], out[2], out[3]);
}

// ---------------------------------------------------------------------------------------
// round begin
// ---------------------------------------------------------------------------------------
__global__ void k_round_begin(const FedbFed* __restrict__ feds, FedbFedState* __restrict__ st, int n_feds,
                              int32_t* __restrict__ conf, int n_conf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_feds) {
    if (st[i].stop_pending) st[i].stop = 1;
    if (!st[i].stop) st[i].round += 1;
  }
  for (int k = i; k < n_conf; k += gridDim.x * blockDim.x) conf[k] = 0;
}

// ---------------------------------------------------------------------------------------
// shuffle: keyed Feistel bijection on [0, n) by cycle walking
// ---------------------------------------------------------------------------------------
__global__ void k_shuffle(const FedbGroup* __restrict__ groups, const int32_t* __restrict__ idx_in,
                          int32_t* __restrict__ idx_out, const int32_t* __restrict__ epoch,
                          const int32_t* __restrict__ active, uint32_t s0, uint32_t s1) {
  const int g = blockIdx.y;
  if (active && !active[g]) return;
  const FedbGroup& G = groups[g];
  const uint32_t n = (uint32_t)G.n_rows;
  int bits = 1;
  while ((1u << bits) < n) ++bits;
  const int half = (bits + 1) >> 1;
  // keyed by the client's index inside its federation, not by its position in the packed table: a federation shuffles
  // the same way whichever other federations share the launch (or the GPU)
  const uint32_t ka = s0 ^ (uint32_t)(G.client * 0x9E3779B1u), kb = s1 + (uint32_t)epoch[g];
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    idx_out[G.idx_off + i] = idx_in[G.idx_off + feistel_perm(i, n, half, ka, kb)];
}

// ---------------------------------------------------------------------------------------
// K2 part 1: forward + loss + backward, one 32-row tile per CTA
// ---------------------------------------------------------------------------------------
// One 32-row tile of one mini-batch of group g: forward, loss, backward; the tile's partial gradients + loss go to its record.
// COH: parameters / row indices may have been rewritten by other CTAs earlier in this kernel (persistent fit kernel).
template <bool COH>
__device__ void fwd_bwd_tile(const FedbGroup& G, int g, int tile, const float* params, const float* __restrict__ X, int ldx,
                             const int32_t* __restrict__ y, const int32_t* idx, int batch, float* scratch, int32_t* step,
                             int smem_floats, float* smem) {
  const int n = G.n_rows;
  const int bs = G.batch_size > 0 ? min(G.batch_size, n) : n;
  const int b0 = batch * bs;
  if (b0 >= n) return;
  const int nb = min(bs, n - b0);
  const int r0 = tile * FEDB_TILE_ROWS;
  if (r0 >= nb) return;
  const int nvalid = min(FEDB_TILE_ROWS, nb - r0);
  const int L = G.n_layers;
  const SlabLayout lay = slab_layout(G.dims, L);
  const SmemPlan sp = smem_plan(G.dims, L);
  const bool ws = sp.total_floats + lay.end <= smem_floats;  // weights fit next to the activations
  const int spare = smem_floats - sp.total_floats;           // otherwise: what is left streams the matrices block by block
  float* wbuf = smem + sp.total_floats;
  const float* P = params + G.slab_off;
  float* rec = scratch + G.scratch_off + (size_t)tile * (G.slab_size + FEDB_LOSS_SLOTS);
  int* ylab = reinterpret_cast<int*>(smem + sp.misc_off);
  const int lane = threadIdx.x & 31;

  if (tile == 0 && threadIdx.x == 0) step[g] += 1;  // optimizer step counter (t in Adam)

  if (ws) {
    stage_weights<COH>(P, smem + sp.total_floats, lay.end);
    P = smem + sp.total_floats;
  }
  load_tile<COH>(X, ldx, y, idx, G.idx_off + b0 + r0, nvalid, lay.ld[0], smem + sp.a_off[0], ylab);
  __syncthreads();

  for (int l = 0; l < L; ++l) {
    // columns of this layer's matrix per block: two blocks of (all outputs x kc) fit in the shared memory left over
    const int kc = min(lay.ld[l], (spare / (2 * ((G.dims[l + 1] + 7) & ~7))) & ~3);
    if (ws) layer_forward<1>(P + lay.w_off[l], P + lay.b_off[l], lay.ld[l], G.dims[l + 1], smem + sp.a_off[l],
                                smem + sp.a_off[l + 1], G.act, l + 1 < L);
    else if (kc >= 16 && G.dims[l + 1] <= 8 * FWD_TMAX * (int)(blockDim.x >> 5))
      layer_forward_chunked<COH ? 2 : 0>(P + lay.w_off[l], P + lay.b_off[l], lay.ld[l], G.dims[l + 1], smem + sp.a_off[l],
                                         smem + sp.a_off[l + 1], G.act, l + 1 < L, wbuf, kc);
    else layer_forward<COH ? 2 : 0>(P + lay.w_off[l], P + lay.b_off[l], lay.ld[l], G.dims[l + 1], smem + sp.a_off[l],
                              smem + sp.a_off[l + 1], G.act, l + 1 < L);
    __syncthreads();
  }

  // head: loss and dz of the output layer, warp 0, lane = row
  float* zL = smem + sp.a_off[L];
  float* dz = smem + sp.dz_off[0];
  const float inv_n = 1.f / (float)nb;
  const int C = G.dims[L];
  if (threadIdx.x < 32) {
    float loss = 0.f;
    const bool valid = lane < nvalid;
    const int yl = ylab[lane];
    if (G.head == FEDB_HEAD_SOFTMAX_CE) {
      float mx = -INFINITY;
      for (int c = 0; c < C; ++c) mx = fmaxf(mx, zL[c * LDS + lane]);
      float se = 0.f;
      for (int c = 0; c < C; ++c) se += expf(zL[c * LDS + lane] - mx);
      const float lse = mx + logf(se);
      loss = valid ? (lse - zL[yl * LDS + lane]) : 0.f;
      for (int c = 0; c < ru4(C); ++c) {
        float d = 0.f;
        if (c < C && valid) d = (expf(zL[c * LDS + lane] - lse) - (c == yl ? 1.f : 0.f)) * inv_n;
        dz[c * LDS + lane] = d;
      }
    } else {
      const float z = zL[lane];
      const float p = 1.f / (1.f + expf(-z));
      const float eps = 1.1920929e-07f;  // SKB:273-274 clips probabilities at dtype eps
      const float pc = fminf(fmaxf(p, eps), 1.f - eps);
      const float yt = (float)yl;
      loss = valid ? -(yt * logf(pc) + (1.f - yt) * logf(1.f - pc)) : 0.f;
      dz[lane] = valid ? (p - yt) * inv_n : 0.f;
      for (int c = 1; c < 4; ++c) dz[c * LDS + lane] = 0.f;
    }
    loss = warp_sum(loss);
    if (lane == 0) rec[G.slab_size] = loss;  // sum (not mean) of row losses of this tile
  }
  __syncthreads();

  int cur = 0;
  for (int l = L - 1; l >= 0; --l) {
    const float* dzc = smem + sp.dz_off[cur];
    layer_wgrad(dzc, smem + sp.a_off[l], lay.ld[l], G.dims[l + 1], rec + lay.w_off[l], rec + lay.b_off[l]);
    if (l > 0) {
      const int rpc = min(G.dims[l + 1], spare / (2 * lay.ld[l]));   // rows per block, two blocks resident
      if (ws) layer_dgrad<1>(P + lay.w_off[l], lay.ld[l], G.dims[l], G.dims[l + 1], dzc, smem + sp.a_off[l],
                                smem + sp.dz_off[cur ^ 1], G.act);
      else if (rpc >= 8 && (lay.ld[l] >> 2) <= DGRAD_TMAX * (int)(blockDim.x >> 5))
        layer_dgrad_chunked<COH ? 2 : 0>(P + lay.w_off[l], lay.ld[l], G.dims[l], G.dims[l + 1], dzc, smem + sp.a_off[l],
                                         smem + sp.dz_off[cur ^ 1], G.act, wbuf, rpc);
      else layer_dgrad<COH ? 2 : 0>(P + lay.w_off[l], lay.ld[l], G.dims[l], G.dims[l + 1], dzc, smem + sp.a_off[l],
                              smem + sp.dz_off[cur ^ 1], G.act);
    }
    __syncthreads();
    cur ^= 1;
  }
}

__global__ void __launch_bounds__(TILE_THREADS)
k_fwd_bwd(const FedbGroup* __restrict__ groups, const FedbTile* __restrict__ tiles, const float* __restrict__ params,
          const float* __restrict__ X, int ldx, const int32_t* __restrict__ y, const int32_t* __restrict__ idx,
          int batch, float* __restrict__ scratch, int32_t* __restrict__ step, const int32_t* __restrict__ active,
          const FedbFedState* __restrict__ fed_state, int smem_floats) {
  extern __shared__ __align__(16) float smem[];
  const FedbTile T = tiles[blockIdx.x];
  const int g = T.owner;
  const FedbGroup& G = groups[g];
  if (active && !active[g]) return;
  if (fed_state && fed_state[G.fed].stop) return;
  fwd_bwd_tile<false>(G, g, T.tile, params, X, ldx, y, idx, batch, scratch, step, smem_floats, smem);
}

// ---------------------------------------------------------------------------------------
// K2 part 2: fixed-order reduction over tiles + L2 + optimizer
// ---------------------------------------------------------------------------------------
// Block `vblock` of `nvblocks` (256 slab float4s each) of group g's reduction + optimizer step for mini-batch `batch`.
// COH: the gradient records and the step counter were written by other CTAs earlier in this kernel.
template <bool COH>
__device__ void reduce_opt_block(const FedbGroup& G, int g, int vblock, int nvblocks, float* params, float* m, float* v,
                                 const float* scratch, int batch, const int32_t* step, float* loss_out, double* batch_stats,
                                 int max_batches) {
  const int n = G.n_rows;
  const int bs = G.batch_size > 0 ? min(G.batch_size, n) : n;
  const int b0 = batch * bs;
  if (b0 >= n) return;
  const int nb = min(bs, n - b0);
  const int ntiles = (nb + FEDB_TILE_ROWS - 1) / FEDB_TILE_ROWS;
  const int P = G.slab_size, stride = P + FEDB_LOSS_SLOTS;
  const SlabLayout lay = slab_layout(G.dims, G.n_layers);
  const float* rec = scratch + G.scratch_off;

  __shared__ StepScalars sc_s;
  __shared__ float red[NTHREADS / 32];
  if (threadIdx.x == 0) sc_s = step_scalars(G, COH ? __ldcg(step + g) : step[g]);
  __syncthreads();
  const StepScalars sc = sc_s;

  float sumsq = 0.f;
  const int q = vblock * NTHREADS + threadIdx.x;   // the 256-float4 partition is part of the result (per-block sums of squares)
  if (threadIdx.x < NTHREADS && q * 4 < lay.end) {
    const int e = q * 4;
    bool is_w = false;
    for (int l = 0; l < G.n_layers; ++l) is_w |= (e >= lay.w_off[l] && e < lay.b_off[l]);
    float4 gsum = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int t = 0; t < ntiles; ++t) {
      const float4 p = COH ? __ldcg(reinterpret_cast<const float4*>(rec + (size_t)t * stride + e))
                           : *reinterpret_cast<const float4*>(rec + (size_t)t * stride + e);
      gsum.x += p.x; gsum.y += p.y; gsum.z += p.z; gsum.w += p.w;
    }
    float4 w = *reinterpret_cast<float4*>(params + G.slab_off + e);
    if (is_w) {
      sumsq = w.x * w.x + w.y * w.y + w.z * w.z + w.w * w.w;
      if (G.alpha != 0.f) {
        const float a = G.alpha / (float)nb;  // SK:235: (a^T delta + alpha W)/n
        gsum.x = fmaf(a, w.x, gsum.x); gsum.y = fmaf(a, w.y, gsum.y);
        gsum.z = fmaf(a, w.z, gsum.z); gsum.w = fmaf(a, w.w, gsum.w);
      }
    }
    float4 mm = make_float4(0.f, 0.f, 0.f, 0.f), vv = mm;
    if (m) mm = *reinterpret_cast<float4*>(m + G.slab_off + e);
    if (v && G.opt != FEDB_OPT_SGD) vv = *reinterpret_cast<float4*>(v + G.slab_off + e);
    w.x = opt_update(G, sc, w.x, gsum.x, mm.x, vv.x);
    w.y = opt_update(G, sc, w.y, gsum.y, mm.y, vv.y);
    w.z = opt_update(G, sc, w.z, gsum.z, mm.z, vv.z);
    w.w = opt_update(G, sc, w.w, gsum.w, mm.w, vv.w);
    *reinterpret_cast<float4*>(params + G.slab_off + e) = w;
    if (m) *reinterpret_cast<float4*>(m + G.slab_off + e) = mm;
    if (v && G.opt != FEDB_OPT_SGD) *reinterpret_cast<float4*>(v + G.slab_off + e) = vv;
  }

  // loss + sum of squares bookkeeping (deterministic: fixed tree, fixed block order)
  if (batch_stats) {
    sumsq = warp_sum(sumsq);
    if ((threadIdx.x & 31) == 0 && threadIdx.x < NTHREADS) red[threadIdx.x >> 5] = sumsq;
    __syncthreads();
    if (threadIdx.x == 0) {
      float s = 0.f;
      for (int i = 0; i < NTHREADS / 32; ++i) s += red[i];
      // per (group, batch): slot 0 = data loss sum, slots 1.. = per-block sumsq
      double* bsrow = batch_stats + ((size_t)g * max_batches + batch) * (2 + nvblocks);
      bsrow[2 + vblock] = (double)s;
    }
  }
  if (vblock == 0 && threadIdx.x == 0) {
    float ls = 0.f;
    for (int t = 0; t < ntiles; ++t) ls += COH ? __ldcg(rec + (size_t)t * stride + P) : rec[(size_t)t * stride + P];
    if (loss_out) loss_out[g] = ls / (float)nb;
    if (batch_stats) {
      double* bsrow = batch_stats + ((size_t)g * max_batches + batch) * (2 + nvblocks);
      bsrow[0] = (double)ls;
      bsrow[1] = (double)nb;
    }
  }
}

__global__ void __launch_bounds__(NTHREADS)
k_reduce_opt(const FedbGroup* __restrict__ groups, float* __restrict__ params, float* __restrict__ m,
             float* __restrict__ v, const float* __restrict__ scratch, int batch, const int32_t* __restrict__ step,
             const int32_t* __restrict__ active, const FedbFedState* __restrict__ fed_state,
             float* __restrict__ loss_out, double* __restrict__ batch_stats, int max_batches) {
  const int g = blockIdx.y;
  const FedbGroup& G = groups[g];
  if (active && !active[g]) return;
  if (fed_state && fed_state[G.fed].stop) return;
  reduce_opt_block<false>(G, g, blockIdx.x, gridDim.x, params, m, v, scratch, batch, step, loss_out, batch_stats, max_batches);
}

// ---------------------------------------------------------------------------------------
// sklearn epoch end: loss curve + tol / n_iter_no_change (SK:750-777)
// ---------------------------------------------------------------------------------------
// COH: batch_stats rows were written by other CTAs earlier in this kernel
template <bool COH>
__device__ void epoch_end_group(const FedbGroup& G, int g, const double* batch_stats, int max_batches, int nblk, double tol,
                                int n_iter_no_change, int max_iter, int32_t* active, int32_t* epoch, int32_t* no_improve,
                                double* best_loss, double* last_loss) {
  const int n = G.n_rows;
  const int bs = G.batch_size > 0 ? min(G.batch_size, n) : n;
  const int nbatches = (n + bs - 1) / bs;
  double accumulated = 0.0;
  for (int b = 0; b < nbatches; ++b) {
    const double* row = batch_stats + ((size_t)g * max_batches + b) * (2 + nblk);
    double ss = 0.0;
    for (int k = 0; k < nblk; ++k) ss += COH ? __ldcg(row + 2 + k) : row[2 + k];
    // batch_loss = mean logloss + 0.5*alpha*||W||^2 / n_b ; accumulated += batch_loss * n_b
    accumulated += (COH ? __ldcg(row) : row[0]) + 0.5 * (double)G.alpha * ss;
  }
  const double loss = accumulated / (double)n;
  last_loss[g] = loss;
  epoch[g] += 1;
  if (loss > best_loss[g] - tol) no_improve[g] += 1; else no_improve[g] = 0;
  if (loss < best_loss[g]) best_loss[g] = loss;
  if (no_improve[g] > n_iter_no_change || epoch[g] >= max_iter) active[g] = 0;
}

__global__ void k_epoch_end(const FedbGroup* __restrict__ groups, int n_groups, const double* __restrict__ batch_stats,
                            int max_batches, int nblk, double tol, int n_iter_no_change, int max_iter,
                            int32_t* __restrict__ active, int32_t* __restrict__ epoch, int32_t* __restrict__ no_improve,
                            double* __restrict__ best_loss, double* __restrict__ last_loss) {
  // A WARP per client: the lanes fetch 128 block sums at a time (coalesced, all in flight) and every lane adds them in
  // block order through shuffles — the same sequence of double additions as epoch_end_group, so the loss curve is
  // bit-identical, without one thread's chain of ~860 dependent global loads (86 us per epoch for an 86 K-parameter model).
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (g >= n_groups || !active[g]) return;     // warp-uniform
  const FedbGroup& G = groups[g];
  const int n = G.n_rows;
  const int bs = G.batch_size > 0 ? min(G.batch_size, n) : n;
  const int nbatches = (n + bs - 1) / bs;
  double accumulated = 0.0;
  for (int b = 0; b < nbatches; ++b) {
    const double* row = batch_stats + ((size_t)g * max_batches + b) * (2 + nblk);
    double ss = 0.0;
    for (int k0 = 0; k0 < nblk; k0 += 128) {
      double v[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) v[c] = (k0 + 32 * c + lane < nblk) ? row[2 + k0 + 32 * c + lane] : 0.0;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int cnt = min(32, nblk - k0 - 32 * c);
        for (int i = 0; i < cnt; ++i) ss += __shfl_sync(0xffffffffu, v[c], i);
      }
    }
    accumulated += row[0] + 0.5 * (double)G.alpha * ss;
  }
  if (lane == 0) {
    const double loss = accumulated / (double)n;
    last_loss[g] = loss;
    epoch[g] += 1;
    if (loss > best_loss[g] - tol) no_improve[g] += 1; else no_improve[g] = 0;
    if (loss < best_loss[g]) best_loss[g] = loss;
    if (no_improve[g] > n_iter_no_change || epoch[g] >= max_iter) active[g] = 0;
  }
}

// ---------------------------------------------------------------------------------------
// K2b, persistent form: a client's whole local fit (SK:703-777) in ONE launch.  A thread-block cluster owns a client and
// loops over epochs and mini-batches on the device: per-epoch shuffle, one tile of the mini-batch per CTA (fwd_bwd_tile),
// cluster barrier, fixed-order reduction + optimizer over the slab split in the same 256-float4 blocks k_reduce_opt uses
// (so the per-block sums of squares, and with them the loss curve and the stopping epoch, are bit-identical to the
// launch-per-mini-batch path), cluster barrier, and at the end of the epoch the tol / n_iter_no_change rule.  Clients are
// independent until the aggregation, so clusters never wait for each other and need not be co-resident.  Data another CTA
// wrote earlier in the kernel (weights, row order, gradient records, statistics) is read with L2-coherent loads.
// ---------------------------------------------------------------------------------------
struct FitArgs {
  const FedbGroup* groups;
  float *params, *m, *v, *scratch, *loss_out;
  const float* X;
  const int32_t *y, *idx_in;
  int32_t *idx_epoch, *step, *active, *epoch, *no_improve;
  const FedbFedState* fed_state;
  double *batch_stats, *best_loss, *last_loss;
  double tol;
  int ldx, n_batches, max_batches, nvblocks, n_iter_no_change, max_iter, max_epochs, shuffle, smem_floats;
  uint32_t s0, s1;
};

__global__ void __launch_bounds__(TILE_THREADS, 1) k_fit_cluster(const FitArgs a) {
  extern __shared__ __align__(16) float smem[];
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const int CS = (int)cluster.num_blocks(), q = (int)cluster.block_rank();
  const int g = blockIdx.x / CS;
  const FedbGroup& G = a.groups[g];
  if (a.fed_state && a.fed_state[G.fed].stop) return;      // whole cluster: the federation has stopped
  const int n = G.n_rows;
  const int bs = G.batch_size > 0 ? min(G.batch_size, n) : n;
  const int ntiles = (bs + FEDB_TILE_ROWS - 1) / FEDB_TILE_ROWS;
  const int32_t* idx = a.shuffle ? a.idx_epoch : a.idx_in;
  int bits = 1;
  while ((1u << bits) < (uint32_t)n) ++bits;
  const int half = (bits + 1) >> 1;
  for (int ep = 0; ep < a.max_epochs; ++ep) {
    if (!__ldcg(a.active + g)) break;                       // same value in every CTA: written before the last cluster barrier
    if (a.shuffle) {
      const uint32_t ka = a.s0 ^ (uint32_t)(G.client * 0x9E3779B1u), kb = a.s1 + (uint32_t)__ldcg(a.epoch + g);
      for (uint32_t i = q * blockDim.x + threadIdx.x; i < (uint32_t)n; i += CS * blockDim.x)
        a.idx_epoch[G.idx_off + i] = a.idx_in[G.idx_off + feistel_perm(i, n, half, ka, kb)];
      __threadfence();
      cluster.sync();
    }
    for (int b = 0; b < a.n_batches; ++b) {
      for (int tile = q; tile < ntiles; tile += CS) {
        fwd_bwd_tile<true>(G, g, tile, a.params, a.X, a.ldx, a.y, idx, b, a.scratch, a.step, a.smem_floats, smem);
        __syncthreads();
      }
      __threadfence();
      cluster.sync();
      for (int vb = q; vb < a.nvblocks; vb += CS) {
        reduce_opt_block<true>(G, g, vb, a.nvblocks, a.params, a.m, a.v, a.scratch, b, a.step, a.loss_out, a.batch_stats, a.max_batches);
        __syncthreads();
      }
      __threadfence();
      cluster.sync();
    }
    if (q == 0 && threadIdx.x == 0)
      epoch_end_group<true>(G, g, a.batch_stats, a.max_batches, a.nvblocks, a.tol, a.n_iter_no_change, a.max_iter, a.active, a.epoch,
                            a.no_improve, a.best_loss, a.last_loss);
    __threadfence();
    cluster.sync();
  }
}

// ---------------------------------------------------------------------------------------
// K3: eval
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TILE_THREADS)
k_eval(const FedbGroup* __restrict__ groups, const FedbEvalJob* __restrict__ jobs, const FedbTile* __restrict__ tiles,
       const float* __restrict__ params, const float* __restrict__ global, const float* __restrict__ X, int ldx,
       const int32_t* __restrict__ y, const int32_t* __restrict__ idx, int32_t* __restrict__ conf,
       const FedbFedState* __restrict__ fed_state, int smem_floats) {
  extern __shared__ __align__(16) float smem[];
  const FedbTile T = tiles[blockIdx.x];
  const FedbEvalJob J = jobs[T.owner];
  const FedbGroup& G = groups[J.group];
  if (fed_state && fed_state[G.fed].stop) return;
  const int r0 = T.tile * FEDB_TILE_ROWS;
  if (r0 >= J.n_rows) return;
  const int nvalid = min(FEDB_TILE_ROWS, J.n_rows - r0);
  const int L = G.n_layers;
  const SlabLayout lay = slab_layout(G.dims, L);
  const SmemPlan sp = smem_plan(G.dims, L);
  const float* P = (J.from_global ? global : params) + J.slab_off;
  const bool ws = sp.total_floats + lay.end <= smem_floats;
  int* ylab = reinterpret_cast<int*>(smem + sp.misc_off);
  if (ws) {
    stage_weights(P, smem + sp.total_floats, lay.end);
    P = smem + sp.total_floats;
  }
  load_tile(X, ldx, y, idx, J.idx_off + r0, nvalid, lay.ld[0], smem + sp.a_off[0], ylab);
  __syncthreads();
  for (int l = 0; l < L; ++l) {
    const int kc = min(lay.ld[l], ((smem_floats - sp.total_floats) / (2 * ((G.dims[l + 1] + 7) & ~7))) & ~3);
    if (ws) layer_forward<true>(P + lay.w_off[l], P + lay.b_off[l], lay.ld[l], G.dims[l + 1], smem + sp.a_off[l],
                                smem + sp.a_off[l + 1], G.act, l + 1 < L);
    else if (kc >= 16 && G.dims[l + 1] <= 8 * FWD_TMAX * (int)(blockDim.x >> 5))
      layer_forward_chunked<0>(P + lay.w_off[l], P + lay.b_off[l], lay.ld[l], G.dims[l + 1], smem + sp.a_off[l],
                               smem + sp.a_off[l + 1], G.act, l + 1 < L, smem + sp.total_floats, kc);
    else layer_forward<false>(P + lay.w_off[l], P + lay.b_off[l], lay.ld[l], G.dims[l + 1], smem + sp.a_off[l],
                              smem + sp.a_off[l + 1], G.act, l + 1 < L);
    __syncthreads();
  }
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    const float* zL = smem + sp.a_off[L];
    int pred;
    if (G.head == FEDB_HEAD_SOFTMAX_CE) {
      // torch.max(outputs, 1): first maximal index (C:81)
      pred = 0;
      float best = zL[lane];
      for (int c = 1; c < G.dims[L]; ++c) {
        const float z = zL[c * LDS + lane];
        if (z > best) { best = z; pred = c; }
      }
    } else {
      pred = zL[lane] > 0.f ? 1 : 0;  // sigmoid(z) > 0.5 (SK:1283-1290)
    }
    if (lane < nvalid) {
      const int yt = ylab[lane];
      if (yt >= 0 && yt < FEDB_MAX_CLASSES && pred < FEDB_MAX_CLASSES)
        atomicAdd(conf + (size_t)J.conf_slot * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES + yt * FEDB_MAX_CLASSES + pred, 1);
    }
  }
}

// ---------------------------------------------------------------------------------------
// K1 + K5: FedAvg, metrics, early stop
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(NTHREADS)
k_fedavg(const FedbGroup* __restrict__ groups, const FedbFed* __restrict__ feds, FedbFedState* __restrict__ state,
         int n_groups, float* __restrict__ params, float* __restrict__ global, const int32_t* __restrict__ conf,
         float* __restrict__ hist_local, float* __restrict__ hist_global, int write_back, int do_metrics, int begin_next,
         int32_t* __restrict__ conf_rw, uint16_t* __restrict__ shadow) {
  const int f = blockIdx.y;
  const FedbFed& F = feds[f];
  if (state[f].stop) return;
  // The clients' slab offsets and weights are staged in shared memory a chunk at a time, so that the loop over clients
  // is plain address arithmetic and eight independent 16-byte loads are in flight per thread: with 128 clients the
  // pass is a pure stream (4 bytes read, 4 + 2 written per client and element) and wants all the memory parallelism
  // it can get.  The arithmetic and its order are unchanged.
  constexpr int CH = 128, UR = 8;
  __shared__ long long s_off[CH];
  __shared__ float s_num[CH], s_den[CH];
  const int e = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const bool active = e < F.slab_size;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int k0 = 0; k0 < F.n_clients; k0 += CH) {
    const int cnt = min(CH, F.n_clients - k0);
    if (k0) __syncthreads();
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
      const FedbGroup& G = groups[F.first_group + k0 + i];
      s_off[i] = G.slab_off; s_num[i] = G.agg_num; s_den[i] = G.agg_den;
    }
    __syncthreads();
    if (active) {
      for (int k = 0; k < cnt; k += UR) {
        float4 w[UR];
#pragma unroll
        for (int u = 0; u < UR; ++u)
          if (k + u < cnt) w[u] = __ldcs(reinterpret_cast<const float4*>(params + s_off[k + u] + e));
#pragma unroll
        for (int u = 0; u < UR; ++u)
          if (k + u < cnt) {
            // reference arithmetic, term by term: (w * n_k) / N, accumulated in client order
            const float num = s_num[k + u], den = s_den[k + u];
            acc.x += __fdiv_rn(__fmul_rn(w[u].x, num), den);
            acc.y += __fdiv_rn(__fmul_rn(w[u].y, num), den);
            acc.z += __fdiv_rn(__fmul_rn(w[u].z, num), den);
            acc.w += __fdiv_rn(__fmul_rn(w[u].w, num), den);
          }
      }
    }
  }
  if (active) {
    *reinterpret_cast<float4*>(global + F.global_off + e) = acc;
    if (write_back) {
      const __nv_bfloat162 lo = __floats2bfloat162_rn(acc.x, acc.y), hi = __floats2bfloat162_rn(acc.z, acc.w);
      const uint2 pk = make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
      if (F.n_clients <= CH) {                         // offsets are still staged
        for (int k = 0; k < F.n_clients; ++k) {
          *reinterpret_cast<float4*>(params + s_off[k] + e) = acc;
          // wide path: the bf16 GEMM operands of every client follow the averaged weights in the same pass
          if (shadow) *reinterpret_cast<uint2*>(shadow + s_off[k] + e) = pk;
        }
      } else {
        for (int k = 0; k < F.n_clients; ++k) {
          const long long off = groups[F.first_group + k].slab_off;
          *reinterpret_cast<float4*>(params + off + e) = acc;
          if (shadow) *reinterpret_cast<uint2*>(shadow + off + e) = pk;
        }
      }
    }
  }
  if (blockIdx.x == 0) {
    __shared__ int pool[FEDB_MAX_CLASSES * FEDB_MAX_CLASSES];
    if (threadIdx.x < 32) {
      if (do_metrics) round_metrics_warp(F, state[f], f, n_groups, conf, hist_local, hist_global, pool, begin_next != 0);
      else if (begin_next && threadIdx.x == 0) begin_next_round(state[f]);
    }
    if (begin_next) {  // clear this federation's confusion slots for the next round (after the metrics read them)
      __syncthreads();
      for (int i = threadIdx.x; i < F.n_clients * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES; i += blockDim.x)
        conf_rw[(size_t)F.first_group * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES + i] = 0;
    }
  }
}

}  // namespace fedb

// =========================================================================================
// C ABI
// =========================================================================================
using namespace fedb;

extern "C" const char* fedb_last_error(void) { return g_err; }
extern "C" int fedb_version(void) { return 100; }
extern "C" void fedb_struct_sizes(int32_t* out) {
  out[0] = (int32_t)sizeof(FedbGroup); out[1] = (int32_t)sizeof(FedbFed); out[2] = (int32_t)sizeof(FedbFedState);
  out[3] = (int32_t)sizeof(FedbEvalJob); out[4] = (int32_t)sizeof(FedbTile);
}

extern "C" int64_t fedb_simt_smem_bytes(const int32_t* dims, int32_t n_layers) {
  if (n_layers < 1 || n_layers > FEDB_MAX_LAYERS) return -1;
  const SmemPlan sp = smem_plan(dims, n_layers);
  const SlabLayout lay = slab_layout(dims, n_layers);
  const int64_t bytes = (int64_t)sp.total_floats * 4, with_w = bytes + (int64_t)lay.end * 4;
  const int64_t cap = 227 * 1024 - 64;     // the kernels' few static __shared__ variables count against the same limit
  if (with_w <= cap) return with_w;  // weights staged in shared memory
  return bytes <= cap ? cap : -1;    // everything, so that the matrices can stream through what the activations leave
}

// Opt in to >48 KB of dynamic shared memory once per (device, function) — the attribute belongs to the device's context, so
// a process that drives several GPUs needs it on each (always the 227 KB maximum, so the call is not repeated inside a
// stream capture after the first, eager, launch).  The table is guarded: entry points may be called from several host threads.
static int set_smem(const void* fn, int64_t bytes) {
  constexpr int kDev = 64, kFn = 8;
  static std::mutex mu;
  static const void* done[kDev][kFn] = {{nullptr}};
  if (bytes <= 48 * 1024) return FEDB_OK;
  int dev = 0;
  FEDB_CUDA_OK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  const bool tracked = dev >= 0 && dev < kDev;
  if (tracked)
    for (int i = 0; i < kFn; ++i) if (done[dev][i] == fn) return FEDB_OK;
  cudaFuncAttributes fa;
  FEDB_CUDA_OK(cudaFuncGetAttributes(&fa, fn));   // static shared memory counts against the same 227 KB
  FEDB_CUDA_OK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - (int)fa.sharedSizeBytes));
  if (tracked)
    for (int i = 0; i < kFn; ++i) if (done[dev][i] == nullptr) { done[dev][i] = fn; break; }
  return FEDB_OK;
}

extern "C" int32_t fedb_batch_stats_stride(int32_t max_slab) { return 2 + (max_slab / 4 + NTHREADS - 1) / NTHREADS; }

extern "C" int fedb_init_params(const FedbGroup* groups, int32_t n_groups, int32_t max_slab, float* params,
                                const int32_t* streams, int32_t kind, uint64_t seed, void* stream) {
  FEDB_REQUIRE(n_groups > 0 && max_slab > 0, "bad sizes");
  dim3 grid((max_slab / 4 + 127) / 128, n_groups);
  k_init<<<grid, 128, 0, (cudaStream_t)stream>>>(groups, params, streams, kind, (uint32_t)seed, (uint32_t)(seed >> 32));
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}

extern "C" int fedb_round_begin(const FedbFed* feds, FedbFedState* state, int32_t n_feds, int32_t* conf,
                                int32_t conf_slots_total, void* stream) {
  const int n_conf = conf_slots_total * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES;
  const int work = n_conf > n_feds ? n_conf : n_feds;
  int blocks = (work + 255) / 256;
  if (blocks > 1024) blocks = 1024;
  if (blocks * 256 < n_feds) return set_err(FEDB_E_ARG, "too many federations for fedb_round_begin");
  k_round_begin<<<blocks, 256, 0, (cudaStream_t)stream>>>(feds, state, n_feds, conf, n_conf);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}

extern "C" int fedb_shuffle(const FedbGroup* groups, int32_t n_groups, const int32_t* idx_in, int32_t* idx_out,
                            const int32_t* epoch, const int32_t* active, uint64_t seed, void* stream) {
  dim3 grid(8, n_groups);
  k_shuffle<<<grid, 256, 0, (cudaStream_t)stream>>>(groups, idx_in, idx_out, epoch, active, (uint32_t)seed,
                                                    (uint32_t)(seed >> 32));
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}

extern "C" int fedb_fwd_bwd(const FedbGroup* groups, const FedbTile* tiles, int32_t n_tiles, const float* params,
                            const float* X, int32_t ldx, const int32_t* y, const int32_t* idx, int32_t batch,
                            float* scratch, int32_t* step, const int32_t* active, const FedbFedState* fed_state,
                            int64_t smem_bytes, void* stream) {
  FEDB_REQUIRE(n_tiles > 0, "n_tiles must be positive");
  FEDB_REQUIRE(ldx % 4 == 0, "ldx must be a multiple of 4");
  FEDB_REQUIRE(smem_bytes > 0 && smem_bytes <= 227 * 1024, "model too wide for the SIMT kernel (shared memory)");
  if (int rc = set_smem((const void*)k_fwd_bwd, smem_bytes)) return rc;
  k_fwd_bwd<<<n_tiles, TILE_THREADS, smem_bytes, (cudaStream_t)stream>>>(groups, tiles, params, X, ldx, y, idx, batch,
                                                                     scratch, step, active, fed_state,
                                                                     (int)(smem_bytes / 4));
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}

extern "C" int fedb_reduce_opt(const FedbGroup* groups, int32_t n_groups, int32_t max_slab, float* params, float* m,
                               float* v, const float* scratch, int32_t batch, const int32_t* step,
                               const int32_t* active, const FedbFedState* fed_state, float* loss_out,
                               double* batch_stats, int32_t max_batches, void* stream) {
  FEDB_REQUIRE(n_groups > 0 && max_slab > 0, "bad sizes");
  dim3 grid((max_slab / 4 + NTHREADS - 1) / NTHREADS, n_groups);
  k_reduce_opt<<<grid, NTHREADS, 0, (cudaStream_t)stream>>>(groups, params, m, v, scratch, batch, step, active,
                                                            fed_state, loss_out, batch_stats, max_batches);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}

extern "C" int fedb_epoch_end(const FedbGroup* groups, int32_t n_groups, const double* batch_stats, int32_t max_batches,
                              int32_t max_slab, double tol, int32_t n_iter_no_change, int32_t max_iter, int32_t* active,
                              int32_t* epoch, int32_t* no_improve, double* best_loss, double* last_loss,
                              void* stream) {
  const int nblk = (max_slab / 4 + NTHREADS - 1) / NTHREADS;
  k_epoch_end<<<(n_groups + 3) / 4, 128, 0, (cudaStream_t)stream>>>(groups, n_groups, batch_stats, max_batches, nblk,
                                                                      tol, n_iter_no_change, max_iter, active, epoch,
                                                                      no_improve, best_loss, last_loss);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}

// The whole local fit of every client in one launch (see k_fit_cluster).  Runs up to max_epochs further epochs per client (clients
// whose active flag is 0, or that converge, stop by themselves); the arguments are those of fedb_shuffle / fedb_fwd_bwd /
// fedb_reduce_opt / fedb_epoch_end, which this replaces call for call.  cluster_size CTAs (1..8) share a client's mini-batch tiles.
extern "C" int fedb_fit_cluster(const FedbGroup* groups, int32_t n_groups, int32_t max_slab, int32_t cluster_size, float* params,
                                float* m, float* v, const float* X, int32_t ldx, const int32_t* y, const int32_t* idx_in,
                                int32_t* idx_epoch, float* scratch, int32_t* step, int32_t* active,
                                const FedbFedState* fed_state, int32_t* epoch, int32_t* no_improve, double* best_loss, double* last_loss, float* loss_out, double* batch_stats,
                                int32_t max_batches, int32_t n_batches, double tol, int32_t n_iter_no_change, int32_t max_iter,
                                int32_t max_epochs, int32_t shuffle, uint64_t seed, int64_t smem_bytes, void* stream) {
  FEDB_REQUIRE(n_groups > 0 && max_slab > 0 && cluster_size >= 1 && cluster_size <= 8, "bad sizes");
  FEDB_REQUIRE(ldx % 4 == 0, "ldx must be a multiple of 4");
  FEDB_REQUIRE(smem_bytes > 0 && smem_bytes <= 227 * 1024, "model too wide for the SIMT kernel (shared memory)");
  FEDB_REQUIRE(batch_stats && active && epoch && no_improve && best_loss && last_loss, "fit state arrays required");
  if (int rc = set_smem((const void*)k_fit_cluster, smem_bytes)) return rc;
  FitArgs a{};
  a.groups = groups; a.params = params; a.m = m; a.v = v; a.scratch = scratch; a.loss_out = loss_out; a.X = X; a.y = y;
  a.idx_in = idx_in; a.idx_epoch = idx_epoch; a.step = step; a.active = active; a.fed_state = fed_state; a.epoch = epoch; a.no_improve = no_improve;
  a.batch_stats = batch_stats; a.best_loss = best_loss; a.last_loss = last_loss; a.tol = tol; a.ldx = ldx; a.n_batches = n_batches;
  a.max_batches = max_batches; a.nvblocks = (max_slab / 4 + NTHREADS - 1) / NTHREADS; a.n_iter_no_change = n_iter_no_change;
  a.max_iter = max_iter; a.max_epochs = max_epochs; a.shuffle = shuffle; a.smem_floats = (int)(smem_bytes / 4);
  a.s0 = (uint32_t)seed; a.s1 = (uint32_t)(seed >> 32);
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = (unsigned)cluster_size; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.gridDim = dim3((unsigned)(n_groups * cluster_size)); cfg.blockDim = dim3(TILE_THREADS); cfg.dynamicSmemBytes = (size_t)smem_bytes;
  cfg.stream = (cudaStream_t)stream; cfg.attrs = at; cfg.numAttrs = 1;
  FEDB_CUDA_OK(cudaLaunchKernelEx(&cfg, k_fit_cluster, a));
  return FEDB_OK;
}

extern "C" int fedb_eval(const FedbGroup* groups, const FedbEvalJob* jobs, const FedbTile* tiles, int32_t n_tiles,
                         const float* params, const float* global, const float* X, int32_t ldx, const int32_t* y,
                         const int32_t* idx, int32_t* conf, const FedbFedState* fed_state, int64_t smem_bytes,
                         void* stream) {
  FEDB_REQUIRE(n_tiles > 0, "n_tiles must be positive");
  FEDB_REQUIRE(smem_bytes > 0 && smem_bytes <= 227 * 1024, "model too wide for the SIMT kernel (shared memory)");
  if (int rc = set_smem((const void*)k_eval, smem_bytes)) return rc;
  k_eval<<<n_tiles, TILE_THREADS, smem_bytes, (cudaStream_t)stream>>>(groups, jobs, tiles, params, global, X, ldx, y, idx,
                                                                  conf, fed_state, (int)(smem_bytes / 4));
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}

extern "C" int fedb_fedavg(const FedbGroup* groups, const FedbFed* feds, FedbFedState* state, int32_t n_feds,
                           int32_t n_groups, int32_t max_slab, float* params, float* global, const int32_t* conf,
                           float* hist_local, float* hist_global, int32_t write_back, int32_t do_metrics,
                           int32_t begin_next, void* shadow_bf16, void* stream) {
  FEDB_REQUIRE(n_feds > 0 && max_slab > 0, "bad sizes");
  dim3 grid((max_slab / 4 + NTHREADS - 1) / NTHREADS, n_feds);
  // begin_next raises the stop flag at the end of block (0, f); with more than one block per federation a
  // block scheduled late could see it one round early
  FEDB_REQUIRE(!begin_next || grid.x == 1, "begin_next needs a single block per federation (slab <= 1024 floats)");
  k_fedavg<<<grid, NTHREADS, 0, (cudaStream_t)stream>>>(groups, feds, state, n_groups, params, global, conf, hist_local,
                                                        hist_global, write_back, do_metrics, begin_next,
                                                        const_cast<int32_t*>(conf), (uint16_t*)shadow_bf16);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}

// =========================================================================================
// Allocation-free local step
// =========================================================================================
// Scratch for fedb_local_step_noalloc.  It lives in the library image, so every GPU that loads
// libfedb200.so carries these 16 MiB from load until the process exits, whether or not the entry
// point is ever called.  In host code `g_noalloc_scratch` is only a handle for the runtime: the
// address kernels may dereference is the one cudaGetSymbolAddress returns for the current device.
#define FEDB_NOALLOC_SCRATCH_FLOATS (4u * 1024u * 1024u)
static __device__ __align__(256) float g_noalloc_scratch[FEDB_NOALLOC_SCRATCH_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
float* g_noalloc_ptr[kMaxDevices] = {nullptr};
std::mutex g_noalloc_mu;   // guards g_noalloc_ptr: the entry point may be entered from several host threads
}  // namespace

extern "C" int64_t fedb_noalloc_scratch_floats(void) { return (int64_t)FEDB_NOALLOC_SCRATCH_FLOATS; }

extern "C" int fedb_local_step_noalloc(const FedbGroup* groups_host, const FedbGroup* groups_dev, int32_t n_groups,
                                       const FedbTile* tiles, int32_t n_tiles, int32_t max_slab, float* params, float* m,
                                       float* v, const float* X, int32_t ldx, const int32_t* y, const int32_t* idx,
                                       int32_t batch, int32_t* step, const int32_t* active,
                                       const FedbFedState* fed_state, float* loss_out, int64_t smem_bytes,
                                       void* stream) {
  FEDB_REQUIRE(groups_host && groups_dev && n_groups > 0 && n_tiles > 0 && max_slab > 0, "bad arguments");
  FEDB_REQUIRE(ldx % 4 == 0, "ldx must be a multiple of 4");
  FEDB_REQUIRE(smem_bytes > 0 && smem_bytes <= 227 * 1024, "model too wide for the SIMT kernel (shared memory)");
  // The records of all groups must fit in the fixed array: check the host copy of the table.
  // (The kernels index the array from the DEVICE table: groups_host must be byte-identical to it.)  A group's records must
  // also cover every tile the kernels will touch: k_reduce_opt derives the tile count from n_rows / batch_size, not from
  // max_tiles, so a table that reserves too few records would run past the group's range — and, for the last group, past
  // the array.
  int64_t need = 0, tiles_reserved = 0;
  for (int g = 0; g < n_groups; ++g) {
    const FedbGroup& G = groups_host[g];
    if (G.scratch_off < 0 || G.max_tiles < 0 || G.slab_size <= 0 || G.n_rows <= 0)
      return set_err(FEDB_E_ARG, "fedb_local_step_noalloc: malformed group table");
    const int rows_per_step = G.batch_size > 0 ? (G.batch_size < G.n_rows ? G.batch_size : G.n_rows) : G.n_rows;
    if (G.max_tiles < (rows_per_step + FEDB_TILE_ROWS - 1) / FEDB_TILE_ROWS)
      return set_err(FEDB_E_ARG, "fedb_local_step_noalloc: a group's max_tiles is smaller than the tiles of one step");
    const int64_t end = G.scratch_off + (int64_t)G.max_tiles * (G.slab_size + FEDB_LOSS_SLOTS);
    if (end > need) need = end;
    tiles_reserved += G.max_tiles;
  }
  if ((int64_t)n_tiles > tiles_reserved) return set_err(FEDB_E_ARG, "fedb_local_step_noalloc: more tiles than the groups reserve records for");
  if (need > (int64_t)FEDB_NOALLOC_SCRATCH_FLOATS) {
    snprintf(g_err, sizeof(g_err),
             "fedb_local_step_noalloc: this step needs %lld floats of gradient records but the built-in scratch holds "
             "%lld; use fedb_fwd_bwd + fedb_reduce_opt with a caller-provided buffer",
             (long long)need, (long long)FEDB_NOALLOC_SCRATCH_FLOATS);
    return FEDB_E_UNSUPPORTED;
  }
  int dev = -1;
  FEDB_CUDA_OK(cudaGetDevice(&dev));
  FEDB_REQUIRE(dev >= 0 && dev < kMaxDevices, "device ordinal out of range");
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_ptr[dev] == nullptr) {
      // Once per device, with that device current: translate the symbol into a device address.
      float* p = nullptr;
      FEDB_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
      FEDB_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
      g_noalloc_ptr[dev] = p;
    }
    scratch = g_noalloc_ptr[dev];
  }
  // Nothing in the scratch has to start at zero: k_fwd_bwd overwrites every float of every record
  // (and its loss slot) that k_reduce_opt reads for the same step, so stale contents of an earlier
  // call are never observed.
  if (int rc = set_smem((const void*)k_fwd_bwd, smem_bytes)) return rc;
  k_fwd_bwd<<<n_tiles, TILE_THREADS, smem_bytes, (cudaStream_t)stream>>>(groups_dev, tiles, params, X, ldx, y, idx, batch,
                                                                     scratch, step, active, fed_state,
                                                                     (int)(smem_bytes / 4));
  FEDB_LAUNCH_OK();
  dim3 grid((max_slab / 4 + NTHREADS - 1) / NTHREADS, n_groups);
  k_reduce_opt<<<grid, NTHREADS, 0, (cudaStream_t)stream>>>(groups_dev, params, m, v, scratch, batch, step, active,
                                                            fed_state, loss_out, nullptr, 0);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}
