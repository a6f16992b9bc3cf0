// K2a/K3 on the 5th-generation tensor cores: one launch = the whole local part of a federated
// round for every client resident on this GPU —
//     n_steps x [ forward -> loss -> backward -> cross-CTA gradient reduction -> optimizer ]
//     -> local evaluation (confusion counts)
// i.e. C:63-73 (train_one_epoch) + C:75-83 (evaluate_local) without leaving the SM.
//
// Mapping.  One thread-block CLUSTER per client, one CTA per 128-row tile of the client's batch
// (UMMA M = 128: data row r of the tile <-> TMEM lane r <-> thread r).  All GEMMs are
// tcgen05.mma kind::f16 (bf16 operands, fp32 accumulation in TMEM); master weights, Adam moments
// and the update itself stay fp32.  Operands live in shared memory as 128B-swizzled [rows][64]
// bf16 panels; the SAME panel is used K-major by the forward / dgrad GEMMs and MN-major by the
// weight-gradient GEMMs (conventions validated on hardware by tests/test_gpu_tc.py), so no
// transposed copies exist.  Biases ride along as an extra K column that is held at 1.0
// (column 15 of X, column 63 of H1): the bias add is part of the forward MMA and the bias
// gradient falls out of the weight-gradient MMA.  The 2..8-wide head runs on CUDA cores inside
// the layer-2 epilogue (a tcgen05 N of 2 would waste the instruction).
//
// Per tile: 6 MMA phases, 4 epilogues, all operands produced in shared memory by the epilogue
// of the previous phase.  Weight-gradient accumulators stay in TMEM across the tiles of a CTA;
// after the last tile every CTA spills them as fp32 to shared memory, the cluster reduces them
// through distributed shared memory in a fixed order (bit-reproducible), each CTA applies the
// optimizer to its slice of the slab, and all CTAs rebuild their bf16 weight panels.
//
// Covers MLPs with 2 hidden ReLU layers, in <= 15, h1 <= 63, h2 <= 256, out <= 8 — the custom
// script's 14-50-200-2 (C:40) and the (50,50)/(50,100)/(50,200) grid points (H:73).  Anything
// else runs on the SIMT kernels (fedb_simt.cu), which are also this kernel's numerics oracle.
#include <cooperative_groups.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace cg = cooperative_groups;

namespace fedb {
namespace {

constexpr int TC_THREADS = 128;
constexpr int TC_ROWS = 128;
constexpr int PANEL = 128 * 128;  // bytes of a [128][64] bf16 panel
// shared-memory map (bytes from the 1024-aligned base)
constexpr int SM_X = 0;
constexpr int SM_H1 = SM_X + PANEL;
constexpr int SM_DZ1 = SM_H1 + PANEL;
constexpr int SM_DZ3 = SM_DZ1 + PANEL;        // directly after DZ1: it is the (ignored) upper M half of the wgrad-1 A operand
constexpr int SM_Z = SM_DZ3 + PANEL;          // 4 panels: H2, then dZ2
constexpr int SM_W1 = SM_Z + 4 * PANEL;       // [64][64] bf16
constexpr int SM_W2 = SM_W1 + 64 * 128;       // [256][64] bf16
constexpr int SM_W3 = SM_W2 + 256 * 128;      // fp32 [8][256]
constexpr int SM_B3 = SM_W3 + 8 * 256 * 4;    // fp32 [8], then scalars
constexpr int SM_END = SM_B3 + 256;
constexpr int SM_STAGE = SM_H1;               // fp32 gradient staging overlays H1..Z once the MMAs are done
// TMEM columns
constexpr uint32_t TM_D1 = 0, TM_D2 = 64, TM_DW2 = 320, TM_DW1 = 448, TM_DW3 = 464;

struct TcArgs {
  const FedbGroup* groups;
  const float* X;
  const int32_t* y;
  const int32_t* idx;
  float* params;
  float* m;
  float* v;
  int32_t* step;
  const int32_t* active;
  const FedbFedState* fed_state;
  float* loss_out;
  int32_t* conf;
  int ldx, batch, n_steps, do_eval;
};

struct Shape {
  int d0, h1, h2, C, N2, NP2, MT2;
  int ld0, ld1, ld2;
};

__device__ __forceinline__ void st16(uint8_t* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  *reinterpret_cast<uint4*>(p) = make_uint4(a, b, c, d);
}

// fp32 slab (global) -> bf16 panels W1, W2 (bias in the spare K column) + fp32 W3 / b3.
// Loads bypass L1 (__ldcg): peers on other SMs rewrite the slab between two calls.
__device__ void load_weights(uint8_t* sm, const float* P, const SlabLayout& lay, const Shape& s) {
  const int tid = threadIdx.x;
  for (int i = tid; i < (SM_B3 + 64 - SM_W1) / 16; i += TC_THREADS) st16(sm + SM_W1 + i * 16, 0, 0, 0, 0);
  __syncthreads();
  for (int task = tid; task < s.h1 * 2; task += TC_THREADS) {
    const int j = task >> 1, c = task & 1;
    float f[8];
#pragma unroll
    for (int ii = 0; ii < 8; ++ii) {
      const int i = 8 * c + ii;
      f[ii] = i < s.d0 ? __ldcg(P + lay.w_off[0] + j * s.ld0 + i) : (i == 15 ? __ldcg(P + lay.b_off[0] + j) : 0.f);
    }
    st16(sm + SM_W1 + tc::panel_off(j, 8 * c), tc::pack_bf16(f[0], f[1]), tc::pack_bf16(f[2], f[3]),
         tc::pack_bf16(f[4], f[5]), tc::pack_bf16(f[6], f[7]));
  }
  for (int task = tid; task < s.h2 * 8; task += TC_THREADS) {
    const int j = task >> 3, c = task & 7;
    float f[8];
#pragma unroll
    for (int ii = 0; ii < 8; ++ii) {
      const int i = 8 * c + ii;
      f[ii] = i < s.h1 ? __ldcg(P + lay.w_off[1] + j * s.ld1 + i) : (i == 63 ? __ldcg(P + lay.b_off[1] + j) : 0.f);
    }
    st16(sm + SM_W2 + tc::panel_off(j, 8 * c), tc::pack_bf16(f[0], f[1]), tc::pack_bf16(f[2], f[3]),
         tc::pack_bf16(f[4], f[5]), tc::pack_bf16(f[6], f[7]));
  }
  float* W3s = reinterpret_cast<float*>(sm + SM_W3);
  float* b3s = reinterpret_cast<float*>(sm + SM_B3);
  for (int e = tid; e < s.C * s.h2; e += TC_THREADS) {
    const int k = e / s.h2, c = e - k * s.h2;
    W3s[k * 256 + c] = __ldcg(P + lay.w_off[2] + k * s.ld2 + c);
  }
  if (tid < s.C) b3s[tid] = __ldcg(P + lay.b_off[2] + tid);
  tc::fence_async_smem();
  __syncthreads();
}

// One data row per thread -> X panel (bf16; column 15 = 1 on valid rows). Returns the label.
__device__ __forceinline__ int load_x_row(uint8_t* sm, const TcArgs& a, int64_t pos, bool valid, const Shape& s) {
  const int t = threadIdx.x;
  float f[16];
  int label = 0;
  if (valid) {
    const int row = __ldg(a.idx + pos);
    const float4* src = reinterpret_cast<const float4*>(a.X + (size_t)row * a.ldx);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 v = __ldg(src + q);
      f[4 * q] = v.x; f[4 * q + 1] = v.y; f[4 * q + 2] = v.z; f[4 * q + 3] = v.w;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if (i >= s.d0) f[i] = 0.f;
    f[15] = 1.f;
    label = __ldg(a.y + row);
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) f[i] = 0.f;
  }
#pragma unroll
  for (int c = 0; c < 2; ++c)
    st16(sm + SM_X + tc::panel_off(t, 8 * c), tc::pack_bf16(f[8 * c], f[8 * c + 1]), tc::pack_bf16(f[8 * c + 2], f[8 * c + 3]),
         tc::pack_bf16(f[8 * c + 4], f[8 * c + 5]), tc::pack_bf16(f[8 * c + 6], f[8 * c + 7]));
  return label;
}

// epilogue -> next MMA hand-off: my TMEM reads are done, my smem writes are visible to the tensor core
__device__ __forceinline__ void handoff() {
  tc::tmem_ld_wait();
  tc::tc_fence_before();
  tc::fence_async_smem();
  __syncthreads();
}

struct Issuer {
  uint32_t sm, tmem;
  uint64_t* bar;
  __device__ __forceinline__ uint64_t kd(int off) const { return tc::make_desc(sm + off, 0, 1024); }          // K-major
  __device__ __forceinline__ uint64_t md(int off) const { return tc::make_desc(sm + off, PANEL, 1024); }      // MN-major
  __device__ void fwd1(const Shape&) const {
    tc::mma_bf16(tmem + TM_D1, kd(SM_X), kd(SM_W1), tc::make_idesc_bf16(128, 64, false, false), false);
  }
  __device__ void fwd2(const Shape& s) const {
    const uint32_t id = tc::make_idesc_bf16(128, s.N2, false, false);
    for (int k = 0; k < 4; ++k) tc::mma_bf16(tmem + TM_D2, kd(SM_H1 + 32 * k), kd(SM_W2 + 32 * k), id, k > 0);
  }
  __device__ void wgrad3(const Shape& s, bool acc) const {
    const uint32_t id = tc::make_idesc_bf16(128, 16, true, true);
    for (int mt = 0; mt < s.MT2; ++mt)
      for (int k = 0; k < 8; ++k)
        tc::mma_bf16(tmem + TM_DW3 + 16 * mt, md(SM_Z + 2 * mt * PANEL + 2048 * k), md(SM_DZ3 + 2048 * k), id, acc || k > 0);
  }
  __device__ void dgrad2_wgrad2(const Shape& s, bool acc) const {
    const uint32_t id4 = tc::make_idesc_bf16(128, 64, false, true);
    for (int p = 0; p < s.NP2; ++p)
      for (int k = 0; k < min(4, (s.N2 - 64 * p) >> 4); ++k)  // never feed unwritten pad columns of the last panel
        tc::mma_bf16(tmem + TM_D1, kd(SM_Z + p * PANEL + 32 * k), md(SM_W2 + (p * 64 + 16 * k) * 128), id4, (p | k) > 0);
    const uint32_t id5 = tc::make_idesc_bf16(128, 64, true, true);
    for (int mt = 0; mt < s.MT2; ++mt)
      for (int k = 0; k < 8; ++k)
        tc::mma_bf16(tmem + TM_DW2 + 64 * mt, md(SM_Z + 2 * mt * PANEL + 2048 * k), md(SM_H1 + 2048 * k), id5, acc || k > 0);
  }
  __device__ void wgrad1(const Shape&, bool acc) const {
    const uint32_t id = tc::make_idesc_bf16(128, 16, true, true);
    for (int k = 0; k < 8; ++k) tc::mma_bf16(tmem + TM_DW1, md(SM_DZ1 + 2048 * k), md(SM_X + 2048 * k), id, acc || k > 0);
  }
};

}  // namespace

__global__ void __launch_bounds__(TC_THREADS, 1) k_local_round_tc(TcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float red_f[4][9];   // per-warp partial sums: loss + db3[8]
  __shared__ float part_out[9];   // this CTA's loss sum + db3 (read by peers over DSMEM)
  __shared__ int cnt[FEDB_MAX_CLASSES * FEDB_MAX_CLASSES];
  __shared__ StepScalars sc_s;

  cg::cluster_group cluster = cg::this_cluster();
  const int CS = (int)cluster.num_blocks(), q = (int)cluster.block_rank();
  const int g = blockIdx.x / CS;
  const FedbGroup& G = a.groups[g];
  if (a.active && !a.active[g]) return;                     // cluster-uniform exits
  if (a.fed_state && a.fed_state[G.fed].stop) return;
  const int n = G.n_rows;
  const int bs = G.batch_size > 0 ? min(G.batch_size, n) : n;
  const int b0 = a.batch * bs;
  if (b0 >= n) return;
  const int nb = min(bs, n - b0);
  const int ntiles = (nb + TC_ROWS - 1) / TC_ROWS;

  // 1024-byte aligned operand area (swizzled panels require it)
  const uint32_t raw = tc::smem_u32(smem_raw);
  uint8_t* sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  const uint32_t sm_u = tc::smem_u32(sm);

  Shape s;
  s.d0 = G.dims[0]; s.h1 = G.dims[1]; s.h2 = G.dims[2]; s.C = G.dims[3];
  s.N2 = (s.h2 + 15) & ~15; s.NP2 = (s.N2 + 63) >> 6; s.MT2 = (s.N2 + 127) >> 7;
  s.ld0 = ru4(s.d0); s.ld1 = ru4(s.h1); s.ld2 = ru4(s.h2);
  const SlabLayout lay = slab_layout(G.dims, 3);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int step0 = a.step[g];
  float* Pg = a.params + G.slab_off;

  if (warp == 0) tc::tmem_alloc<512>(&tmem_slot);
  if (tid == 0) {
    tc::mbar_init(&bar, 1);
    tc::fence_mbar_init();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t tl = tmem + ((uint32_t)(warp * 32) << 16);  // this thread's TMEM lane
  const Issuer mma{sm_u, tmem, &bar};
  uint32_t ph = 0;
  float* W3s = reinterpret_cast<float*>(sm + SM_W3);
  float* b3s = reinterpret_cast<float*>(sm + SM_B3);
  float* stage = reinterpret_cast<float*>(sm + SM_STAGE);
  const float inv_n = 1.f / (float)nb;

  load_weights(sm, Pg, lay, s);

#define FEDB_MMA_PHASE(ISSUE)            \
  do {                                   \
    if (tid == 0) {                      \
      tc::tc_fence_after();              \
      ISSUE;                             \
      tc::mma_commit(&bar);              \
    }                                    \
    tc::mbar_wait(&bar, ph);             \
    ph ^= 1;                             \
    tc::tc_fence_after();                \
  } while (0)

  for (int st = 0; st < a.n_steps; ++st) {
    float loss_acc = 0.f, db3_acc[FEDB_MAX_CLASSES];
#pragma unroll
    for (int k = 0; k < FEDB_MAX_CLASSES; ++k) db3_acc[k] = 0.f;
    int my_tiles = 0;

    for (int tile = q; tile < ntiles; tile += CS, ++my_tiles) {
      const bool acc = my_tiles > 0;
      const int r = tile * TC_ROWS + tid;
      const bool valid = r < nb;
      const int label = load_x_row(sm, a, G.idx_off + b0 + r, valid, s);
      handoff();
      // ---- layer 1 -----------------------------------------------------------------------
      FEDB_MMA_PHASE(mma.fwd1(s));
      uint64_t mask1 = 0;
      {
        float v[16];
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 16) {
          tc::tmem_ld16(tl + TM_D1 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            if (v[i] > 0.f) mask1 |= 1ull << (c0 + i); else v[i] = 0.f;
          }
          if (c0 == 48) { v[15] = valid ? 1.f : 0.f; mask1 &= ~(1ull << 63); }
#pragma unroll
          for (int h = 0; h < 2; ++h)
            st16(sm + SM_H1 + tc::panel_off(tid, c0 + 8 * h), tc::pack_bf16(v[8 * h], v[8 * h + 1]),
                 tc::pack_bf16(v[8 * h + 2], v[8 * h + 3]), tc::pack_bf16(v[8 * h + 4], v[8 * h + 5]),
                 tc::pack_bf16(v[8 * h + 6], v[8 * h + 7]));
        }
      }
      handoff();
      // ---- layer 2 + head ----------------------------------------------------------------
      FEDB_MMA_PHASE(mma.fwd2(s));
      float z[FEDB_MAX_CLASSES], dz3[FEDB_MAX_CLASSES];
#pragma unroll
      for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = k < s.C ? b3s[k] : 0.f;
      for (int c0 = 0; c0 < s.N2; c0 += 16) {
        float v[16];
        tc::tmem_ld16(tl + TM_D2 + c0, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
#pragma unroll
        for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
          if (k < s.C) {
#pragma unroll
            for (int i = 0; i < 16; ++i) z[k] = fmaf(v[i], W3s[k * 256 + c0 + i], z[k]);
          }
        uint8_t* pz = sm + SM_Z + (c0 >> 6) * PANEL;
#pragma unroll
        for (int h = 0; h < 2; ++h)
          st16(pz + tc::panel_off(tid, (c0 & 63) + 8 * h), tc::pack_bf16(v[8 * h], v[8 * h + 1]),
               tc::pack_bf16(v[8 * h + 2], v[8 * h + 3]), tc::pack_bf16(v[8 * h + 4], v[8 * h + 5]),
               tc::pack_bf16(v[8 * h + 6], v[8 * h + 7]));
      }
      {
        float loss = 0.f;
#pragma unroll
        for (int k = 0; k < FEDB_MAX_CLASSES; ++k) dz3[k] = 0.f;
        if (valid) {
          if (G.head == FEDB_HEAD_SOFTMAX_CE) {
            float mx = -INFINITY;
#pragma unroll
            for (int k = 0; k < FEDB_MAX_CLASSES; ++k) if (k < s.C) mx = fmaxf(mx, z[k]);
            float se = 0.f;
#pragma unroll
            for (int k = 0; k < FEDB_MAX_CLASSES; ++k) if (k < s.C) se += expf(z[k] - mx);
            const float lse = mx + logf(se);
#pragma unroll
            for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
              if (k < s.C) {
                if (k == label) loss = lse - z[k];
                dz3[k] = (expf(z[k] - lse) - (k == label ? 1.f : 0.f)) * inv_n;
              }
          } else {
            const float p = 1.f / (1.f + expf(-z[0]));
            const float eps = 1.1920929e-07f;
            const float pc = fminf(fmaxf(p, eps), 1.f - eps), yt = (float)label;
            loss = -(yt * logf(pc) + (1.f - yt) * logf(1.f - pc));
            dz3[0] = (p - yt) * inv_n;
          }
        }
        loss_acc += loss;
#pragma unroll
        for (int k = 0; k < FEDB_MAX_CLASSES; ++k) db3_acc[k] += dz3[k];
        st16(sm + SM_DZ3 + tc::panel_off(tid, 0), tc::pack_bf16(dz3[0], dz3[1]), tc::pack_bf16(dz3[2], dz3[3]),
             tc::pack_bf16(dz3[4], dz3[5]), tc::pack_bf16(dz3[6], dz3[7]));
        st16(sm + SM_DZ3 + tc::panel_off(tid, 8), 0, 0, 0, 0);
      }
      handoff();
      // ---- head weight gradient (needs H2), then dZ2 overwrites H2 --------------------------
      FEDB_MMA_PHASE(mma.wgrad3(s, acc));
      for (int c0 = 0; c0 < s.N2; c0 += 16) {
        float v[16], d[16];
        tc::tmem_ld16(tl + TM_D2 + c0, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) d[i] = 0.f;
#pragma unroll
        for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
          if (k < s.C) {
#pragma unroll
            for (int i = 0; i < 16; ++i) d[i] = fmaf(dz3[k], W3s[k * 256 + c0 + i], d[i]);
          }
#pragma unroll
        for (int i = 0; i < 16; ++i) d[i] = v[i] > 0.f ? d[i] : 0.f;
        uint8_t* pz = sm + SM_Z + (c0 >> 6) * PANEL;
#pragma unroll
        for (int h = 0; h < 2; ++h)
          st16(pz + tc::panel_off(tid, (c0 & 63) + 8 * h), tc::pack_bf16(d[8 * h], d[8 * h + 1]),
               tc::pack_bf16(d[8 * h + 2], d[8 * h + 3]), tc::pack_bf16(d[8 * h + 4], d[8 * h + 5]),
               tc::pack_bf16(d[8 * h + 6], d[8 * h + 7]));
      }
      handoff();
      // ---- dH1 = dZ2 W2 ; dW2 += dZ2^T H1 ------------------------------------------------------
      FEDB_MMA_PHASE(mma.dgrad2_wgrad2(s, acc));
      {
        float v[16];
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 16) {
          tc::tmem_ld16(tl + TM_D1 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = ((mask1 >> (c0 + i)) & 1ull) ? v[i] : 0.f;
#pragma unroll
          for (int h = 0; h < 2; ++h)
            st16(sm + SM_DZ1 + tc::panel_off(tid, c0 + 8 * h), tc::pack_bf16(v[8 * h], v[8 * h + 1]),
                 tc::pack_bf16(v[8 * h + 2], v[8 * h + 3]), tc::pack_bf16(v[8 * h + 4], v[8 * h + 5]),
                 tc::pack_bf16(v[8 * h + 6], v[8 * h + 7]));
        }
      }
      handoff();
      // ---- dW1 += dZ1^T X ------------------------------------------------------------------------
      FEDB_MMA_PHASE(mma.wgrad1(s, acc));
    }  // tiles

    // ---- spill this CTA's gradients (fp32, slab layout) to the staging area -------------------
    __syncthreads();
    for (int i = tid; i < (lay.total >> 2); i += TC_THREADS) reinterpret_cast<float4*>(stage)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    if (my_tiles > 0) {
      float v[16];
      for (int mt = 0; mt < s.MT2; ++mt) {
        const int j = 128 * mt + tid;
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 16) {
          tc::tmem_ld16(tl + TM_DW2 + 64 * mt + c0, v);
          tc::tmem_ld_wait();
          if (j < s.h2) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int col = c0 + i;
              if (col < s.h1) stage[lay.w_off[1] + j * s.ld1 + col] = v[i];
              else if (col == 63) stage[lay.b_off[1] + j] = v[i];
            }
          }
        }
        tc::tmem_ld16(tl + TM_DW3 + 16 * mt, v);
        tc::tmem_ld_wait();
        if (j < s.h2) {
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
            if (k < s.C) stage[lay.w_off[2] + k * s.ld2 + j] = v[k];
        }
      }
      tc::tmem_ld16(tl + TM_DW1, v);
      tc::tmem_ld_wait();
      if (tid < s.h1) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          if (i < s.d0) stage[lay.w_off[0] + tid * s.ld0 + i] = v[i];
          else if (i == 15) stage[lay.b_off[0] + tid] = v[i];
        }
      }
    }
    // loss and db3: warp shuffle, then the 4 warps in a fixed order
    {
      float vals[9];
      vals[0] = loss_acc;
#pragma unroll
      for (int k = 0; k < 8; ++k) vals[k + 1] = db3_acc[k];
#pragma unroll
      for (int i = 0; i < 9; ++i) {
        float x = vals[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) red_f[warp][i] = x;
      }
      __syncthreads();
      if (tid < 9) {
        const float x = red_f[0][tid] + red_f[1][tid] + red_f[2][tid] + red_f[3][tid];
        part_out[tid] = x;
        if (tid >= 1 && tid - 1 < s.C) stage[lay.b_off[2] + tid - 1] = x;
      }
    }
    if (tid == 0) sc_s = step_scalars(G, step0 + st + 1);
    tc::tc_fence_before();
    cluster.sync();

    // ---- cluster reduction over DSMEM (fixed rank order) + optimizer on my slice ----------------
    {
      const StepScalars sc = sc_s;
      const int n4 = lay.end >> 2, per = (n4 + CS - 1) / CS;
      const int lo = q * per, hi = min(n4, lo + per);
      for (int e4 = lo + tid; e4 < hi; e4 += TC_THREADS) {
        float4 gsum = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int rk = 0; rk < CS; ++rk) {
          const float4 p = reinterpret_cast<const float4*>(cluster.map_shared_rank(stage, rk))[e4];
          gsum.x += p.x; gsum.y += p.y; gsum.z += p.z; gsum.w += p.w;
        }
        const int e = e4 * 4;
        bool is_w = false;
        for (int l = 0; l < 3; ++l) is_w |= (e >= lay.w_off[l] && e < lay.b_off[l]);
        float4 w = *reinterpret_cast<float4*>(Pg + e);
        if (is_w && G.alpha != 0.f) {
          const float al = G.alpha * inv_n;
          gsum.x = fmaf(al, w.x, gsum.x); gsum.y = fmaf(al, w.y, gsum.y);
          gsum.z = fmaf(al, w.z, gsum.z); gsum.w = fmaf(al, w.w, gsum.w);
        }
        float4 mm = make_float4(0.f, 0.f, 0.f, 0.f), vv = mm;
        if (a.m) mm = *reinterpret_cast<float4*>(a.m + G.slab_off + e);
        if (a.v && G.opt != FEDB_OPT_SGD) vv = *reinterpret_cast<float4*>(a.v + G.slab_off + e);
        w.x = opt_update(G, sc, w.x, gsum.x, mm.x, vv.x);
        w.y = opt_update(G, sc, w.y, gsum.y, mm.y, vv.y);
        w.z = opt_update(G, sc, w.z, gsum.z, mm.z, vv.z);
        w.w = opt_update(G, sc, w.w, gsum.w, mm.w, vv.w);
        *reinterpret_cast<float4*>(Pg + e) = w;
        if (a.m) *reinterpret_cast<float4*>(a.m + G.slab_off + e) = mm;
        if (a.v && G.opt != FEDB_OPT_SGD) *reinterpret_cast<float4*>(a.v + G.slab_off + e) = vv;
      }
      if (q == 0 && tid == 0 && a.loss_out) {
        float ls = 0.f;
        for (int rk = 0; rk < CS; ++rk) ls += cluster.map_shared_rank(part_out, rk)[0];
        a.loss_out[g] = ls * inv_n;
      }
    }
    __threadfence();
    cluster.sync();   // all slices written and visible; nobody still reads my staging area
    load_weights(sm, Pg, lay, s);
  }  // steps

  // ---- local evaluation with the updated weights (C:75-83) ----------------------------------
  if (a.do_eval) {
    if (tid < FEDB_MAX_CLASSES * FEDB_MAX_CLASSES) cnt[tid] = 0;
    __syncthreads();
    for (int tile = q; tile < ntiles; tile += CS) {
      const int r = tile * TC_ROWS + tid;
      const bool valid = r < nb;
      const int label = load_x_row(sm, a, G.idx_off + b0 + r, valid, s);
      handoff();
      FEDB_MMA_PHASE(mma.fwd1(s));
      {
        float v[16];
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 16) {
          tc::tmem_ld16(tl + TM_D1 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
          if (c0 == 48) v[15] = valid ? 1.f : 0.f;
#pragma unroll
          for (int h = 0; h < 2; ++h)
            st16(sm + SM_H1 + tc::panel_off(tid, c0 + 8 * h), tc::pack_bf16(v[8 * h], v[8 * h + 1]),
                 tc::pack_bf16(v[8 * h + 2], v[8 * h + 3]), tc::pack_bf16(v[8 * h + 4], v[8 * h + 5]),
                 tc::pack_bf16(v[8 * h + 6], v[8 * h + 7]));
        }
      }
      handoff();
      FEDB_MMA_PHASE(mma.fwd2(s));
      float z[FEDB_MAX_CLASSES];
#pragma unroll
      for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = k < s.C ? b3s[k] : 0.f;
      for (int c0 = 0; c0 < s.N2; c0 += 16) {
        float v[16];
        tc::tmem_ld16(tl + TM_D2 + c0, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
          if (k < s.C) {
#pragma unroll
            for (int i = 0; i < 16; ++i) z[k] = fmaf(fmaxf(v[i], 0.f), W3s[k * 256 + c0 + i], z[k]);
          }
      }
      int pred = 0;
      if (G.head == FEDB_HEAD_SOFTMAX_CE) {
        float best = z[0];
#pragma unroll
        for (int k = 1; k < FEDB_MAX_CLASSES; ++k)
          if (k < s.C && z[k] > best) { best = z[k]; pred = k; }
      } else {
        pred = z[0] > 0.f ? 1 : 0;
      }
      if (valid && label >= 0 && label < FEDB_MAX_CLASSES) atomicAdd(&cnt[label * FEDB_MAX_CLASSES + pred], 1);
    }
    tc::tmem_ld_wait();
    __syncthreads();
    if (tid < FEDB_MAX_CLASSES * FEDB_MAX_CLASSES && cnt[tid] != 0)
      atomicAdd(a.conf + (size_t)g * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES + tid, cnt[tid]);
  }
#undef FEDB_MMA_PHASE

  if (q == 0 && tid == 0) a.step[g] = step0 + a.n_steps;
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(tmem);
}

}  // namespace fedb

using namespace fedb;

extern "C" int64_t fedb_tc_smem_bytes(void) { return SM_END + 1024; }

// One launch: n_steps local optimizer steps (+ optional local evaluation) for every group, on the
// tensor cores.  cluster_size CTAs cooperate on one client; every group must satisfy
// 3 layers, relu, dims[0] <= 15, dims[1] <= 63, dims[2] <= 256, dims[3] <= 8 (checked by the
// caller on its host table: ops/tc.py: tc_eligible).
extern "C" int fedb_local_round_tc(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X,
                                   int32_t ldx, const int32_t* y, const int32_t* idx, float* params, float* m, float* v,
                                   int32_t* step, const int32_t* active, const FedbFedState* fed_state, float* loss_out,
                                   int32_t* conf, int32_t batch, int32_t n_steps, int32_t do_eval, void* stream) {
  FEDB_REQUIRE(n_groups > 0 && cluster_size >= 1 && cluster_size <= 8, "bad grid");
  FEDB_REQUIRE(ldx >= 16 && ldx % 4 == 0, "feature rows must be padded to >= 16 floats");
  static bool attr_set = false;
  const int smem = SM_END + 1024;
  if (!attr_set) {
    FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_local_round_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  TcArgs a{groups, X, y, idx, params, m, v, step, active, fed_state, loss_out, conf, ldx, batch, n_steps, do_eval};
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(n_groups * cluster_size);
  cfg.blockDim = dim3(TC_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster_size;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  FEDB_CUDA_OK(cudaLaunchKernelEx(&cfg, k_local_round_tc, a));
  return FEDB_OK;
}
