// K2b on the 5th-generation tensor cores: the sklearn-style local fit (SK:703-777 `_fit_stochastic` as driven by S:101 and
// H:90-91: per-epoch reshuffle, mini-batches of 200, forward / log-loss (+ L2) / backward / Adam with sklearn's bias
// correction, training-loss curve, tol / n_iter_no_change stopping rule) of every client in ONE launch, with the epoch
// loop on the device and the weights resident on chip — `mlp_persistent_train<sigmoid_logloss, adam_sklearn>` of SURVEY §8.
//
// Arithmetic as in csrc/fedb_tcx.cu: every MMA operand is split into two fp16 numbers (22 significant bits) and every GEMM
// is three tcgen05.mma products into fp32 TMEM accumulators, so the loss curve — which decides the stopping epoch — is
// fp32-grade.  What differs is the PARALLELISATION, because the scripts' default model 14-50-400-1 does not fit the
// round kernel's scheme (one CTA = a 64-row tile of the whole network: 102 KB of split W2 panels + 112 KB of H2 / dZ2 panels):
//
//   a cluster of NC = 4 CTAs owns a client and splits the SECOND HIDDEN LAYER, not the rows: CTA q holds rows
//   [q S, (q+1) S) of W2 (S = ceil(h2 / 4); with their b2 and W3 entries) and pushes EVERY row of the mini-batch through
//   them, 128 rows per tile (UMMA M = 128, all four SM sub-partitions busy in the epilogues).
//
// Consequences: layer 1 is replicated (K = 16, negligible); the logits need one exchange per tile — every CTA writes its
// partial row sums into all peers' shared memory, one cluster barrier, everyone adds the four partials in rank order —;
// dW2 / db2 / dW3 of a CTA's slice are COMPLETE in its own TMEM (it saw all rows), so the bulk of the model (96 %) never
// crosses a CTA boundary: the owner applies Adam straight from the accumulators and rewrites its own operand panels; the
// fp32 master weights and both moments of the slice live in TENSOR MEMORY for the whole fit (3 x 64 columns in the same
// lane / column geometry as the dW2 accumulator: a thread reads gradient, weight, m and v of its 16 elements with four
// tcgen05.ld and puts them back with three tcgen05.st, no global memory in the step); the backward pass is linear in dH1,
// so each CTA turns its PARTIAL dH1 = dZ2_slice W2_slice into a partial dW1 / db1 (50 x 16 numbers), which are exchanged like
// the logits and applied redundantly by all four CTAs to their replicated copy of W1 (fp32 master and moments in shared
// memory).  Three cluster barriers per mini-batch of 200 rows, no other synchronisation between CTAs, none between clients.
//
// Shapes: 2 hidden ReLU layers, in <= 15, h1 <= 63, h2 <= 512, out <= 8 (sigmoid: 1); any batch size.  Other models run on
// the CUDA-core kernels (csrc/fedb_simt.cu), which are also this kernel's numerics oracle (tests/test_gpu_fitx.py).
#include <cooperative_groups.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "tcx_common.cuh"

namespace cg = cooperative_groups;

namespace fedb {
namespace {

constexpr int FT = 512;            // threads per CTA: 128 rows x 4 column parts
constexpr int FR = 128;            // data rows per tile
constexpr int FP = FR * 128;       // bytes of a [128][64] fp16 panel
constexpr int FNC = 4;             // CTAs per client
// shared-memory map (bytes from the 1024-aligned base)
constexpr int FX_X = 0;                         // cols 0-15 X hi (col 15 = 1), 16-31 X lo, 32-39 dZ3 hi, 40-47 dZ3 lo
constexpr int FX_H1 = FX_X + FP;                // H1 hi, H1 lo; later dZ1 hi, dZ1 lo
constexpr int FX_Z = FX_H1 + 2 * FP;            // hi panels 0, 1; lo panels 2, 3: H2 slice, then dZ2 slice
constexpr int FX_W1 = FX_Z + 4 * FP;            // [64][64]: cols 0-15 W1 hi (col 15 = b1), cols 16-31 lo
constexpr int FX_W2 = FX_W1 + 64 * 128;         // my slice, hi [128][64] (col 63 = b2)
constexpr int FX_W2L = FX_W2 + 128 * 128;       // lo
constexpr int FX_W3 = FX_W2L + 128 * 128;       // fp32 [128 units of my slice][8 classes]
constexpr int FX_B3 = FX_W3 + 128 * 8 * 4;      // fp32 b3[8], then its moments m[8], v[8]
constexpr int FX_M1 = FX_B3 + 128;              // fp32 replicated layer-1 state: w[64][16], m[64][16], v[64][16] (col 15 = b1)
constexpr int FX_ZP = FX_M1 + 3 * 64 * 16 * 4;  // partial logits [2 parities][NC][8][128] floats
constexpr int FX_G1 = FX_ZP + 2 * FNC * 8 * FR * 4;   // partial layer-1 gradients [NC][64 * 16 + 16] floats (+ sum of squares of the slice)
constexpr int G1_FLOATS = 64 * 16 + 16;
constexpr int FX_END = FX_G1 + FNC * G1_FLOATS * 4;
static_assert(FX_END + 1024 + 4096 <= 227 * 1024, "fit kernel: shared memory map too large");
// TMEM columns
// TMEM columns: accumulators D1 (layer-1 pre-activations, later the partial dH1), D2 (my slice of layer 2), DW2 / DW1 / DW3
// (weight gradients); ZX (the exchange of a row's four partial logits) reuses D1 while it is dead.  Then the optimizer state
// of the slice, resident for the whole fit: W / M / V [128 units][64] (column 63 = b2), S3 = {W3[0], W3[1], m, m, v, v} per unit.
constexpr uint32_t TF_D1 = 0, TF_D2 = 64, TF_DW2 = 192, TF_DW1 = 256, TF_DW3 = 288, TF_ZX = TF_D1;
constexpr uint32_t TF_W = 304, TF_M = 368, TF_V = 432, TF_S3 = 496;
// Adam's square root and division as single MUFU approximations (2 ulp) instead of the IEEE sequences: the error of a step is
// 2^-22 of lr, far below the rounding of the fp32 weight it is added to; the step is instruction-issue bound otherwise.
constexpr bool FITX_FAST = true;

struct FitXArgs {
  const FedbGroup* groups;
  const float* X;
  const int32_t* y;
  const int32_t* idx_in;
  int32_t* idx_epoch;
  float* params;
  float* m;
  float* v;
  int32_t* step;
  int32_t* active;
  const FedbFedState* fed_state;
  int32_t* epoch;
  int32_t* no_improve;
  double* best_loss;
  double* last_loss;
  float* loss_out;
  double tol;
  int ldx, n_iter_no_change, max_iter, max_epochs, shuffle;
  uint32_t s0, s1;
  long long* prof;   // optional: globaltimer stamps of thread 0 of every CTA of client 0 during the first mini-batch, [NC][32]
};

struct ShapeF {
  int d0, h1, h2, C, ld0, ld1, ld2;
  int J0, nj, N2;      // my slice of hidden layer 2: first unit, units, MMA N (multiple of 16)
  int w_off0, b_off0, w_off1, b_off1, w_off2, b_off2;
};

__device__ __forceinline__ void st_cluster_f1(uint32_t a, const float x) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(a), "f"(x) : "memory");
}
__device__ __forceinline__ void st_cluster_f4(uint32_t a, const float4 v) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void handoff_f() {
  tc::tmem_ld_wait();
  tc::tc_fence_before();
  tc::fence_async_smem();
  __syncthreads();
}

__device__ __forceinline__ float sqrt_approx(float x) {
  float y;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// opt_update_t (common.cuh) with the fast square root / division
template <int OPT>
__device__ __forceinline__ float opt_upd_x(const FedbGroup& G, const StepScalars& sc, float w, float g, float& m, float& v) {
  if (!FITX_FAST || OPT == FEDB_OPT_SGD) return opt_update_t<OPT>(G, sc, w, g, m, v);
  m = fmaf(G.beta1, m, (1.f - G.beta1) * g);
  v = fmaf(G.beta2, v, (1.f - G.beta2) * g * g);
  if (OPT == FEDB_OPT_ADAM_TORCH) return w - sc.lr_eff * __fdividef(m, sqrt_approx(v) * sc.inv_sqrt_bc2 + G.eps);
  return w - __fdividef(sc.lr_eff * m, sqrt_approx(v) + G.eps);
}
__device__ __forceinline__ float opt_upd_any(const FedbGroup& G, const StepScalars& sc, float w, float g, float& m, float& v) {
  if (G.opt == FEDB_OPT_ADAM_SKLEARN) return opt_upd_x<FEDB_OPT_ADAM_SKLEARN>(G, sc, w, g, m, v);
  if (G.opt == FEDB_OPT_ADAM_TORCH) return opt_upd_x<FEDB_OPT_ADAM_TORCH>(G, sc, w, g, m, v);
  return opt_upd_x<FEDB_OPT_SGD>(G, sc, w, g, m, v);
}

// split-fp16 writers into THIS CTA's panels (weights are held scaled by WS)
__device__ __forceinline__ void put_w2f(uint8_t* sm, int j, int col, const float4 w) {
  uint32_t h0, l0, h1, l1;
  split2(w.x * WS, w.y * WS, h0, l0);
  split2(w.z * WS, w.w * WS, h1, l1);
  const uint32_t off = tc::panel_off(j, col);
  *reinterpret_cast<uint2*>(sm + FX_W2 + off) = make_uint2(h0, h1);
  *reinterpret_cast<uint2*>(sm + FX_W2L + off) = make_uint2(l0, l1);
}
__device__ __forceinline__ void put_w1f(uint8_t* sm, int i, int col, const float4 w) {   // col in {0, 4, 8, 12}
  uint32_t h0, l0, h1, l1;
  split2(w.x * WS, w.y * WS, h0, l0);
  split2(w.z * WS, w.w * WS, h1, l1);
  *reinterpret_cast<uint2*>(sm + FX_W1 + tc::panel_off(i, col)) = make_uint2(h0, h1);
  *reinterpret_cast<uint2*>(sm + FX_W1 + tc::panel_off(i, 16 + col)) = make_uint2(l0, l1);
}

// All MMAs are issued by one thread.  Inside a group of three products the two small cross terms go first.
struct IssuerF {
  uint64_t kbase, mbase;  // K-major / MN-major descriptor of shared-memory offset 0
  uint32_t tmem;
  __device__ __forceinline__ uint64_t kd(int off) const { return kbase + (uint64_t)(off >> 4); }
  __device__ __forceinline__ uint64_t md(int off) const { return mbase + (uint64_t)(off >> 4); }
  __device__ __forceinline__ void fwd1() const {
    const uint32_t id = idesc_f16(128, 64, false, false);
    tc::mma_bf16(tmem + TF_D1, kd(FX_X + 32), kd(FX_W1), id, false);
    tc::mma_bf16(tmem + TF_D1, kd(FX_X), kd(FX_W1 + 32), id, true);
    tc::mma_bf16(tmem + TF_D1, kd(FX_X), kd(FX_W1), id, true);
  }
  __device__ __forceinline__ void fwd2(int N2) const {   // D2 = H1 W2_slice^T
    const uint32_t id = idesc_f16(128, N2, false, false);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      tc::mma_bf16(tmem + TF_D2, kd(FX_H1 + FP + 32 * k), kd(FX_W2 + 32 * k), id, k > 0);
      tc::mma_bf16(tmem + TF_D2, kd(FX_H1 + 32 * k), kd(FX_W2L + 32 * k), id, true);
      tc::mma_bf16(tmem + TF_D2, kd(FX_H1 + 32 * k), kd(FX_W2 + 32 * k), id, true);
    }
  }
  __device__ __forceinline__ void wgrad3(bool acc) const {   // DW3[j][0..7 | 8..15] += H2^T [dZ3 hi | dZ3 lo]
    const uint32_t id = idesc_f16(128, 16, true, true);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint64_t b = md(FX_X + 64 + 2048 * k);
      tc::mma_bf16(tmem + TF_DW3, md(FX_Z + 2 * FP + 2048 * k), b, id, acc || k > 0);
      tc::mma_bf16(tmem + TF_DW3, md(FX_Z + 2048 * k), b, id, true);
    }
  }
  __device__ __forceinline__ void dgrad2_wgrad2(int N2, bool acc) const {
    const uint32_t id4 = idesc_f16(128, 64, false, true);   // partial dH1 = dZ2_slice W2_slice
    // (fully unrolled under a uniform predicate: every descriptor is a compile-time constant; as run-time values each costs
    //  the single issuing thread a move into the uniform registers, ~25 cycles per MMA)
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      if (c < (N2 >> 4)) {
        const int p = c >> 2, k = c & 3, wrow = c * 16 * 128;
        tc::mma_bf16(tmem + TF_D1, kd(FX_Z + (2 + p) * FP + 32 * k), md(FX_W2 + wrow), id4, c > 0);
        tc::mma_bf16(tmem + TF_D1, kd(FX_Z + p * FP + 32 * k), md(FX_W2L + wrow), id4, true);
        tc::mma_bf16(tmem + TF_D1, kd(FX_Z + p * FP + 32 * k), md(FX_W2 + wrow), id4, true);
      }
    }
    const uint32_t id5 = idesc_f16(128, 64, true, true);    // dW2_slice += dZ2_slice^T H1
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      tc::mma_bf16(tmem + TF_DW2, md(FX_Z + 2 * FP + 2048 * k), md(FX_H1 + 2048 * k), id5, acc || k > 0);
      tc::mma_bf16(tmem + TF_DW2, md(FX_Z + 2048 * k), md(FX_H1 + FP + 2048 * k), id5, true);
      tc::mma_bf16(tmem + TF_DW2, md(FX_Z + 2048 * k), md(FX_H1 + 2048 * k), id5, true);
    }
  }
  __device__ __forceinline__ void wgrad1(bool acc) const {   // partial DW1[i][0..15 | 16..31] += dZ1p^T [X hi | X lo]
    const uint32_t id = idesc_f16(128, 32, true, true);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      tc::mma_bf16(tmem + TF_DW1, md(FX_H1 + FP + 2048 * k), md(FX_X + 2048 * k), id, acc || k > 0);
      tc::mma_bf16(tmem + TF_DW1, md(FX_H1 + 2048 * k), md(FX_X + 2048 * k), id, true);
    }
  }
};

}  // namespace

__global__ void __launch_bounds__(FT, 1) k_fit_tcx(FitXArgs a) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float red_f[4][9];        // per-warp partial sums of the four part-0 warps: loss + db3[8]
  __shared__ float red_ss[FT / 32];    // per-warp partial sums of squares: my slice (after its update) ...
  __shared__ float red_ss1[FT / 32];   // ... and layer 1 (before its update)
  __shared__ float s_ss_slice;         // ||W||^2 of my slice as it stands: the L2 term of the NEXT step's loss
  __shared__ StepScalars sc_s;
  __shared__ FedbGroup Gs;
  __shared__ int s_alive;

  cg::cluster_group cluster = cg::this_cluster();
  const int q = (int)cluster.block_rank();
  const int g = blockIdx.x / FNC;
  if (threadIdx.x < sizeof(FedbGroup) / 4)
    reinterpret_cast<int32_t*>(&Gs)[threadIdx.x] = reinterpret_cast<const int32_t*>(a.groups + g)[threadIdx.x];
  __syncthreads();
  const FedbGroup& G = Gs;
  if (a.fed_state && a.fed_state[G.fed].stop) return;      // cluster-uniform exits
  if (!a.active[g]) return;
  const int n = G.n_rows;
  const int bs = G.batch_size > 0 ? min(G.batch_size, n) : n;
  const int nbatches = (n + bs - 1) / bs;

  const uint32_t raw = tc::smem_u32(smem_raw);
  uint8_t* sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  const uint32_t sm_u = tc::smem_u32(sm);

  ShapeF s;
  s.d0 = G.dims[0]; s.h1 = G.dims[1]; s.h2 = G.dims[2]; s.C = G.dims[3];
  s.ld0 = ru4(s.d0); s.ld1 = ru4(s.h1); s.ld2 = ru4(s.h2);
  {
    const SlabLayout lay = slab_layout(G.dims, 3);
    s.w_off0 = lay.w_off[0]; s.b_off0 = lay.b_off[0]; s.w_off1 = lay.w_off[1]; s.b_off1 = lay.b_off[1];
    s.w_off2 = lay.w_off[2]; s.b_off2 = lay.b_off[2];
    const int S = (s.h2 + FNC - 1) / FNC;
    s.J0 = q * S;
    s.nj = max(0, min(S, s.h2 - s.J0));
    s.N2 = max(16, (s.nj + 15) & ~15);
  }
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int lane128 = (warp & 3) * 32 + lane;   // data row inside the tile = TMEM lane
  const int part = warp >> 2;                   // column quarter of every epilogue
  const int row = lane128;
  const int c1 = 16 * part;
  float* Pg = a.params + G.slab_off;
  float* Mg = a.m + G.slab_off;
  float* Vg = a.v + G.slab_off;
  const bool use_v = G.opt != FEDB_OPT_SGD;

  if (warp == 0) tc::tmem_alloc<512>(&tmem_slot);
  if (tid == 0) {
    tc::mbar_init(&bar, 1);
    tc::fence_mbar_init();
  }
  for (int i = tid; i < (FX_ZP - FX_X) / 16; i += FT) *reinterpret_cast<uint4*>(sm + FX_X + i * 16) = make_uint4(0, 0, 0, 0);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  const IssuerF mma{tc::make_desc(sm_u, 0, 1024), tc::make_desc(sm_u, FP, 1024), tmem};
  uint32_t ph = 0;
  float* W3s = reinterpret_cast<float*>(sm + FX_W3);
  float* b3s = reinterpret_cast<float*>(sm + FX_B3);     // [0..7] b3, [8..15] m, [16..23] v
  float* M1 = reinterpret_cast<float*>(sm + FX_M1);      // [0] w, [1024] m, [2048] v ; element i * 16 + col
  float* zpart = reinterpret_cast<float*>(sm + FX_ZP);
  float* g1part = reinterpret_cast<float*>(sm + FX_G1);
  uint8_t* const h1_hi = sm + FX_H1;
  uint8_t* const h1_lo = sm + FX_H1 + FP;

  // ---- my slice of W2 / b2 / W3 from the master copy: fp32 weights and moments into tensor memory, split operands into
  //      the panels (thread = unit j of the slice x a 16-column part); layer 1 and b3 replicated in shared memory
  {
    const int j = lane128;
    const bool mine = j < s.nj;
    const int e0 = s.w_off1 + (s.J0 + (mine ? j : 0)) * s.ld1;
    float ss0 = 0.f;
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      const int c0 = c1 + 8 * hf;
      float w8[8], m8[8], v8[8];
#pragma unroll
      for (int i4 = 0; i4 < 2; ++i4) {
        const int col = c0 + 4 * i4;
        const bool in = mine && col < s.ld1;
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 w = in ? __ldcg(reinterpret_cast<const float4*>(Pg + e0 + col)) : z4;
        const float4 mm = in ? __ldcg(reinterpret_cast<const float4*>(Mg + e0 + col)) : z4;
        const float4 vv = (in && use_v) ? __ldcg(reinterpret_cast<const float4*>(Vg + e0 + col)) : z4;
        w8[4 * i4] = w.x; w8[4 * i4 + 1] = w.y; w8[4 * i4 + 2] = w.z; w8[4 * i4 + 3] = w.w;
        m8[4 * i4] = mm.x; m8[4 * i4 + 1] = mm.y; m8[4 * i4 + 2] = mm.z; m8[4 * i4 + 3] = mm.w;
        v8[4 * i4] = vv.x; v8[4 * i4 + 1] = vv.y; v8[4 * i4 + 2] = vv.z; v8[4 * i4 + 3] = vv.w;
        ss0 += w.x * w.x + w.y * w.y + w.z * w.z + w.w * w.w;
      }
      if (part == 3 && hf == 1) {   // column 63 (beyond h1 <= 63, a zero so far) carries b2: it multiplies H1's constant-1 column
        const int e = s.b_off1 + s.J0 + (mine ? j : 0);
        w8[7] = mine ? __ldcg(Pg + e) : 0.f;
        m8[7] = mine ? __ldcg(Mg + e) : 0.f;
        v8[7] = (mine && use_v) ? __ldcg(Vg + e) : 0.f;
      }
      tc::tmem_st8(tl + TF_W + c0, w8);
      tc::tmem_st8(tl + TF_M + c0, m8);
      tc::tmem_st8(tl + TF_V + c0, v8);
      put_w2f(sm, j, c0, make_float4(w8[0], w8[1], w8[2], w8[3]));
      put_w2f(sm, j, c0 + 4, make_float4(w8[4], w8[5], w8[6], w8[7]));
    }
    if (part == 0) {   // head weights of classes 0, 1 of unit j with their moments
      float s3[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) s3[i] = 0.f;
#pragma unroll
      for (int k = 0; k < 2; ++k)
        if (mine && k < s.C) {
          const int e = s.w_off2 + k * s.ld2 + s.J0 + j;
          s3[k] = __ldcg(Pg + e);
          s3[2 + k] = __ldcg(Mg + e);
          s3[4 + k] = use_v ? __ldcg(Vg + e) : 0.f;
          ss0 += s3[k] * s3[k];
        }
      tc::tmem_st8(tl + TF_S3, s3);
    }
    if (part == 2 && mine)
      for (int k = 2; k < s.C; ++k) {
        const float w = __ldcg(Pg + s.w_off2 + k * s.ld2 + s.J0 + j);
        ss0 += w * w;
      }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss0 += __shfl_xor_sync(0xffffffffu, ss0, o);
    if (lane == 0) red_ss[warp] = ss0;
    tc::tmem_st_wait();
    for (int jj = tid; jj < s.nj; jj += FT)
      for (int k = 0; k < s.C; ++k) W3s[jj * 8 + k] = __ldcg(Pg + s.w_off2 + k * s.ld2 + s.J0 + jj);
    for (int t = tid; t < s.h1 * 16; t += FT) {
      const int i = t >> 4, col = t & 15;
      int e = -1;
      if (col < s.d0) e = s.w_off0 + i * s.ld0 + col;
      else if (col == 15) e = s.b_off0 + i;
      M1[t] = e >= 0 ? __ldcg(Pg + e) : 0.f;
      M1[1024 + t] = e >= 0 ? __ldcg(Mg + e) : 0.f;
      M1[2048 + t] = (e >= 0 && use_v) ? __ldcg(Vg + e) : 0.f;
    }
    if (tid < s.C) {
      b3s[tid] = __ldcg(Pg + s.b_off2 + tid);
      b3s[8 + tid] = __ldcg(Mg + s.b_off2 + tid);
      b3s[16 + tid] = use_v ? __ldcg(Vg + s.b_off2 + tid) : 0.f;
    }
    __syncthreads();
    if (tid == 0) {
      float t = 0.f;
      for (int w = 0; w < FT / 32; ++w) t += red_ss[w];
      s_ss_slice = t;
    }
    for (int t = tid; t < s.h1 * 4; t += FT) {
      const int i = t >> 2, c4 = t & 3;
      put_w1f(sm, i, 4 * c4, *reinterpret_cast<const float4*>(M1 + i * 16 + 4 * c4));
    }
    tc::fence_async_smem();
  }
  __syncthreads();
  cluster_barrier();   // every CTA of the cluster is running: its shared memory may be written from here on

#define FSTAMP(i) do { if (a.prof && g == 0 && tid == 0 && ep == 0 && b == 0) { long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); a.prof[q * 32 + (i)] = t_; } } while (0)
#define FIT_MMA_PHASE(ISSUE)             \
  do {                                   \
    if (warp == 0 && tc::elect_one()) {      \
      tc::tc_fence_after();              \
      ISSUE;                             \
      tc::mma_commit(&bar);              \
    }                                    \
    tc::mbar_wait(&bar, ph);             \
    ph ^= 1;                             \
    tc::tc_fence_after();                \
  } while (0)

  const int nch2 = s.N2 >> 4;
  const int ch_lo = part * (nch2 >> 2) + min(part, nch2 & 3), ch_hi = ch_lo + (nch2 >> 2) + (part < (nch2 & 3) ? 1 : 0);
  int t_step = a.step[g];
  const int ep_base = a.epoch[g];   // epochs completed before this launch (also keys the reshuffle)
  int eps_run = 0, no_imp = a.no_improve[g];
  double best = a.best_loss[g], last = 0.0;
  float last_batch_loss = 0.f;
  unsigned tile_ctr = 0;        // parity of the partial-logit buffers
  int bits = 1;
  while ((1u << bits) < (uint32_t)n) ++bits;
  const int half_bits = (bits + 1) >> 1;
  const int32_t* idxsrc = a.shuffle ? a.idx_epoch : a.idx_in;
  bool alive = true;

  for (int ep = 0; ep < a.max_epochs && alive; ++ep) {
    // ---- per-epoch reshuffle (SK:704-718): a keyed bijection of the shard, a quarter per CTA, visible after the barrier
    if (a.shuffle) {
      const uint32_t ka = a.s0 ^ (uint32_t)(G.client * 0x9E3779B1u), kb = a.s1 + (uint32_t)(ep_base + ep);
      for (uint32_t i = q * FT + tid; i < (uint32_t)n; i += FNC * FT)
        a.idx_epoch[G.idx_off + i] = a.idx_in[G.idx_off + feistel_perm(i, (uint32_t)n, half_bits, ka, kb)];
      __threadfence();
      cluster_barrier();
    }
    double acc_loss = 0.0;   // thread 0: sum over mini-batches of (row losses + 0.5 alpha ||W||^2), SK:353-361
    // Input rows run two tiles ahead of the math: the row INDEX of the tile after next and the DATA of the next tile are in
    // flight while a tile computes, so neither of the two dependent loads is ever waited for.
    auto pos_of = [&](int bb, int tt) -> int {   // position in the shard of this thread's row of tile tt of mini-batch bb, or -1
      const int b0_ = bb * bs;
      if (b0_ >= n) return -1;
      const int r_ = tt * FR + row;
      return r_ < min(bs, n - b0_) ? b0_ + r_ : -1;
    };
    auto advance = [&](int& bb, int& tt) {
      const int nt_ = (min(bs, max(0, n - bb * bs)) + FR - 1) / FR;
      if (++tt >= nt_) { tt = 0; ++bb; }
    };
    float4 nx = make_float4(0.f, 0.f, 0.f, 0.f);
    int nlabel = 0, ngrow = -1, pb = 0, pt = 0;   // (pb, pt): the tile whose row index is in ngrow
    if (row < min(bs, n)) {
      const int grow = __ldcg(idxsrc + G.idx_off + row);
      nx = __ldg(reinterpret_cast<const float4*>(a.X + (size_t)grow * a.ldx) + part);
      nlabel = __ldg(a.y + grow);
    }
    advance(pb, pt);
    {
      const int p1 = pos_of(pb, pt);
      if (p1 >= 0) ngrow = __ldcg(idxsrc + G.idx_off + p1);
    }

    for (int b = 0; b < nbatches; ++b) {
      const int b0 = b * bs, nb = min(bs, n - b0);
      const int ntiles = (nb + FR - 1) / FR;
      const float inv_n = 1.f / (float)nb, gscale = inv_n * (1.f / DZS);
      t_step += 1;
      if (tid == 0) sc_s = step_scalars_fast(G, t_step);   // (read after the block barrier that follows the tiles)
      float loss_acc = 0.f, db3_acc[FEDB_MAX_CLASSES];
#pragma unroll
      for (int k = 0; k < FEDB_MAX_CLASSES; ++k) db3_acc[k] = 0.f;

      FSTAMP(0);
      for (int tile = 0; tile < ntiles; ++tile) {
        const bool acc = tile > 0;
        const int r = tile * FR + row;
        const bool valid = r < nb;
        // ---- this thread's four input columns of data row r (fetched one tile ahead) ---------------------------------
        int label = 0;
        {
          float f[4] = {nx.x, nx.y, nx.z, nx.w};
          label = nlabel;
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (4 * part + i >= s.d0 || !valid) f[i] = 0.f;
          if (part == 3) f[3] = valid ? 1.f : 0.f;   // column 15: the bias input
          uint32_t h0, l0, h1, l1;
          split2(f[0], f[1], h0, l0);
          split2(f[2], f[3], h1, l1);
          *reinterpret_cast<uint2*>(sm + FX_X + tc::panel_off(row, 4 * part)) = make_uint2(h0, h1);
          *reinterpret_cast<uint2*>(sm + FX_X + tc::panel_off(row, 16 + 4 * part)) = make_uint2(l0, l1);
        }
        {   // the next tile's data (its index arrived a tile ago) and the index of the tile after it
          nx = make_float4(0.f, 0.f, 0.f, 0.f);
          nlabel = 0;
          if (ngrow >= 0) {
            nx = __ldg(reinterpret_cast<const float4*>(a.X + (size_t)ngrow * a.ldx) + part);
            nlabel = __ldg(a.y + ngrow);
          }
          advance(pb, pt);
          const int p2 = pos_of(pb, pt);
          ngrow = -1;
          if (p2 >= 0) ngrow = __ldcg(idxsrc + G.idx_off + p2);
        }
        handoff_f();
        if (tile == 0) FSTAMP(1);
        // ---- layer 1 (replicated in the four CTAs) ----------------------------------------------------------
        FIT_MMA_PHASE(mma.fwd1());
        if (tile == 0) FSTAMP(2);
        uint32_t mask1 = 0;
        {
          float v[16];
          tc::tmem_ld16(tl + TF_D1 + c1, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            if (v[i] > 0.f) { mask1 |= 1u << i; v[i] = fminf(v[i] * INV_WS, HMAX); } else v[i] = 0.f;
          }
          if (part == 3) { v[15] = valid ? 1.f : 0.f; mask1 &= ~(1u << 15); }   // column 63: the bias input of layer 2
          st_panel8(h1_hi, h1_lo, row, c1, v);
          st_panel8(h1_hi, h1_lo, row, c1 + 8, v + 8);
        }
        handoff_f();
        if (tile == 0) FSTAMP(3);
        // ---- my slice of layer 2, partial logits -------------------------------------------------------------
        FIT_MMA_PHASE(mma.fwd2(s.N2));
        if (tile == 0) FSTAMP(4);
        float z[FEDB_MAX_CLASSES], dz3[FEDB_MAX_CLASSES];
#pragma unroll
        for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = dz3[k] = 0.f;
#pragma unroll 1
        for (int ch = ch_lo; ch < ch_hi; ++ch) {
          const int c0 = ch << 4;
          float v[16];
          tc::tmem_ld16(tl + TF_D2 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = fminf(fmaxf(v[i], 0.f) * INV_WS, HMAX);
          if (s.C == 1) {          // the scripts' sigmoid head: one output
#pragma unroll
            for (int i = 0; i < 16; ++i) z[0] = fmaf(v[i], W3s[(c0 + i) * 8], z[0]);
          } else if (s.C == 2) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float2 w = *reinterpret_cast<const float2*>(W3s + (c0 + i) * 8);
              z[0] = fmaf(v[i], w.x, z[0]);
              z[1] = fmaf(v[i], w.y, z[1]);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float4 w0 = *reinterpret_cast<const float4*>(W3s + (c0 + i) * 8), w1 = *reinterpret_cast<const float4*>(W3s + (c0 + i) * 8 + 4);
              z[0] = fmaf(v[i], w0.x, z[0]); z[1] = fmaf(v[i], w0.y, z[1]); z[2] = fmaf(v[i], w0.z, z[2]); z[3] = fmaf(v[i], w0.w, z[3]);
              z[4] = fmaf(v[i], w1.x, z[4]); z[5] = fmaf(v[i], w1.y, z[5]); z[6] = fmaf(v[i], w1.z, z[6]); z[7] = fmaf(v[i], w1.w, z[7]);
            }
          }
          uint8_t* pz = sm + FX_Z + (c0 >> 6) * FP;
          st_panel8(pz, pz + 2 * FP, row, c0 & 63, v);
          st_panel8(pz, pz + 2 * FP, row, (c0 & 63) + 8, v + 8);
        }
        // the row's four parts add up through 32 spare TMEM columns; part 0 sends the slice's partial logits to every CTA
        tc::tmem_st8(tl + TF_ZX + 8 * part, z);
        tc::tmem_st_wait();
        tc::tc_fence_before();
        __syncthreads();
        tc::tc_fence_after();
        if (tile == 0) FSTAMP(5);
        const int par = (int)(tile_ctr & 1u);
        ++tile_ctr;
        if (part == 0) {
          float z0[16], z1[16];
          tc::tmem_ld16(tl + TF_ZX, z0);
          tc::tmem_ld16(tl + TF_ZX + 16, z1);
          tc::tmem_ld_wait();
          const uint32_t dst = sm_u + FX_ZP + (uint32_t)(((par * FNC + q) * 8) * FR + row) * 4u;
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
            if (k < s.C) {
              const float zk = (z0[k] + z0[8 + k]) + (z1[k] + z1[8 + k]);
#pragma unroll
              for (int rk = 0; rk < FNC; ++rk) st_cluster_f1(mapa(dst + (uint32_t)(k * FR) * 4u, rk), zk);
            }
        }
        cluster_barrier();
        if (tile == 0) FSTAMP(6);
        {
          const float* zp = zpart + (size_t)par * FNC * 8 * FR + row;
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
            z[k] = k < s.C ? ((zp[(0 * 8 + k) * FR] + zp[(1 * 8 + k) * FR]) + (zp[(2 * 8 + k) * FR] + zp[(3 * 8 + k) * FR])) + b3s[k] : 0.f;
        }
        float loss = 0.f;
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
                dz3[k] = (expf(z[k] - lse) - (k == label ? 1.f : 0.f)) * DZS;
              }
          } else {
            const float p = 1.f / (1.f + expf(-z[0]));
            const float eps = 1.1920929e-07f;   // SKB:273-274 clips probabilities at dtype eps
            const float pc = fminf(fmaxf(p, eps), 1.f - eps), yt = (float)label;
            if (part == 0) loss = -(yt * logf(pc) + (1.f - yt) * logf(1.f - pc));   // (only part 0 books the loss)
            dz3[0] = (p - yt) * DZS;
          }
        }
        if (part == 0) {   // one thread per row books the loss / db3 and publishes dZ3 as [hi | lo] next to X
          loss_acc += loss;
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k) db3_acc[k] += dz3[k];
          uint4 dh, dl;
          split8(dz3, dh, dl);
          *reinterpret_cast<uint4*>(sm + FX_X + tc::panel_off(row, 32)) = dh;
          *reinterpret_cast<uint4*>(sm + FX_X + tc::panel_off(row, 40)) = dl;
        }
        handoff_f();
        if (tile == 0) FSTAMP(7);
        // ---- head weight gradient of my slice (needs H2), then dZ2 overwrites H2 ---------------------------------
        // (measured: running the first chunk of the next epilogue under these MMAs and waiting only before its stores made the
        //  FOLLOWING MMA phases slower by more than it saved: 18.3 against 17.2 us for the two tiles)
        FIT_MMA_PHASE(mma.wgrad3(acc));
        if (tile == 0) FSTAMP(8);
#pragma unroll 1
        for (int ch = ch_lo; ch < ch_hi; ++ch) {
          const int c0 = ch << 4;
          float v[16], d[16];
          tc::tmem_ld16(tl + TF_D2 + c0, v);
          tc::tmem_ld_wait();
          if (s.C == 1) {
#pragma unroll
            for (int i = 0; i < 16; ++i) d[i] = dz3[0] * W3s[(c0 + i) * 8];
          } else if (s.C == 2) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float2 w = *reinterpret_cast<const float2*>(W3s + (c0 + i) * 8);
              d[i] = fmaf(dz3[1], w.y, dz3[0] * w.x);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float4 w0 = *reinterpret_cast<const float4*>(W3s + (c0 + i) * 8), w1 = *reinterpret_cast<const float4*>(W3s + (c0 + i) * 8 + 4);
              d[i] = dz3[0] * w0.x + dz3[1] * w0.y + dz3[2] * w0.z + dz3[3] * w0.w + dz3[4] * w1.x + dz3[5] * w1.y + dz3[6] * w1.z + dz3[7] * w1.w;
            }
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) d[i] = v[i] > 0.f ? d[i] : 0.f;
          uint8_t* pz = sm + FX_Z + (c0 >> 6) * FP;
          st_panel8(pz, pz + 2 * FP, row, c0 & 63, d);
          st_panel8(pz, pz + 2 * FP, row, (c0 & 63) + 8, d + 8);
        }
        handoff_f();
        if (tile == 0) FSTAMP(9);
        // ---- partial dH1 = dZ2_slice W2_slice ; dW2_slice += dZ2_slice^T H1 -----------------------------------------
        FIT_MMA_PHASE(mma.dgrad2_wgrad2(s.N2, acc));
        if (tile == 0) FSTAMP(10);
        {
          float v[16];
          tc::tmem_ld16(tl + TF_D1 + c1, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = ((mask1 >> i) & 1u) ? v[i] * INV_WS : 0.f;
          st_panel8(h1_hi, h1_lo, row, c1, v);        // (partial) dZ1 takes H1's place
          st_panel8(h1_hi, h1_lo, row, c1 + 8, v + 8);
        }
        handoff_f();
        if (tile == 0) FSTAMP(11);
        // (measured: a second X panel, so that these MMAs go out with the next tile's first group and its rows are staged here,
        //  saves a wait per tile, but staging the next tile here and the longer first MMA group cost more than that:
        //  16.5 against 16.0 us for the two tiles)
        FIT_MMA_PHASE(mma.wgrad1(acc));
        if (tile == 0) FSTAMP(12);
      }  // tiles
      FSTAMP(13);

      // ---- layer-1 partial gradients -> every CTA of the cluster (they are summed after the barrier below) -----------------
      if (part == 1) {
        float v[16], w[16];
        tc::tmem_ld16(tl + TF_DW1, v);
        tc::tmem_ld16(tl + TF_DW1 + 16, w);
        tc::tmem_ld_wait();
        if (lane128 < s.h1) {
          const uint32_t dst = sm_u + FX_G1 + (uint32_t)(q * G1_FLOATS + lane128 * 16) * 4u;
#pragma unroll
          for (int i4 = 0; i4 < 4; ++i4) {
            const float4 gq = make_float4((v[4 * i4] + w[4 * i4]) * gscale, (v[4 * i4 + 1] + w[4 * i4 + 1]) * gscale,
                                          (v[4 * i4 + 2] + w[4 * i4 + 2]) * gscale, (v[4 * i4 + 3] + w[4 * i4 + 3]) * gscale);
#pragma unroll
            for (int rk = 0; rk < FNC; ++rk) st_cluster_f4(mapa(dst + 16u * i4, rk), gq);
          }
        }
      }
      // loss and db3 (identical in the four CTAs: every CTA saw every row): warp shuffle, then the four part-0 warps in order
      if (part == 0) {
        float vals[9];
        vals[0] = loss_acc;
#pragma unroll
        for (int k = 0; k < 8; ++k) vals[k + 1] = db3_acc[k] * gscale;
#pragma unroll
        for (int i = 0; i < 9; ++i) {
          float x = vals[i];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
          if (lane == 0) red_f[warp][i] = x;
        }
      }
      if (tid == 0) {
        const float t = s_ss_slice;   // my slice's ||W||^2 before this step's update (kept from the previous update)
#pragma unroll
        for (int rk = 0; rk < FNC; ++rk) st_cluster_f1(mapa(sm_u + FX_G1 + (uint32_t)(q * G1_FLOATS + 64 * 16) * 4u, rk), t);
      }
      __syncthreads();
      cluster_arrive();   // my layer-1 partials and slice sum are on their way; the wait follows the slice optimizer
      FSTAMP(14);
      const StepScalars sc = sc_s;
      const float al = G.alpha * inv_n;

      // ---- my slice: Adam straight from the accumulators (complete: this CTA saw every row) on the state in tensor memory,
      //      new operands into my panels.  The optimizer kind is decided ONCE around straight-line code, so the independent
      //      chains of a thread overlap.
      float ss = 0.f;   // sum of squares of the NEW weights: the L2 term of the next step's loss
      {
        const int j = lane128;
        const bool mine = j < s.nj;
        auto update_all = [&](auto tag) {
          constexpr int OPT = decltype(tag)::value;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const int c0 = c1 + 8 * hf;
            float g8[8], w8[8], m8[8], v8[8];
            tc::tmem_ld8(tl + TF_DW2 + c0, g8);
            tc::tmem_ld8(tl + TF_W + c0, w8);
            tc::tmem_ld8(tl + TF_M + c0, m8);
            tc::tmem_ld8(tl + TF_V + c0, v8);
            tc::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const bool is_b2 = hf == 1 && i == 7 && part == 3;   // column 63: no L2 term, not part of ||W||^2
              const float w = w8[i], gr = g8[i] * gscale;
              const float wn = opt_upd_x<OPT>(G, sc, w, is_b2 ? gr : fmaf(al, w, gr), m8[i], v8[i]);
              ss += (is_b2 || !mine) ? 0.f : wn * wn;
              // (columns h1 .. 62 are padding: gradient and weight are exact zeros and stay so; lanes beyond the slice are
              //  kept at zero whatever the accumulator rows there hold)
              w8[i] = mine ? wn : 0.f;
              m8[i] = mine ? m8[i] : 0.f;
              v8[i] = mine ? v8[i] : 0.f;
            }
            tc::tmem_st8(tl + TF_W + c0, w8);
            tc::tmem_st8(tl + TF_M + c0, m8);
            tc::tmem_st8(tl + TF_V + c0, v8);
            put_w2f(sm, j, c0, make_float4(w8[0], w8[1], w8[2], w8[3]));
            put_w2f(sm, j, c0 + 4, make_float4(w8[4], w8[5], w8[6], w8[7]));
          }
          if (part == 0) {   // head weights of classes 0, 1: dW3[k][j] = DW3[j][k] + DW3[j][8 + k]
            float s3[8], g3[16];
            tc::tmem_ld8(tl + TF_S3, s3);
            tc::tmem_ld16(tl + TF_DW3, g3);
            tc::tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              const float w = s3[k];
              const float wn = opt_upd_x<OPT>(G, sc, w, fmaf(al, w, (g3[k] + g3[8 + k]) * gscale), s3[2 + k], s3[4 + k]);
              const bool on = mine && k < s.C;
              ss += on ? wn * wn : 0.f;
              s3[k] = on ? wn : 0.f;
              s3[2 + k] = on ? s3[2 + k] : 0.f;
              s3[4 + k] = on ? s3[4 + k] : 0.f;
              if (on) W3s[j * 8 + k] = wn;
            }
            tc::tmem_st8(tl + TF_S3, s3);
          }
        };
        FSTAMP(18);
        if (G.opt == FEDB_OPT_ADAM_SKLEARN) update_all(std::integral_constant<int, FEDB_OPT_ADAM_SKLEARN>{});
        else if (G.opt == FEDB_OPT_ADAM_TORCH) update_all(std::integral_constant<int, FEDB_OPT_ADAM_TORCH>{});
        else update_all(std::integral_constant<int, FEDB_OPT_SGD>{});
        FSTAMP(19);
        if (part == 2 && s.C > 2) {   // head weights of further classes (rare): state in global memory
          // (tcgen05.ld is warp-collective: every lane of the part-2 warps executes it, whether or not its unit is in the slice)
          float g3[16];
          tc::tmem_ld16(tl + TF_DW3, g3);
          tc::tmem_ld_wait();
          for (int k = 2; k < s.C && mine; ++k) {
            const int e = s.w_off2 + k * s.ld2 + s.J0 + j;
            float w0 = Pg[e], mm = Mg[e], vv = use_v ? Vg[e] : 0.f, gk = 0.f;
#pragma unroll
            for (int t = 2; t < 8; ++t) gk = (t == k) ? (g3[t] + g3[8 + t]) * gscale : gk;
            const float w = opt_upd_any(G, sc, w0, fmaf(al, w0, gk), mm, vv);
            ss += w * w;
            Pg[e] = w; Mg[e] = mm;
            if (use_v) Vg[e] = vv;
            W3s[j * 8 + k] = w;
          }
        }
        tc::tmem_st_wait();
        FSTAMP(20);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
      if (lane == 0) red_ss[warp] = ss;    // (summed by thread 0 after the block barrier of the layer-1 phase)
      FSTAMP(15);
      cluster_wait();      // every CTA's layer-1 partials and slice sums have landed everywhere
      FSTAMP(16);

      // ---- layer 1 and b3: the same update in all four CTAs, on their replicated fp32 copies in shared memory ------------------
      float ss1 = 0.f;
      if (tid < s.h1 * 4) {
        const int i = tid >> 2, c4 = tid & 3, o = i * 16 + 4 * c4;
        float4 gs = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int rk = 0; rk < FNC; ++rk) {
          const float4 p = *reinterpret_cast<const float4*>(g1part + rk * G1_FLOATS + o);
          gs.x += p.x; gs.y += p.y; gs.z += p.z; gs.w += p.w;
        }
        const float gk[4] = {gs.x, gs.y, gs.z, gs.w};
        float wn[4];
#pragma unroll
        for (int l = 0; l < 4; ++l) {
          const int col = 4 * c4 + l;
          float w = M1[o + l], mm = M1[1024 + o + l], vv = M1[2048 + o + l];
          if (col < s.d0) {
            ss1 += w * w;
            w = opt_upd_any(G, sc, w, fmaf(al, w, gk[l]), mm, vv);
          } else if (col == 15) {
            w = opt_upd_any(G, sc, w, gk[l], mm, vv);
          }
          M1[o + l] = w; M1[1024 + o + l] = mm; M1[2048 + o + l] = vv;
          wn[l] = w;
        }
        put_w1f(sm, i, 4 * c4, make_float4(wn[0], wn[1], wn[2], wn[3]));
      }
      if (tid >= FT - 32 && lane < s.C) {   // b3 (its gradient is the same in every CTA)
        float w = b3s[lane], mm = b3s[8 + lane], vv = b3s[16 + lane];
        const float gr = (red_f[0][1 + lane] + red_f[1][1 + lane]) + (red_f[2][1 + lane] + red_f[3][1 + lane]);
        w = opt_upd_any(G, sc, w, gr, mm, vv);
        b3s[lane] = w; b3s[8 + lane] = mm; b3s[16 + lane] = vv;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ss1 += __shfl_xor_sync(0xffffffffu, ss1, o);
      if (lane == 0) red_ss1[warp] = ss1;
      tc::fence_async_smem();
      __syncthreads();
      if (tid == 0) {
        float t1 = 0.f, t2 = 0.f;
        for (int w = 0; w < FT / 32; ++w) { t1 += red_ss1[w]; t2 += red_ss[w]; }
        s_ss_slice = t2;
        const float ss_all = ((g1part[0 * G1_FLOATS + 1024] + g1part[1 * G1_FLOATS + 1024]) + (g1part[2 * G1_FLOATS + 1024] + g1part[3 * G1_FLOATS + 1024])) + t1;
        const float rowloss = (red_f[0][0] + red_f[1][0]) + (red_f[2][0] + red_f[3][0]);
        acc_loss += (double)rowloss + 0.5 * (double)G.alpha * (double)ss_all;
        last_batch_loss = rowloss * inv_n;
      }
      FSTAMP(17);
    }  // mini-batches

    // ---- end of the epoch: loss curve and the tol / n_iter_no_change rule (SK:750-777), identical in every CTA ----------
    eps_run = ep + 1;
    if (tid == 0) {
      const double loss = acc_loss / (double)n;
      last = loss;
      if (loss > best - a.tol) no_imp += 1; else no_imp = 0;
      if (loss < best) best = loss;
      s_alive = (no_imp > a.n_iter_no_change || ep_base + eps_run >= a.max_iter) ? 0 : 1;
    }
    __syncthreads();
    alive = s_alive != 0;
    __syncthreads();
  }
#undef FIT_MMA_PHASE
#undef FSTAMP

  // ---- write back what lived on chip: every CTA its slice (weights and moments from tensor memory); layer 1, b3 and their
  //      moments, the counters and the loss curve state by CTA 0 ----------------------------------------------------------------
  {
    const int j = lane128;
    const bool mine = j < s.nj;
    const int e0 = s.w_off1 + (s.J0 + (mine ? j : 0)) * s.ld1;
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      const int c0 = c1 + 8 * hf;
      float w8[8], m8[8], v8[8];
      tc::tmem_ld8(tl + TF_W + c0, w8);
      tc::tmem_ld8(tl + TF_M + c0, m8);
      tc::tmem_ld8(tl + TF_V + c0, v8);
      tc::tmem_ld_wait();
      if (part == 3 && hf == 1) {
        if (mine) {
          const int e = s.b_off1 + s.J0 + j;
          Pg[e] = w8[7]; Mg[e] = m8[7];
          if (use_v) Vg[e] = v8[7];
        }
        w8[7] = m8[7] = v8[7] = 0.f;   // column 63 of the master row (present when ld1 = 64) is padding
      }
#pragma unroll
      for (int i4 = 0; i4 < 2; ++i4) {
        const int col = c0 + 4 * i4;
        if (mine && col < s.ld1) {
          *reinterpret_cast<float4*>(Pg + e0 + col) = make_float4(w8[4 * i4], w8[4 * i4 + 1], w8[4 * i4 + 2], w8[4 * i4 + 3]);
          *reinterpret_cast<float4*>(Mg + e0 + col) = make_float4(m8[4 * i4], m8[4 * i4 + 1], m8[4 * i4 + 2], m8[4 * i4 + 3]);
          if (use_v) *reinterpret_cast<float4*>(Vg + e0 + col) = make_float4(v8[4 * i4], v8[4 * i4 + 1], v8[4 * i4 + 2], v8[4 * i4 + 3]);
        }
      }
    }
    if (part == 0) {
      float s3[8];
      tc::tmem_ld8(tl + TF_S3, s3);
      tc::tmem_ld_wait();
#pragma unroll
      for (int k = 0; k < 2; ++k)
        if (mine && k < s.C) {
          const int e = s.w_off2 + k * s.ld2 + s.J0 + j;
          Pg[e] = s3[k]; Mg[e] = s3[2 + k];
          if (use_v) Vg[e] = s3[4 + k];
        }
    }
  }
  if (q == 0) {
    for (int t = tid; t < s.h1 * 16; t += FT) {
      const int i = t >> 4, col = t & 15;
      int e = -1;
      if (col < s.d0) e = s.w_off0 + i * s.ld0 + col;
      else if (col == 15) e = s.b_off0 + i;
      if (e >= 0) {
        Pg[e] = M1[t]; Mg[e] = M1[1024 + t];
        if (use_v) Vg[e] = M1[2048 + t];
      }
    }
    if (tid < s.C) {
      Pg[s.b_off2 + tid] = b3s[tid]; Mg[s.b_off2 + tid] = b3s[8 + tid];
      if (use_v) Vg[s.b_off2 + tid] = b3s[16 + tid];
    }
    if (tid == 0) {
      a.step[g] = t_step;
      a.epoch[g] = ep_base + eps_run;
      a.no_improve[g] = no_imp;
      a.best_loss[g] = best;
      a.last_loss[g] = last;
      if (!alive) a.active[g] = 0;
      if (a.loss_out) a.loss_out[g] = last_batch_loss;
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(tmem);
  cluster_barrier();   // nobody leaves while a peer may still write its shared memory
}

}  // namespace fedb

using namespace fedb;

extern "C" int64_t fedb_fitx_smem_bytes(void) { return FX_END + 1024; }
// 1 if the tensor-core fit kernel runs this model: 2 hidden ReLU layers, in <= 15, h1 <= 63, h2 <= 512, out <= 8
extern "C" int32_t fedb_fitx_fits(int32_t d0, int32_t h1, int32_t h2, int32_t C) {
  return (d0 >= 1 && d0 <= 15 && h1 >= 1 && h1 <= 63 && h2 >= 1 && h2 <= 128 * FNC && C >= 1 && C <= 8) ? 1 : 0;
}

// The whole local fit of every client (up to max_epochs further epochs) in one launch; arguments as fedb_fit_cluster
// (include/fedb200.h).  Clients whose active flag is 0 or whose federation has stopped return at once.
extern "C" int fedb_fit_tcx(const FedbGroup* groups, int32_t n_groups, float* params, float* m, float* v, const float* X, int32_t ldx,
                            const int32_t* y, const int32_t* idx_in, int32_t* idx_epoch, int32_t* step, int32_t* active,
                            const FedbFedState* fed_state, int32_t* epoch, int32_t* no_improve, double* best_loss, double* last_loss,
                            float* loss_out, double tol, int32_t n_iter_no_change, int32_t max_iter, int32_t max_epochs,
                            int32_t shuffle, uint64_t seed, long long* prof, void* stream) {
  FEDB_REQUIRE(n_groups > 0 && max_epochs >= 0, "bad sizes");
  FEDB_REQUIRE(ldx >= 16 && ldx % 4 == 0, "feature rows must be padded to >= 16 floats");
  FEDB_REQUIRE(m != nullptr && v != nullptr, "moment arenas required");
  FitXArgs a{};
  a.groups = groups; a.X = X; a.y = y; a.idx_in = idx_in; a.idx_epoch = idx_epoch; a.params = params; a.m = m; a.v = v;
  a.step = step; a.active = active; a.fed_state = fed_state; a.epoch = epoch; a.no_improve = no_improve; a.best_loss = best_loss;
  a.last_loss = last_loss; a.loss_out = loss_out; a.tol = tol; a.ldx = ldx; a.n_iter_no_change = n_iter_no_change;
  a.max_iter = max_iter; a.max_epochs = max_epochs; a.shuffle = shuffle; a.s0 = (uint32_t)seed; a.s1 = (uint32_t)(seed >> 32);
  a.prof = prof;
  const int smem = FX_END + 1024;
  FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_fit_tcx, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(n_groups * FNC);
  cfg.blockDim = dim3(FT);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = FNC;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  FEDB_CUDA_OK(cudaLaunchKernelEx(&cfg, k_fit_tcx, a));
  return FEDB_OK;
}
