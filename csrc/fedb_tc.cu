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
  float* scratch;   // per-group gradient records (FedbGroup::scratch_off; >= cluster_size records each)
  int ldx, batch, n_steps, do_eval;
  long long* prof;  // optional: clock64 stamps of CTA 0 / thread 0 (debug, tools/tc_timeline.py)
  // ---- fused round tail (tail_mode 1): the last cluster to finish aggregates every federation ----
  int tail_mode, n_groups, n_feds, n_conf;
  const FedbFed* feds;
  FedbFedState* fed_state_rw;
  float* global;
  float* hist_local;
  float* hist_global;
  unsigned int* tail_count;   // wraps to 0 by itself (atomicInc)
};

struct Shape {
  int d0, h1, h2, C, N2, NP2, MT2;
  int ld0, ld1, ld2;
};

__device__ __forceinline__ void st16(uint8_t* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  *reinterpret_cast<uint4*>(p) = make_uint4(a, b, c, d);
}

// fp32 slab (global) -> bf16 panels W1, W2 (bias in the spare K column) + fp32 W3 / b3.
// Step 1: ONE bulk async copy (cp.async.bulk, the TMA engine's 1-D mode) lands the whole slab in the
// fp32 landing zone (the area that otherwise holds H1/dZ panels, idle at these points); step 2:
// the threads convert from shared memory.  This replaces ~11 dependent L2 round trips by one.
// `zero_first` clears the panels' padding (rows / columns no slab element maps to), once per kernel.
__device__ __forceinline__ void bulk_load_begin(uint8_t* dst, const float* src, uint32_t bytes, uint64_t* wbar) {
  asm volatile("fence.proxy.async;" ::: "memory");
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc::smem_u32(wbar)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(tc::smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(tc::smem_u32(wbar))
               : "memory");
}

__device__ void convert_weights(uint8_t* sm, const float* L, const SlabLayout& lay, const Shape& s) {
  const int tid = threadIdx.x;
  {  // W2 [h2][ld1] (+ b2 into column 63): 16 lanes walk one row
    const int q = tid & 15, jr = tid >> 4, nq = s.ld1 >> 2;
    for (int j = jr; j < s.h2; j += TC_THREADS / 16) {
      if (q < nq) {
        const float4 w = *reinterpret_cast<const float4*>(L + lay.w_off[1] + j * s.ld1 + 4 * q);
        *reinterpret_cast<uint2*>(sm + SM_W2 + tc::panel_off(j, 4 * q)) = make_uint2(tc::pack_bf16(w.x, w.y), tc::pack_bf16(w.z, w.w));
      }
      if (q == 15) *reinterpret_cast<__nv_bfloat16*>(sm + SM_W2 + tc::panel_off(j, 63)) = __float2bfloat16_rn(L[lay.b_off[1] + j]);
    }
  }
  {  // W1 [h1][ld0] (+ b1 into column 15): 4 lanes per row
    const int q = tid & 3, nq = s.ld0 >> 2;
    for (int j = tid >> 2; j < s.h1; j += TC_THREADS / 4) {
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      if (q < nq) w = *reinterpret_cast<const float4*>(L + lay.w_off[0] + j * s.ld0 + 4 * q);
      if (q == 3) w.w = L[lay.b_off[0] + j];   // column 15 carries the bias (d0 <= 15: the slab has padding there)
      *reinterpret_cast<uint2*>(sm + SM_W1 + tc::panel_off(j, 4 * q)) = make_uint2(tc::pack_bf16(w.x, w.y), tc::pack_bf16(w.z, w.w));
    }
  }
  {  // W3 [C][ld2] -> fp32 [C][256], b3
    const int nq = s.ld2 >> 2;
    for (int k = 0; k < s.C; ++k)
      for (int q = tid; q < nq; q += TC_THREADS)
        *reinterpret_cast<float4*>(sm + SM_W3 + (k * 256 + 4 * q) * 4) = *reinterpret_cast<const float4*>(L + lay.w_off[2] + k * s.ld2 + 4 * q);
    if (tid < s.C) reinterpret_cast<float*>(sm + SM_B3)[tid] = L[lay.b_off[2] + tid];
  }
  tc::fence_async_smem();
  __syncthreads();
}

// One data row per thread: global -> registers (issued early so the latency overlaps other work)
struct RowRegs {
  float4 q[4];
  int label;
};
__device__ __forceinline__ RowRegs fetch_row(const TcArgs& a, int64_t pos, bool valid) {
  RowRegs r;
  r.label = 0;
  if (valid) {
    const int row = __ldg(a.idx + pos);
    const float4* src = reinterpret_cast<const float4*>(a.X + (size_t)row * a.ldx);
#pragma unroll
    for (int q = 0; q < 4; ++q) r.q[q] = __ldg(src + q);
    r.label = __ldg(a.y + row);
  } else {
#pragma unroll
    for (int q = 0; q < 4; ++q) r.q[q] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  return r;
}
// registers -> X panel (bf16; column 15 = 1 on valid rows). Returns the label.
__device__ __forceinline__ int store_row(uint8_t* sm, const RowRegs& r, bool valid, const Shape& s) {
  const int t = threadIdx.x;
  float f[16];
#pragma unroll
  for (int q = 0; q < 4; ++q) { f[4 * q] = r.q[q].x; f[4 * q + 1] = r.q[q].y; f[4 * q + 2] = r.q[q].z; f[4 * q + 3] = r.q[q].w; }
#pragma unroll
  for (int i = 0; i < 16; ++i)
    if (i >= s.d0) f[i] = 0.f;
  f[15] = valid ? 1.f : 0.f;
#pragma unroll
  for (int c = 0; c < 2; ++c)
    st16(sm + SM_X + tc::panel_off(t, 8 * c), tc::pack_bf16(f[8 * c], f[8 * c + 1]), tc::pack_bf16(f[8 * c + 2], f[8 * c + 3]),
         tc::pack_bf16(f[8 * c + 4], f[8 * c + 5]), tc::pack_bf16(f[8 * c + 6], f[8 * c + 7]));
  return r.label;
}

// epilogue -> next MMA hand-off: my TMEM reads are done, my smem writes are visible to the tensor core
__device__ __forceinline__ void handoff() {
  tc::tmem_ld_wait();
  tc::tc_fence_before();
  tc::fence_async_smem();
  __syncthreads();
}

// All MMAs are issued by one thread; descriptors are a 64-bit base plus (byte offset >> 4).
struct Issuer {
  uint64_t kbase, mbase;  // K-major / MN-major descriptor of shared-memory offset 0
  uint32_t tmem;
  __device__ __forceinline__ uint64_t kd(int off) const { return kbase + (uint64_t)(off >> 4); }
  __device__ __forceinline__ uint64_t md(int off) const { return mbase + (uint64_t)(off >> 4); }
  __device__ __forceinline__ void fwd1(const Shape&) const {
    tc::mma_bf16(tmem + TM_D1, kd(SM_X), kd(SM_W1), tc::make_idesc_bf16(128, 64, false, false), false);
  }
  __device__ __forceinline__ void fwd2(const Shape& s) const {
    const uint32_t id = tc::make_idesc_bf16(128, s.N2, false, false);
#pragma unroll
    for (int k = 0; k < 4; ++k) tc::mma_bf16(tmem + TM_D2, kd(SM_H1 + 32 * k), kd(SM_W2 + 32 * k), id, k > 0);
  }
  __device__ __forceinline__ void wgrad3(const Shape& s, bool acc) const {
    const uint32_t id = tc::make_idesc_bf16(128, 16, true, true);
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
      if (mt < s.MT2) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          tc::mma_bf16(tmem + TM_DW3 + 16 * mt, md(SM_Z + 2 * mt * PANEL + 2048 * k), md(SM_DZ3 + 2048 * k), id, acc || k > 0);
      }
  }
  __device__ __forceinline__ void dgrad2_wgrad2(const Shape& s, bool acc) const {
    const uint32_t id4 = tc::make_idesc_bf16(128, 64, false, true);
#pragma unroll
    for (int p = 0; p < 4; ++p) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (64 * p + 16 * k < s.N2)  // never feed unwritten pad columns of the last panel
          tc::mma_bf16(tmem + TM_D1, kd(SM_Z + p * PANEL + 32 * k), md(SM_W2 + (p * 64 + 16 * k) * 128), id4, (p | k) > 0);
    }
    const uint32_t id5 = tc::make_idesc_bf16(128, 64, true, true);
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
      if (mt < s.MT2) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          tc::mma_bf16(tmem + TM_DW2 + 64 * mt, md(SM_Z + 2 * mt * PANEL + 2048 * k), md(SM_H1 + 2048 * k), id5, acc || k > 0);
      }
  }
  __device__ __forceinline__ void wgrad1(const Shape&, bool acc) const {
    const uint32_t id = tc::make_idesc_bf16(128, 16, true, true);
#pragma unroll
    for (int k = 0; k < 8; ++k) tc::mma_bf16(tmem + TM_DW1, md(SM_DZ1 + 2048 * k), md(SM_X + 2048 * k), id, acc || k > 0);
  }
};

}  // namespace

__device__ __forceinline__ void local_round_body(const TcArgs& a, uint8_t* smem_raw) {
  __shared__ uint64_t bar, wbar;
  __shared__ uint32_t tmem_slot;
  __shared__ float red_f[4][9];   // per-warp partial sums: loss + db3[8]
  __shared__ int cnt[FEDB_MAX_CLASSES * FEDB_MAX_CLASSES];
  __shared__ StepScalars sc_s;

  cg::cluster_group cluster = cg::this_cluster();
  const int CS = (int)cluster.num_blocks(), q = (int)cluster.block_rank();
  const int g = blockIdx.x / CS;
  // Private copy of the descriptor: read through the global pointer, every field access after a
  // store would be a fresh (possibly aliased) L2 round trip on the critical path.
  __shared__ FedbGroup Gs;
  if (threadIdx.x < sizeof(FedbGroup) / 4)
    reinterpret_cast<int32_t*>(&Gs)[threadIdx.x] = reinterpret_cast<const int32_t*>(a.groups + g)[threadIdx.x];
  __syncthreads();
  const FedbGroup& G = Gs;
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
  float* landing = reinterpret_cast<float*>(sm + SM_STAGE);   // fp32 slab landing zone (idle panel area)
  if (tid == 0) {
    tc::mbar_init(&bar, 1);
    tc::mbar_init(&wbar, 1);
    tc::fence_mbar_init();
    bulk_load_begin(reinterpret_cast<uint8_t*>(landing), Pg, (uint32_t)lay.end * 4u, &wbar);
  }
  for (int i = tid; i < (SM_B3 + 64 - SM_W1) / 16; i += TC_THREADS) st16(sm + SM_W1 + i * 16, 0, 0, 0, 0);  // panel padding
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t tl = tmem + ((uint32_t)(warp * 32) << 16);  // this thread's TMEM lane
  const Issuer mma{tc::make_desc(sm_u, 0, 1024), tc::make_desc(sm_u, PANEL, 1024), tmem};
  uint32_t ph = 0, wph = 0;
  float* W3s = reinterpret_cast<float*>(sm + SM_W3);
  float* b3s = reinterpret_cast<float*>(sm + SM_B3);
  const float inv_n = 1.f / (float)nb;

#define STAMP(i) do { if (a.prof && blockIdx.x == 0 && tid == 0) a.prof[i] = clock64(); } while (0)
  STAMP(0);
  // first tile's row: issue its loads before the weight conversion so both latencies overlap
  RowRegs rr = fetch_row(a, G.idx_off + b0 + q * TC_ROWS + tid, q < ntiles && q * TC_ROWS + tid < nb);
  tc::mbar_wait(&wbar, wph); wph ^= 1;
  convert_weights(sm, landing, lay, s);
  STAMP(1);

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
      if (my_tiles > 0 || st > 0) rr = fetch_row(a, G.idx_off + b0 + r, valid);
      const int label = store_row(sm, rr, valid, s);
      handoff();
      STAMP(2);
      // ---- layer 1 -----------------------------------------------------------------------
      FEDB_MMA_PHASE(mma.fwd1(s));
      STAMP(3);
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
      STAMP(4);
      // ---- layer 2 + head ----------------------------------------------------------------
      FEDB_MMA_PHASE(mma.fwd2(s));
      STAMP(5);
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
      STAMP(6);
      // ---- head weight gradient (needs H2), then dZ2 overwrites H2 --------------------------
      FEDB_MMA_PHASE(mma.wgrad3(s, acc));
      STAMP(7);
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
      STAMP(8);
      // ---- dH1 = dZ2 W2 ; dW2 += dZ2^T H1 ------------------------------------------------------
      FEDB_MMA_PHASE(mma.dgrad2_wgrad2(s, acc));
      STAMP(9);
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
      STAMP(10);
      // ---- dW1 += dZ1^T X ------------------------------------------------------------------------
      FEDB_MMA_PHASE(mma.wgrad1(s, acc));
      STAMP(11);
    }  // tiles

    // ---- spill this CTA's gradients (fp32, slab layout) straight from TMEM to its record in the global
    // scratch (L2-resident; the cluster reduction reads it back from there — measured 3x faster than
    // reducing through distributed shared memory for this size).  Every slab position below lay.end is
    // written (padding explicitly as 0).
    float* rec = a.scratch + G.scratch_off + (size_t)q * (G.slab_size + FEDB_LOSS_SLOTS);
    if (my_tiles == 0) {
      for (int i = tid; i < (lay.total >> 2); i += TC_THREADS) reinterpret_cast<float4*>(rec)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      float v[16];
      for (int mt = 0; mt < s.MT2; ++mt) {
        const int j = 128 * mt + tid;
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 16) {
          tc::tmem_ld16(tl + TM_DW2 + 64 * mt + c0, v);
          tc::tmem_ld_wait();
          if (j < s.h2) {
#pragma unroll
            for (int i4 = 0; i4 < 4; ++i4) {
              const int col = c0 + 4 * i4;
              if (col < s.ld1)
                *reinterpret_cast<float4*>(rec + lay.w_off[1] + j * s.ld1 + col) =
                    make_float4(col < s.h1 ? v[4 * i4] : 0.f, col + 1 < s.h1 ? v[4 * i4 + 1] : 0.f, col + 2 < s.h1 ? v[4 * i4 + 2] : 0.f,
                                col + 3 < s.h1 ? v[4 * i4 + 3] : 0.f);
            }
            if (c0 == 48) rec[lay.b_off[1] + j] = v[15];
          } else if (j < ru4(s.h2)) {
            if (c0 == 0) rec[lay.b_off[1] + j] = 0.f;
          }
        }
        tc::tmem_ld16(tl + TM_DW3 + 16 * mt, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
          if (k < s.C && j < s.ld2) rec[lay.w_off[2] + k * s.ld2 + j] = j < s.h2 ? v[k] : 0.f;
      }
      tc::tmem_ld16(tl + TM_DW1, v);
      tc::tmem_ld_wait();
      if (tid < s.h1) {
#pragma unroll
        for (int i4 = 0; i4 < 4; ++i4) {
          const int col = 4 * i4;
          if (col < s.ld0)
            *reinterpret_cast<float4*>(rec + lay.w_off[0] + tid * s.ld0 + col) =
                make_float4(col < s.d0 ? v[col] : 0.f, col + 1 < s.d0 ? v[col + 1] : 0.f, col + 2 < s.d0 ? v[col + 2] : 0.f,
                            col + 3 < s.d0 ? v[col + 3] : 0.f);
        }
        rec[lay.b_off[0] + tid] = v[15];
      } else if (tid < ru4(s.h1)) {
        rec[lay.b_off[0] + tid] = 0.f;
      }
      if (tid >= s.C && tid < ru4(s.C)) rec[lay.b_off[2] + tid] = 0.f;
    }
    STAMP(12);
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
        if (tid == 0) rec[G.slab_size] = x;                             // loss sum of this CTA's rows
        else if (tid - 1 < s.C) rec[lay.b_off[2] + tid - 1] = x;        // db3
      }
    }
    if (tid == 0) sc_s = step_scalars_fast(G, step0 + st + 1);
    // my slice of the slab and its moments do not depend on the peers: fetch them before the barrier
    constexpr int PF = 3;
    const int n4 = lay.end >> 2, per = (n4 + CS - 1) / CS;
    const int lo = q * per, hi = min(n4, lo + per);
    const bool use_v = a.v && G.opt != FEDB_OPT_SGD;
    float4 w_pf[PF], m_pf[PF], v_pf[PF];
#pragma unroll
    for (int it = 0; it < PF; ++it) {
      const int e4 = lo + tid + it * TC_THREADS;
      w_pf[it] = m_pf[it] = v_pf[it] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e4 < hi) {
        w_pf[it] = *reinterpret_cast<float4*>(Pg + 4 * e4);
        if (a.m) m_pf[it] = *reinterpret_cast<float4*>(a.m + G.slab_off + 4 * e4);
        if (use_v) v_pf[it] = *reinterpret_cast<float4*>(a.v + G.slab_off + 4 * e4);
      }
    }
    tc::tc_fence_before();
    __threadfence();
    cluster.sync();
    STAMP(13);

    // ---- cluster reduction over DSMEM (fixed rank order) + optimizer on my slice ----------------
    {
      const StepScalars sc = sc_s;
      const float* rec0 = a.scratch + G.scratch_off;
      const size_t rstride = (size_t)(G.slab_size + FEDB_LOSS_SLOTS);
      int it = 0;
      for (int e4 = lo + tid; e4 < hi; e4 += TC_THREADS, ++it) {
        const int e = e4 * 4;
        float4 pr[8];   // every load of this element group is issued before the first use
#pragma unroll
        for (int rk = 0; rk < 8; ++rk)
          if (rk < CS) pr[rk] = __ldcg(reinterpret_cast<const float4*>(rec0 + rk * rstride + e));
        float4 w, mm, vv;
        if (it < PF) {
          // (static indexing keeps the prefetch arrays in registers)
          w = it == 0 ? w_pf[0] : (it == 1 ? w_pf[1] : w_pf[2]);
          mm = it == 0 ? m_pf[0] : (it == 1 ? m_pf[1] : m_pf[2]);
          vv = it == 0 ? v_pf[0] : (it == 1 ? v_pf[1] : v_pf[2]);
        } else {
          w = *reinterpret_cast<float4*>(Pg + e);
          mm = vv = make_float4(0.f, 0.f, 0.f, 0.f);
          if (a.m) mm = *reinterpret_cast<float4*>(a.m + G.slab_off + e);
          if (use_v) vv = *reinterpret_cast<float4*>(a.v + G.slab_off + e);
        }
        float4 gsum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int rk = 0; rk < 8; ++rk)
          if (rk < CS) { gsum.x += pr[rk].x; gsum.y += pr[rk].y; gsum.z += pr[rk].z; gsum.w += pr[rk].w; }
        const bool is_w = (e < lay.b_off[0]) || (e >= lay.w_off[1] && e < lay.b_off[1]) || (e >= lay.w_off[2] && e < lay.b_off[2]);
        if (is_w && G.alpha != 0.f) {
          const float al = G.alpha * inv_n;
          gsum.x = fmaf(al, w.x, gsum.x); gsum.y = fmaf(al, w.y, gsum.y);
          gsum.z = fmaf(al, w.z, gsum.z); gsum.w = fmaf(al, w.w, gsum.w);
        }
        w.x = opt_update(G, sc, w.x, gsum.x, mm.x, vv.x);
        w.y = opt_update(G, sc, w.y, gsum.y, mm.y, vv.y);
        w.z = opt_update(G, sc, w.z, gsum.z, mm.z, vv.z);
        w.w = opt_update(G, sc, w.w, gsum.w, mm.w, vv.w);
        *reinterpret_cast<float4*>(Pg + e) = w;
        if (a.m) *reinterpret_cast<float4*>(a.m + G.slab_off + e) = mm;
        if (use_v) *reinterpret_cast<float4*>(a.v + G.slab_off + e) = vv;
      }
      if (q == 0 && warp == 3 && a.loss_out) {   // 8 lanes fetch the 8 partial losses at once; fixed-order sum
        float x = lane < CS ? __ldcg(rec0 + lane * rstride + G.slab_size) : 0.f;
        float ls = 0.f;
#pragma unroll
        for (int rk = 0; rk < 8; ++rk) ls += __shfl_sync(0xffffffffu, x, rk);
        if (lane == 0) a.loss_out[g] = ls * inv_n;
      }
    }
    STAMP(14);
    __threadfence();
    asm volatile("fence.proxy.async;" ::: "memory");   // my slice must be visible to the peers' bulk copies
    cluster.sync();   // all slices written; all records consumed
    STAMP(15);
    if (tid == 0) bulk_load_begin(reinterpret_cast<uint8_t*>(landing), Pg, (uint32_t)lay.end * 4u, &wbar);
    tc::mbar_wait(&wbar, wph); wph ^= 1;
    convert_weights(sm, landing, lay, s);
    STAMP(16);
  }  // steps

  // ---- local evaluation with the updated weights (C:75-83) ----------------------------------
  if (a.do_eval) {
    if (tid < FEDB_MAX_CLASSES * FEDB_MAX_CLASSES) cnt[tid] = 0;
    __syncthreads();
    for (int tile = q; tile < ntiles; tile += CS) {
      const int r = tile * TC_ROWS + tid;
      const bool valid = r < nb;
      int label;
      if (ntiles <= CS && a.n_steps > 0) {
        label = rr.label;   // single tile per CTA: the X panel from the training step is still in place
      } else {
        rr = fetch_row(a, G.idx_off + b0 + r, valid);
        label = store_row(sm, rr, valid, s);
      }
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
    STAMP(17);
  }
#undef FEDB_MMA_PHASE
#undef STAMP

  if (q == 0 && tid == 0) a.step[g] = step0 + a.n_steps;
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(tmem);
}

// K1 + K5 for small problems, executed by the last cluster to finish: FedAvg of every federation in
// client order with the reference's arithmetic (C:108-116), write-back (C:119-120), metrics and early
// stop (C:165-192), then the next round's begin bookkeeping — same results as k_fedavg.
constexpr int TAIL_MAX_FEDS = 256, TAIL_MAX_CLIENTS = 32;

__device__ void round_tail(const TcArgs& a, int crank, int csize, const unsigned char* s_stop) {
  const int tid = threadIdx.x;
  __shared__ long long t_off[TAIL_MAX_CLIENTS];
  __shared__ float t_num[TAIL_MAX_CLIENTS], t_den[TAIL_MAX_CLIENTS];
  const int nthr = csize * TC_THREADS, gt = crank * TC_THREADS + tid;
  for (int f = 0; f < a.n_feds; ++f) {
    if (s_stop[f]) continue;
    const FedbFed F = a.feds[f];
    __syncthreads();
    if (tid < F.n_clients) {
      const FedbGroup* G = a.groups + F.first_group + tid;
      t_off[tid] = G->slab_off; t_num[tid] = G->agg_num; t_den[tid] = G->agg_den;
    }
    __syncthreads();
    const int n4 = F.slab_size >> 2;
    for (int base = gt; base < n4; base += 4 * nthr) {
      float4 acc[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
      for (int k = 0; k < F.n_clients; ++k) {
        const float num = t_num[k], den = t_den[k];
        const float4* src = reinterpret_cast<const float4*>(a.params + t_off[k]);
        float4 w[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (base + u * nthr < n4) w[u] = __ldcg(src + base + u * nthr);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (base + u * nthr < n4) {
            acc[u].x += __fdiv_rn(__fmul_rn(w[u].x, num), den);
            acc[u].y += __fdiv_rn(__fmul_rn(w[u].y, num), den);
            acc[u].z += __fdiv_rn(__fmul_rn(w[u].z, num), den);
            acc[u].w += __fdiv_rn(__fmul_rn(w[u].w, num), den);
          }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int e4 = base + u * nthr;
        if (e4 < n4) {
          reinterpret_cast<float4*>(a.global + F.global_off)[e4] = acc[u];
          for (int k = 0; k < F.n_clients; ++k) reinterpret_cast<float4*>(a.params + t_off[k])[e4] = acc[u];
        }
      }
    }
  }
  if (a.prof && crank == 0 && tid == 0) a.prof[20] = clock64();
  if (crank == 0) {   // metrics read only the confusion counts: they overlap the averaging above
    __shared__ int pool[FEDB_MAX_CLASSES * FEDB_MAX_CLASSES];
    if (tid < 32)
      for (int f = 0; f < a.n_feds; ++f) {
        if (s_stop[f]) continue;
        round_metrics_warp(a.feds[f], a.fed_state_rw[f], f, a.n_groups, a.conf, a.hist_local, a.hist_global, pool, true);
      }
    __syncthreads();
    if (a.prof && tid == 0) a.prof[21] = clock64();
    for (int i = tid; i < a.n_conf; i += TC_THREADS) a.conf[i] = 0;
  }
}

__global__ void __launch_bounds__(TC_THREADS, 1) k_local_round_tc(TcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  local_round_body(a, smem_raw);
  if (a.tail_mode == 0) return;
  // Every cluster — including those that had nothing to do — checks in once all of its CTAs are
  // done; the cluster that observes the last check-in runs the tail.  No grid-wide barrier and no
  // co-residency assumption: the other clusters have simply finished.
  __shared__ unsigned int s_last;
  __shared__ unsigned char s_stop[TAIL_MAX_FEDS];
  cg::cluster_group cluster = cg::this_cluster();
  const int CS = (int)cluster.num_blocks(), crank = (int)cluster.block_rank();
  // stop flags as they stand for THIS round (the tail raises them for the next one)
  for (int f = threadIdx.x; f < a.n_feds; f += TC_THREADS) s_stop[f] = a.fed_state_rw[f].stop ? 1 : 0;
  __threadfence();
  cluster.sync();
  if (crank == 0 && threadIdx.x == 0) {
    const unsigned int n_clusters = gridDim.x / CS;
    const unsigned int last = (atomicInc(a.tail_count, n_clusters - 1) == n_clusters - 1) ? 1u : 0u;
    for (int r = 0; r < CS; ++r) *cluster.map_shared_rank(&s_last, r) = last;
  }
  cluster.sync();
  if (a.prof && blockIdx.x == 0 && threadIdx.x == 0) a.prof[18] = clock64();
  if (!s_last) return;
  __threadfence();
  round_tail(a, crank, CS, s_stop);
  if (a.prof && crank == 0 && threadIdx.x == 0) a.prof[19] = clock64();
}

}  // namespace fedb

using namespace fedb;

extern "C" int64_t fedb_tc_smem_bytes(void) { return SM_END + 1024; }

extern "C" int fedb_round_tc_prof(const FedbGroup*, int32_t, int32_t, const float*, int32_t, const int32_t*, const int32_t*, float*, float*,
                                  float*, int32_t*, const int32_t*, FedbFedState*, float*, int32_t*, float*, int32_t, const FedbFed*,
                                  int32_t, float*, float*, float*, int32_t, uint32_t*, long long*, void*);

static int launch_tc(TcArgs a, int32_t n_groups, int32_t cluster_size, void* stream) {
  FEDB_REQUIRE(n_groups > 0 && cluster_size >= 1 && cluster_size <= 8, "bad grid");
  FEDB_REQUIRE(a.ldx >= 16 && a.ldx % 4 == 0, "feature rows must be padded to >= 16 floats");
  FEDB_REQUIRE(a.scratch != nullptr, "scratch (gradient records) is required");
  static bool attr_set = false;
  const int smem = SM_END + 1024;
  if (!attr_set) {
    FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_local_round_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
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

// One launch: n_steps local optimizer steps (+ optional local evaluation) for every group, on the
// tensor cores.  cluster_size CTAs cooperate on one client; every group must satisfy
// 3 layers, relu, dims[0] <= 15, dims[1] <= 63, dims[2] <= 256, dims[3] <= 8 (checked by the
// caller on its host table: ops/tc.py: tc_eligible).
extern "C" int fedb_local_round_tc_prof(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X,
                                        int32_t ldx, const int32_t* y, const int32_t* idx, float* params, float* m, float* v,
                                        int32_t* step, const int32_t* active, const FedbFedState* fed_state, float* loss_out,
                                        int32_t* conf, float* scratch, int32_t batch, int32_t n_steps, int32_t do_eval,
                                        long long* prof, void* stream) {
  TcArgs a{};
  a.groups = groups; a.X = X; a.y = y; a.idx = idx; a.params = params; a.m = m; a.v = v; a.step = step; a.active = active;
  a.fed_state = fed_state; a.loss_out = loss_out; a.conf = conf; a.scratch = scratch; a.ldx = ldx; a.batch = batch;
  a.n_steps = n_steps; a.do_eval = do_eval; a.prof = prof; a.tail_mode = 0;
  return launch_tc(a, n_groups, cluster_size, stream);
}

extern "C" int fedb_local_round_tc(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X,
                                   int32_t ldx, const int32_t* y, const int32_t* idx, float* params, float* m, float* v,
                                   int32_t* step, const int32_t* active, const FedbFedState* fed_state, float* loss_out,
                                   int32_t* conf, float* scratch, int32_t batch, int32_t n_steps, int32_t do_eval, void* stream) {
  return fedb_local_round_tc_prof(groups, n_groups, cluster_size, X, ldx, y, idx, params, m, v, step, active, fed_state, loss_out,
                                  conf, scratch, batch, n_steps, do_eval, nullptr, stream);
}

// The whole federated round in ONE launch (small problems): local steps + local evaluation as above,
// then the last cluster to finish runs FedAvg + write-back + metrics + early stop + the next
// round's begin bookkeeping for every federation (what fedb_fedavg(begin_next=1) would do).
// conf_ints = number of int32 counters to clear for the next round; tail_count: one zeroed uint32.
extern "C" int fedb_round_tc(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X, int32_t ldx,
                             const int32_t* y, const int32_t* idx, float* params, float* m, float* v, int32_t* step,
                             const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf, float* scratch,
                             int32_t n_steps, const FedbFed* feds, int32_t n_feds, float* global, float* hist_local,
                             float* hist_global, int32_t conf_ints, uint32_t* tail_count, void* stream) {
  return fedb_round_tc_prof(groups, n_groups, cluster_size, X, ldx, y, idx, params, m, v, step, active, fed_state, loss_out, conf,
                            scratch, n_steps, feds, n_feds, global, hist_local, hist_global, conf_ints, tail_count, nullptr, stream);
}

extern "C" int fedb_round_tc_prof(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X, int32_t ldx,
                                  const int32_t* y, const int32_t* idx, float* params, float* m, float* v, int32_t* step,
                                  const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf, float* scratch,
                                  int32_t n_steps, const FedbFed* feds, int32_t n_feds, float* global, float* hist_local,
                                  float* hist_global, int32_t conf_ints, uint32_t* tail_count, long long* prof, void* stream) {
  FEDB_REQUIRE(feds && fed_state && global && tail_count && n_feds > 0, "bad tail arguments");
  TcArgs a{};
  a.groups = groups; a.X = X; a.y = y; a.idx = idx; a.params = params; a.m = m; a.v = v; a.step = step; a.active = active;
  a.fed_state = fed_state; a.loss_out = loss_out; a.conf = conf; a.scratch = scratch; a.ldx = ldx; a.batch = 0;
  a.n_steps = n_steps; a.do_eval = 1; a.prof = prof;
  a.tail_mode = 1; a.n_groups = n_groups; a.n_feds = n_feds; a.n_conf = conf_ints; a.feds = feds; a.fed_state_rw = fed_state;
  a.global = global; a.hist_local = hist_local; a.hist_global = hist_global; a.tail_count = tail_count;
  return launch_tc(a, n_groups, cluster_size, stream);
}
