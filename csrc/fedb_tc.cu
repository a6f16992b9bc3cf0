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
#include "p2p_common.cuh"
#include "tc_common.cuh"
#include "tc_round_common.cuh"

namespace cg = cooperative_groups;

namespace fedb {
namespace {

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
constexpr int SM_SLOTS = SM_END;               // [CS][per] float4 partial-gradient slices + CS loss floats (never overlaps live panels)
constexpr int SLOTS_MAX_BYTES = 227 * 1024 - 4096 - SM_SLOTS;   // what is left of the SM's shared memory
constexpr int SM_STAGE = SM_H1;               // fp32 gradient staging overlays H1..Z once the MMAs are done
// TMEM columns
constexpr uint32_t TM_D1 = 0, TM_D2 = 64, TM_DW2 = 320, TM_DW1 = 448, TM_DW3 = 464, TM_ZX = 496;   // ZX: 2 x 8 logit-exchange columns


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

// The small part of convert_weights: W1 (+ b1 in column 15) and b2 (column 63 of W2) from the fp32 landing
// zone.  W2 itself and the head arrive already converted (publish_weights in the round body).
__device__ void convert_small(uint8_t* sm, const float* L, const SlabLayout& lay, const Shape& s) {
  const int tid = threadIdx.x;
  {
    const int q = tid & 3, nq = s.ld0 >> 2;
    for (int j = tid >> 2; j < s.h1; j += TC_THREADS / 4) {
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      if (q < nq) w = *reinterpret_cast<const float4*>(L + lay.w_off[0] + j * s.ld0 + 4 * q);
      if (q == 3) w.w = L[lay.b_off[0] + j];
      *reinterpret_cast<uint2*>(sm + SM_W1 + tc::panel_off(j, 4 * q)) = make_uint2(tc::pack_bf16(w.x, w.y), tc::pack_bf16(w.z, w.w));
    }
  }
  for (int j = tid; j < s.h2; j += TC_THREADS)
    *reinterpret_cast<__nv_bfloat16*>(sm + SM_W2 + tc::panel_off(j, 63)) = __float2bfloat16_rn(L[lay.b_off[1] + j]);
  tc::fence_async_smem();
  __syncthreads();
}

// Two threads per data row: half h moves columns 8h..8h+7 (global -> registers, issued early so the
// latency overlaps other work; then registers -> X panel as bf16, column 15 = 1 on valid rows).
struct RowRegs {
  float4 q[2];
  int label;
};
__device__ __forceinline__ RowRegs fetch_row(const TcArgs& a, int64_t pos, bool valid, int half) {
  RowRegs r;
  r.label = 0;
  r.q[0] = r.q[1] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (valid) {
    const int row = __ldg(a.idx + pos);
    const float4* src = reinterpret_cast<const float4*>(a.X + (size_t)row * a.ldx) + 2 * half;
    r.q[0] = __ldg(src);
    r.q[1] = __ldg(src + 1);
    r.label = __ldg(a.y + row);
  }
  return r;
}
// Host-input mode (TcArgs::x_host): the tile's rows come from pinned host memory.  Over the host link the request size is
// what matters (tools/host_pull_bench.cu: 1000 rows as 2 x 16 B per thread 13.7 us, as one 64-byte request per row 6.2 us,
// against 3.7 us from device memory), so four lanes fetch one whole row with a single instruction: warp w takes tile rows
// 16w .. 16w+15, two instructions of eight rows.  host_deliver() later passes the rows through shared memory (the H1 panel,
// idle at that point) to the threads that own them, and writes them through to the device-resident X / y, 64 bytes per row.
struct HostRegs {
  float4 v[2];
  int grow[2], label[2];   // global row index (-1: none) and, on the row's first lane, its label
};
__device__ __forceinline__ HostRegs host_fetch(const TcArgs& a, int64_t pos0, int nvalid, int tid) {
  HostRegs h;
  const int w = tid >> 5, lane = tid & 31, c = lane & 3;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int rl = 16 * w + 8 * j + (lane >> 2);
    h.v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    h.grow[j] = -1;
    h.label[j] = 0;
    if (rl < nvalid) {
      const int row = __ldg(a.idx + pos0 + rl);
      h.grow[j] = row;
      h.v[j] = __ldcv(reinterpret_cast<const float4*>(a.x_host + (size_t)row * a.ldx) + c);
      if (c == 0) h.label[j] = __ldcv(a.y_host + row);
    }
  }
  return h;
}
__device__ __forceinline__ RowRegs host_deliver(uint8_t* sm, const TcArgs& a, const HostRegs& h, int tid, int row, int half) {
  float* stage = reinterpret_cast<float*>(sm + SM_H1);           // 128 rows x 16 floats, then 128 labels
  int* stage_lab = reinterpret_cast<int*>(sm + SM_H1 + TC_ROWS * 64);
  const int w = tid >> 5, lane = tid & 31, c = lane & 3;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int rl = 16 * w + 8 * j + (lane >> 2);
    *reinterpret_cast<float4*>(stage + rl * 16 + 4 * c) = h.v[j];
    if (c == 0) stage_lab[rl] = h.label[j];
    if (h.grow[j] >= 0) {
      *(reinterpret_cast<float4*>(const_cast<float*>(a.X) + (size_t)h.grow[j] * a.ldx) + c) = h.v[j];
      if (c == 0) const_cast<int32_t*>(a.y)[h.grow[j]] = h.label[j];
    }
  }
  __syncthreads();
  RowRegs r;
  r.q[0] = *reinterpret_cast<const float4*>(stage + row * 16 + 8 * half);
  r.q[1] = *reinterpret_cast<const float4*>(stage + row * 16 + 8 * half + 4);
  r.label = stage_lab[row];
  return r;
}
__device__ __forceinline__ int store_row(uint8_t* sm, const RowRegs& r, bool valid, const Shape& s, int row, int half) {
  float f[8] = {r.q[0].x, r.q[0].y, r.q[0].z, r.q[0].w, r.q[1].x, r.q[1].y, r.q[1].z, r.q[1].w};
#pragma unroll
  for (int i = 0; i < 8; ++i)
    if (8 * half + i >= s.d0) f[i] = 0.f;
  if (half == 1) f[7] = valid ? 1.f : 0.f;
  st16(sm + SM_X + tc::panel_off(row, 8 * half), tc::pack_bf16(f[0], f[1]), tc::pack_bf16(f[2], f[3]), tc::pack_bf16(f[4], f[5]),
       tc::pack_bf16(f[6], f[7]));
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

// Row <-> thread mapping: 256 threads, two per data row.  Thread t serves row (t & 127) — TMEM lane
// (t & 127), reachable from warps w and w + 4 alike — and the column half (t >> 7) of every epilogue,
// so each scheduler has two warps to interleave and every epilogue takes half as long.
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
  const int row = tid & 127, half = tid >> 7;
  const int step0 = a.step[g];
  float* Pg = a.params + G.slab_off;
  // gradient exchange geometry: CTA o owns float4 indices [o*per, (o+1)*per) of the slab
  const int n4 = lay.end >> 2, per = (n4 + CS - 1) / CS;
  const uint32_t inv_per = 0xFFFFFFFFu / (uint32_t)per + 1u;   // e4 / per == __umulhi(e4, inv_per) for e4 < 2^16
  const int lo = q * per, hi = min(n4, lo + per);

  // fused cross-GPU aggregation: where this rank publishes its weighted slab for the current round
  uint64_t p2p_seq = 0;
  float* p2p_partial = nullptr;
  if (a.tail_mode == 2) {
    p2p_seq = *reinterpret_cast<volatile uint64_t*>(a.p2p->seq);
    p2p_partial = reinterpret_cast<float*>(a.p2p->sym[a.p2p->rank] + a.p2p->sym_partial_off) + (p2p_seq & 1) * a.p2p->global_total;
  }
  float* landing = reinterpret_cast<float*>(sm + SM_STAGE);   // fp32 slab landing zone (idle panel area)
  float4* slots = reinterpret_cast<float4*>(sm + SM_SLOTS);   // [CS][per] partial-gradient slices sent by every CTA
  if (warp == 0) tc::tmem_alloc<512>(&tmem_slot);
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
  const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);  // this thread's TMEM lane
  const Issuer mma{tc::make_desc(sm_u, 0, 1024), tc::make_desc(sm_u, PANEL, 1024), tmem};
  uint32_t ph = 0, wph = 0;
  float* W3s = reinterpret_cast<float*>(sm + SM_W3);
  float* b3s = reinterpret_cast<float*>(sm + SM_B3);
  const float inv_n = 1.f / (float)nb;
  // peers' slot arrays / landing zones over distributed shared memory
  float4* peer_slots[8];
  uint8_t* peer_sm[8];   // every CTA's operand area: updated weights are stored straight into the peers' bf16 panels
#pragma unroll
  for (int rk = 0; rk < 8; ++rk) {
    peer_slots[rk] = cluster.map_shared_rank(slots, rk < CS ? rk : 0);
    peer_sm[rk] = cluster.map_shared_rank(sm, rk < CS ? rk : 0);
  }
  // One updated float4 of the slab (element index e) -> the operand image of every CTA of the cluster:
  // bf16 into the swizzled W1 / W2 panels (biases into their spare K column), fp32 into the head arrays.
  // This is the all-gather of the new weights AND their conversion in one step.
  auto publish_weights = [&](int e, const float4 w) {
    // W2 (88 % of the slab): 4 columns of one row = 8 contiguous bytes of the swizzled panel
    if (e >= lay.w_off[1] && e < lay.b_off[1]) {
      const int r = e - lay.w_off[1];
      const int j = r / s.ld1, i = r - j * s.ld1;
      const int off = SM_W2 + (int)tc::panel_off(j, i);
      const uint2 pk = make_uint2(tc::pack_bf16(w.x, w.y), tc::pack_bf16(w.z, w.w));
#pragma unroll
      for (int rk = 0; rk < 8; ++rk)
        if (rk < CS) *reinterpret_cast<uint2*>(peer_sm[rk] + off) = pk;
      return;
    }
    int off;
    if (e >= lay.w_off[2] && e < lay.b_off[2]) {          // W3 [C][ld2] -> fp32 [C][256]
      const int r = e - lay.w_off[2];
      const int k = r / s.ld2, c = r - k * s.ld2;
      off = SM_W3 + (k * 256 + c) * 4;
    } else if (e >= lay.b_off[2]) {                       // b3
      off = SM_B3 + (e - lay.b_off[2]) * 4;
    } else {                                              // W1, b1, b2: fp32 into the landing zone, converted locally (convert_small)
      off = SM_STAGE + e * 4;
    }
#pragma unroll
    for (int rk = 0; rk < 8; ++rk)
      if (rk < CS) *reinterpret_cast<float4*>(peer_sm[rk] + off) = w;
  };
  // one gradient float4 (slab index e4) -> the owner CTA's slot for this CTA
  auto send_grad = [&](int e4, const float4 v) {
    const int o = (int)__umulhi((uint32_t)e4, inv_per);
    float4* dst = peer_slots[0];
#pragma unroll
    for (int rk = 1; rk < 8; ++rk) dst = (o == rk) ? peer_slots[rk] : dst;
    dst[q * per + (e4 - o * per)] = v;
  };

#define STAMP(i) do { if (a.prof && blockIdx.x == 0 && tid == 0) a.prof[i] = clock64(); } while (0)
  STAMP(0);
  // first tile's row: issue its loads before the weight conversion so both latencies overlap
  RowRegs rr;
  HostRegs hr;
  if (a.x_host) hr = host_fetch(a, G.idx_off + b0 + q * TC_ROWS, q < ntiles ? min(TC_ROWS, nb - q * TC_ROWS) : 0, tid);
  else rr = fetch_row(a, G.idx_off + b0 + q * TC_ROWS + row, q < ntiles && q * TC_ROWS + row < nb, half);
  tc::mbar_wait(&wbar, wph); wph ^= 1;
  convert_weights(sm, landing, lay, s);
  STAMP(1);

#define FEDB_MMA_PHASE(ISSUE)            \
  do {                                   \
    if (warp == 0 && tc::elect_one()) {  \
      tc::tc_fence_after();              \
      ISSUE;                             \
      tc::mma_commit(&bar);              \
    }                                    \
    tc::mbar_wait(&bar, ph);             \
    ph ^= 1;                             \
    tc::tc_fence_after();                \
  } while (0)

  // split of the layer-2 columns between the two halves (in 16-column chunks)
  const int nch2 = s.N2 >> 4, ch_lo = half ? (nch2 + 1) >> 1 : 0, ch_hi = half ? nch2 : (nch2 + 1) >> 1;

  for (int st = 0; st < a.n_steps; ++st) {
    float loss_acc = 0.f, db3_acc[FEDB_MAX_CLASSES];
#pragma unroll
    for (int k = 0; k < FEDB_MAX_CLASSES; ++k) db3_acc[k] = 0.f;
    int my_tiles = 0;

    for (int tile = q; tile < ntiles; tile += CS, ++my_tiles) {
      const bool acc = my_tiles > 0;
      const int r = tile * TC_ROWS + row;
      const bool valid = r < nb;
      if (a.x_host) {
        if (my_tiles > 0 || st > 0) hr = host_fetch(a, G.idx_off + b0 + tile * TC_ROWS, min(TC_ROWS, nb - tile * TC_ROWS), tid);
        rr = host_deliver(sm, a, hr, tid, row, half);
      } else if (my_tiles > 0 || st > 0) {
        rr = fetch_row(a, G.idx_off + b0 + r, valid, half);
      }
      const int label = store_row(sm, rr, valid, s, row, half);
      handoff();
      STAMP(2);
      // ---- layer 1 -----------------------------------------------------------------------
      FEDB_MMA_PHASE(mma.fwd1(s));
      STAMP(3);
      uint32_t mask1 = 0;   // ReLU mask of this half's 32 columns of H1
      {
        float v[16];
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c0 = 32 * half + 16 * cc;
          tc::tmem_ld16(tl + TM_D1 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            if (v[i] > 0.f) mask1 |= 1u << (16 * cc + i); else v[i] = 0.f;
          }
          if (c0 == 48) { v[15] = valid ? 1.f : 0.f; mask1 &= ~(1u << 31); }
#pragma unroll
          for (int h = 0; h < 2; ++h)
            st16(sm + SM_H1 + tc::panel_off(row, c0 + 8 * h), tc::pack_bf16(v[8 * h], v[8 * h + 1]),
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
      for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = 0.f;
      for (int ch = ch_lo; ch < ch_hi; ++ch) {
        const int c0 = ch << 4;
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
          st16(pz + tc::panel_off(row, (c0 & 63) + 8 * h), tc::pack_bf16(v[8 * h], v[8 * h + 1]),
               tc::pack_bf16(v[8 * h + 2], v[8 * h + 3]), tc::pack_bf16(v[8 * h + 4], v[8 * h + 5]),
               tc::pack_bf16(v[8 * h + 6], v[8 * h + 7]));
      }
      // the two halves swap their partial logits through 16 spare TMEM columns of the row's lane
      tc::tmem_st8(tl + TM_ZX + 8 * half, z);
      tc::tmem_st_wait();
      tc::tc_fence_before();
      __syncthreads();
      tc::tc_fence_after();
      {
        float zo[8];
        tc::tmem_ld8(tl + TM_ZX + 8 * (half ^ 1), zo);
        tc::tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = (half ? zo[k] + z[k] : z[k] + zo[k]) + (k < s.C ? b3s[k] : 0.f);
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
        if (half == 0) {   // one thread per row books the loss / db3 and publishes dZ3
          loss_acc += loss;
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k) db3_acc[k] += dz3[k];
          st16(sm + SM_DZ3 + tc::panel_off(row, 0), tc::pack_bf16(dz3[0], dz3[1]), tc::pack_bf16(dz3[2], dz3[3]),
               tc::pack_bf16(dz3[4], dz3[5]), tc::pack_bf16(dz3[6], dz3[7]));
        } else {
          st16(sm + SM_DZ3 + tc::panel_off(row, 8), 0, 0, 0, 0);
        }
      }
      handoff();
      STAMP(6);
      // ---- head weight gradient (needs H2), then dZ2 overwrites H2 --------------------------
      FEDB_MMA_PHASE(mma.wgrad3(s, acc));
      STAMP(7);
      for (int ch = ch_lo; ch < ch_hi; ++ch) {
        const int c0 = ch << 4;
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
          st16(pz + tc::panel_off(row, (c0 & 63) + 8 * h), tc::pack_bf16(d[8 * h], d[8 * h + 1]),
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
        for (int cc = 0; cc < 2; ++cc) {
          const int c0 = 32 * half + 16 * cc;
          tc::tmem_ld16(tl + TM_D1 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = ((mask1 >> (16 * cc + i)) & 1u) ? v[i] : 0.f;
#pragma unroll
          for (int h = 0; h < 2; ++h)
            st16(sm + SM_DZ1 + tc::panel_off(row, c0 + 8 * h), tc::pack_bf16(v[8 * h], v[8 * h + 1]),
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

    // ---- send this CTA's partial gradients, slice by slice, straight from TMEM into the owning CTA's
    // shared memory (distributed shared memory stores): a reduce-scatter with no trip through L2.
    // Every slab position below lay.end is sent (padding explicitly as 0).
    {
      const bool have = my_tiles > 0;
      float v[16];
      for (int mt = 0; mt < s.MT2; ++mt) {
        const int j = 128 * mt + row;
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c0 = 32 * half + 16 * cc;
          if (have) { tc::tmem_ld16(tl + TM_DW2 + 64 * mt + c0, v); tc::tmem_ld_wait(); }
          else {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = 0.f;
          }
          if (j < s.h2) {
#pragma unroll
            for (int i4 = 0; i4 < 4; ++i4) {
              const int col = c0 + 4 * i4;
              if (col < s.ld1)
                send_grad((lay.w_off[1] + j * s.ld1 + col) >> 2,
                          make_float4(col < s.h1 ? v[4 * i4] : 0.f, col + 1 < s.h1 ? v[4 * i4 + 1] : 0.f,
                                      col + 2 < s.h1 ? v[4 * i4 + 2] : 0.f, col + 3 < s.h1 ? v[4 * i4 + 3] : 0.f));
            }
          }
          // bias gradients sit in column 63: one float each, gathered four rows at a time by the lane that
          // owns the aligned group (warp shuffle keeps it a single float4 store)
          if (c0 == 48) {
            const float b0v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 0), b1v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 1);
            const float b2v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 2), b3v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 3);
            if ((lane & 3) == 0 && j < ru4(s.h2))
              send_grad((lay.b_off[1] + j) >> 2, make_float4(j < s.h2 ? b0v : 0.f, j + 1 < s.h2 ? b1v : 0.f, j + 2 < s.h2 ? b2v : 0.f,
                                                             j + 3 < s.h2 ? b3v : 0.f));
          }
        }
        if (half == 0) {   // head weights: dW3[k][j] for this thread's j, four neighbouring j per store
          if (have) { tc::tmem_ld16(tl + TM_DW3 + 16 * mt, v); tc::tmem_ld_wait(); }
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
            if (k < s.C) {
              const float x0 = __shfl_sync(0xffffffffu, v[k], (lane & ~3) + 0), x1 = __shfl_sync(0xffffffffu, v[k], (lane & ~3) + 1);
              const float x2 = __shfl_sync(0xffffffffu, v[k], (lane & ~3) + 2), x3 = __shfl_sync(0xffffffffu, v[k], (lane & ~3) + 3);
              if ((lane & 3) == 0 && j < s.ld2)
                send_grad((lay.w_off[2] + k * s.ld2 + j) >> 2, make_float4(j < s.h2 ? x0 : 0.f, j + 1 < s.h2 ? x1 : 0.f,
                                                                           j + 2 < s.h2 ? x2 : 0.f, j + 3 < s.h2 ? x3 : 0.f));
            }
        }
      }
      if (half == 0) {     // layer 1: dW1[row][0..15] (+ db1 in column 15)
        if (have) { tc::tmem_ld16(tl + TM_DW1, v); tc::tmem_ld_wait(); }
        else {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = 0.f;
        }
        if (row < s.h1) {
#pragma unroll
          for (int i4 = 0; i4 < 4; ++i4) {
            const int col = 4 * i4;
            if (col < s.ld0)
              send_grad((lay.w_off[0] + row * s.ld0 + col) >> 2,
                        make_float4(col < s.d0 ? v[col] : 0.f, col + 1 < s.d0 ? v[col + 1] : 0.f, col + 2 < s.d0 ? v[col + 2] : 0.f,
                                    col + 3 < s.d0 ? v[col + 3] : 0.f));
          }
        }
        const float b0v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 0), b1v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 1);
        const float b2v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 2), b3v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 3);
        if ((lane & 3) == 0 && row < ru4(s.h1))
          send_grad((lay.b_off[0] + row) >> 2, make_float4(row < s.h1 ? b0v : 0.f, row + 1 < s.h1 ? b1v : 0.f,
                                                           row + 2 < s.h1 ? b2v : 0.f, row + 3 < s.h1 ? b3v : 0.f));
      }
    }
    STAMP(12);
    // loss and db3: warp shuffle, then the 4 row-warps in a fixed order; sent like any other gradient
    {
      float vals[9];
      vals[0] = loss_acc;
#pragma unroll
      for (int k = 0; k < 8; ++k) vals[k + 1] = db3_acc[k];
      if (half == 0) {
#pragma unroll
        for (int i = 0; i < 9; ++i) {
          float x = vals[i];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
          if (lane == 0) red_f[warp][i] = x;
        }
      }
      __syncthreads();
      if (tid == 0) {
        float t[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) t[i] = red_f[0][i] + red_f[1][i] + red_f[2][i] + red_f[3][i];
        send_grad(lay.b_off[2] >> 2, make_float4(0 < s.C ? t[1] : 0.f, 1 < s.C ? t[2] : 0.f, 2 < s.C ? t[3] : 0.f, 3 < s.C ? t[4] : 0.f));
        if (s.C > 4) send_grad((lay.b_off[2] >> 2) + 1, make_float4(t[5], 5 < s.C ? t[6] : 0.f, 6 < s.C ? t[7] : 0.f, 7 < s.C ? t[8] : 0.f));
        // loss partial: slot n4 of rank 0's extra row (index per*CS + q)
        reinterpret_cast<float*>(peer_slots[0] + CS * per)[q] = t[0];
        sc_s = step_scalars_fast(G, step0 + st + 1);
      }
    }
    // my slice of the slab and its moments do not depend on the peers: fetch them before the barrier
    constexpr int PF = 2;
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
    cluster.sync();     // every partial slice has landed in its owner's shared memory
    STAMP(13);

    // ---- owner: fixed-order sum of the CS partials (local shared memory), optimizer, then all-gather +
    // bf16 conversion of the updated slice into every CTA's operand panels (distributed shared memory) ------
    {
      const StepScalars sc = sc_s;
      int it = 0;
      for (int e4 = lo + tid; e4 < hi; e4 += TC_THREADS, ++it) {
        const int e = e4 * 4, i = e4 - lo;
        float4 gsum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int rk = 0; rk < 8; ++rk)
          if (rk < CS) { const float4 p = slots[rk * per + i]; gsum.x += p.x; gsum.y += p.y; gsum.z += p.z; gsum.w += p.w; }
        float4 w, mm, vv;
        if (it < PF) {
          w = it == 0 ? w_pf[0] : w_pf[1];
          mm = it == 0 ? m_pf[0] : m_pf[1];
          vv = it == 0 ? v_pf[0] : v_pf[1];
        } else {
          w = *reinterpret_cast<float4*>(Pg + e);
          mm = vv = make_float4(0.f, 0.f, 0.f, 0.f);
          if (a.m) mm = *reinterpret_cast<float4*>(a.m + G.slab_off + e);
          if (use_v) vv = *reinterpret_cast<float4*>(a.v + G.slab_off + e);
        }
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
        publish_weights(e, w);
        if (p2p_partial && st == a.n_steps - 1)          // this rank's FedAvg term, (w * n_k) / N, for the peers to fetch (C:110-115)
          reinterpret_cast<float4*>(p2p_partial)[e4] =
              make_float4(__fdiv_rn(__fmul_rn(w.x, G.agg_num), G.agg_den), __fdiv_rn(__fmul_rn(w.y, G.agg_num), G.agg_den),
                          __fdiv_rn(__fmul_rn(w.z, G.agg_num), G.agg_den), __fdiv_rn(__fmul_rn(w.w, G.agg_num), G.agg_den));
        *reinterpret_cast<float4*>(Pg + e) = w;          // master copy (off the critical path)
        if (a.m) *reinterpret_cast<float4*>(a.m + G.slab_off + e) = mm;
        if (use_v) *reinterpret_cast<float4*>(a.v + G.slab_off + e) = vv;
      }
      if (q == 0 && tid == 32 && a.loss_out) {
        const float* lp = reinterpret_cast<const float*>(slots + CS * per);
        float ls = 0.f;
        for (int rk = 0; rk < CS; ++rk) ls += lp[rk];
        a.loss_out[g] = ls * inv_n;
      }
    }
    STAMP(14);
    cluster.sync();   // every CTA holds the complete updated operand image (bf16 panels + fp32 head)
    if (p2p_partial && st == a.n_steps - 1 && q == 0 && tid < a.p2p->world) {
      // arrival flag A ("my partial is complete") into every peer's buffer; the local evaluation below
      // overlaps its flight over NVLink and the peers' own progress.  The cluster barrier ordered every
      // CTA's partial stores before this thread; the system-scope fence + release store make them
      // visible to whoever acquires the flag (cumulativity) — no per-thread system fence needed.
      __threadfence_system();
      uint64_t* pf = reinterpret_cast<uint64_t*>(a.p2p->sym[tid] + a.p2p->sym_flag_off);
      st_release_sys_u64(pf + a.p2p->rank, p2p_seq + 1);
    }
    STAMP(15);
    convert_small(sm, landing, lay, s);   // W1 / b1 / b2 from the landing zone (+ proxy fence and block barrier)
    STAMP(16);
  }  // steps

  // ---- local evaluation with the updated weights (C:75-83) ----------------------------------
  if (a.do_eval) {
    if (tid < FEDB_MAX_CLASSES * FEDB_MAX_CLASSES) cnt[tid] = 0;
    __syncthreads();
    for (int tile = q; tile < ntiles; tile += CS) {
      const int r = tile * TC_ROWS + row;
      const bool valid = r < nb;
      int label;
      if (ntiles <= CS && a.n_steps > 0) {
        label = rr.label;   // single tile per CTA: the X panel from the training step is still in place
      } else {
        if (a.x_host) {
          hr = host_fetch(a, G.idx_off + b0 + tile * TC_ROWS, min(TC_ROWS, nb - tile * TC_ROWS), tid);
          rr = host_deliver(sm, a, hr, tid, row, half);
        } else {
          rr = fetch_row(a, G.idx_off + b0 + r, valid, half);
        }
        label = store_row(sm, rr, valid, s, row, half);
      }
      handoff();
      FEDB_MMA_PHASE(mma.fwd1(s));
      {
        float v[16];
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c0 = 32 * half + 16 * cc;
          tc::tmem_ld16(tl + TM_D1 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
          if (c0 == 48) v[15] = valid ? 1.f : 0.f;
#pragma unroll
          for (int h = 0; h < 2; ++h)
            st16(sm + SM_H1 + tc::panel_off(row, c0 + 8 * h), tc::pack_bf16(v[8 * h], v[8 * h + 1]),
                 tc::pack_bf16(v[8 * h + 2], v[8 * h + 3]), tc::pack_bf16(v[8 * h + 4], v[8 * h + 5]),
                 tc::pack_bf16(v[8 * h + 6], v[8 * h + 7]));
        }
      }
      handoff();
      FEDB_MMA_PHASE(mma.fwd2(s));
      float z[FEDB_MAX_CLASSES];
#pragma unroll
      for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = 0.f;
      for (int ch = ch_lo; ch < ch_hi; ++ch) {
        const int c0 = ch << 4;
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
      tc::tmem_st8(tl + TM_ZX + 8 * half, z);
      tc::tmem_st_wait();
      tc::tc_fence_before();
      __syncthreads();
      tc::tc_fence_after();
      if (half == 0) {
        float zo[8];
        tc::tmem_ld8(tl + TM_ZX + 8, zo);
        tc::tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = z[k] + zo[k] + (k < s.C ? b3s[k] : 0.f);
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


__global__ void __launch_bounds__(TC_THREADS, 1) k_local_round_tc(TcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  local_round_body(a, smem_raw);
  if (a.tail_mode == 0) return;
  if (a.tail_mode == 2) { p2p_tail(a); return; }
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
  const unsigned int n_clusters = gridDim.x / CS;
  if (n_clusters > 1) {
    if (crank == 0 && threadIdx.x == 0) {
      const unsigned int last = (atomicInc(a.tail_count, n_clusters - 1) == n_clusters - 1) ? 1u : 0u;
      for (int r = 0; r < CS; ++r) *cluster.map_shared_rank(&s_last, r) = last;
    }
    cluster.sync();
    if (!s_last) return;
    __threadfence();
  }
  if (a.prof && blockIdx.x == 0 && threadIdx.x == 0) a.prof[18] = clock64();
  round_tail(a, crank, CS, s_stop);
  if (a.prof && crank == CS - 1 && threadIdx.x == 0) a.prof[19] = clock64();
}

}  // namespace fedb

using namespace fedb;

extern "C" int64_t fedb_tc_smem_bytes(void) { return SM_SLOTS + SLOTS_MAX_BYTES + 1024; }
// largest slab (floats) whose gradient slices fit the per-CTA exchange area
extern "C" int32_t fedb_tc_max_slab(void) { return (SLOTS_MAX_BYTES - 64) / 4 - 32; }

extern "C" int fedb_round_tc_prof(const FedbGroup*, int32_t, int32_t, const float*, int32_t, const int32_t*, const int32_t*, float*, float*,
                                  float*, int32_t*, const int32_t*, FedbFedState*, float*, int32_t*, float*, int32_t, const FedbFed*,
                                  int32_t, float*, float*, float*, int32_t, uint32_t*, FedbFedState*, long long*, void*);

static int launch_tc(TcArgs a, int32_t n_groups, int32_t cluster_size, void* stream) {
  FEDB_REQUIRE(n_groups > 0 && cluster_size >= 1 && cluster_size <= 8, "bad grid");
  FEDB_REQUIRE(a.ldx >= 16 && a.ldx % 4 == 0, "feature rows must be padded to >= 16 floats");
  static bool attr_set = false;
  const int smem = SM_SLOTS + SLOTS_MAX_BYTES + 1024;
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
                             float* hist_global, int32_t conf_ints, uint32_t* tail_count, FedbFedState* state_mirror, void* stream) {
  return fedb_round_tc_prof(groups, n_groups, cluster_size, X, ldx, y, idx, params, m, v, step, active, fed_state, loss_out, conf,
                            scratch, n_steps, feds, n_feds, global, hist_local, hist_global, conf_ints, tail_count, state_mirror, nullptr,
                            stream);
}

// The whole round of a multi-GPU federation with ONE resident client per rank in one launch per rank:
// local steps + (partial published, flag raised) + local evaluation + cross-GPU aggregation over NVLink
// peer memory + metrics + early stop (see p2p_tail).  rank_dev: this rank's FedbP2PRank record.
extern "C" int fedb_round_tc_p2p_hostin(const FedbGroup* groups, int32_t cluster_size, float* X, int32_t ldx, int32_t* y,
                                        const float* x_host, const int32_t* y_host, const int32_t* idx, float* params, float* m,
                                        float* v, int32_t* step, const int32_t* active, FedbFedState* fed_state, float* loss_out,
                                        int32_t* conf, float* scratch, int32_t n_steps, const FedbP2PRank* rank_dev,
                                        FedbFedState* state_mirror, void* stream) {
  FEDB_REQUIRE(rank_dev != nullptr, "rank record required");
  FEDB_REQUIRE((x_host == nullptr) == (y_host == nullptr), "x_host and y_host come together");
  TcArgs a{};
  a.x_host = x_host; a.y_host = y_host;
  a.groups = groups; a.X = X; a.y = y; a.idx = idx; a.params = params; a.m = m; a.v = v; a.step = step; a.active = active;
  a.fed_state = fed_state; a.loss_out = loss_out; a.conf = conf; a.scratch = scratch; a.ldx = ldx; a.batch = 0;
  a.n_steps = n_steps; a.do_eval = 1; a.prof = nullptr; a.tail_mode = 2; a.n_groups = 1; a.n_feds = 1; a.p2p = rank_dev;
  a.state_mirror = state_mirror;
  return launch_tc(a, 1, cluster_size, stream);
}

extern "C" int fedb_round_tc_p2p(const FedbGroup* groups, int32_t cluster_size, const float* X, int32_t ldx, const int32_t* y,
                                 const int32_t* idx, float* params, float* m, float* v, int32_t* step, const int32_t* active,
                                 FedbFedState* fed_state, float* loss_out, int32_t* conf, float* scratch, int32_t n_steps,
                                 const FedbP2PRank* rank_dev, FedbFedState* state_mirror, void* stream) {
  return fedb_round_tc_p2p_hostin(groups, cluster_size, const_cast<float*>(X), ldx, const_cast<int32_t*>(y), nullptr, nullptr, idx, params,
                                  m, v, step, active, fed_state, loss_out, conf, scratch, n_steps, rank_dev, state_mirror, stream);
}

extern "C" int fedb_round_tc_hostin(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, float* X, int32_t ldx, int32_t* y,
                                    const float* x_host, const int32_t* y_host, const int32_t* idx, float* params, float* m, float* v,
                                    int32_t* step, const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf,
                                    float* scratch, int32_t n_steps, const FedbFed* feds, int32_t n_feds, float* global,
                                    float* hist_local, float* hist_global, int32_t conf_ints, uint32_t* tail_count,
                                    FedbFedState* state_mirror, long long* prof, void* stream);

extern "C" int fedb_round_tc_prof(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X, int32_t ldx,
                                  const int32_t* y, const int32_t* idx, float* params, float* m, float* v, int32_t* step,
                                  const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf, float* scratch,
                                  int32_t n_steps, const FedbFed* feds, int32_t n_feds, float* global, float* hist_local,
                                  float* hist_global, int32_t conf_ints, uint32_t* tail_count, FedbFedState* state_mirror,
                                  long long* prof, void* stream) {
  return fedb_round_tc_hostin(groups, n_groups, cluster_size, const_cast<float*>(X), ldx, const_cast<int32_t*>(y), nullptr, nullptr, idx,
                              params, m, v, step, active, fed_state, loss_out, conf, scratch, n_steps, feds, n_feds, global, hist_local,
                              hist_global, conf_ints, tail_count, state_mirror, prof, stream);
}

// fedb_round_tc with this step's inputs taken from pinned host memory by the kernel itself (x_host / y_host: device-visible
// addresses of host arrays laid out like X / y; rows are written through to X / y).  Both null: identical to fedb_round_tc.
extern "C" int fedb_round_tc_hostin(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, float* X, int32_t ldx, int32_t* y,
                                    const float* x_host, const int32_t* y_host, const int32_t* idx, float* params, float* m, float* v,
                                    int32_t* step, const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf,
                                    float* scratch, int32_t n_steps, const FedbFed* feds, int32_t n_feds, float* global,
                                    float* hist_local, float* hist_global, int32_t conf_ints, uint32_t* tail_count,
                                    FedbFedState* state_mirror, long long* prof, void* stream) {
  FEDB_REQUIRE(feds && fed_state && global && tail_count && n_feds > 0, "bad tail arguments");
  FEDB_REQUIRE((x_host == nullptr) == (y_host == nullptr), "x_host and y_host come together");
  TcArgs a{};
  a.x_host = x_host; a.y_host = y_host;
  a.groups = groups; a.X = X; a.y = y; a.idx = idx; a.params = params; a.m = m; a.v = v; a.step = step; a.active = active;
  a.fed_state = fed_state; a.loss_out = loss_out; a.conf = conf; a.scratch = scratch; a.ldx = ldx; a.batch = 0;
  a.n_steps = n_steps; a.do_eval = 1; a.prof = prof;
  a.tail_mode = 1; a.n_groups = n_groups; a.n_feds = n_feds; a.n_conf = conf_ints; a.feds = feds; a.fed_state_rw = fed_state;
  a.global = global; a.hist_local = hist_local; a.hist_global = hist_global; a.tail_count = tail_count;
  a.state_mirror = state_mirror;
  return launch_tc(a, n_groups, cluster_size, stream);
}
