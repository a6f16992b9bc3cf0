// K2a/K3 on the 5th-generation tensor cores at fp32-grade accuracy: the whole local part of a federated round
//     n_steps x [ forward -> loss -> backward -> cross-CTA gradient reduction -> optimizer ] -> local evaluation
// (C:63-73 train_one_epoch + C:75-83 evaluate_local; with a tail: + FedAvg C:101-120, metrics and early stop C:165-192)
// in one launch, like csrc/fedb_tc.cu — but where that kernel rounds every MMA operand to bf16 (8 significant bits),
// this one feeds the tensor cores SPLIT operands: every fp32 value a is stored as two fp16 numbers
//     hi = fp16(a),  lo = fp16(a - hi)           (22 significant bits together, exponent range of fp16)
// and every GEMM is issued as three tcgen05.mma (kind::f16) products accumulated in the same fp32 TMEM columns:
//     A B  ~=  A_lo B_hi + A_hi B_lo + A_hi B_hi  (the dropped A_lo B_lo term is 2^-22 of the product).
// The reference trains in fp32 (C:65, C:69-72) and its early-stop rule compares metrics at atol 1e-4 (C:181-192); with
// bf16 operands the accuracy keeps jittering and the rule fires ~100 rounds late.  tests/test_gpu_tcx.py bounds this
// kernel's gradients against float64 and the fp32 SIMT kernels at 1e-5 of each layer's scale (measured: 2e-6).
//
// Range management (fp16 has 5 exponent bits): weights are held scaled by WS = 16 and the accumulators descaled on
// read-out; the loss gradient travels through the backward GEMMs multiplied by DZS = 32 and WITHOUT the 1/n factor
// (|dz| <= 32), which the weight-gradient read-out applies.  Activations are clamped to 6e4.  Values below 6e-5 lose
// relative (not absolute) precision: the absolute floor is 3e-8 against operand scales of O(1).
//
// Mapping.  Twice the bytes per operand element means half the rows per CTA: a tile is 64 data rows (TMEM lanes
// 0..63; the MMAs still run at M = 128 and the upper 64 accumulator lanes hold don't-care values computed from
// whatever follows the 64-row panel in shared memory), a cluster of up to 16 CTAs (non-portable size, one GPC) owns a
// client.  Panels are [64 rows][64 fp16] with the hardware 128-byte swizzle (csrc/tc_common.cuh); hi and lo panels of
// an operand are separate, except where an operand is <= 16 columns wide and both halves fit one panel side by side
// (X, W1, dZ3 — the last one shares X's panel) — there a single MMA with N = 32 (or 16) yields both products at once.
// 16 warps: warp w reaches TMEM lanes 32 (w % 4) .. +31, so the data rows (lanes 0..63) are served by the eight warps
// with w % 4 < 2, four per row ("parts": each takes a quarter of the columns of every epilogue).
//
// Cross-CTA exchange (per optimizer step):
//   reduce-scatter  every CTA drains its weight-gradient accumulators from TMEM into a staging area laid out by OWNER
//                   (CTA o owns W2 rows [o R2, (o+1) R2) with their b2 / W3 entries, W1 rows of block CS-1-o and, for
//                   o = CS-1, b3 and the loss) and sends block o to CTA o with ONE bulk asynchronous copy between the two
//                   shared memories (cp.async.bulk.shared::cluster.shared::cta, completion counted on the receiver's
//                   mbarrier, receivers visited in a rotated order); no cluster barrier;
//   optimizer       the owner adds the CS partials in rank order (bit-reproducible) and updates master weights / moments
//                   in global memory, one scalar per thread and pass;
//   all-gather      through L2: one cluster barrier, then every CTA bulk-loads the whole fp32 slab (cp.async.bulk, the TMA
//                   engine's 1-D mode) and rebuilds its split-fp16 operand image.  Measured on B200 the SM-to-SM fabric
//                   moves ~6-20 B/clk per SM, so pushing 16 copies of the weights through it cost twice what L2 does.
// Hazards are excluded by the protocol itself: a CTA sends partials only after its MMAs are done, and nobody reuses the
// staging area before the barrier that follows every owner's wait.
//
// Shapes: 2 hidden ReLU layers, in <= 15, h1 <= 63, h2 <= 256, out <= 8 (the custom script's 14-50-200-2, C:40).
#include <cooperative_groups.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "p2p_common.cuh"
#include "tc_common.cuh"
#include "tc_round_common.cuh"
#include "tcx_common.cuh"

namespace cg = cooperative_groups;

namespace fedb {
namespace {

constexpr int XT = 512;           // threads per CTA
constexpr int XR = 64;            // data rows per tile
constexpr int XP = XR * 128;      // bytes of a [64][64] fp16 panel
// shared-memory map (bytes from the 1024-aligned base)
constexpr int SX_X = 0;                      // cols 0-15 X hi (col 15 = 1), 16-31 X lo, 32-39 dZ3 hi, 40-47 dZ3 lo
constexpr int SX_H1 = SX_X + XP;             // H1 hi, H1 lo; later dZ1 hi, dZ1 lo
constexpr int SX_Z = SX_H1 + 2 * XP;         // hi panels 0..3, lo panels 0..3: H2, then dZ2
constexpr int SX_W1 = SX_Z + 8 * XP;         // [64][64]: cols 0-15 W1 hi (col 15 = b1), cols 16-31 lo
constexpr int SX_W2 = SX_W1 + 64 * 128;      // hi [256][64] (col 63 = b2)
constexpr int SX_W2L = SX_W2 + 256 * 128;    // lo
constexpr int SX_W3 = SX_W2L + 256 * 128;    // fp32 [256 units][8 classes]
constexpr int SX_B3 = SX_W3 + 256 * 8 * 4;   // fp32 [8]
constexpr int SX_END = SX_B3 + 256;
constexpr int SX_SLOTS = SX_END;              // [CS senders][share] floats: the partials this CTA owns
constexpr int SX_SLOTS_MAX = 227 * 1024 - 7168 - SX_SLOTS;   // (static shared memory of the kernel and its tails: < 6 KB, + 1 KB alignment)
constexpr int SX_STAGE = SX_H1;               // outgoing partials [CS owners][share] (and the fp32 slab at launch) overlay H1..Z
constexpr int SX_STAGE_MAX = 10 * XP;
constexpr int DZ3_COL = 32;                   // first column of [dZ3 hi | dZ3 lo] inside the X panel
// TMEM columns
constexpr uint32_t TX_D1 = 0, TX_D2 = 64, TX_DW2 = 320, TX_DW1 = 448, TX_DW3 = 480;
constexpr uint32_t TX_ZX = TX_D1;             // 4 x 8 partial logits per row (D1 is dead at that point)

struct ShapeX {
  int d0, h1, h2, C, N2, MT2;
  int ld0, ld1, ld2;
  int w_off0, b_off0, w_off1, b_off1, w_off2, b_off2, end;   // slab offsets (floats)
};

// Who owns what in the cross-CTA exchange, and the layout of one (sender, owner) block:
//   [R2 records of recf floats: W2 row (ld1) | b2, dW3[0..2] (C <= 3)  or  b2, 0, 0, 0 | dW3[0..7] (C > 3)]
//   [R1 records of 16 floats: W1 row, column 15 = b1]   [16 floats: db3[0..7], loss sum, 0...]   (only owner CS-1 uses the last)
struct OwnX {
  int R2, R1, recf, share;   // share: floats per block (multiple of 4)
};
__host__ __device__ inline OwnX own_geometry(int h1, int h2, int C, int CS) {
  OwnX o;
  o.R2 = (h2 + CS - 1) / CS;
  o.R1 = (h1 + CS - 1) / CS;
  o.recf = ru4(h1) + 4 + (C > 3 ? 8 : 0);
  o.share = o.R2 * o.recf + o.R1 * 16 + 16;
  return o;
}

// kind::f16 instruction descriptor with fp16 (not bf16) operands, fp32 accumulation

__device__ __forceinline__ void bulk_load_begin_x(uint8_t* dst, const float* src, uint32_t bytes, uint64_t* wbar) {
  asm volatile("fence.proxy.async;" ::: "memory");
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc::smem_u32(wbar)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(tc::smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(tc::smem_u32(wbar))
               : "memory");
}

// One W2 row segment (4 columns) / one W1 row segment / one b2 entry -> split fp16 in this CTA's panels
__device__ __forceinline__ void put_w2(uint8_t* sm, int j, int col, const float4 w) {
  uint32_t h0, l0, h1, l1;
  split2(w.x * WS, w.y * WS, h0, l0);
  split2(w.z * WS, w.w * WS, h1, l1);
  const uint32_t off = tc::panel_off(j, col);
  *reinterpret_cast<uint2*>(sm + SX_W2 + off) = make_uint2(h0, h1);
  *reinterpret_cast<uint2*>(sm + SX_W2L + off) = make_uint2(l0, l1);
}
__device__ __forceinline__ void put_w1(uint8_t* sm, int i, int col, const float4 w) {   // col in {0, 4, 8, 12}
  uint32_t h0, l0, h1, l1;
  split2(w.x * WS, w.y * WS, h0, l0);
  split2(w.z * WS, w.w * WS, h1, l1);
  *reinterpret_cast<uint2*>(sm + SX_W1 + tc::panel_off(i, col)) = make_uint2(h0, h1);
  *reinterpret_cast<uint2*>(sm + SX_W1 + tc::panel_off(i, 16 + col)) = make_uint2(l0, l1);
}
__device__ __forceinline__ void put_b2(uint8_t* sm, int j, float b) {
  b *= WS;
  const __half h = __float2half_rn(b);
  *reinterpret_cast<__half*>(sm + SX_W2 + tc::panel_off(j, 63)) = h;
  *reinterpret_cast<__half*>(sm + SX_W2L + tc::panel_off(j, 63)) = __float2half_rn(b - __half2float(h));
}

// whole slab (fp32 landing zone) -> operand image; once per launch
__device__ void convert_weights_x(uint8_t* sm, const float* L, const ShapeX& s) {
  const int tid = threadIdx.x;
  {  // W2 [h2][ld1]: 16 lanes walk one row
    const int q = tid & 15, nq = s.ld1 >> 2;
    for (int j = tid >> 4; j < s.h2; j += XT / 16)
      if (q < nq) put_w2(sm, j, 4 * q, *reinterpret_cast<const float4*>(L + s.w_off1 + j * s.ld1 + 4 * q));
  }
  {  // W1 [h1][ld0] (+ b1 into column 15)
    const int q = tid & 3, nq = s.ld0 >> 2;
    for (int i = tid >> 2; i < s.h1; i += XT / 4) {
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      if (q < nq) w = *reinterpret_cast<const float4*>(L + s.w_off0 + i * s.ld0 + 4 * q);
      if (q == 3) w.w = L[s.b_off0 + i];   // d0 <= 15: column 15 is free for the bias
      put_w1(sm, i, 4 * q, w);
    }
  }
  for (int j = tid; j < s.h2; j += XT) {
    put_b2(sm, j, L[s.b_off1 + j]);
    float* w3 = reinterpret_cast<float*>(sm + SX_W3) + j * 8;
    for (int k = 0; k < s.C; ++k) w3[k] = L[s.w_off2 + k * s.ld2 + j];
  }
  if (tid < s.C) reinterpret_cast<float*>(sm + SX_B3)[tid] = L[s.b_off2 + tid];
  tc::fence_async_smem();
  __syncthreads();
}

// Row threads: columns 4 part .. 4 part + 3 of data row `pos` (global -> registers; issued early)
struct RowRegsX {
  float4 q;
  int label;
};
__device__ __forceinline__ RowRegsX fetch_row_x(const TcArgs& a, int64_t pos, bool valid, int part) {
  RowRegsX r;
  r.label = 0;
  r.q = make_float4(0.f, 0.f, 0.f, 0.f);
  if (valid) {
    const int row = __ldg(a.idx + pos);
    r.q = __ldg(reinterpret_cast<const float4*>(a.X + (size_t)row * a.ldx) + part);
    r.label = __ldg(a.y + row);
  }
  return r;
}
// Host-input mode (TcArgs::x_host): the tile's rows come from pinned host memory, 64 bytes per row in one request by four
// neighbouring lanes (tools/host_pull_bench.cu: the request size is what matters over the host link), written through
// to the device-resident X / y.  Threads 0..255 fetch (row t / 4, chunk t % 4) — which is exactly what row thread
// (row, part) needs, so the hand-over goes through shared memory (the H1 panel, idle at that point).
struct HostRegsX {
  float4 v;
  int grow, label;
};
// part 1 (issued as early as possible: the loads cross the host link, ~2 us): threads 0..255 fetch (row t / 4, chunk t % 4)
__device__ __forceinline__ HostRegsX host_fetch_x(const TcArgs& a, int64_t pos0, int nvalid, int tid) {
  HostRegsX h;
  h.v = make_float4(0.f, 0.f, 0.f, 0.f);
  h.grow = -1;
  h.label = 0;
  const int rl = tid >> 2, c = tid & 3;
  if (tid < 4 * XR && rl < nvalid) {
    h.grow = __ldg(a.idx + pos0 + rl);
    h.v = __ldcv(reinterpret_cast<const float4*>(a.x_host + (size_t)h.grow * a.ldx) + c);
    if (c == 0) h.label = __ldcv(a.y_host + h.grow);
  }
  return h;
}
// part 2: write-through to X / y and hand-over to the row threads through shared memory
__device__ __forceinline__ RowRegsX host_deliver_x(uint8_t* sm, const TcArgs& a, const HostRegsX& h, int tid, int row, int part, bool rowt) {
  float* stage = reinterpret_cast<float*>(sm + SX_H1);           // 64 rows x 16 floats, then 64 labels
  int* stage_lab = reinterpret_cast<int*>(sm + SX_H1 + XR * 64);
  if (tid < 4 * XR) {
    const int rl = tid >> 2, c = tid & 3;
    if (h.grow >= 0) {
      *(reinterpret_cast<float4*>(const_cast<float*>(a.X) + (size_t)h.grow * a.ldx) + c) = h.v;
      if (c == 0) const_cast<int32_t*>(a.y)[h.grow] = h.label;
    }
    *reinterpret_cast<float4*>(stage + rl * 16 + 4 * c) = h.v;
    if (c == 0) stage_lab[rl] = h.label;
  }
  __syncthreads();
  RowRegsX r;
  r.label = 0;
  r.q = make_float4(0.f, 0.f, 0.f, 0.f);
  if (rowt) {
    r.q = *reinterpret_cast<const float4*>(stage + row * 16 + 4 * part);
    r.label = stage_lab[row];
  }
  __syncthreads();   // the stage area is the H1 panel: everyone has read before anyone's epilogue writes it
  return r;
}
__device__ __forceinline__ void store_row_x(uint8_t* sm, const RowRegsX& r, bool valid, const ShapeX& s, int row, int part) {
  float f[4] = {r.q.x, r.q.y, r.q.z, r.q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (4 * part + i >= s.d0) f[i] = 0.f;
  if (part == 3) f[3] = valid ? 1.f : 0.f;   // column 15: the bias input
  uint32_t h0, l0, h1, l1;
  split2(f[0], f[1], h0, l0);
  split2(f[2], f[3], h1, l1);
  *reinterpret_cast<uint2*>(sm + SX_X + tc::panel_off(row, 4 * part)) = make_uint2(h0, h1);
  *reinterpret_cast<uint2*>(sm + SX_X + tc::panel_off(row, 16 + 4 * part)) = make_uint2(l0, l1);
}

// epilogue -> next MMA hand-off: my TMEM reads are done, my smem writes are visible to the tensor core
__device__ __forceinline__ void handoff_x() {
  tc::tmem_ld_wait();
  tc::tc_fence_before();
  tc::fence_async_smem();
  __syncthreads();
}

// All MMAs are issued by one thread.  Order inside a group of three products: the two small cross terms first.
struct IssuerX {
  uint64_t kbase, mbase;  // K-major / MN-major descriptor of shared-memory offset 0
  uint32_t tmem;
  __device__ __forceinline__ uint64_t kd(int off) const { return kbase + (uint64_t)(off >> 4); }
  __device__ __forceinline__ uint64_t md(int off) const { return mbase + (uint64_t)(off >> 4); }
  __device__ __forceinline__ void fwd1() const {   // D1 = X W1^T: K = 16, hi / lo side by side in both panels
    const uint32_t id = idesc_f16(128, 64, false, false);
    tc::mma_bf16(tmem + TX_D1, kd(SX_X + 32), kd(SX_W1), id, false);
    tc::mma_bf16(tmem + TX_D1, kd(SX_X), kd(SX_W1 + 32), id, true);
    tc::mma_bf16(tmem + TX_D1, kd(SX_X), kd(SX_W1), id, true);
  }
  __device__ __forceinline__ void fwd2(const ShapeX& s) const {   // D2 = H1 W2^T
    const uint32_t id = idesc_f16(128, s.N2, false, false);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      tc::mma_bf16(tmem + TX_D2, kd(SX_H1 + XP + 32 * k), kd(SX_W2 + 32 * k), id, k > 0);
      tc::mma_bf16(tmem + TX_D2, kd(SX_H1 + 32 * k), kd(SX_W2L + 32 * k), id, true);
      tc::mma_bf16(tmem + TX_D2, kd(SX_H1 + 32 * k), kd(SX_W2 + 32 * k), id, true);
    }
  }
  __device__ __forceinline__ void wgrad3(const ShapeX& s, bool acc) const {   // DW3[j][0..7 | 8..15] = H2^T [dZ3 hi | dZ3 lo]
    const uint32_t id = idesc_f16(128, 16, true, true);
#pragma unroll 1
    for (int mt = 0; mt < s.MT2; ++mt) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t b = md(SX_X + 2 * DZ3_COL + 2048 * k);
        tc::mma_bf16(tmem + TX_DW3 + 16 * mt, md(SX_Z + (4 + 2 * mt) * XP + 2048 * k), b, id, acc || k > 0);
        tc::mma_bf16(tmem + TX_DW3 + 16 * mt, md(SX_Z + 2 * mt * XP + 2048 * k), b, id, true);
      }
    }
  }
  __device__ __forceinline__ void dgrad2_wgrad2(const ShapeX& s, bool acc) const {
    const uint32_t id4 = idesc_f16(128, 64, false, true);   // dH1 = dZ2 W2
#pragma unroll 1
    for (int c = 0; c < (s.N2 >> 4); ++c) {   // 16 columns of dZ2 per step; never feeds unwritten pad columns
      const int p = c >> 2, k = c & 3, wrow = c * 16 * 128;
      tc::mma_bf16(tmem + TX_D1, kd(SX_Z + (4 + p) * XP + 32 * k), md(SX_W2 + wrow), id4, c > 0);
      tc::mma_bf16(tmem + TX_D1, kd(SX_Z + p * XP + 32 * k), md(SX_W2L + wrow), id4, true);
      tc::mma_bf16(tmem + TX_D1, kd(SX_Z + p * XP + 32 * k), md(SX_W2 + wrow), id4, true);
    }
    const uint32_t id5 = idesc_f16(128, 64, true, true);    // dW2 += dZ2^T H1
#pragma unroll 1
    for (int mt = 0; mt < s.MT2; ++mt) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        tc::mma_bf16(tmem + TX_DW2 + 64 * mt, md(SX_Z + (4 + 2 * mt) * XP + 2048 * k), md(SX_H1 + 2048 * k), id5, acc || k > 0);
        tc::mma_bf16(tmem + TX_DW2 + 64 * mt, md(SX_Z + 2 * mt * XP + 2048 * k), md(SX_H1 + XP + 2048 * k), id5, true);
        tc::mma_bf16(tmem + TX_DW2 + 64 * mt, md(SX_Z + 2 * mt * XP + 2048 * k), md(SX_H1 + 2048 * k), id5, true);
      }
    }
  }
  __device__ __forceinline__ void wgrad1(bool acc) const {   // DW1[i][0..15 | 16..31] = dZ1^T [X hi | X lo]
    const uint32_t id = idesc_f16(128, 32, true, true);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      tc::mma_bf16(tmem + TX_DW1, md(SX_H1 + XP + 2048 * k), md(SX_X + 2048 * k), id, acc || k > 0);
      tc::mma_bf16(tmem + TX_DW1, md(SX_H1 + 2048 * k), md(SX_X + 2048 * k), id, true);
    }
  }
};

}  // namespace

// Returns true when the optimizer phase has already produced the FedAvg result (single-client federation on one GPU:
// global = (w n) / n written by the owners, C:110-115), so the tail only scores the round.
__device__ __forceinline__ bool round_body_x(const TcArgs& a, uint8_t* smem_raw, TailPre* pre) {
  __shared__ uint64_t bar, wbar, gbar;   // MMA phase done | fp32 slab landed | every sender's partials for my rows are in
  __shared__ uint32_t tmem_slot;
  __shared__ float red_f[2][9];   // per-warp partial sums: loss + db3[8]
  __shared__ int cnt[FEDB_MAX_CLASSES * FEDB_MAX_CLASSES];
  __shared__ StepScalars sc_s;
  __shared__ FedbGroup Gs;

  cg::cluster_group cluster = cg::this_cluster();
  const int CS = (int)cluster.num_blocks(), q = (int)cluster.block_rank();
  const int g = blockIdx.x / CS;
  if (threadIdx.x < sizeof(FedbGroup) / 4)
    reinterpret_cast<int32_t*>(&Gs)[threadIdx.x] = reinterpret_cast<const int32_t*>(a.groups + g)[threadIdx.x];
  __syncthreads();
  const FedbGroup& G = Gs;
  if (a.active && !a.active[g]) return false;                     // cluster-uniform exits
  if (a.fed_state && a.fed_state[G.fed].stop) return false;
  const int n = G.n_rows;
  const int bs = G.batch_size > 0 ? min(G.batch_size, n) : n;
  const int b0 = a.batch * bs;
  if (b0 >= n) return false;
  const int nb = min(bs, n - b0);
  const int ntiles = (nb + XR - 1) / XR;

  const uint32_t raw = tc::smem_u32(smem_raw);
  uint8_t* sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  const uint32_t sm_u = tc::smem_u32(sm);

  ShapeX s;
  s.d0 = G.dims[0]; s.h1 = G.dims[1]; s.h2 = G.dims[2]; s.C = G.dims[3];
  s.N2 = (s.h2 + 15) & ~15; s.MT2 = (s.N2 + 127) >> 7;
  s.ld0 = ru4(s.d0); s.ld1 = ru4(s.h1); s.ld2 = ru4(s.h2);
  {
    const SlabLayout lay = slab_layout(G.dims, 3);
    s.w_off0 = lay.w_off[0]; s.b_off0 = lay.b_off[0]; s.w_off1 = lay.w_off[1]; s.b_off1 = lay.b_off[1];
    s.w_off2 = lay.w_off[2]; s.b_off2 = lay.b_off[2]; s.end = lay.end;
  }
  const OwnX own = own_geometry(s.h1, s.h2, s.C, CS);
  const int share4 = own.share >> 2;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int lane128 = (warp & 3) * 32 + lane;   // this thread's TMEM lane
  const int part = warp >> 2;                   // column quarter
  const bool rowt = (warp & 2) == 0;            // owns data row `lane128` (< 64) in the forward / backward epilogues
  const int row = lane128;
  const int step0 = a.step[g];
  float* Pg = a.params + G.slab_off;

  // single-GPU, single-client federation: the owners write the FedAvg result themselves
  const bool inline_avg = a.tail_mode == 1 && gridDim.x == (unsigned)CS && a.n_feds == 1 && a.feds[0].n_clients == 1;
  uint64_t p2p_seq = 0;
  float* avg_out = nullptr;    // where (w n_k) / N goes on the last step: the peers' fetch area, or the global arena
  if (a.tail_mode == 2) {
    p2p_seq = *reinterpret_cast<volatile uint64_t*>(a.p2p->seq);
    avg_out = reinterpret_cast<float*>(a.p2p->sym[a.p2p->rank] + a.p2p->sym_partial_off) + (p2p_seq & 1) * a.p2p->global_total;
  } else if (inline_avg) {
    avg_out = a.global + a.feds[0].global_off;
  }
  float* landing = reinterpret_cast<float*>(sm + SX_STAGE);   // fp32 slab (at launch and after every step); outgoing partials in between
  float* stage = landing;
  const float* slots = reinterpret_cast<const float*>(sm + SX_SLOTS);
  if (warp == 0) tc::tmem_alloc<512>(&tmem_slot);
  if (tid == 0) {
    tc::mbar_init(&bar, 1);
    tc::mbar_init(&wbar, 1);
    tc::mbar_init(&gbar, 1);
    tc::fence_mbar_init();
    bulk_load_begin_x(reinterpret_cast<uint8_t*>(landing), Pg, (uint32_t)s.end * 4u, &wbar);
  }
  for (int i = tid; i < (SX_B3 + 64 - SX_W1) / 16; i += XT) *reinterpret_cast<uint4*>(sm + SX_W1 + i * 16) = make_uint4(0, 0, 0, 0);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  // every CTA's mbarriers exist before any peer's copy can name them: arrive now, wait before the first send
  if (CS > 1) asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  bool cluster_waited = CS == 1;
  const uint32_t tmem = tmem_slot;
  const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  const IssuerX mma{tc::make_desc(sm_u, 0, 1024), tc::make_desc(sm_u, XP, 1024), tmem};
  uint32_t ph = 0, gph = 0, wph = 0;
  float* W3s = reinterpret_cast<float*>(sm + SX_W3);
  float* b3s = reinterpret_cast<float*>(sm + SX_B3);
  const float inv_n = 1.f / (float)nb;
  const float gscale = inv_n * (1.f / DZS);
  const int my_j0 = q * own.R2, my_i0 = (CS - 1 - q) * own.R1;   // first W2 row / W1 row this CTA owns

#define STAMP(i) do { if (a.prof && g == 0 && tid == 0) tc_stamp(a, q, i); } while (0)
  STAMP(0);
  RowRegsX rr;
  rr.label = 0;
  rr.q = make_float4(0.f, 0.f, 0.f, 0.f);
  HostRegsX hr;
  hr.v = make_float4(0.f, 0.f, 0.f, 0.f); hr.grow = -1; hr.label = 0;
  if (a.x_host) hr = host_fetch_x(a, G.idx_off + b0 + q * XR, q < ntiles ? min(XR, nb - q * XR) : 0, tid);
  else if (rowt) rr = fetch_row_x(a, G.idx_off + b0 + q * XR + row, q < ntiles && q * XR + row < nb, part);
  tc::mbar_wait(&wbar, wph); wph ^= 1;
  convert_weights_x(sm, landing, s);
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

  // split of the layer-2 columns between the four parts (in 16-column chunks)
  const int nch2 = s.N2 >> 4;
  const int ch_lo = part * (nch2 >> 2) + min(part, nch2 & 3), ch_hi = ch_lo + (nch2 >> 2) + (part < (nch2 & 3) ? 1 : 0);
  uint8_t* const h1_hi = sm + SX_H1;
  uint8_t* const h1_lo = sm + SX_H1 + XP;
  const int c1 = 16 * part;   // this part's 16 columns of the 64-wide layer-1 accumulators
  bool first_fetch = true;

  for (int st = 0; st < a.n_steps; ++st) {
    if (tid == 0 && CS > 1) mbar_expect(&gbar, (uint32_t)(CS * own.share * 4));   // (a peer's bytes may arrive first: the count goes negative)
    float loss_acc = 0.f, db3_acc[FEDB_MAX_CLASSES];
#pragma unroll
    for (int k = 0; k < FEDB_MAX_CLASSES; ++k) db3_acc[k] = 0.f;
    int my_tiles = 0;

    for (int tile = q; tile < ntiles; tile += CS, ++my_tiles) {
      const bool acc = my_tiles > 0;
      const int r = tile * XR + row;
      const bool valid = rowt && r < nb;
      if (a.x_host) {
        if (!first_fetch) hr = host_fetch_x(a, G.idx_off + b0 + tile * XR, min(XR, nb - tile * XR), tid);
        rr = host_deliver_x(sm, a, hr, tid, row, part, rowt);
      } else if (rowt && !first_fetch) {
        rr = fetch_row_x(a, G.idx_off + b0 + r, valid, part);
      }
      first_fetch = false;
      const int label = rr.label;
      if (rowt) store_row_x(sm, rr, valid, s, row, part);
      handoff_x();
      STAMP(2);
      // ---- layer 1 -----------------------------------------------------------------------
      FEDB_MMA_PHASE(mma.fwd1());
      STAMP(3);
      uint32_t mask1 = 0;   // ReLU mask of this part's 16 columns of H1
      if (rowt) {
        float v[16];
        tc::tmem_ld16(tl + TX_D1 + c1, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          if (v[i] > 0.f) { mask1 |= 1u << i; v[i] = fminf(v[i] * INV_WS, HMAX); } else v[i] = 0.f;
        }
        if (part == 3) { v[15] = valid ? 1.f : 0.f; mask1 &= ~(1u << 15); }   // column 63: the bias input of layer 2
        st_panel8(h1_hi, h1_lo, row, c1, v);
        st_panel8(h1_hi, h1_lo, row, c1 + 8, v + 8);
      }
      handoff_x();
      STAMP(4);
      // ---- layer 2 + head ----------------------------------------------------------------
      FEDB_MMA_PHASE(mma.fwd2(s));
      STAMP(5);
      float z[FEDB_MAX_CLASSES], dz3[FEDB_MAX_CLASSES];
#pragma unroll
      for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = dz3[k] = 0.f;
      if (rowt) {
#pragma unroll 1
        for (int ch = ch_lo; ch < ch_hi; ++ch) {
          const int c0 = ch << 4;
          float v[16];
          tc::tmem_ld16(tl + TX_D2 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = fminf(fmaxf(v[i], 0.f) * INV_WS, HMAX);
          if (s.C <= 2) {
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
          uint8_t* pz = sm + SX_Z + (c0 >> 6) * XP;
          st_panel8(pz, pz + 4 * XP, row, c0 & 63, v);
          st_panel8(pz, pz + 4 * XP, row, (c0 & 63) + 8, v + 8);
        }
        // the four parts swap their partial logits through 32 spare TMEM columns of the row's lane
        tc::tmem_st8(tl + TX_ZX + 8 * part, z);
        tc::tmem_st_wait();
      }
      tc::tc_fence_before();
      __syncthreads();
      tc::tc_fence_after();
      if (rowt) {
        {
          float z0[16], z1[16];
          tc::tmem_ld16(tl + TX_ZX, z0);
          tc::tmem_ld16(tl + TX_ZX + 16, z1);
          tc::tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = ((z0[k] + z0[8 + k]) + (z1[k] + z1[8 + k])) + (k < s.C ? b3s[k] : 0.f);
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
            const float eps = 1.1920929e-07f;
            const float pc = fminf(fmaxf(p, eps), 1.f - eps), yt = (float)label;
            loss = -(yt * logf(pc) + (1.f - yt) * logf(1.f - pc));
            dz3[0] = (p - yt) * DZS;
          }
        }
        if (part == 0) {   // one thread per row books the loss / db3 and publishes dZ3 as [hi | lo] next to X
          loss_acc += loss;
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k) db3_acc[k] += dz3[k];
          uint4 dh, dl;
          split8(dz3, dh, dl);
          *reinterpret_cast<uint4*>(sm + SX_X + tc::panel_off(row, DZ3_COL)) = dh;
          *reinterpret_cast<uint4*>(sm + SX_X + tc::panel_off(row, DZ3_COL + 8)) = dl;
        }
      }
      handoff_x();
      STAMP(6);
      // ---- head weight gradient (needs H2), then dZ2 overwrites H2 --------------------------
      FEDB_MMA_PHASE(mma.wgrad3(s, acc));
      STAMP(7);
      if (rowt) {
#pragma unroll 1
        for (int ch = ch_lo; ch < ch_hi; ++ch) {
          const int c0 = ch << 4;
          float v[16], d[16];
          tc::tmem_ld16(tl + TX_D2 + c0, v);
          tc::tmem_ld_wait();
          if (s.C <= 2) {
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
          uint8_t* pz = sm + SX_Z + (c0 >> 6) * XP;
          st_panel8(pz, pz + 4 * XP, row, c0 & 63, d);
          st_panel8(pz, pz + 4 * XP, row, (c0 & 63) + 8, d + 8);
        }
      }
      handoff_x();
      STAMP(8);
      // ---- dH1 = dZ2 W2 ; dW2 += dZ2^T H1 ------------------------------------------------------
      FEDB_MMA_PHASE(mma.dgrad2_wgrad2(s, acc));
      STAMP(9);
      if (rowt) {
        float v[16];
        tc::tmem_ld16(tl + TX_D1 + c1, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = ((mask1 >> i) & 1u) ? v[i] * INV_WS : 0.f;
        st_panel8(h1_hi, h1_lo, row, c1, v);        // dZ1 takes H1's place
        st_panel8(h1_hi, h1_lo, row, c1 + 8, v + 8);
      }
      handoff_x();
      STAMP(10);
      // ---- dW1 += dZ1^T X ------------------------------------------------------------------------
      FEDB_MMA_PHASE(mma.wgrad1(acc));
      STAMP(11);
    }  // tiles

    // ---- reduce-scatter, sender side: weight-gradient accumulators (TMEM) -> staging blocks by owner -> one bulk copy per owner
    {
      const bool have = my_tiles > 0;
      float v[16];
#pragma unroll 1
      for (int mt = 0; mt < s.MT2; ++mt) {
        const int j = 128 * mt + lane128;
        const int o = j / own.R2;
        float* rec = stage + o * own.share + (j - o * own.R2) * own.recf;
        if (have) { tc::tmem_ld16(tl + TX_DW2 + 64 * mt + c1, v); tc::tmem_ld_wait(); }
        else {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = 0.f;
        }
        if (j < s.h2) {
#pragma unroll
          for (int i4 = 0; i4 < 4; ++i4)
            if (c1 + 4 * i4 < s.ld1)   // columns h1 .. ld1-1 of the accumulator are exact zeros (H1's pad columns are)
              *reinterpret_cast<float4*>(rec + c1 + 4 * i4) =
                  make_float4(v[4 * i4] * gscale, v[4 * i4 + 1] * gscale, v[4 * i4 + 2] * gscale, v[4 * i4 + 3] * gscale);
          if (part == 3) rec[s.ld1] = v[15] * gscale;   // column 63 multiplied the constant-1 input: db2
        }
        if (part == 0) {   // head weights: dW3[k][j] = DW3[j][k] + DW3[j][8 + k]
          if (have) { tc::tmem_ld16(tl + TX_DW3 + 16 * mt, v); tc::tmem_ld_wait(); }
          if (j < s.h2) {
            float g3[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) g3[k] = (v[k] + v[8 + k]) * gscale;
            if (s.C <= 3) {
              rec[s.ld1 + 1] = g3[0]; rec[s.ld1 + 2] = g3[1]; rec[s.ld1 + 3] = g3[2];
            } else {
              rec[s.ld1 + 1] = 0.f; rec[s.ld1 + 2] = 0.f; rec[s.ld1 + 3] = 0.f;
              *reinterpret_cast<float4*>(rec + s.ld1 + 4) = make_float4(g3[0], g3[1], g3[2], g3[3]);
              *reinterpret_cast<float4*>(rec + s.ld1 + 8) = make_float4(g3[4], g3[5], g3[6], g3[7]);
            }
          }
        }
      }
      if (part == 1) {     // layer 1: dW1[i][c] = DW1[i][c] + DW1[i][16 + c] (column 15: db1)
        float w[16];
        if (have) {
          tc::tmem_ld16(tl + TX_DW1, v);
          tc::tmem_ld16(tl + TX_DW1 + 16, w);
          tc::tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = w[i] = 0.f;
        }
        if (lane128 < s.h1) {
          const int blk = lane128 / own.R1;
          float* rec = stage + (CS - 1 - blk) * own.share + own.R2 * own.recf + (lane128 - blk * own.R1) * 16;
#pragma unroll
          for (int i4 = 0; i4 < 4; ++i4)
            *reinterpret_cast<float4*>(rec + 4 * i4) = make_float4((v[4 * i4] + w[4 * i4]) * gscale, (v[4 * i4 + 1] + w[4 * i4 + 1]) * gscale,
                                                                   (v[4 * i4 + 2] + w[4 * i4 + 2]) * gscale, (v[4 * i4 + 3] + w[4 * i4 + 3]) * gscale);
        }
      }
      // loss and db3: warp shuffle, then the two row warps of part 0 in a fixed order
      if (rowt && part == 0) {
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
    }
    STAMP(12);
    // ---- owner, before the partials arrive: the floats of my share that are parameters, and their weights / moments (these
    // do not depend on the peers).  One scalar per task, TK tasks per thread and pass, so every thread of the CTA takes part.
    constexpr int TK = 2;
    const int w2f = own.R2 * own.recf, w1f = own.R1 * 16;
    const bool use_m = a.m != nullptr, use_v = a.v && G.opt != FEDB_OPT_SGD;
    float* Mw = a.m + G.slab_off;
    float* Vw = a.v + G.slab_off;
    // float f of my share -> slab index (-1: padding / not mine; -2: the loss sum); bit 30 marks a weight (L2 applies)
    auto plan = [&](int f) -> int {
      if (f < w2f) {
        const int jr = f / own.recf, c = f - jr * own.recf, j = my_j0 + jr;
        if (j >= s.h2) return -1;
        if (c < s.h1) return (s.w_off1 + j * s.ld1 + c) | (1 << 30);
        if (c == s.ld1) return s.b_off1 + j;
        const int k = c - s.ld1 - (s.C > 3 ? 4 : 1);
        return (c > s.ld1 && k >= 0 && k < s.C) ? ((s.w_off2 + k * s.ld2 + j) | (1 << 30)) : -1;
      }
      if (f < w2f + w1f) {
        const int t = f - w2f, i = my_i0 + (t >> 4), col = t & 15;
        if (i >= s.h1) return -1;
        if (col < s.d0) return (s.w_off0 + i * s.ld0 + col) | (1 << 30);
        return col == 15 ? s.b_off0 + i : -1;
      }
      if (q != CS - 1 || f >= own.share) return -1;
      const int t = f - w2f - w1f;
      return t < s.C ? s.b_off2 + t : (t == 8 ? -2 : -1);
    };
    int te[TK];
    float uw[TK], um[TK], uv[TK];
    auto fetch = [&](int f0) {
#pragma unroll
      for (int i = 0; i < TK; ++i) {
        te[i] = plan(f0 + i * XT);
        uw[i] = um[i] = uv[i] = 0.f;
        if (te[i] >= 0) {
          const int e = te[i] & ~(1 << 30);
          uw[i] = __ldcg(Pg + e);
          if (use_m) um[i] = __ldcg(Mw + e);
          if (use_v) uv[i] = __ldcg(Vw + e);
        }
      }
    };
    fetch(tid);
    tc::tc_fence_before();
    tc::fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      float* misc = stage + (CS - 1) * own.share + w2f + w1f;
#pragma unroll
      for (int i = 0; i < 9; ++i) misc[i == 0 ? 8 : i - 1] = red_f[0][i] + red_f[1][i];   // [db3[0..7], loss]
      sc_s = step_scalars_fast(G, step0 + st + 1);
      tc::fence_async_smem();
    }
    if (!cluster_waited) {
      asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
      cluster_waited = true;
    }
    __syncthreads();
    if (CS > 1) {
      if (tid < CS) {   // my staging block for owner o -> slot `q` of owner o; the owners are visited in a rotated order so that
                        // no receiver is the target of all senders at once
        const int o = (q + 1 + tid) % CS;
        bulk_s2c(mapa(sm_u + SX_SLOTS + (uint32_t)(q * own.share) * 4u, o), sm_u + SX_STAGE + (uint32_t)(o * own.share) * 4u,
                 (uint32_t)own.share * 4u, mapa(tc::smem_u32(&gbar), o));
      }
      mbar_wait_cluster(&gbar, gph);
      gph ^= 1;
    }
    STAMP(13);

    // ---- owner: fixed-order sum of the CS partials, optimizer; master weights / moments (and, on the last step, this client's
    // FedAvg term) go to global memory, from where every CTA rebuilds its operand image after the cluster barrier below
    {
      const StepScalars sc = sc_s;
      const float* parts = CS > 1 ? slots : stage;
      const bool emit_avg = avg_out != nullptr && st == a.n_steps - 1;
      const float al = G.alpha * inv_n;
#pragma unroll 1
      for (int f0 = tid; f0 < own.share; f0 += TK * XT) {
        if (f0 != tid) fetch(f0);
        float gs[TK];
#pragma unroll
        for (int i = 0; i < TK; ++i) {
          gs[i] = 0.f;
          if (te[i] != -1) {
#pragma unroll 4
            for (int rk = 0; rk < CS; ++rk) gs[i] += parts[rk * own.share + f0 + i * XT];
          }
        }
#pragma unroll
        for (int i = 0; i < TK; ++i) {
          if (te[i] == -2) {
            if (a.loss_out) a.loss_out[g] = gs[i] * inv_n;
          } else if (te[i] >= 0) {
            const int e = te[i] & ~(1 << 30);
            const float gr = (te[i] >> 30) ? fmaf(al, uw[i], gs[i]) : gs[i];
            const float w = opt_update(G, sc, uw[i], gr, um[i], uv[i]);
            if (use_m) Mw[e] = um[i];
            if (use_v) Vw[e] = uv[i];
            Pg[e] = w;
            if (emit_avg) avg_out[e] = __fdiv_rn(__fmul_rn(w, G.agg_num), G.agg_den);   // (w n_k) / N, C:110-115
          }
        }
      }
    }
    STAMP(14);
    // ---- all-gather through L2 (its path into an SM is several times wider than the SM-to-SM fabric): every owner's stores
    // are ordered before the barrier, the bulk load of the whole slab after it; the staging area is free again by then (all
    // owners have passed their wait, so every outgoing copy has completed)
    if (CS > 1) cluster.sync(); else __syncthreads();
    if (tid == 0) bulk_load_begin_x(reinterpret_cast<uint8_t*>(landing), Pg, (uint32_t)s.end * 4u, &wbar);
    if (a.tail_mode == 2 && st == a.n_steps - 1 && q == 0 && tid < a.p2p->world) {
      // arrival flag A ("my partial is complete") into every peer GPU's buffer; the local evaluation below overlaps its
      // flight.  The cluster barrier ordered every owner's stores before this thread: fence + release make them visible
      // to whoever acquires the flag.
      __threadfence_system();
      uint64_t* pf = reinterpret_cast<uint64_t*>(a.p2p->sym[tid] + a.p2p->sym_flag_off);
      st_release_sys_u64(pf + a.p2p->rank, p2p_seq + 1);
    }
    STAMP(15);
    tc::mbar_wait(&wbar, wph); wph ^= 1;
    convert_weights_x(sm, landing, s);
    STAMP(16);
  }  // steps

  // ---- local evaluation with the updated weights (C:75-83) ----------------------------------
  if (a.do_eval) {
    if (tid < FEDB_MAX_CLASSES * FEDB_MAX_CLASSES) cnt[tid] = 0;
    __syncthreads();
    for (int tile = q; tile < ntiles; tile += CS) {
      const int r = tile * XR + row;
      const bool valid = rowt && r < nb;
      if (!(ntiles <= CS && a.n_steps > 0)) {   // (single tile per CTA: the X panel from the training step is still in place)
        if (a.x_host) {
          if (!first_fetch) hr = host_fetch_x(a, G.idx_off + b0 + tile * XR, min(XR, nb - tile * XR), tid);
          rr = host_deliver_x(sm, a, hr, tid, row, part, rowt);
        } else if (rowt && !first_fetch) {
          rr = fetch_row_x(a, G.idx_off + b0 + r, valid, part);
        }
        first_fetch = false;
        if (rowt) store_row_x(sm, rr, valid, s, row, part);
      }
      const int label = rr.label;
      handoff_x();
      FEDB_MMA_PHASE(mma.fwd1());
      if (rowt) {
        float v[16];
        tc::tmem_ld16(tl + TX_D1 + c1, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = fminf(fmaxf(v[i], 0.f) * INV_WS, HMAX);
        if (part == 3) v[15] = valid ? 1.f : 0.f;
        st_panel8(h1_hi, h1_lo, row, c1, v);
        st_panel8(h1_hi, h1_lo, row, c1 + 8, v + 8);
      }
      handoff_x();
      FEDB_MMA_PHASE(mma.fwd2(s));
      float z[FEDB_MAX_CLASSES];
#pragma unroll
      for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = 0.f;
      if (rowt) {
#pragma unroll 1
        for (int ch = ch_lo; ch < ch_hi; ++ch) {
          const int c0 = ch << 4;
          float v[16];
          tc::tmem_ld16(tl + TX_D2 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = fminf(fmaxf(v[i], 0.f) * INV_WS, HMAX);
          if (s.C <= 2) {
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
        }
        tc::tmem_st8(tl + TX_ZX + 8 * part, z);
        tc::tmem_st_wait();
      }
      tc::tc_fence_before();
      __syncthreads();
      tc::tc_fence_after();
      if (rowt && part == 0) {
        float z0[16], z1[16];
        tc::tmem_ld16(tl + TX_ZX, z0);
        tc::tmem_ld16(tl + TX_ZX + 16, z1);
        tc::tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = ((z0[k] + z0[8 + k]) + (z1[k] + z1[8 + k])) + (k < s.C ? b3s[k] : 0.f);
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
    if (!cluster_waited) {   // (evaluation-only launch: the peers' shared memory is initialised once this completes)
      asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
      cluster_waited = true;
    }
    if (tid < FEDB_MAX_CLASSES * FEDB_MAX_CLASSES && cnt[tid] != 0) {
      if ((a.tail_mode == 1 || a.tail_mode == 2) && gridDim.x == (unsigned)CS)   // single cluster: straight into the scoring CTA's shared memory
        asm volatile("red.relaxed.cluster.shared::cluster.add.u32 [%0], %1;" ::"r"(mapa(tc::smem_u32(&pre->cnt[tid]), CS - 1)), "r"((uint32_t)cnt[tid]) : "memory");
      else
        atomicAdd(a.conf + (size_t)g * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES + tid, cnt[tid]);
    }
    STAMP(17);
  }
#undef FEDB_MMA_PHASE
#undef STAMP

  if (q == 0 && tid == 0) a.step[g] = step0 + a.n_steps;
  if (!cluster_waited) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");   // (n_steps == 0: complete the pending phase)
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<512>(tmem);
  // a CTA's shared memory is the source of its outgoing copies and the target of its peers': nobody leaves before all have
  // received everything (the tails synchronise the cluster themselves)
  if (a.tail_mode == 0 && CS > 1) cluster.sync();
  return inline_avg && a.n_steps > 0;
}

__global__ void __launch_bounds__(XT, 1) k_round_tcx(TcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ TailPre pre;
  __shared__ unsigned char s_stop[TAIL_MAX_FEDS];   // stop flags as they stand for THIS round (only a tail ever writes them)
  if (a.tail_mode == 1)
    for (int f = threadIdx.x; f < a.n_feds; f += XT) s_stop[f] = a.fed_state_rw[f].stop ? 1 : 0;
  {
    cg::cluster_group cl = cg::this_cluster();
    const bool single = gridDim.x == cl.num_blocks();
    if (threadIdx.x < FEDB_MAX_CLASSES * FEDB_MAX_CLASSES) pre.cnt[threadIdx.x] = 0;
    if (threadIdx.x == 0) { pre.have_fs = 0; pre.have_cnt = ((a.tail_mode == 1 || a.tail_mode == 2) && single) ? 1 : 0; }
    __syncthreads();
    // the scoring CTA's last warp (idle in every epilogue) fetches the federation record / state while the round runs
    if ((a.tail_mode == 1 || a.tail_mode == 2) && cl.block_rank() == cl.num_blocks() - 1 && (threadIdx.x >> 5) == XT / 32 - 1) tail_prefetch(a, &pre);
  }
  const bool avg_done = round_body_x(a, smem_raw, &pre);
  if (a.tail_mode == 0) return;
  if (a.tail_mode == 2) { p2p_tail(a, &pre); return; }
  if (a.tail_mode == 3) {   // many resident clients, cross-GPU: the rank's last cluster to finish runs the exchange
    __shared__ unsigned int s_last3;
    cg::cluster_group cl = cg::this_cluster();
    const int CSz = (int)cl.num_blocks();
    __threadfence();
    cl.sync();
    const unsigned int n_cl = gridDim.x / CSz;
    if (n_cl > 1) {
      if (cl.block_rank() == 0 && threadIdx.x == 0) {
        const unsigned int last = (atomicInc(a.tail_count, n_cl - 1) == n_cl - 1) ? 1u : 0u;
        for (int r = 0; r < CSz; ++r) *cl.map_shared_rank(&s_last3, r) = last;
      }
      cl.sync();
      if (!s_last3) return;
      __threadfence();
    }
    p2p_tail_many(a, &pre);
    return;
  }
  // the cluster that observes the last check-in runs the single-GPU aggregation tail (see k_local_round_tc)
  __shared__ unsigned int s_last;
  cg::cluster_group cluster = cg::this_cluster();
  const int CS = (int)cluster.num_blocks(), crank = (int)cluster.block_rank();
  const unsigned int n_clusters = gridDim.x / CS;
  if (n_clusters > 1) __threadfence();   // (single cluster: nothing of this round went through global atomics; the barrier orders the rest)
  cluster.sync();
  if (n_clusters > 1) {
    if (crank == 0 && threadIdx.x == 0) {
      const unsigned int last = (atomicInc(a.tail_count, n_clusters - 1) == n_clusters - 1) ? 1u : 0u;
      for (int r = 0; r < CS; ++r) *cluster.map_shared_rank(&s_last, r) = last;
    }
    cluster.sync();
    if (!s_last) return;
    __threadfence();
  }
  if (a.prof && threadIdx.x == 0) tc_stamp(a, crank, 18);
  round_tail(a, crank, CS, s_stop, avg_done, &pre);
  if (a.prof && threadIdx.x == 0) tc_stamp(a, crank, 19);
}

}  // namespace fedb

using namespace fedb;

static constexpr int kTcxSmem = SX_SLOTS + SX_SLOTS_MAX + 1024;

extern "C" int64_t fedb_tcx_smem_bytes(void) { return kTcxSmem; }
extern "C" int32_t fedb_tcx_rows(void) { return XR; }
// 1 if a model with these layer sizes runs with `cluster` CTAs per client: shape limits, and the exchange blocks of all
// senders / owners must fit the receive area and the staging area.
extern "C" int32_t fedb_tcx_fits(int32_t d0, int32_t h1, int32_t h2, int32_t C, int32_t cluster) {
  if (cluster < 1 || cluster > 16 || d0 < 1 || d0 > 15 || h1 < 1 || h1 > 63 || h2 < 1 || h2 > 256 || C < 1 || C > 8) return 0;
  const OwnX o = own_geometry(h1, h2, C, cluster);
  const int64_t bytes = (int64_t)cluster * o.share * 4;
  int32_t dims[4] = {d0, h1, h2, C};
  const SlabLayout lay = slab_layout(dims, 3);
  if ((int64_t)lay.end * 4 > SX_STAGE_MAX) return 0;                         // the fp32 slab lands in the staging area at launch
  return bytes <= SX_STAGE_MAX && (cluster == 1 || bytes <= SX_SLOTS_MAX) ? 1 : 0;
}

// Largest cluster size (<= 16) the device can co-schedule for this kernel; 16 needs the non-portable opt-in and a GPC
// with 16 free SMs.  Cached per device.
extern "C" int32_t fedb_tcx_max_cluster(void) {
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 8;
  if (cached[dev]) return cached[dev];
  if (cudaFuncSetAttribute((const void*)k_round_tcx, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcxSmem) != cudaSuccess ||
      cudaFuncSetAttribute((const void*)k_round_tcx, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
    cudaGetLastError();
    return cached[dev] = 8;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(16);
  cfg.blockDim = dim3(XT);
  cfg.dynamicSmemBytes = kTcxSmem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 16;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, k_round_tcx, &cfg) != cudaSuccess) { cudaGetLastError(); n = 0; }
  return cached[dev] = (n >= 1 ? 16 : 8);
}

static int launch_tcx(TcArgs a, int32_t n_groups, int32_t cluster_size, void* stream) {
  FEDB_REQUIRE(n_groups > 0 && cluster_size >= 1 && cluster_size <= 16, "bad grid");
  FEDB_REQUIRE(a.ldx >= 16 && a.ldx % 4 == 0, "feature rows must be padded to >= 16 floats");
  // per-device attributes: cheap, legal during stream capture, and correct for a process that drives several GPUs
  FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_round_tcx, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcxSmem));
  if (cluster_size > 8) FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_round_tcx, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(n_groups * cluster_size);
  cfg.blockDim = dim3(XT);
  cfg.dynamicSmemBytes = kTcxSmem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster_size;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  FEDB_CUDA_OK(cudaLaunchKernelEx(&cfg, k_round_tcx, a));
  return FEDB_OK;
}

// The three general entry points of the split-fp16 round kernel; arguments exactly as for the bf16 kernel's
// fedb_local_round_tc_prof / fedb_round_tc_hostin / fedb_round_tc_p2p_hostin (include/fedb200.h), cluster_size 1..16.
// The caller checks the model against fedb_tcx_fits (the kernel reads shapes from the device table).
extern "C" int fedb_local_round_tcx(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X, int32_t ldx,
                                    const int32_t* y, const int32_t* idx, float* params, float* m, float* v, int32_t* step,
                                    const int32_t* active, const FedbFedState* fed_state, float* loss_out, int32_t* conf,
                                    float* scratch, int32_t batch, int32_t n_steps, int32_t do_eval, long long* prof, void* stream) {
  TcArgs a{};
  a.groups = groups; a.X = X; a.y = y; a.idx = idx; a.params = params; a.m = m; a.v = v; a.step = step; a.active = active;
  a.fed_state = fed_state; a.loss_out = loss_out; a.conf = conf; a.scratch = scratch; a.ldx = ldx; a.batch = batch;
  a.n_steps = n_steps; a.do_eval = do_eval; a.prof = prof; a.prof_stride = 32; a.tail_mode = 0;
  return launch_tcx(a, n_groups, cluster_size, stream);
}

extern "C" int fedb_round_tcx(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, float* X, int32_t ldx, int32_t* y,
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
  a.n_steps = n_steps; a.do_eval = 1; a.prof = prof; a.prof_stride = 32;
  a.tail_mode = 1; a.n_groups = n_groups; a.n_feds = n_feds; a.n_conf = conf_ints; a.feds = feds; a.fed_state_rw = fed_state;
  a.global = global; a.hist_local = hist_local; a.hist_global = hist_global; a.tail_count = tail_count;
  a.state_mirror = state_mirror;
  return launch_tcx(a, n_groups, cluster_size, stream);
}

extern "C" int fedb_round_tcx_p2p(const FedbGroup* groups, int32_t cluster_size, float* X, int32_t ldx, int32_t* y,
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
  return launch_tcx(a, 1, cluster_size, stream);
}

// Multi-GPU round with ANY number of resident clients per rank in one launch per rank (p2p_tail_many): groups / n_groups are
// this rank's resident clients, tail_count one zeroed uint32; everything else as fedb_round_tcx_p2p.
extern "C" int fedb_round_tcx_p2p_many(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, float* X, int32_t ldx, int32_t* y,
                                       const float* x_host, const int32_t* y_host, const int32_t* idx, float* params, float* m,
                                       float* v, int32_t* step, const int32_t* active, FedbFedState* fed_state, float* loss_out,
                                       int32_t* conf, float* scratch, int32_t n_steps, const FedbP2PRank* rank_dev,
                                       uint32_t* tail_count, FedbFedState* state_mirror, void* stream) {
  FEDB_REQUIRE(rank_dev != nullptr && tail_count != nullptr && n_groups > 0, "rank record and tail counter required");
  FEDB_REQUIRE((x_host == nullptr) == (y_host == nullptr), "x_host and y_host come together");
  TcArgs a{};
  a.x_host = x_host; a.y_host = y_host;
  a.groups = groups; a.X = X; a.y = y; a.idx = idx; a.params = params; a.m = m; a.v = v; a.step = step; a.active = active;
  a.fed_state = fed_state; a.loss_out = loss_out; a.conf = conf; a.scratch = scratch; a.ldx = ldx; a.batch = 0;
  a.n_steps = n_steps; a.do_eval = 1; a.prof = nullptr; a.tail_mode = 3; a.n_groups = n_groups; a.n_feds = 1; a.p2p = rank_dev;
  a.tail_count = tail_count; a.state_mirror = state_mirror;
  return launch_tcx(a, n_groups, cluster_size, stream);
}
