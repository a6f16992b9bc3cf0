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
// kernel's gradients against the fp32 SIMT kernels at 1e-5 of each layer's scale.
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
// an operand are separate, except where an operand is <= 16 columns wide and both fit one panel side by side (X, W1,
// dZ3) — there a single MMA with N = 32 (or 16) against the [hi | lo] panel yields both products at once.
//
// Shapes: 2 hidden ReLU layers, in <= 15, h1 <= 63, h2 <= 256, out <= 8 (the custom script's 14-50-200-2, C:40).
#include <cooperative_groups.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "p2p_common.cuh"
#include "tc_common.cuh"
#include "tc_round_common.cuh"

namespace cg = cooperative_groups;

namespace fedb {
namespace {

constexpr int XR = 64;            // data rows per tile
constexpr int XP = XR * 128;      // bytes of a [64][64] fp16 panel
constexpr float WS = 16.f, INV_WS = 1.f / 16.f;   // weight panels hold WS * w
constexpr float DZS = 32.f;                        // backward operands hold DZS * n * (true gradient)
constexpr float HMAX = 60000.f;                    // activation clamp (fp16 max 65504)
// shared-memory map (bytes from the 1024-aligned base)
constexpr int SX_X = 0;                      // cols 0-15 X hi (col 15 = 1), cols 16-31 X lo
constexpr int SX_DZ3 = SX_X + XP;            // cols 0-7 dZ3 hi, cols 8-15 dZ3 lo
constexpr int SX_H1 = SX_DZ3 + XP;           // H1 hi, H1 lo; later dZ1 hi, dZ1 lo
constexpr int SX_Z = SX_H1 + 2 * XP;         // hi panels 0..3, lo panels 0..3: H2, then dZ2
constexpr int SX_W1 = SX_Z + 8 * XP;         // [64][64]: cols 0-15 W1 hi (col 15 = b1), cols 16-31 lo
constexpr int SX_W2 = SX_W1 + 64 * 128;      // hi [256][64] (col 63 = b2)
constexpr int SX_W2L = SX_W2 + 256 * 128;    // lo
constexpr int SX_W3 = SX_W2L + 256 * 128;    // fp32 [8][256]
constexpr int SX_B3 = SX_W3 + 8 * 256 * 4;   // fp32 [8]
constexpr int SX_END = SX_B3 + 256;
constexpr int SX_SLOTS = SX_END;              // [CS][per] float4 partial-gradient slices + CS loss floats
constexpr int SX_SLOTS_MAX = 227 * 1024 - 4096 - SX_SLOTS;
constexpr int SX_STAGE = SX_H1;               // fp32 slab landing zone overlays H1..Z (idle at those points)
// TMEM columns
constexpr uint32_t TX_D1 = 0, TX_D2 = 64, TX_DW2 = 320, TX_DW1 = 448, TX_DW3 = 480;
constexpr uint32_t TX_ZX = TX_D1;             // logit exchange between the two column halves (D1 is dead at that point)

struct ShapeX {
  int d0, h1, h2, C, N2, MT2;
  int ld0, ld1, ld2;
};

// kind::f16 instruction descriptor with fp16 (not bf16) operands, fp32 accumulation
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N, bool a_mn, bool b_mn) {
  return (1u << 4) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ uint32_t h2u(const __half2 h) { return *reinterpret_cast<const uint32_t*>(&h); }
// (a, b) -> packed fp16 pair of the leading parts and of the remainders
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
  const __half2 h = __floats2half2_rn(a, b);
  const float2 f = __half22float2(h);
  hi = h2u(h);
  lo = h2u(__floats2half2_rn(a - f.x, b - f.y));
}
__device__ __forceinline__ void split8(const float* v, uint4& hi, uint4& lo) {
  split2(v[0], v[1], hi.x, lo.x);
  split2(v[2], v[3], hi.y, lo.y);
  split2(v[4], v[5], hi.z, lo.z);
  split2(v[6], v[7], hi.w, lo.w);
}
__device__ __forceinline__ void st_panel8(uint8_t* hi_panel, uint8_t* lo_panel, int row, int col, const float* v) {
  uint4 hi, lo;
  split8(v, hi, lo);
  const uint32_t off = tc::panel_off(row, col);
  *reinterpret_cast<uint4*>(hi_panel + off) = hi;
  *reinterpret_cast<uint4*>(lo_panel + off) = lo;
}

// distributed shared memory by 32-bit shared::cluster addresses (no 64-bit pointer per peer in registers)
__device__ __forceinline__ uint32_t mapa(uint32_t saddr, int rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_f4(uint32_t a, const float4 v) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_cluster_u2(uint32_t a, const uint32_t x, const uint32_t y) {
  asm volatile("st.shared::cluster.v2.u32 [%0], {%1,%2};" ::"r"(a), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void st_cluster_f1(uint32_t a, const float x) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(a), "f"(x) : "memory");
}

__device__ __forceinline__ void bulk_load_begin_x(uint8_t* dst, const float* src, uint32_t bytes, uint64_t* wbar) {
  asm volatile("fence.proxy.async;" ::: "memory");
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc::smem_u32(wbar)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(tc::smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(tc::smem_u32(wbar))
               : "memory");
}

// W1 (+ b1 in column 15) from the fp32 landing zone into the [hi | lo] panel; b2 into column 63 of the W2 panels
__device__ __forceinline__ void convert_w1_b2(uint8_t* sm, const float* L, const SlabLayout& lay, const ShapeX& s, int tid) {
  {
    const int q = tid & 3, nq = s.ld0 >> 2;
    for (int j = tid >> 2; j < s.h1; j += TC_THREADS / 4) {
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      if (q < nq) w = *reinterpret_cast<const float4*>(L + lay.w_off[0] + j * s.ld0 + 4 * q);
      if (q == 3) w.w = L[lay.b_off[0] + j];   // column 15 carries the bias (d0 <= 15: the slab has padding there)
      uint32_t h0, l0, h1, l1;
      split2(w.x * WS, w.y * WS, h0, l0);
      split2(w.z * WS, w.w * WS, h1, l1);
      *reinterpret_cast<uint2*>(sm + SX_W1 + tc::panel_off(j, 4 * q)) = make_uint2(h0, h1);
      *reinterpret_cast<uint2*>(sm + SX_W1 + tc::panel_off(j, 16 + 4 * q)) = make_uint2(l0, l1);
    }
  }
  for (int j = tid; j < s.h2; j += TC_THREADS) {
    const float b = L[lay.b_off[1] + j] * WS;
    const __half h = __float2half_rn(b);
    *reinterpret_cast<__half*>(sm + SX_W2 + tc::panel_off(j, 63)) = h;
    *reinterpret_cast<__half*>(sm + SX_W2L + tc::panel_off(j, 63)) = __float2half_rn(b - __half2float(h));
  }
}

// whole slab (fp32 landing zone) -> operand image; once per launch
__device__ void convert_weights_x(uint8_t* sm, const float* L, const SlabLayout& lay, const ShapeX& s) {
  const int tid = threadIdx.x;
  {  // W2 [h2][ld1]: 16 lanes walk one row
    const int q = tid & 15, jr = tid >> 4, nq = s.ld1 >> 2;
    for (int j = jr; j < s.h2; j += TC_THREADS / 16) {
      if (q < nq) {
        const float4 w = *reinterpret_cast<const float4*>(L + lay.w_off[1] + j * s.ld1 + 4 * q);
        uint32_t h0, l0, h1, l1;
        split2(w.x * WS, w.y * WS, h0, l0);
        split2(w.z * WS, w.w * WS, h1, l1);
        const uint32_t off = tc::panel_off(j, 4 * q);
        *reinterpret_cast<uint2*>(sm + SX_W2 + off) = make_uint2(h0, h1);
        *reinterpret_cast<uint2*>(sm + SX_W2L + off) = make_uint2(l0, l1);
      }
    }
  }
  convert_w1_b2(sm, L, lay, s, tid);
  {  // W3 [C][ld2] -> fp32 [C][256], b3
    const int nq = s.ld2 >> 2;
    for (int k = 0; k < s.C; ++k)
      for (int q = tid; q < nq; q += TC_THREADS)
        *reinterpret_cast<float4*>(sm + SX_W3 + (k * 256 + 4 * q) * 4) = *reinterpret_cast<const float4*>(L + lay.w_off[2] + k * s.ld2 + 4 * q);
    if (tid < s.C) reinterpret_cast<float*>(sm + SX_B3)[tid] = L[lay.b_off[2] + tid];
  }
  tc::fence_async_smem();
  __syncthreads();
}

struct RowRegsX {
  float4 q[2];
  int label;
};
// Row threads: column half h of data row `pos` (global -> registers; issued early so the latency overlaps other work)
__device__ __forceinline__ RowRegsX fetch_row_x(const TcArgs& a, int64_t pos, bool valid, int half) {
  RowRegsX r;
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
// Host-input mode (TcArgs::x_host): four lanes fetch one whole 64-byte row from pinned host memory with one instruction
// (tools/host_pull_bench.cu: the request size is what matters over the host link); 64 rows x 4 lanes = the whole CTA.
struct HostRegsX {
  float4 v;
  int grow, label;
};
__device__ __forceinline__ HostRegsX host_fetch_x(const TcArgs& a, int64_t pos0, int nvalid, int tid) {
  HostRegsX h;
  const int rl = tid >> 2, c = tid & 3;
  h.v = make_float4(0.f, 0.f, 0.f, 0.f);
  h.grow = -1;
  h.label = 0;
  if (rl < nvalid) {
    const int row = __ldg(a.idx + pos0 + rl);
    h.grow = row;
    h.v = __ldcv(reinterpret_cast<const float4*>(a.x_host + (size_t)row * a.ldx) + c);
    if (c == 0) h.label = __ldcv(a.y_host + row);
  }
  return h;
}
// ... handed to the row threads through shared memory (the H1 panels, idle at that point) and written through to X / y
__device__ __forceinline__ RowRegsX host_deliver_x(uint8_t* sm, const TcArgs& a, const HostRegsX& h, int tid, int row, int half, bool rowt) {
  float* stage = reinterpret_cast<float*>(sm + SX_H1);           // 64 rows x 16 floats, then 64 labels
  int* stage_lab = reinterpret_cast<int*>(sm + SX_H1 + XR * 64);
  const int rl = tid >> 2, c = tid & 3;
  *reinterpret_cast<float4*>(stage + rl * 16 + 4 * c) = h.v;
  if (c == 0) stage_lab[rl] = h.label;
  if (h.grow >= 0) {
    *(reinterpret_cast<float4*>(const_cast<float*>(a.X) + (size_t)h.grow * a.ldx) + c) = h.v;
    if (c == 0) const_cast<int32_t*>(a.y)[h.grow] = h.label;
  }
  __syncthreads();
  RowRegsX r;
  r.label = 0;
  r.q[0] = r.q[1] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (rowt) {
    r.q[0] = *reinterpret_cast<const float4*>(stage + row * 16 + 8 * half);
    r.q[1] = *reinterpret_cast<const float4*>(stage + row * 16 + 8 * half + 4);
    r.label = stage_lab[row];
  }
  __syncthreads();   // the stage area is the H1 panel: everyone has read before anyone's epilogue writes it
  return r;
}
__device__ __forceinline__ int store_row_x(uint8_t* sm, const RowRegsX& r, bool valid, const ShapeX& s, int row, int half) {
  float f[8] = {r.q[0].x, r.q[0].y, r.q[0].z, r.q[0].w, r.q[1].x, r.q[1].y, r.q[1].z, r.q[1].w};
#pragma unroll
  for (int i = 0; i < 8; ++i)
    if (8 * half + i >= s.d0) f[i] = 0.f;
  if (half == 1) f[7] = valid ? 1.f : 0.f;   // column 15: the bias input
  uint4 hi, lo;
  split8(f, hi, lo);
  *reinterpret_cast<uint4*>(sm + SX_X + tc::panel_off(row, 8 * half)) = hi;
  *reinterpret_cast<uint4*>(sm + SX_X + tc::panel_off(row, 16 + 8 * half)) = lo;
  return r.label;
}

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
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
      if (mt < s.MT2) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          tc::mma_bf16(tmem + TX_DW3 + 16 * mt, md(SX_Z + (4 + 2 * mt) * XP + 2048 * k), md(SX_DZ3 + 2048 * k), id, acc || k > 0);
          tc::mma_bf16(tmem + TX_DW3 + 16 * mt, md(SX_Z + 2 * mt * XP + 2048 * k), md(SX_DZ3 + 2048 * k), id, true);
        }
      }
  }
  __device__ __forceinline__ void dgrad2_wgrad2(const ShapeX& s, bool acc) const {
    const uint32_t id4 = idesc_f16(128, 64, false, true);   // dH1 = dZ2 W2
#pragma unroll
    for (int p = 0; p < 4; ++p) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (64 * p + 16 * k < s.N2) {   // never feed unwritten pad columns of the last panel
          const int wrow = (p * 64 + 16 * k) * 128;
          tc::mma_bf16(tmem + TX_D1, kd(SX_Z + (4 + p) * XP + 32 * k), md(SX_W2 + wrow), id4, (p | k) > 0);
          tc::mma_bf16(tmem + TX_D1, kd(SX_Z + p * XP + 32 * k), md(SX_W2L + wrow), id4, true);
          tc::mma_bf16(tmem + TX_D1, kd(SX_Z + p * XP + 32 * k), md(SX_W2 + wrow), id4, true);
        }
    }
    const uint32_t id5 = idesc_f16(128, 64, true, true);    // dW2 += dZ2^T H1
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
      if (mt < s.MT2) {
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

// Thread mapping: 8 warps.  Warp w reaches TMEM lanes 32 (w % 4) .. +31.  Data rows live in lanes 0..63, so the "row
// threads" of the forward / backward epilogues are warps 0, 1 (column half 0) and 4, 5 (column half 1); all eight warps
// take part in the weight-gradient read-out (lanes = output features), the optimizer and the weight conversion.
__device__ __forceinline__ void round_body_x(const TcArgs& a, uint8_t* smem_raw) {
  __shared__ uint64_t bar, wbar;
  __shared__ uint32_t tmem_slot;
  __shared__ float red_f[2][9];   // per-warp partial sums: loss + db3[8]
  __shared__ int cnt[FEDB_MAX_CLASSES * FEDB_MAX_CLASSES];
  __shared__ StepScalars sc_s;

  cg::cluster_group cluster = cg::this_cluster();
  const int CS = (int)cluster.num_blocks(), q = (int)cluster.block_rank();
  const int g = blockIdx.x / CS;
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
  const int ntiles = (nb + XR - 1) / XR;

  const uint32_t raw = tc::smem_u32(smem_raw);
  uint8_t* sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  const uint32_t sm_u = tc::smem_u32(sm);

  ShapeX s;
  s.d0 = G.dims[0]; s.h1 = G.dims[1]; s.h2 = G.dims[2]; s.C = G.dims[3];
  s.N2 = (s.h2 + 15) & ~15; s.MT2 = (s.N2 + 127) >> 7;
  s.ld0 = ru4(s.d0); s.ld1 = ru4(s.h1); s.ld2 = ru4(s.h2);
  const SlabLayout lay = slab_layout(G.dims, 3);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int lane128 = (warp & 3) * 32 + lane;   // this thread's TMEM lane
  const int half = warp >> 2;
  const bool rowt = (warp & 2) == 0;            // owns data row `lane128` (< 64) in the forward / backward epilogues
  const int row = lane128;
  const int step0 = a.step[g];
  float* Pg = a.params + G.slab_off;
  // gradient exchange geometry: CTA o owns float4 indices [o*per, (o+1)*per) of the slab
  const int n4 = lay.end >> 2, per = (n4 + CS - 1) / CS;
  const uint32_t inv_per = 0xFFFFFFFFu / (uint32_t)per + 1u;   // e4 / per == __umulhi(e4, inv_per) for e4 < 2^16
  const int lo = q * per, hi = min(n4, lo + per);

  uint64_t p2p_seq = 0;
  float* p2p_partial = nullptr;
  if (a.tail_mode == 2) {
    p2p_seq = *reinterpret_cast<volatile uint64_t*>(a.p2p->seq);
    p2p_partial = reinterpret_cast<float*>(a.p2p->sym[a.p2p->rank] + a.p2p->sym_partial_off) + (p2p_seq & 1) * a.p2p->global_total;
  }
  float* landing = reinterpret_cast<float*>(sm + SX_STAGE);
  float4* slots = reinterpret_cast<float4*>(sm + SX_SLOTS);
  const uint32_t slots_u = sm_u + SX_SLOTS;
  if (warp == 0) tc::tmem_alloc<512>(&tmem_slot);
  if (tid == 0) {
    tc::mbar_init(&bar, 1);
    tc::mbar_init(&wbar, 1);
    tc::fence_mbar_init();
    bulk_load_begin_x(reinterpret_cast<uint8_t*>(landing), Pg, (uint32_t)lay.end * 4u, &wbar);
  }
  for (int i = tid; i < (SX_B3 + 64 - SX_W1) / 16; i += TC_THREADS) *reinterpret_cast<uint4*>(sm + SX_W1 + i * 16) = make_uint4(0, 0, 0, 0);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  const IssuerX mma{tc::make_desc(sm_u, 0, 1024), tc::make_desc(sm_u, XP, 1024), tmem};
  uint32_t ph = 0, wph = 0;
  const float* W3s = reinterpret_cast<const float*>(sm + SX_W3);
  const float* b3s = reinterpret_cast<const float*>(sm + SX_B3);
  const float inv_n = 1.f / (float)nb;
  const float gscale = inv_n * (1.f / DZS);

  // One updated float4 of the slab (element index e) -> the operand image of every CTA of the cluster: W2 as split fp16
  // into the hi / lo panels, the head as fp32, W1 / b1 / b2 as fp32 into the landing zone (converted locally afterwards).
  auto publish_weights = [&](int e, const float4 w) {
    if (e >= lay.w_off[1] && e < lay.b_off[1]) {
      const int r = e - lay.w_off[1];
      const int j = r / s.ld1, i = r - j * s.ld1;
      const uint32_t off = tc::panel_off(j, i);
      uint32_t h0, l0, h1, l1;
      split2(w.x * WS, w.y * WS, h0, l0);
      split2(w.z * WS, w.w * WS, h1, l1);
#pragma unroll 4
      for (int rk = 0; rk < CS; ++rk) {
        const uint32_t base = mapa(sm_u, rk);
        st_cluster_u2(base + SX_W2 + off, h0, h1);
        st_cluster_u2(base + SX_W2L + off, l0, l1);
      }
      return;
    }
    int off;
    if (e >= lay.w_off[2] && e < lay.b_off[2]) {          // W3 [C][ld2] -> fp32 [C][256]
      const int r = e - lay.w_off[2];
      const int k = r / s.ld2, c = r - k * s.ld2;
      off = SX_W3 + (k * 256 + c) * 4;
    } else if (e >= lay.b_off[2]) {                       // b3
      off = SX_B3 + (e - lay.b_off[2]) * 4;
    } else {                                              // W1, b1, b2
      off = SX_STAGE + e * 4;
    }
#pragma unroll 4
    for (int rk = 0; rk < CS; ++rk) st_cluster_f4(mapa(sm_u, rk) + off, w);
  };
  // one gradient float4 (slab index e4) -> the owner CTA's slot for this CTA
  auto send_grad = [&](int e4, const float4 v) {
    const int o = (int)__umulhi((uint32_t)e4, inv_per);
    st_cluster_f4(mapa(slots_u + (uint32_t)(q * per + (e4 - o * per)) * 16u, o), v);
  };

#define STAMP(i) do { if (a.prof && blockIdx.x == 0 && tid == 0) a.prof[i] = clock64(); } while (0)
  STAMP(0);
  RowRegsX rr;
  HostRegsX hr;
  rr.label = 0;
  rr.q[0] = rr.q[1] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (a.x_host) hr = host_fetch_x(a, G.idx_off + b0 + q * XR, q < ntiles ? min(XR, nb - q * XR) : 0, tid);
  else if (rowt) rr = fetch_row_x(a, G.idx_off + b0 + q * XR + row, q < ntiles && q * XR + row < nb, half);
  tc::mbar_wait(&wbar, wph); wph ^= 1;
  convert_weights_x(sm, landing, lay, s);
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

  // split of the layer-2 columns between the two halves (in 16-column chunks)
  const int nch2 = s.N2 >> 4, ch_lo = half ? (nch2 + 1) >> 1 : 0, ch_hi = half ? nch2 : (nch2 + 1) >> 1;
  uint8_t* const h1_hi = sm + SX_H1;
  uint8_t* const h1_lo = sm + SX_H1 + XP;

  for (int st = 0; st < a.n_steps; ++st) {
    float loss_acc = 0.f, db3_acc[FEDB_MAX_CLASSES];
#pragma unroll
    for (int k = 0; k < FEDB_MAX_CLASSES; ++k) db3_acc[k] = 0.f;
    int my_tiles = 0;

    for (int tile = q; tile < ntiles; tile += CS, ++my_tiles) {
      const bool acc = my_tiles > 0;
      const int r = tile * XR + row;
      const bool valid = rowt && r < nb;
      if (a.x_host) {
        if (my_tiles > 0 || st > 0) hr = host_fetch_x(a, G.idx_off + b0 + tile * XR, min(XR, nb - tile * XR), tid);
        rr = host_deliver_x(sm, a, hr, tid, row, half, rowt);
      } else if (rowt && (my_tiles > 0 || st > 0)) {
        rr = fetch_row_x(a, G.idx_off + b0 + r, valid, half);
      }
      int label = 0;
      if (rowt) label = store_row_x(sm, rr, valid, s, row, half);
      handoff_x();
      STAMP(2);
      // ---- layer 1 -----------------------------------------------------------------------
      FEDB_MMA_PHASE(mma.fwd1());
      STAMP(3);
      uint32_t mask1 = 0;   // ReLU mask of this half's 32 columns of H1
      if (rowt) {
        float v[16];
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c0 = 32 * half + 16 * cc;
          tc::tmem_ld16(tl + TX_D1 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            if (v[i] > 0.f) { mask1 |= 1u << (16 * cc + i); v[i] = fminf(v[i] * INV_WS, HMAX); } else v[i] = 0.f;
          }
          if (c0 == 48) { v[15] = valid ? 1.f : 0.f; mask1 &= ~(1u << 31); }
          st_panel8(h1_hi, h1_lo, row, c0, v);
          st_panel8(h1_hi, h1_lo, row, c0 + 8, v + 8);
        }
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
        for (int ch = ch_lo; ch < ch_hi; ++ch) {
          const int c0 = ch << 4;
          float v[16];
          tc::tmem_ld16(tl + TX_D2 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = fminf(fmaxf(v[i], 0.f) * INV_WS, HMAX);
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
            if (k < s.C) {
#pragma unroll
              for (int i = 0; i < 16; ++i) z[k] = fmaf(v[i], W3s[k * 256 + c0 + i], z[k]);
            }
          uint8_t* pz = sm + SX_Z + (c0 >> 6) * XP;
          st_panel8(pz, pz + 4 * XP, row, c0 & 63, v);
          st_panel8(pz, pz + 4 * XP, row, (c0 & 63) + 8, v + 8);
        }
        // the two halves swap their partial logits through 16 spare TMEM columns of the row's lane
        tc::tmem_st8(tl + TX_ZX + 8 * half, z);
        tc::tmem_st_wait();
      }
      tc::tc_fence_before();
      __syncthreads();
      tc::tc_fence_after();
      if (rowt) {
        {
          float zo[8];
          tc::tmem_ld8(tl + TX_ZX + 8 * (half ^ 1), zo);
          tc::tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = (half ? zo[k] + z[k] : z[k] + zo[k]) + (k < s.C ? b3s[k] : 0.f);
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
        if (half == 0) {   // one thread per row books the loss / db3 and publishes dZ3 as [hi | lo]
          loss_acc += loss;
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k) db3_acc[k] += dz3[k];
          uint4 dh, dl;
          split8(dz3, dh, dl);
          *reinterpret_cast<uint4*>(sm + SX_DZ3 + tc::panel_off(row, 0)) = dh;
          *reinterpret_cast<uint4*>(sm + SX_DZ3 + tc::panel_off(row, 8)) = dl;
        }
      }
      handoff_x();
      STAMP(6);
      // ---- head weight gradient (needs H2), then dZ2 overwrites H2 --------------------------
      FEDB_MMA_PHASE(mma.wgrad3(s, acc));
      STAMP(7);
      if (rowt) {
        for (int ch = ch_lo; ch < ch_hi; ++ch) {
          const int c0 = ch << 4;
          float v[16], d[16];
          tc::tmem_ld16(tl + TX_D2 + c0, v);
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
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c0 = 32 * half + 16 * cc;
          tc::tmem_ld16(tl + TX_D1 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = ((mask1 >> (16 * cc + i)) & 1u) ? v[i] * INV_WS : 0.f;
          st_panel8(h1_hi, h1_lo, row, c0, v);        // dZ1 takes H1's place
          st_panel8(h1_hi, h1_lo, row, c0 + 8, v + 8);
        }
      }
      handoff_x();
      STAMP(10);
      // ---- dW1 += dZ1^T X ------------------------------------------------------------------------
      FEDB_MMA_PHASE(mma.wgrad1(acc));
      STAMP(11);
    }  // tiles

    // ---- send this CTA's partial gradients, slice by slice, straight from TMEM into the owning CTA's shared memory
    // (a reduce-scatter over distributed shared memory).  Every slab position below lay.end is sent (padding as 0).
    {
      const bool have = my_tiles > 0;
      float v[16];
      for (int mt = 0; mt < s.MT2; ++mt) {
        const int j = 128 * mt + lane128;
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c0 = 32 * half + 16 * cc;
          if (have) { tc::tmem_ld16(tl + TX_DW2 + 64 * mt + c0, v); tc::tmem_ld_wait(); }
          else {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = 0.f;
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] *= gscale;
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
          if (c0 == 48) {   // bias gradients sit in column 63: gathered four rows at a time by the lane that owns the aligned group
            const float b0v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 0), b1v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 1);
            const float b2v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 2), b3v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 3);
            if ((lane & 3) == 0 && j < ru4(s.h2))
              send_grad((lay.b_off[1] + j) >> 2, make_float4(j < s.h2 ? b0v : 0.f, j + 1 < s.h2 ? b1v : 0.f, j + 2 < s.h2 ? b2v : 0.f,
                                                             j + 3 < s.h2 ? b3v : 0.f));
          }
        }
        if (half == 0) {   // head weights: dW3[k][j] = DW3[j][k] + DW3[j][8 + k]
          if (have) { tc::tmem_ld16(tl + TX_DW3 + 16 * mt, v); tc::tmem_ld_wait(); }
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
            if (k < s.C) {
              const float x = (v[k] + v[8 + k]) * gscale;
              const float x0 = __shfl_sync(0xffffffffu, x, (lane & ~3) + 0), x1 = __shfl_sync(0xffffffffu, x, (lane & ~3) + 1);
              const float x2 = __shfl_sync(0xffffffffu, x, (lane & ~3) + 2), x3 = __shfl_sync(0xffffffffu, x, (lane & ~3) + 3);
              if ((lane & 3) == 0 && j < s.ld2)
                send_grad((lay.w_off[2] + k * s.ld2 + j) >> 2, make_float4(j < s.h2 ? x0 : 0.f, j + 1 < s.h2 ? x1 : 0.f,
                                                                           j + 2 < s.h2 ? x2 : 0.f, j + 3 < s.h2 ? x3 : 0.f));
            }
        }
      }
      if (half == 0) {     // layer 1: dW1[i][c] = DW1[i][c] + DW1[i][16 + c] (+ db1 in column 15)
        float w[16];
        if (have) {
          tc::tmem_ld16(tl + TX_DW1, v);
          tc::tmem_ld16(tl + TX_DW1 + 16, w);
          tc::tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = w[i] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = (v[i] + w[i]) * gscale;
        if (lane128 < s.h1) {
#pragma unroll
          for (int i4 = 0; i4 < 4; ++i4) {
            const int col = 4 * i4;
            if (col < s.ld0)
              send_grad((lay.w_off[0] + lane128 * s.ld0 + col) >> 2,
                        make_float4(col < s.d0 ? v[col] : 0.f, col + 1 < s.d0 ? v[col + 1] : 0.f, col + 2 < s.d0 ? v[col + 2] : 0.f,
                                    col + 3 < s.d0 ? v[col + 3] : 0.f));
          }
        }
        const float b0v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 0), b1v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 1);
        const float b2v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 2), b3v = __shfl_sync(0xffffffffu, v[15], (lane & ~3) + 3);
        if ((lane & 3) == 0 && lane128 < ru4(s.h1))
          send_grad((lay.b_off[0] + lane128) >> 2, make_float4(lane128 < s.h1 ? b0v : 0.f, lane128 + 1 < s.h1 ? b1v : 0.f,
                                                               lane128 + 2 < s.h1 ? b2v : 0.f, lane128 + 3 < s.h1 ? b3v : 0.f));
      }
    }
    STAMP(12);
    // loss and db3: warp shuffle, then the 2 row-warps of half 0 in a fixed order; sent like any other gradient
    {
      if (rowt && half == 0) {
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
      __syncthreads();
      if (tid == 0) {
        float t[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) t[i] = red_f[0][i] + red_f[1][i];
        send_grad(lay.b_off[2] >> 2, make_float4(0 < s.C ? t[1] : 0.f, 1 < s.C ? t[2] : 0.f, 2 < s.C ? t[3] : 0.f, 3 < s.C ? t[4] : 0.f));
        if (s.C > 4) send_grad((lay.b_off[2] >> 2) + 1, make_float4(t[5], 5 < s.C ? t[6] : 0.f, 6 < s.C ? t[7] : 0.f, 7 < s.C ? t[8] : 0.f));
        // loss partial: float q of the extra row behind rank 0's slots
        st_cluster_f1(mapa(slots_u + (uint32_t)(CS * per) * 16u + 4u * (uint32_t)q, 0), t[0]);
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

    // ---- owner: fixed-order sum of the CS partials (local shared memory), optimizer, then all-gather + conversion of the
    // updated slice into every CTA's operand image (distributed shared memory) ------
    {
      const StepScalars sc = sc_s;
      int it = 0;
      for (int e4 = lo + tid; e4 < hi; e4 += TC_THREADS, ++it) {
        const int e = e4 * 4, i = e4 - lo;
        float4 gsum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
        for (int rk = 0; rk < CS; ++rk) { const float4 p = slots[rk * per + i]; gsum.x += p.x; gsum.y += p.y; gsum.z += p.z; gsum.w += p.w; }
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
    cluster.sync();   // every CTA holds the complete updated operand image
    if (p2p_partial && st == a.n_steps - 1 && q == 0 && tid < a.p2p->world) {
      // arrival flag A ("my partial is complete") into every peer's buffer; the local evaluation below overlaps its flight.
      // The cluster barrier ordered every CTA's partial stores before this thread; fence + release make them visible to
      // whoever acquires the flag.
      __threadfence_system();
      uint64_t* pf = reinterpret_cast<uint64_t*>(a.p2p->sym[tid] + a.p2p->sym_flag_off);
      st_release_sys_u64(pf + a.p2p->rank, p2p_seq + 1);
    }
    STAMP(15);
    convert_w1_b2(sm, landing, lay, s, tid);
    tc::fence_async_smem();
    __syncthreads();
    STAMP(16);
  }  // steps

  // ---- local evaluation with the updated weights (C:75-83) ----------------------------------
  if (a.do_eval) {
    if (tid < FEDB_MAX_CLASSES * FEDB_MAX_CLASSES) cnt[tid] = 0;
    __syncthreads();
    for (int tile = q; tile < ntiles; tile += CS) {
      const int r = tile * XR + row;
      const bool valid = rowt && r < nb;
      int label = 0;
      if (ntiles <= CS && a.n_steps > 0) {
        label = rr.label;   // single tile per CTA: the X panel from the training step is still in place
      } else {
        if (a.x_host) {
          hr = host_fetch_x(a, G.idx_off + b0 + tile * XR, min(XR, nb - tile * XR), tid);
          rr = host_deliver_x(sm, a, hr, tid, row, half, rowt);
        } else if (rowt) {
          rr = fetch_row_x(a, G.idx_off + b0 + r, valid, half);
        }
        if (rowt) label = store_row_x(sm, rr, valid, s, row, half);
      }
      handoff_x();
      FEDB_MMA_PHASE(mma.fwd1());
      if (rowt) {
        float v[16];
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c0 = 32 * half + 16 * cc;
          tc::tmem_ld16(tl + TX_D1 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = fminf(fmaxf(v[i], 0.f) * INV_WS, HMAX);
          if (c0 == 48) v[15] = valid ? 1.f : 0.f;
          st_panel8(h1_hi, h1_lo, row, c0, v);
          st_panel8(h1_hi, h1_lo, row, c0 + 8, v + 8);
        }
      }
      handoff_x();
      FEDB_MMA_PHASE(mma.fwd2(s));
      float z[FEDB_MAX_CLASSES];
#pragma unroll
      for (int k = 0; k < FEDB_MAX_CLASSES; ++k) z[k] = 0.f;
      if (rowt) {
        for (int ch = ch_lo; ch < ch_hi; ++ch) {
          const int c0 = ch << 4;
          float v[16];
          tc::tmem_ld16(tl + TX_D2 + c0, v);
          tc::tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < FEDB_MAX_CLASSES; ++k)
            if (k < s.C) {
#pragma unroll
              for (int i = 0; i < 16; ++i) z[k] = fmaf(fminf(fmaxf(v[i], 0.f) * INV_WS, HMAX), W3s[k * 256 + c0 + i], z[k]);
            }
        }
        tc::tmem_st8(tl + TX_ZX + 8 * half, z);
        tc::tmem_st_wait();
      }
      tc::tc_fence_before();
      __syncthreads();
      tc::tc_fence_after();
      if (rowt && half == 0) {
        float zo[8];
        tc::tmem_ld8(tl + TX_ZX + 8, zo);
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

__global__ void __launch_bounds__(TC_THREADS, 1) k_round_tcx(TcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  round_body_x(a, smem_raw);
  if (a.tail_mode == 0) return;
  if (a.tail_mode == 2) { p2p_tail(a); return; }
  // the cluster that observes the last check-in runs the single-GPU aggregation tail (see k_local_round_tc)
  __shared__ unsigned int s_last;
  __shared__ unsigned char s_stop[TAIL_MAX_FEDS];
  cg::cluster_group cluster = cg::this_cluster();
  const int CS = (int)cluster.num_blocks(), crank = (int)cluster.block_rank();
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

extern "C" int64_t fedb_tcx_smem_bytes(void) { return SX_SLOTS + SX_SLOTS_MAX + 1024; }
// largest slab (floats) whose gradient slices (rounded up to a multiple of 16 CTAs) fit the per-CTA exchange area
extern "C" int32_t fedb_tcx_max_slab(void) { return (SX_SLOTS_MAX - 64 - 16 * 16) / 4; }
extern "C" int32_t fedb_tcx_rows(void) { return XR; }

// Largest cluster size (<= 16) the device can co-schedule for this kernel; 16 needs the non-portable opt-in and a GPC
// with 16 free SMs.  Cached per device.
extern "C" int32_t fedb_tcx_max_cluster(void) {
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 8;
  if (cached[dev]) return cached[dev];
  const int smem = SX_SLOTS + SX_SLOTS_MAX + 1024;
  if (cudaFuncSetAttribute((const void*)k_round_tcx, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess ||
      cudaFuncSetAttribute((const void*)k_round_tcx, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
    cudaGetLastError();
    return cached[dev] = 8;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(16);
  cfg.blockDim = dim3(TC_THREADS);
  cfg.dynamicSmemBytes = smem;
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
  const int smem = SX_SLOTS + SX_SLOTS_MAX + 1024;
  // per-device attributes: cheap, legal during stream capture, and correct for a process that drives several GPUs
  FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_round_tcx, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  if (cluster_size > 8) FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_round_tcx, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
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
  FEDB_CUDA_OK(cudaLaunchKernelEx(&cfg, k_round_tcx, a));
  return FEDB_OK;
}

// The three general entry points of the split-fp16 round kernel; arguments exactly as for the bf16 kernel's
// fedb_local_round_tc_prof / fedb_round_tc_hostin / fedb_round_tc_p2p_hostin (include/fedb200.h), cluster_size 1..16.
extern "C" int fedb_local_round_tcx(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X, int32_t ldx,
                                    const int32_t* y, const int32_t* idx, float* params, float* m, float* v, int32_t* step,
                                    const int32_t* active, const FedbFedState* fed_state, float* loss_out, int32_t* conf,
                                    float* scratch, int32_t batch, int32_t n_steps, int32_t do_eval, long long* prof, void* stream) {
  TcArgs a{};
  a.groups = groups; a.X = X; a.y = y; a.idx = idx; a.params = params; a.m = m; a.v = v; a.step = step; a.active = active;
  a.fed_state = fed_state; a.loss_out = loss_out; a.conf = conf; a.scratch = scratch; a.ldx = ldx; a.batch = batch;
  a.n_steps = n_steps; a.do_eval = do_eval; a.prof = prof; a.tail_mode = 0;
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
  a.n_steps = n_steps; a.do_eval = 1; a.prof = prof;
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
