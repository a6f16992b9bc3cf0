// K2d, 2-CTA variant: the grouped GEMM of fedb_gemm_tma.cu on CTA PAIRS (thread-block cluster of 2,
// tcgen05 cta_group::2).  One pair computes a 256 x 256 output tile: each CTA owns 128 rows of A and
// HALF of the B tile, a single tcgen05.mma.cta_group::2 issued by the leader CTA reads both halves and
// writes 128 x 256 fp32 accumulators into EACH CTA's tensor memory.  Per SM and k-step that is 32 KB of
// operands for 128 x 256 x 64 MACs — twice the arithmetic intensity of the single-CTA 128 x 128 tile,
// which is what the L2 -> SM operand stream limits.  Both CTAs' TMA loads complete on the LEADER's
// "full" barrier (cp.async.bulk.tensor ... cta_group::2); tcgen05.commit ... multicast::cluster releases
// the stage in both CTAs.  Epilogues and arguments are those of fedb_gemm_tma (FedbGemm).
// Requires M % 256 == 0 and N % 256 == 0; other shapes use the single-CTA kernel.
#include <cooperative_groups.h>
#include <cuda.h>

#include "common.cuh"
#include "pair_common.cuh"
#include "tma_map.cuh"
#include "tc_common.cuh"

namespace cg = cooperative_groups;

namespace fedb {
namespace {

constexpr int T_THREADS = 192;
constexpr int BM = 128, BN = 256, BNH = 128, BK = 64;      // per CTA: 128 rows x 256 cols; it loads 128 of the 256 B columns
constexpr int STAGES_MAX = 3;                             // fwd / dgrad / plain store: 3 stages; optimizer epilogue: 2 stages + a deep ring
constexpr int A_BYTES = BM * BK * 2, B_BYTES = BNH * BK * 2;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;             // 32 KB per CTA
constexpr int TILE_LD = 128 + 4;                           // the epilogue stages 128 columns at a time
// The optimizer epilogue streams W / momentum / second moment through per-warp rings in shared memory filled by
// cp.async.bulk: with 12 warps per SM, register loads keep too few bytes in flight (and every outstanding load
// pins an L1 line, of which little is left beside 2 x 97 KB of shared memory): 5.1 TB/s against 6.2 TB/s for the
// ring (tools/adam_pattern_bench.cu).  The rings live behind the fp32 staging tile, inside the idle operand stages.
// One CTA alone must keep its SM's share of HBM busy while its neighbour runs a main loop (~45 GB/s x ~1 us), so the
// optimizer variant runs 2 operand stages (64 KB, reused by a 64 KB XOR-swizzled staging tile) and spends the rest
// of its shared memory on rings that are already filling during the main loop.
constexpr int RING_D = 5, RING_SLOT = 3 * 512;             // rows in flight per warp; one row = 512 B of each array
constexpr int RING_BYTES = (T_THREADS / 32) * RING_D * RING_SLOT;
template <bool OPT> struct Cfg {
  static constexpr int STAGES = OPT ? 2 : 3;
  static constexpr int RING_OFF = STAGES * STAGE_BYTES;
  static constexpr int SMEM = (OPT ? RING_OFF + RING_BYTES : (STAGES * STAGE_BYTES > BM * TILE_LD * 4 ? STAGES * STAGE_BYTES : BM * TILE_LD * 4)) + 1024;
};
static_assert(Cfg<true>::RING_OFF >= BM * 128 * 4, "the swizzled staging tile must fit in the stages");
// fp32 staging tile: padded rows for the plain epilogues, 16-byte XOR swizzle (exactly 64 KB) for the optimizer one
template <bool OPT> __device__ __forceinline__ int tile_at(int r, int c4) { return OPT ? r * 128 + ((c4 ^ (r & 7)) << 2) : r * TILE_LD + (c4 << 2); }
using namespace pair;

}  // namespace

template <bool OPT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(T_THREADS, 2)
k_gemm_tma2(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, FedbGemm p, int* __restrict__ err,
            unsigned long long* __restrict__ prof) {
  extern __shared__ uint8_t smem_raw[];
  constexpr int STAGES = Cfg<OPT>::STAGES, RING_OFF = Cfg<OPT>::RING_OFF;
  __shared__ uint64_t full_bar[STAGES_MAX], empty_bar[STAGES_MAX], done_bar, ring_bar[T_THREADS / 32][RING_D];
  __shared__ uint32_t tmem_slot;
  __shared__ StepScalars sc_s;
  __shared__ FedbGroup Gs;
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();          // 0 = leader (issues the MMAs)
  const uint32_t raw = tc::smem_u32(smem_raw);
  const uint32_t sm = (raw + 1023u) & ~1023u;
  uint8_t* smp = smem_raw + (sm - raw);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform: ring addresses stay in uniform registers
  const int g = blockIdx.z, m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int nh0 = n0 + rank * BNH;                     // the half of the B tile this CTA loads
  const int KT = (p.K + BK - 1) / BK;

  if (OPT && tid < (int)(sizeof(FedbGroup) / 4))
    reinterpret_cast<int32_t*>(&Gs)[tid] = reinterpret_cast<const int32_t*>(p.groups + g)[tid];
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc::smem_u32(&tmem_slot)), "n"(256) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
    tc::mbar_init(&done_bar, 1);
    for (int i = 0; i < (T_THREADS / 32) * RING_D; ++i) tc::mbar_init(&ring_bar[0][0] + i, 1);
    tc::fence_mbar_init();
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  tc::tc_fence_before();
  cluster.sync();                                       // the peer's barriers exist before anything signals them
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  // optional phase stamps (ns, globaltimer) of warp 2: [start, main loop done, pass 0 staged, pass 0 rows done, pass 1 staged, end, smid]
  unsigned long long* stamp = nullptr;
  if (prof && tid == 64) {
    stamp = prof + 8 * ((size_t)blockIdx.x + (size_t)gridDim.x * (blockIdx.y + (size_t)gridDim.y * blockIdx.z));
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    stamp[6] = smid;
  }
  auto mark = [&](int k) {
    if (stamp) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); stamp[k] = t; }
  };
  mark(0);
  if (OPT && tid == 64) sc_s = step_scalars_fast(Gs, p.step[g]);
  bool ok = true;
  // optimizer-state ring of this warp: row i of its sequence is (pass i / rpw, tile row warp + (i % rpw) * NW)
  constexpr int NW = T_THREADS / 32;
  const int rpw = (BM - warp + NW - 1) / NW, ring_total = OPT ? 2 * rpw : 0;
  const bool use_v = OPT && p.Vel && Gs.opt != FEDB_OPT_SGD;
  const bool use_m = OPT && p.Mom && !(Gs.opt == FEDB_OPT_SGD && Gs.momentum == 0.f);   // plain SGD keeps no state: skip its traffic
  const uint32_t ring_sm = sm + RING_OFF + warp * RING_D * RING_SLOT;
  const float* ring = reinterpret_cast<const float*>(smp + RING_OFF + warp * RING_D * RING_SLOT);
  const size_t opt_base = OPT ? (size_t)g * p.w_group_stride + (size_t)m0 * p.ldw + n0 : 0;
  int is_j = 0, is_h = 0, is_d = 0;                     // next row to request: ordinal in the pass, pass, ring slot
  auto ring_issue = [&]() {                              // lane 0 only; no-op once both passes are requested
    if (is_h < 2) {
      const size_t off = opt_base + (size_t)(warp + is_j * NW) * p.ldw + is_h * 128;
      const uint32_t bar = tc::smem_u32(&ring_bar[warp][is_d]), dst = ring_sm + is_d * RING_SLOT;
      mbar_expect_tx(&ring_bar[warp][is_d], 512u * (1 + (int)use_m + (int)use_v));
      bulk_load_512(dst, p.W + off, bar);
      if (use_m) bulk_load_512(dst + 512, p.Mom + off, bar);
      if (use_v) bulk_load_512(dst + 1024, p.Vel + off, bar);
      if (++is_j == rpw) { is_j = 0; ++is_h; }
      if (++is_d == RING_D) is_d = 0;
    }
  };
  if (OPT && lane == 0)                                  // the rings fill while the main loop runs
    for (int i = 0; i < RING_D && i < ring_total; ++i) ring_issue();
  int cs_d = 0;                                          // consumer side: slot and phase parity of the next row
  uint32_t cs_par = 0;

  if (warp == 0) {
    // ===== TMA producer (one lane per CTA): own A rows + own half of B; completion on the LEADER's full barrier =====
    // (elected with elect.sync, not `lane == 0`: under a plain divergent branch the compiler wraps every UTMALDG / UTCHMMA in
    //  a loop over the active lanes)
    __syncwarp();
    if (tc::elect_one()) {
      for (int kt = 0; kt < KT && ok; ++kt) {
        const int s = kt % STAGES, k0 = kt * BK;
        if (kt >= STAGES) ok = mbar_wait_to(&empty_bar[s], ((kt / STAGES) - 1) & 1);
        if (!ok) break;
        if (rank == 0) mbar_expect_tx(&full_bar[s], 2 * STAGE_BYTES);
        const uint32_t lbar = tc::smem_u32(&full_bar[s]) & PEER_MASK;
        const uint32_t a_dst = sm + s * STAGE_BYTES, b_dst = a_dst + A_BYTES;
        if (!p.a_mn) tma2_load_3d(a_dst, &tmA, k0, m0, g, lbar);
        else { tma2_load_3d(a_dst, &tmA, m0, k0, g, lbar); tma2_load_3d(a_dst + BK * 128, &tmA, m0 + 64, k0, g, lbar); }
        if (!p.b_mn) tma2_load_3d(b_dst, &tmB, k0, nh0, g, lbar);
        else { tma2_load_3d(b_dst, &tmB, nh0, k0, g, lbar); tma2_load_3d(b_dst + BK * 128, &tmB, nh0 + 64, k0, g, lbar); }
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ===== MMA issuer: one lane of the leader CTA drives both SMs' tensor cores =====
    __syncwarp();
    if (tc::elect_one()) {
      const uint32_t idesc = tc::make_idesc_bf16(2 * BM, BN, p.a_mn != 0, p.b_mn != 0);
      for (int kt = 0; kt < KT && ok; ++kt) {
        const int s = kt % STAGES;
        ok = mbar_wait_to(&full_bar[s], (kt / STAGES) & 1);
        if (!ok) break;
        tc::tc_fence_after();
        const int k_cnt = min(BK, p.K - kt * BK);
        const uint32_t a_base = sm + s * STAGE_BYTES, b_base = a_base + A_BYTES;
        for (int k = 0; k < (k_cnt >> 4); ++k) {
          const uint64_t ad = p.a_mn ? tc::make_desc(a_base + 2048 * k, BK * 128, 1024) : tc::make_desc(a_base + 32 * k, 0, 1024);
          const uint64_t bd = p.b_mn ? tc::make_desc(b_base + 2048 * k, BK * 128, 1024) : tc::make_desc(b_base + 32 * k, 0, 1024);
          mma2_bf16(tmem, ad, bd, idesc, (kt | k) != 0);
        }
        commit2(&empty_bar[s]);
        if (kt == KT - 1) commit2(&done_bar);
      }
    }
  }
  // ===== epilogue: each CTA drains its own 128 x 256 accumulator, 128 columns at a time =====
  float* tile = reinterpret_cast<float*>(smp);
  if (warp >= 2) ok = mbar_wait_to(&done_bar, 0);
  if (!ok && err) *err = FEDB_E_TIMEOUT;
  tc::tc_fence_after();
  mark(1);
  for (int hc = 0; hc < 2; ++hc) {
    const int nc0 = n0 + hc * 128;                     // global column base of this pass
    if (warp >= 2) {
      const int quad = warp & 3;
      const int r = quad * 32 + lane;
      const uint32_t tl = tmem + ((uint32_t)(quad * 32) << 16) + hc * 128;
      for (int c0 = 0; c0 < 128; c0 += 16) {
        float v[16];
        tc::tmem_ld16(tl + c0, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4)
          *reinterpret_cast<float4*>(tile + tile_at<OPT>(r, (c0 >> 2) + q4)) = make_float4(v[4 * q4], v[4 * q4 + 1], v[4 * q4 + 2], v[4 * q4 + 3]);
      }
    }
    tc::tc_fence_before();
    __syncthreads();
    mark(2 + 2 * hc);
    if constexpr (OPT) {
      const OptRegs o = opt_regs(Gs, sc_s, Gs.alpha * p.l2_scale);
      for (int j = 0; j < rpw; j += 2) {               // two rows per trip: independent chains hide SFU and shared-memory latency
        const int cnt = min(2, rpw - j);
        float4 x[2], mm[2], vv[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (u < cnt) {
            tc::mbar_wait(&ring_bar[warp][cs_d], cs_par);
            const float* slot = ring + cs_d * (RING_SLOT / 4);
            x[u] = *reinterpret_cast<const float4*>(slot + 4 * lane);
            mm[u] = vv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (use_m) mm[u] = *reinterpret_cast<const float4*>(slot + 128 + 4 * lane);
            if (use_v) vv[u] = *reinterpret_cast<const float4*>(slot + 256 + 4 * lane);
            if (++cs_d == RING_D) { cs_d = 0; cs_par ^= 1u; }
          }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (u < cnt) {
            const int r = warp + (j + u) * NW;
            const float4 gr = *reinterpret_cast<const float4*>(tile + tile_at<true>(r, lane));
            x[u].x = opt_update_fast(o, x[u].x, gr.x, mm[u].x, vv[u].x);
            x[u].y = opt_update_fast(o, x[u].y, gr.y, mm[u].y, vv[u].y);
            x[u].z = opt_update_fast(o, x[u].z, gr.z, mm[u].z, vv[u].z);
            x[u].w = opt_update_fast(o, x[u].w, gr.w, mm[u].w, vv[u].w);
            const size_t off = opt_base + (size_t)r * p.ldw + hc * 128 + 4 * lane;
            *reinterpret_cast<float4*>(p.W + off) = x[u];
            if (use_m) *reinterpret_cast<float4*>(p.Mom + off) = mm[u];
            if (use_v) *reinterpret_cast<float4*>(p.Vel + off) = vv[u];
            if (p.Wsh) *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.Wsh) + off) = make_uint2(tc::pack_bf16(x[u].x, x[u].y), tc::pack_bf16(x[u].z, x[u].w));
          }
        }
        __syncwarp();                                  // every lane has consumed its slot values: the slots may be refilled
        if (lane == 0)
          for (int u = 0; u < cnt; ++u) ring_issue();
      }
      if (!ok && err) *err = FEDB_E_TIMEOUT;
    } else if (p.epilogue == FEDB_GEMM_STORE_F32) {
      for (int r = warp; r < BM; r += T_THREADS / 32)
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.C) + (size_t)g * p.c_group_stride + (size_t)(m0 + r) * p.ldc + nc0 + 4 * lane) =
            *reinterpret_cast<const float4*>(tile + r * TILE_LD + 4 * lane);
    } else {
      const float4 bias4 = (p.epilogue == FEDB_GEMM_FWD_BIAS_RELU && p.bias)
                               ? *reinterpret_cast<const float4*>(p.bias + (size_t)g * p.bias_group_stride + nc0 + 4 * lane)
                               : make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = warp; r < BM; r += T_THREADS / 32) {
        float4 v = *reinterpret_cast<const float4*>(tile + r * TILE_LD + 4 * lane);
        if (p.epilogue == FEDB_GEMM_FWD_BIAS_RELU) {
          v.x = fmaxf(v.x + bias4.x, 0.f); v.y = fmaxf(v.y + bias4.y, 0.f); v.z = fmaxf(v.z + bias4.z, 0.f); v.w = fmaxf(v.w + bias4.w, 0.f);
        } else {
          const uint2 a = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(p.act) + (size_t)g * p.act_group_stride +
                                                          (size_t)(m0 + r) * p.ldact + nc0 + 4 * lane);
          const __nv_bfloat16* ab = reinterpret_cast<const __nv_bfloat16*>(&a);
          v.x = __bfloat162float(ab[0]) > 0.f ? v.x : 0.f; v.y = __bfloat162float(ab[1]) > 0.f ? v.y : 0.f;
          v.z = __bfloat162float(ab[2]) > 0.f ? v.z : 0.f; v.w = __bfloat162float(ab[3]) > 0.f ? v.w : 0.f;
        }
        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)g * p.c_group_stride + (size_t)(m0 + r) * p.ldc + nc0 + 4 * lane) =
            make_uint2(tc::pack_bf16(v.x, v.y), tc::pack_bf16(v.z, v.w));
      }
    }
    mark(3 + 2 * hc);
    __syncthreads();                                   // the staging tile is rewritten by the next pass
  }
  tc::tc_fence_before();
  cluster.sync();                                      // both CTAs are done with tensor memory
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(256) : "memory");
}

}  // namespace fedb

using namespace fedb;

// err_dev: optional device int that receives FEDB_E_TIMEOUT if the pair's barrier protocol stalls.
static int gemm_tma2_launch(const FedbGemm* args_host, int32_t n_groups, int32_t* err_dev, unsigned long long* prof_dev, void* stream) {
  const FedbGemm& p = *args_host;
  FEDB_REQUIRE(n_groups > 0 && p.M % 256 == 0 && p.N % 256 == 0 && p.K % 16 == 0 && p.K > 0, "2-CTA GEMM needs M % 256 == 0, N % 256 == 0, K % 16 == 0");
  FEDB_REQUIRE(p.lda % 8 == 0 && p.ldb % 8 == 0 && p.a_group_stride % 8 == 0 && p.b_group_stride % 8 == 0, "strides must be multiples of 8 elements");
  alignas(64) CUtensorMap tmA, tmB;
  if (int rc = make_gemm_maps(p, n_groups, &tmA, &tmB)) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_gemm_tma2<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<false>::SMEM));
    FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_gemm_tma2<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<true>::SMEM));
    attr_set = true;
  }
  dim3 grid(p.M / BM, p.N / BN, n_groups);            // grid.x is even: consecutive x form the CTA pair
  if (p.epilogue == FEDB_GEMM_WGRAD_OPT) k_gemm_tma2<true><<<grid, T_THREADS, Cfg<true>::SMEM, (cudaStream_t)stream>>>(tmA, tmB, p, err_dev, prof_dev);
  else k_gemm_tma2<false><<<grid, T_THREADS, Cfg<false>::SMEM, (cudaStream_t)stream>>>(tmA, tmB, p, err_dev, prof_dev);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}

extern "C" int fedb_gemm_tma2(const FedbGemm* args_host, int32_t n_groups, int32_t* err_dev, void* stream) {
  return gemm_tma2_launch(args_host, n_groups, err_dev, nullptr, stream);
}
// Same launch with phase stamps: prof_dev holds 8 uint64 per CTA (grid order x, y, group): globaltimer ns at start, main loop
// done, pass 0 staged, pass 0 done, pass 1 staged, end; then the SM id.  For tools/gemm_timeline.py.
extern "C" int fedb_gemm_tma2_prof(const FedbGemm* args_host, int32_t n_groups, int32_t* err_dev, uint64_t* prof_dev, void* stream) {
  return gemm_tma2_launch(args_host, n_groups, err_dev, reinterpret_cast<unsigned long long*>(prof_dev), stream);
}
