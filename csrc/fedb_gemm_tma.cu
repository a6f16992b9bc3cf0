// K2d, main variant: the grouped GEMM of fedb_gemm.cu with the canonical Blackwell structure —
// operands fetched by TMA (cp.async.bulk.tensor, 3-D tensor maps: K x rows x client, hardware 128B
// swizzle straight into the UMMA panel layout), a warp-specialised CTA (warp 0 = TMA producer,
// warp 1 = tcgen05.mma issuer + TMEM owner, warps 2-5 = epilogue), a full/empty mbarrier ring, and an
// epilogue that stages the fp32 accumulator tile through shared memory so that every global access of
// the fused bias/ReLU, ReLU-mask and optimizer epilogues is a fully coalesced 512-byte row segment.
// Same arguments, shapes and epilogues as fedb_gemm_tc (include/fedb200.h: FedbGemm).
#include <cuda.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "tma_map.cuh"

namespace fedb {
namespace {

constexpr int T_THREADS = 192;
constexpr int BM = 128, BN = 128, BK = 64;
constexpr int STAGES = 3;                                  // 96 KB: two CTAs per SM overlap epilogue and main loop
constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int TILE_LD = BN + 4;                            // fp32 staging row stride (floats): 16-byte aligned, bank-shifted

// Optimizer epilogue: W / momentum / second moment reach the SM through per-warp rings in shared memory filled
// by cp.async.bulk (see fedb_gemm_tma2.cu and tools/adam_pattern_bench.cu: 6.2 TB/s against 5.1 TB/s for
// register loads at this occupancy).  The rings sit behind the fp32 staging tile, inside the idle operand stages.
constexpr int RING_D = 3, RING_SLOT = 3 * 512;
constexpr int RING_OFF = BM * TILE_LD * 4;
static_assert(RING_OFF % 128 == 0 && RING_OFF + (T_THREADS / 32) * RING_D * RING_SLOT <= STAGES * STAGE_BYTES, "rings must fit in the stages");

__device__ __forceinline__ void bulk_load_512(uint32_t dst, const float* src, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 512, [%2];" ::"r"(dst), "l"(src), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
               "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(tc::smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc::smem_u32(bar)), "r"(bytes) : "memory");
}

}  // namespace

__global__ void __launch_bounds__(T_THREADS) k_gemm_tma(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                                        FedbGemm p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], done_bar, ring_bar[T_THREADS / 32][RING_D];
  __shared__ uint32_t tmem_slot;
  __shared__ StepScalars sc_s;
  __shared__ FedbGroup Gs;
  const uint32_t raw = tc::smem_u32(smem_raw);
  const uint32_t sm = (raw + 1023u) & ~1023u;
  uint8_t* smp = smem_raw + (sm - raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = blockIdx.z, m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int n_cnt = min(BN, p.N - n0);
  const int KT = (p.K + BK - 1) / BK;

  if (p.epilogue == FEDB_GEMM_WGRAD_OPT && tid < (int)(sizeof(FedbGroup) / 4))
    reinterpret_cast<int32_t*>(&Gs)[tid] = reinterpret_cast<const int32_t*>(p.groups + g)[tid];
  if (warp == 1) tc::tmem_alloc<128>(&tmem_slot);
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
    tc::mbar_init(&done_bar, 1);
    for (int i = 0; i < (T_THREADS / 32) * RING_D; ++i) tc::mbar_init(&ring_bar[0][0] + i, 1);
    tc::fence_mbar_init();
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (p.epilogue == FEDB_GEMM_WGRAD_OPT && tid == 64) sc_s = step_scalars_fast(Gs, p.step[g]);

  if (warp == 0) {
    // ===== TMA producer (one lane, elected: see fedb_gemm_tma2.cu) =====
    __syncwarp();
    if (tc::elect_one()) {
      for (int kt = 0; kt < KT; ++kt) {
        const int s = kt % STAGES, k0 = kt * BK;
        if (kt >= STAGES) tc::mbar_wait(&empty_bar[s], ((kt / STAGES) - 1) & 1);
        mbar_expect_tx(&full_bar[s], STAGE_BYTES);
        const uint32_t a_dst = sm + s * STAGE_BYTES, b_dst = a_dst + A_BYTES;
        if (!p.a_mn) tma_load_3d(a_dst, &tmA, k0, m0, g, &full_bar[s]);                       // box {64 k, 128 rows}
        else { tma_load_3d(a_dst, &tmA, m0, k0, g, &full_bar[s]); tma_load_3d(a_dst + BK * 128, &tmA, m0 + 64, k0, g, &full_bar[s]); }
        if (!p.b_mn) tma_load_3d(b_dst, &tmB, k0, n0, g, &full_bar[s]);
        else { tma_load_3d(b_dst, &tmB, n0, k0, g, &full_bar[s]); tma_load_3d(b_dst + BK * 128, &tmB, n0 + 64, k0, g, &full_bar[s]); }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one lane, elected) =====
    __syncwarp();
    if (tc::elect_one()) {
      const uint32_t idesc = tc::make_idesc_bf16(BM, n_cnt, p.a_mn != 0, p.b_mn != 0);
      for (int kt = 0; kt < KT; ++kt) {
        const int s = kt % STAGES;
        tc::mbar_wait(&full_bar[s], (kt / STAGES) & 1);
        tc::tc_fence_after();
        const int k_cnt = min(BK, p.K - kt * BK);
        const uint32_t a_base = sm + s * STAGE_BYTES, b_base = a_base + A_BYTES;
        for (int k = 0; k < (k_cnt >> 4); ++k) {
          const uint64_t ad = p.a_mn ? tc::make_desc(a_base + 2048 * k, BK * 128, 1024) : tc::make_desc(a_base + 32 * k, 0, 1024);
          const uint64_t bd = p.b_mn ? tc::make_desc(b_base + 2048 * k, BK * 128, 1024) : tc::make_desc(b_base + 32 * k, 0, 1024);
          tc::mma_bf16(tmem, ad, bd, idesc, (kt | k) != 0);
        }
        tc::mma_commit(&empty_bar[s]);
        if (kt == KT - 1) tc::mma_commit(&done_bar);
      }
    }
  }
  // ===== epilogue =====
  // (1) warps 2..5: accumulator rows -> fp32 staging tile in shared memory (the operand ring is idle by then)
  float* tile = reinterpret_cast<float*>(smp);
  constexpr int NW = T_THREADS / 32;
  const bool ring_on = p.epilogue == FEDB_GEMM_WGRAD_OPT && n_cnt == BN;
  const bool use_v = p.epilogue == FEDB_GEMM_WGRAD_OPT && p.Vel && Gs.opt != FEDB_OPT_SGD;
  const bool use_m = p.epilogue == FEDB_GEMM_WGRAD_OPT && p.Mom && !(Gs.opt == FEDB_OPT_SGD && Gs.momentum == 0.f);   // plain SGD keeps no state
  const int rpw = (BM - warp + NW - 1) / NW;               // rows of the tile this warp updates: warp, warp + NW, ...
  const uint32_t ring_sm = sm + RING_OFF + warp * RING_D * RING_SLOT;
  const size_t opt_base = (size_t)g * p.w_group_stride + (size_t)m0 * p.ldw + n0;
  auto ring_issue = [&](int i) {
    const size_t off = opt_base + (size_t)(warp + i * NW) * p.ldw;
    const int d = i % RING_D;
    const uint32_t bar = tc::smem_u32(&ring_bar[warp][d]), dst = ring_sm + d * RING_SLOT;
    mbar_expect_tx(&ring_bar[warp][d], 512u * (1 + (int)use_m + (int)use_v));
    bulk_load_512(dst, p.W + off, bar);
    if (use_m) bulk_load_512(dst + 512, p.Mom + off, bar);
    if (use_v) bulk_load_512(dst + 1024, p.Vel + off, bar);
  };
  if (ring_on) {
    tc::mbar_wait(&done_bar, 0);                           // the stages are idle from here on
    if (lane == 0)
      for (int i = 0; i < RING_D && i < rpw; ++i) ring_issue(i);
  }
  if (warp >= 2) {
    tc::mbar_wait(&done_bar, 0);
    tc::tc_fence_after();
    const int quad = warp & 3;                       // TMEM lane quadrant this warp may read
    const int r = quad * 32 + lane;
    const uint32_t tl = tmem + ((uint32_t)(quad * 32) << 16);
    for (int c0 = 0; c0 < n_cnt; c0 += 16) {
      float v[16];
      tc::tmem_ld16(tl + c0, v);
      tc::tmem_ld_wait();
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4)
        *reinterpret_cast<float4*>(tile + r * TILE_LD + c0 + 4 * q4) = make_float4(v[4 * q4], v[4 * q4 + 1], v[4 * q4 + 2], v[4 * q4 + 3]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  // (2) all 6 warps: one warp per row at a time, lane = 4 consecutive columns -> coalesced global traffic
  const int nq = n_cnt >> 2;                         // float4 groups per row (<= 32)
  if (ring_on) {
    const StepScalars sc = sc_s;
    const float l2 = Gs.alpha * p.l2_scale;
    const float* ring = reinterpret_cast<const float*>(smp + RING_OFF + warp * RING_D * RING_SLOT);
    for (int i = 0; i < rpw; ++i) {
      const int d = i % RING_D, r = warp + i * NW;
      tc::mbar_wait(&ring_bar[warp][d], (i / RING_D) & 1);
      const float* slot = ring + d * (RING_SLOT / 4);
      float4 x = *reinterpret_cast<const float4*>(slot + 4 * lane);
      float4 mm = use_m ? *reinterpret_cast<const float4*>(slot + 128 + 4 * lane) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 vv = use_v ? *reinterpret_cast<const float4*>(slot + 256 + 4 * lane) : make_float4(0.f, 0.f, 0.f, 0.f);
      __syncwarp();                                        // every lane has read the slot: it may be refilled
      if (lane == 0 && i + RING_D < rpw) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        ring_issue(i + RING_D);
      }
      const float4 gr = *reinterpret_cast<const float4*>(tile + r * TILE_LD + 4 * lane);
      x.x = opt_update(Gs, sc, x.x, fmaf(l2, x.x, gr.x), mm.x, vv.x);
      x.y = opt_update(Gs, sc, x.y, fmaf(l2, x.y, gr.y), mm.y, vv.y);
      x.z = opt_update(Gs, sc, x.z, fmaf(l2, x.z, gr.z), mm.z, vv.z);
      x.w = opt_update(Gs, sc, x.w, fmaf(l2, x.w, gr.w), mm.w, vv.w);
      const size_t off = opt_base + (size_t)r * p.ldw + 4 * lane;
      *reinterpret_cast<float4*>(p.W + off) = x;
      if (use_m) *reinterpret_cast<float4*>(p.Mom + off) = mm;
      if (use_v) *reinterpret_cast<float4*>(p.Vel + off) = vv;
      if (p.Wsh) *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.Wsh) + off) = make_uint2(tc::pack_bf16(x.x, x.y), tc::pack_bf16(x.z, x.w));
    }
  } else if (p.epilogue == FEDB_GEMM_WGRAD_OPT) {          // narrow tail tile: plain register loads
    const StepScalars sc = sc_s;
    const float l2 = Gs.alpha * p.l2_scale;
    constexpr int UR = 4;
    for (int r0 = warp; r0 < BM; r0 += NW * UR) {
      size_t off[UR];
      float4 w[UR], mm[UR], vv[UR];
      if (lane < nq) {
#pragma unroll
        for (int u = 0; u < UR; ++u) {
          const int r = r0 + u * NW;
          off[u] = (size_t)g * p.w_group_stride + (size_t)(m0 + r) * p.ldw + n0 + 4 * lane;
          if (r < BM) {
            w[u] = *reinterpret_cast<const float4*>(p.W + off[u]);
            mm[u] = use_m ? *reinterpret_cast<const float4*>(p.Mom + off[u]) : make_float4(0.f, 0.f, 0.f, 0.f);
            vv[u] = use_v ? *reinterpret_cast<const float4*>(p.Vel + off[u]) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
        }
#pragma unroll
        for (int u = 0; u < UR; ++u) {
          const int r = r0 + u * NW;
          if (r < BM) {
            const float4 gr = *reinterpret_cast<const float4*>(tile + r * TILE_LD + 4 * lane);
            float4 x = w[u];
            x.x = opt_update(Gs, sc, x.x, fmaf(l2, x.x, gr.x), mm[u].x, vv[u].x);
            x.y = opt_update(Gs, sc, x.y, fmaf(l2, x.y, gr.y), mm[u].y, vv[u].y);
            x.z = opt_update(Gs, sc, x.z, fmaf(l2, x.z, gr.z), mm[u].z, vv[u].z);
            x.w = opt_update(Gs, sc, x.w, fmaf(l2, x.w, gr.w), mm[u].w, vv[u].w);
            *reinterpret_cast<float4*>(p.W + off[u]) = x;
            if (use_m) *reinterpret_cast<float4*>(p.Mom + off[u]) = mm[u];
            if (use_v) *reinterpret_cast<float4*>(p.Vel + off[u]) = vv[u];
            if (p.Wsh) *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.Wsh) + off[u]) = make_uint2(tc::pack_bf16(x.x, x.y), tc::pack_bf16(x.z, x.w));
          }
        }
      }
    }
  } else if (p.epilogue == FEDB_GEMM_STORE_F32) {
    for (int r = warp; r < BM; r += T_THREADS / 32)
      if (lane < nq)
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.C) + (size_t)g * p.c_group_stride + (size_t)(m0 + r) * p.ldc + n0 + 4 * lane) =
            *reinterpret_cast<const float4*>(tile + r * TILE_LD + 4 * lane);
  } else {
    const float4 bias4 = (p.epilogue == FEDB_GEMM_FWD_BIAS_RELU && p.bias && lane < nq)
                             ? *reinterpret_cast<const float4*>(p.bias + (size_t)g * p.bias_group_stride + n0 + 4 * lane)
                             : make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = warp; r < BM; r += T_THREADS / 32) {
      if (lane < nq) {
        float4 v = *reinterpret_cast<const float4*>(tile + r * TILE_LD + 4 * lane);
        if (p.epilogue == FEDB_GEMM_FWD_BIAS_RELU) {
          v.x = fmaxf(v.x + bias4.x, 0.f); v.y = fmaxf(v.y + bias4.y, 0.f); v.z = fmaxf(v.z + bias4.z, 0.f); v.w = fmaxf(v.w + bias4.w, 0.f);
        } else {
          const uint2 a = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(p.act) + (size_t)g * p.act_group_stride +
                                                          (size_t)(m0 + r) * p.ldact + n0 + 4 * lane);
          const __nv_bfloat16* ab = reinterpret_cast<const __nv_bfloat16*>(&a);
          v.x = __bfloat162float(ab[0]) > 0.f ? v.x : 0.f; v.y = __bfloat162float(ab[1]) > 0.f ? v.y : 0.f;
          v.z = __bfloat162float(ab[2]) > 0.f ? v.z : 0.f; v.w = __bfloat162float(ab[3]) > 0.f ? v.w : 0.f;
        }
        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)g * p.c_group_stride + (size_t)(m0 + r) * p.ldc + n0 + 4 * lane) =
            make_uint2(tc::pack_bf16(v.x, v.y), tc::pack_bf16(v.z, v.w));
      }
    }
  }
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc<128>(tmem);
}

}  // namespace fedb

using namespace fedb;

extern "C" int fedb_gemm_tma(const FedbGemm* args_host, int32_t n_groups, void* stream) {
  const FedbGemm& p = *args_host;
  FEDB_REQUIRE(n_groups > 0 && p.M > 0 && p.N > 0 && p.K > 0, "bad GEMM sizes");
  FEDB_REQUIRE(p.M % 128 == 0 && p.N % 16 == 0 && p.K % 16 == 0, "GEMM needs M % 128 == 0, N % 16 == 0, K % 16 == 0");
  FEDB_REQUIRE(p.lda % 8 == 0 && p.ldb % 8 == 0 && p.a_group_stride % 8 == 0 && p.b_group_stride % 8 == 0,
               "operand leading dimensions / group strides must be multiples of 8 elements (16 bytes)");
  FEDB_REQUIRE(((uintptr_t)p.A % 16) == 0 && ((uintptr_t)p.B % 16) == 0, "operands must be 16-byte aligned");
  alignas(64) CUtensorMap tmA, tmB;
  // K-major source [MN rows][K]: inner = K, box rows = 128 (MN).  MN-major source [K rows][MN]: inner = MN, box rows = 64 (K).
  if (int rc = make_gemm_maps(p, n_groups, &tmA, &tmB)) return rc;
  static bool attr_set = false;
  const int tile_bytes = BM * TILE_LD * 4;
  const int smem = (STAGES * STAGE_BYTES > tile_bytes ? STAGES * STAGE_BYTES : tile_bytes) + 1024;
  if (!attr_set) {
    FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_gemm_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  dim3 grid((p.N + BN - 1) / BN, p.M / BM, n_groups);
  k_gemm_tma<<<grid, T_THREADS, smem, (cudaStream_t)stream>>>(tmA, tmB, p);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}
