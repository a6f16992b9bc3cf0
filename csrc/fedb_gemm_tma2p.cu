// K2d, persistent CTA-pair variant.  One cluster of two CTAs per SM pair walks a static list of 256 x 256 output tiles
// (fedb_gemm_tma2.cu computes one such tile per cluster and relies on a second resident CTA for overlap).  Here the
// overlap is explicit, inside the CTA:
//   warp 0      TMA producer: operand stages for tile after tile, never waiting for an epilogue
//   warp 1      tcgen05.mma issuer (leader CTA): alternates between two 256-column accumulators in tensor memory
//   warps 2-9   epilogue: drain accumulator b of tile i while the tensor cores fill accumulator 1-b with tile i+1
// so the eight epilogue warps stream continuously — which is what the optimizer epilogue needs: it moves 26 bytes per
// output element (W, momentum, second moment in and out, bf16 shadow out) and is bound by HBM, not by the MMAs.  Its
// W / m / v rows arrive through per-warp cp.async.bulk rings that run ahead across tile boundaries.
// Requires M % 256 == 0 and N % 256 == 0.  Arguments and epilogues: FedbGemm (include/fedb200.h).
#include <cooperative_groups.h>
#include <cuda.h>

#include "common.cuh"
#include "pair_common.cuh"
#include "tma_map.cuh"
#include "tc_common.cuh"

namespace cg = cooperative_groups;

namespace fedb {
namespace {

using namespace pair;

constexpr int P_THREADS = 320, EPI_WARPS = 8, EPI_THREADS = EPI_WARPS * 32;
constexpr int BM = 128, BN = 256, BNH = 128, BK = 64;
constexpr int A_BYTES = BM * BK * 2, B_BYTES = BNH * BK * 2, STAGE_BYTES = A_BYTES + B_BYTES;   // 32 KB per CTA and stage
constexpr int TILE_BYTES = BM * 128 * 4;                   // fp32 staging tile of one 128-column pass, 16-byte XOR swizzle
constexpr int RING_D = 5, RING_SLOT = 3 * 512;             // optimizer rows in flight per epilogue warp
constexpr int STAGES_MAX = 5;
template <bool OPT> struct PCfg {
  static constexpr int STAGES = OPT ? 3 : 5;               // the optimizer variant is memory-bound: shared memory goes to the rings
  static constexpr int TILE_OFF = STAGES * STAGE_BYTES;
  static constexpr int RING_OFF = TILE_OFF + TILE_BYTES;
  static constexpr int SMEM = RING_OFF + (OPT ? EPI_WARPS * RING_D * RING_SLOT : 0) + 1024;
};
static_assert(PCfg<true>::SMEM <= 227 * 1024 && PCfg<false>::SMEM <= 227 * 1024, "shared memory budget");

__device__ __forceinline__ int tile_at(int r, int c4) { return r * 128 + ((c4 ^ (r & 7)) << 2); }
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory"); }

struct TileCoord { int g, m0, n0; };
__device__ __forceinline__ TileCoord tile_coord(int t, int tiles_m, int tiles_n, int rank) {
  TileCoord c;
  const int mp = t % tiles_m, rest = t / tiles_m;
  c.m0 = mp * 256 + rank * BM;
  c.n0 = (rest % tiles_n) * BN;
  c.g = rest / tiles_n;
  return c;
}

}  // namespace

template <bool OPT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(P_THREADS, 1)
k_gemm_tma2p(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, FedbGemm p, int n_tiles, int tiles_m,
             int tiles_n, int* __restrict__ err) {
  extern __shared__ uint8_t smem_raw[];
  constexpr int STAGES = PCfg<OPT>::STAGES;
  __shared__ uint64_t full_bar[STAGES_MAX], empty_bar[STAGES_MAX], acc_full[2], acc_empty[2], ring_bar[EPI_WARPS][RING_D];
  __shared__ uint32_t tmem_slot;
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();              // 0 = leader (issues the MMAs)
  const uint32_t raw = tc::smem_u32(smem_raw);
  const uint32_t sm = (raw + 1023u) & ~1023u;
  uint8_t* smp = smem_raw + (sm - raw);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int cid = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int KT = (p.K + BK - 1) / BK;

  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc::smem_u32(&tmem_slot)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { tc::mbar_init(&acc_full[b], 1); tc::mbar_init(&acc_empty[b], 2 * EPI_WARPS); }
    for (int i = 0; i < EPI_WARPS * RING_D; ++i) tc::mbar_init(&ring_bar[0][0] + i, 1);
    tc::fence_mbar_init();
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  tc::tc_fence_before();
  cluster.sync();                                          // the peer's barriers exist before anything signals them
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  bool ok = true;                                          // a wait that times out is reported and then ignored: every loop below has
                                                           // a static trip count, so the kernel always terminates
  if (warp == 0) {
    // ===== TMA producer (one lane per CTA): own A rows + own half of B; completion on the LEADER's full barrier =====
    // (elected with elect.sync, not `lane == 0`: under a plain divergent branch the compiler wraps every UTMALDG / UTCHMMA in
    //  a loop over the active lanes)
    __syncwarp();
    if (tc::elect_one()) {
      int it = 0;
      for (int t = cid; t < n_tiles; t += n_clusters) {
        const TileCoord c = tile_coord(t, tiles_m, tiles_n, rank);
        const int nh0 = c.n0 + rank * BNH;
        for (int kt = 0; kt < KT; ++kt, ++it) {
          const int s = it % STAGES, k0 = kt * BK;
          if (ok) ok = mbar_wait_to(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
          if (rank == 0) mbar_expect_tx(&full_bar[s], 2 * STAGE_BYTES);
          const uint32_t lbar = tc::smem_u32(&full_bar[s]) & PEER_MASK;
          const uint32_t a_dst = sm + s * STAGE_BYTES, b_dst = a_dst + A_BYTES;
          if (!p.a_mn) tma2_load_3d(a_dst, &tmA, k0, c.m0, c.g, lbar);
          else { tma2_load_3d(a_dst, &tmA, c.m0, k0, c.g, lbar); tma2_load_3d(a_dst + BK * 128, &tmA, c.m0 + 64, k0, c.g, lbar); }
          if (!p.b_mn) tma2_load_3d(b_dst, &tmB, k0, nh0, c.g, lbar);
          else { tma2_load_3d(b_dst, &tmB, nh0, k0, c.g, lbar); tma2_load_3d(b_dst + BK * 128, &tmB, nh0 + 64, k0, c.g, lbar); }
        }
      }
      if (!ok && err) *err = FEDB_E_TIMEOUT;
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one lane of the leader CTA drives both SMs' tensor cores, two accumulators in turn =====
    __syncwarp();
    if (rank == 0 && tc::elect_one()) {
      const uint32_t idesc = tc::make_idesc_bf16(2 * BM, BN, p.a_mn != 0, p.b_mn != 0);
      int it = 0, tc_n = 0;
      for (int t = cid; t < n_tiles; t += n_clusters, ++tc_n) {
        const int b = tc_n & 1;
        if (ok) ok = mbar_wait_cluster_to(&acc_empty[b], ((tc_n >> 1) & 1) ^ 1);    // both CTAs' epilogues have drained accumulator b
        tc::tc_fence_after();
        for (int kt = 0; kt < KT; ++kt, ++it) {
          const int s = it % STAGES;
          if (ok) ok = mbar_wait_to(&full_bar[s], (it / STAGES) & 1);
          tc::tc_fence_after();
          const int k_cnt = min(BK, p.K - kt * BK);
          const uint32_t a_base = sm + s * STAGE_BYTES, b_base = a_base + A_BYTES;
          for (int k = 0; k < (k_cnt >> 4); ++k) {
            const uint64_t ad = p.a_mn ? tc::make_desc(a_base + 2048 * k, BK * 128, 1024) : tc::make_desc(a_base + 32 * k, 0, 1024);
            const uint64_t bd = p.b_mn ? tc::make_desc(b_base + 2048 * k, BK * 128, 1024) : tc::make_desc(b_base + 32 * k, 0, 1024);
            mma2_bf16(tmem + b * 256, ad, bd, idesc, (kt | k) != 0);
          }
          commit2(&empty_bar[s]);
        }
        commit2(&acc_full[b]);
      }
      if (!ok && err) *err = FEDB_E_TIMEOUT;
    }
  } else {
    // ===== epilogue warps =====
    const int e = warp - 2;                                // 0..7: rows e, e + 8, ... of the staged tile
    const int quad = warp & 3, chalf = e >> 2;             // tensor-memory lane quadrant this warp may read; 64-column half it stages
    float* tile = reinterpret_cast<float*>(smp + PCfg<OPT>::TILE_OFF);
    // --- optimizer-state ring of this warp: requests run ahead of consumption, across passes and tiles ---
    const uint32_t ring_sm = sm + PCfg<OPT>::RING_OFF + e * RING_D * RING_SLOT;
    const float* ring = reinterpret_cast<const float*>(smp + PCfg<OPT>::RING_OFF + e * RING_D * RING_SLOT);
    int is_t = cid, is_h = 0, is_j = 0, is_d = 0;          // next row to request: tile, pass, ordinal, slot
    size_t is_base = 0;
    bool is_m = false, is_v = false;
    auto ring_ctx = [&]() {                                // per-tile constants of the request side
      if (is_t < n_tiles) {
        const TileCoord c = tile_coord(is_t, tiles_m, tiles_n, rank);
        const FedbGroup& G = p.groups[c.g];
        const int opt = __ldg(&G.opt);
        is_v = p.Vel && opt != FEDB_OPT_SGD;
        is_m = p.Mom && !(opt == FEDB_OPT_SGD && __ldg(&G.momentum) == 0.f);   // plain SGD keeps no state: skip its traffic
        is_base = (size_t)c.g * p.w_group_stride + (size_t)c.m0 * p.ldw + c.n0;
      }
    };
    auto ring_issue = [&]() {                              // lane 0 only; no-op once every tile is requested
      if (is_t < n_tiles) {
        const size_t off = is_base + (size_t)(e + is_j * EPI_WARPS) * p.ldw + is_h * 128;
        const uint32_t bar = tc::smem_u32(&ring_bar[e][is_d]), dst = ring_sm + is_d * RING_SLOT;
        mbar_expect_tx(&ring_bar[e][is_d], 512u * (1 + (int)is_m + (int)is_v));
        bulk_load_512(dst, p.W + off, bar);
        if (is_m) bulk_load_512(dst + 512, p.Mom + off, bar);
        if (is_v) bulk_load_512(dst + 1024, p.Vel + off, bar);
        if (++is_d == RING_D) is_d = 0;
        if (++is_j == BM / EPI_WARPS) {
          is_j = 0;
          if (++is_h == 2) { is_h = 0; is_t += n_clusters; ring_ctx(); }
        }
      }
    };
    if (OPT && lane == 0) {
      ring_ctx();
      for (int i = 0; i < RING_D; ++i) ring_issue();
    }
    int cs_d = 0;
    uint32_t cs_par = 0;

    int tc_n = 0;
    for (int t = cid; t < n_tiles; t += n_clusters, ++tc_n) {
      const TileCoord c = tile_coord(t, tiles_m, tiles_n, rank);
      const int b = tc_n & 1;
      OptRegs o;
      bool use_m = false, use_v = false;
      if constexpr (OPT) {
        const FedbGroup& G = p.groups[c.g];
        FedbGroup Gr;                                      // only the optimizer fields are read
        Gr.opt = __ldg(&G.opt); Gr.lr = __ldg(&G.lr); Gr.lr_gamma = __ldg(&G.lr_gamma); Gr.lr_step_size = __ldg(&G.lr_step_size);
        Gr.beta1 = __ldg(&G.beta1); Gr.beta2 = __ldg(&G.beta2); Gr.eps = __ldg(&G.eps); Gr.momentum = __ldg(&G.momentum);
        Gr.nesterov = __ldg(&G.nesterov); Gr.alpha = __ldg(&G.alpha);
        o = opt_regs(Gr, step_scalars_fast(Gr, __ldg(p.step + c.g)), Gr.alpha * p.l2_scale);
        use_v = p.Vel && Gr.opt != FEDB_OPT_SGD;
        use_m = p.Mom && !(Gr.opt == FEDB_OPT_SGD && Gr.momentum == 0.f);
      }
      if (ok) ok = mbar_wait_to(&acc_full[b], (tc_n >> 1) & 1);
      tc::tc_fence_after();
      for (int hc = 0; hc < 2; ++hc) {
        const int nc0 = c.n0 + hc * 128;                   // global column base of this pass
        {                                                  // accumulator rows -> staging tile: this warp's quadrant, 64 columns
          const int r = quad * 32 + lane;
          const uint32_t tl = tmem + ((uint32_t)(quad * 32) << 16) + b * 256 + hc * 128 + chalf * 64;
#pragma unroll
          for (int c0 = 0; c0 < 64; c0 += 16) {
            float v[16];
            tc::tmem_ld16(tl + c0, v);
            tc::tmem_ld_wait();
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4)
              *reinterpret_cast<float4*>(tile + tile_at(r, chalf * 16 + (c0 >> 2) + q4)) = make_float4(v[4 * q4], v[4 * q4 + 1], v[4 * q4 + 2], v[4 * q4 + 3]);
          }
        }
        if (hc == 1) {                                     // this warp is done with accumulator b: hand it back to the MMA issuer
          tc::tc_fence_before();
          if (lane == 0) mbar_arrive_remote(&acc_empty[b], 0);
        }
        epi_bar();
        if constexpr (OPT) {
          const size_t opt_base = (size_t)c.g * p.w_group_stride + (size_t)c.m0 * p.ldw + nc0;
          for (int j = 0; j < BM / EPI_WARPS; j += 2) {    // two rows per trip: independent chains hide SFU and shared-memory latency
            float4 x[2], mm[2], vv[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              if (ok) ok = mbar_wait_to(&ring_bar[e][cs_d], cs_par);
              const float* slot = ring + cs_d * (RING_SLOT / 4);
              x[u] = *reinterpret_cast<const float4*>(slot + 4 * lane);
              mm[u] = vv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (use_m) mm[u] = *reinterpret_cast<const float4*>(slot + 128 + 4 * lane);
              if (use_v) vv[u] = *reinterpret_cast<const float4*>(slot + 256 + 4 * lane);
              if (++cs_d == RING_D) { cs_d = 0; cs_par ^= 1u; }
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int r = e + (j + u) * EPI_WARPS;
              const float4 gr = *reinterpret_cast<const float4*>(tile + tile_at(r, lane));
              x[u].x = opt_update_fast(o, x[u].x, gr.x, mm[u].x, vv[u].x);
              x[u].y = opt_update_fast(o, x[u].y, gr.y, mm[u].y, vv[u].y);
              x[u].z = opt_update_fast(o, x[u].z, gr.z, mm[u].z, vv[u].z);
              x[u].w = opt_update_fast(o, x[u].w, gr.w, mm[u].w, vv[u].w);
              const size_t off = opt_base + (size_t)r * p.ldw + 4 * lane;
              *reinterpret_cast<float4*>(p.W + off) = x[u];
              if (use_m) *reinterpret_cast<float4*>(p.Mom + off) = mm[u];
              if (use_v) *reinterpret_cast<float4*>(p.Vel + off) = vv[u];
              if (p.Wsh) *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.Wsh) + off) = make_uint2(tc::pack_bf16(x[u].x, x[u].y), tc::pack_bf16(x[u].z, x[u].w));
            }
            __syncwarp();                                  // every lane has consumed its slot values: the slots may be refilled
            if (lane == 0) { ring_issue(); ring_issue(); }
          }
        } else if (p.epilogue == FEDB_GEMM_STORE_F32) {
          for (int r = e; r < BM; r += EPI_WARPS)
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.C) + (size_t)c.g * p.c_group_stride + (size_t)(c.m0 + r) * p.ldc + nc0 + 4 * lane) =
                *reinterpret_cast<const float4*>(tile + tile_at(r, lane));
        } else {
          const float4 bias4 = (p.epilogue == FEDB_GEMM_FWD_BIAS_RELU && p.bias)
                                   ? __ldg(reinterpret_cast<const float4*>(p.bias + (size_t)c.g * p.bias_group_stride + nc0 + 4 * lane))
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
          for (int r = e; r < BM; r += EPI_WARPS) {
            float4 v = *reinterpret_cast<const float4*>(tile + tile_at(r, lane));
            if (p.epilogue == FEDB_GEMM_FWD_BIAS_RELU) {
              v.x = fmaxf(v.x + bias4.x, 0.f); v.y = fmaxf(v.y + bias4.y, 0.f); v.z = fmaxf(v.z + bias4.z, 0.f); v.w = fmaxf(v.w + bias4.w, 0.f);
            } else {
              const uint2 a = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(p.act) + (size_t)c.g * p.act_group_stride +
                                                              (size_t)(c.m0 + r) * p.ldact + nc0 + 4 * lane);
              const __nv_bfloat16* ab = reinterpret_cast<const __nv_bfloat16*>(&a);
              v.x = __bfloat162float(ab[0]) > 0.f ? v.x : 0.f; v.y = __bfloat162float(ab[1]) > 0.f ? v.y : 0.f;
              v.z = __bfloat162float(ab[2]) > 0.f ? v.z : 0.f; v.w = __bfloat162float(ab[3]) > 0.f ? v.w : 0.f;
            }
            *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)c.g * p.c_group_stride + (size_t)(c.m0 + r) * p.ldc + nc0 + 4 * lane) =
                make_uint2(tc::pack_bf16(v.x, v.y), tc::pack_bf16(v.z, v.w));
          }
        }
        epi_bar();                                         // the staging tile is rewritten by the next pass
      }
    }
    if (!ok && err) *err = FEDB_E_TIMEOUT;
  }
  tc::tc_fence_before();
  cluster.sync();                                          // both CTAs are done with tensor memory and with each other's barriers
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
}

}  // namespace fedb

using namespace fedb;

// err_dev: optional device int that receives FEDB_E_TIMEOUT if a barrier wait of the pipeline times out.
extern "C" int fedb_gemm_tma2p(const FedbGemm* args_host, int32_t n_groups, int32_t* err_dev, void* stream) {
  const FedbGemm& p = *args_host;
  FEDB_REQUIRE(n_groups > 0 && p.M % 256 == 0 && p.N % 256 == 0 && p.K % 16 == 0 && p.K > 0, "persistent pair GEMM needs M % 256 == 0, N % 256 == 0, K % 16 == 0");
  FEDB_REQUIRE(p.lda % 8 == 0 && p.ldb % 8 == 0 && p.a_group_stride % 8 == 0 && p.b_group_stride % 8 == 0, "strides must be multiples of 8 elements");
  alignas(64) CUtensorMap tmA, tmB;
  if (int rc = make_gemm_maps(p, n_groups, &tmA, &tmB)) return rc;
  // clusters that can be resident at once (a GPC with an odd SM count leaves one SM unpaired): the static tile lists
  // are balanced over exactly that many clusters
  static int max_clusters[2] = {0, 0};
  if (!max_clusters[0]) {
    FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_gemm_tma2p<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, PCfg<false>::SMEM));
    FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_gemm_tma2p<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, PCfg<true>::SMEM));
    int dev = 0, n_sm = 0;
    FEDB_CUDA_OK(cudaGetDevice(&dev));
    FEDB_CUDA_OK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
    for (int v = 0; v < 2; ++v) {
      cudaLaunchConfig_t cfg = {};
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.gridDim = dim3(n_sm & ~1); cfg.blockDim = dim3(P_THREADS);
      cfg.dynamicSmemBytes = v ? PCfg<true>::SMEM : PCfg<false>::SMEM;
      cfg.attrs = at; cfg.numAttrs = 1;
      int n = 0;
      cudaError_t e = cudaOccupancyMaxActiveClusters(&n, v ? (const void*)k_gemm_tma2p<true> : (const void*)k_gemm_tma2p<false>, &cfg);
      if (e != cudaSuccess || n <= 0) { (void)cudaGetLastError(); n = n_sm / 2 - 4; }
      max_clusters[v] = n < n_sm / 2 ? n : n_sm / 2;
    }
  }
  const int tiles_m = p.M / 256, tiles_n = p.N / 256, n_tiles = n_groups * tiles_m * tiles_n;
  const int cap = max_clusters[p.epilogue == FEDB_GEMM_WGRAD_OPT ? 1 : 0];
  const int n_clusters = n_tiles < cap ? n_tiles : cap;
  if (p.epilogue == FEDB_GEMM_WGRAD_OPT)
    k_gemm_tma2p<true><<<2 * n_clusters, P_THREADS, PCfg<true>::SMEM, (cudaStream_t)stream>>>(tmA, tmB, p, n_tiles, tiles_m, tiles_n, err_dev);
  else
    k_gemm_tma2p<false><<<2 * n_clusters, P_THREADS, PCfg<false>::SMEM, (cudaStream_t)stream>>>(tmA, tmB, p, n_tiles, tiles_m, tiles_n, err_dev);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}
