// K2d: grouped GEMM on tcgen05 for the wide-MLP path (BASELINE config 5: 4 x 4096-wide layers, many
// clients per GPU).  One kernel serves the three GEMMs of a Linear layer for every resident client
// (blockIdx.z = client), selected by operand majorness + epilogue:
//
//   forward   A_{l+1} = relu(A_l W_l^T + b_l)      A K-major [B][in],   B K-major [out][in]
//   dgrad     dZ_l    = (dZ_{l+1} W_{l+1}) .* (A_{l+1} > 0)
//                                                   A K-major [B][out],  B MN-major (W stored [out][in])
//   wgrad+opt W_l    <- opt(W_l, dZ_l^T A_{l-1})   A MN-major (dZ stored [B][out]), B MN-major (A stored [B][in])
//
// The weight-gradient tile never reaches memory: its epilogue applies the optimizer (Adam with
// either bias-correction convention, or SGD) to the fp32 master weights and moments and refreshes the
// bf16 shadow copy the next forward pass reads (C:72-73 fused into the GEMM).
//
// 128 x 128 x 64 tiles, UMMA M = 128, N = 16..128, K = 16 per instruction, accumulators in TMEM.
// Operands are staged through a 3-deep ring of 128B-swizzled panels filled with 16-byte cp.async
// (LDGSTS) whose destination address carries the swizzle; tcgen05.commit on a per-stage mbarrier
// releases a stage back to the loaders, so copies of tile k+2 overlap the MMAs of tile k.
// M must be a multiple of 128; N and K multiples of 16.
#include "common.cuh"
#include "tc_common.cuh"

namespace fedb {
namespace {

constexpr int G_THREADS = 128;
constexpr int BM = 128, BN = 128, BK = 64;
constexpr int STAGES = 3;   // 97 KB of shared memory: two CTAs per SM, so one CTA's epilogue overlaps the other's main loop
constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;   // 16 KB each
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;

__device__ __forceinline__ void cp_async16(uint32_t dst_smem, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_smem), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Stage one operand tile.  K-major source [rows = MN][K contiguous]: one panel [128][64].
// MN-major source [rows = K][MN contiguous]: two panels [64 k][64 mn] (mn 0..63 | 64..127).
__device__ __forceinline__ void load_operand(uint32_t dst, const __nv_bfloat16* src, int ld, bool mn_major, int mn0, int mn_cnt,
                                             int k0, int k_cnt) {
  const int tid = threadIdx.x;
  if (!mn_major) {
    const int chunks = k_cnt >> 3;                 // 16-byte chunks per row actually needed
    for (int i = tid; i < BM * 8; i += G_THREADS) {
      const int r = i >> 3, c = i & 7;
      if (c < chunks) {
        if (r < mn_cnt) cp_async16(dst + r * 128 + ((c ^ (r & 7)) << 4), src + (size_t)(mn0 + r) * ld + k0 + 8 * c);
        else *reinterpret_cast<uint4*>(__cvta_shared_to_generic(dst + r * 128 + ((c ^ (r & 7)) << 4))) = make_uint4(0, 0, 0, 0);
      }
    }
  } else {
    const int chunks = (mn_cnt + 7) >> 3;          // chunks along MN actually present
    for (int i = tid; i < BK * 16; i += G_THREADS) {
      const int kk = i >> 4, c = i & 15;
      if (kk < k_cnt) {
        const uint32_t d = dst + (c >> 3) * (BK * 128) + kk * 128 + (((c & 7) ^ (kk & 7)) << 4);
        if (c < chunks) cp_async16(d, src + (size_t)(k0 + kk) * ld + mn0 + 8 * c);
        else *reinterpret_cast<uint4*>(__cvta_shared_to_generic(d)) = make_uint4(0, 0, 0, 0);
      }
    }
  }
}

}  // namespace

__global__ void __launch_bounds__(G_THREADS) k_gemm_tc(FedbGemm p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t empty_bar[STAGES], done_bar;
  __shared__ uint32_t tmem_slot;
  __shared__ StepScalars sc_s;
  const uint32_t raw = tc::smem_u32(smem_raw);
  const uint32_t sm = (raw + 1023u) & ~1023u;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int g = blockIdx.z, m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int n_cnt = min(BN, p.N - n0);             // multiple of 16
  const int m_cnt = min(BM, p.M - m0);
  const __nv_bfloat16* A = reinterpret_cast<const __nv_bfloat16*>(p.A) + (size_t)g * p.a_group_stride;
  const __nv_bfloat16* B = reinterpret_cast<const __nv_bfloat16*>(p.B) + (size_t)g * p.b_group_stride;

  if (warp == 0) tc::tmem_alloc<128>(&tmem_slot);
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) tc::mbar_init(&empty_bar[s], 1);
    tc::mbar_init(&done_bar, 1);
    tc::fence_mbar_init();
    if (p.epilogue == FEDB_GEMM_WGRAD_OPT) sc_s = step_scalars_fast(p.groups[g], p.step[g]);
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t idesc = tc::make_idesc_bf16(BM, n_cnt, p.a_mn != 0, p.b_mn != 0);
  const int KT = (p.K + BK - 1) / BK;

  auto issue_loads = [&](int kt) {
    const int s = kt % STAGES, k0 = kt * BK, k_cnt = min(BK, p.K - k0);
    const uint32_t a_dst = sm + s * STAGE_BYTES, b_dst = a_dst + A_BYTES;
    load_operand(a_dst, A, p.lda, p.a_mn != 0, m0, m_cnt, k0, k_cnt);
    load_operand(b_dst, B, p.ldb, p.b_mn != 0, n0, n_cnt, k0, k_cnt);
  };

  // prologue: fill STAGES-1 stages
  for (int kt = 0; kt < STAGES - 1; ++kt) {
    if (kt < KT) issue_loads(kt);
    cp_async_commit();
  }
  uint32_t empty_phase = 0;   // bit s = parity to wait for on empty_bar[s]
  for (int kt = 0; kt < KT; ++kt) {
    const int s = kt % STAGES;
    cp_async_wait<STAGES - 2>();                   // this thread's copies of tile kt have landed
    tc::fence_async_smem();
    __syncthreads();                               // ... and everybody else's
    if (tid == 0) {
      tc::tc_fence_after();
      const int k_cnt = min(BK, p.K - kt * BK);
      const uint32_t a_base = sm + s * STAGE_BYTES, b_base = a_base + A_BYTES;
      for (int k = 0; k < (k_cnt >> 4); ++k) {
        const uint64_t ad = p.a_mn ? tc::make_desc(a_base + 2048 * k, BK * 128, 1024) : tc::make_desc(a_base + 32 * k, 0, 1024);
        const uint64_t bd = p.b_mn ? tc::make_desc(b_base + 2048 * k, BK * 128, 1024) : tc::make_desc(b_base + 32 * k, 0, 1024);
        tc::mma_bf16(tmem, ad, bd, idesc, (kt | k) != 0);
      }
      tc::mma_commit(&empty_bar[s]);               // stage s is free once these MMAs have read it
      if (kt == KT - 1) tc::mma_commit(&done_bar);
    }
    // refill the stage that tile kt-1 used (its MMAs were committed one iteration ago)
    const int nk = kt + STAGES - 1;
    if (nk < KT) {
      const int ns = nk % STAGES;
      if (nk >= STAGES) {                          // the stage has been used before: wait until its MMAs are done
        tc::mbar_wait(&empty_bar[ns], (empty_phase >> ns) & 1u);
        empty_phase ^= 1u << ns;
      }
      issue_loads(nk);
    }
    cp_async_commit();
  }
  tc::mbar_wait(&done_bar, 0);
  tc::tc_fence_after();

  // ---- epilogue: thread = accumulator row ---------------------------------------------------------
  const int row = m0 + tid;
  const uint32_t tl = tmem + ((uint32_t)(warp * 32) << 16);
  const bool row_ok = tid < m_cnt;
  if (p.epilogue == FEDB_GEMM_FWD_BIAS_RELU || p.epilogue == FEDB_GEMM_DGRAD_MASK) {
    __nv_bfloat16* C = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)g * p.c_group_stride + (size_t)row * p.ldc + n0;
    const float* bias = p.bias ? p.bias + (size_t)g * p.bias_group_stride + n0 : nullptr;
    const __nv_bfloat16* act = p.act ? reinterpret_cast<const __nv_bfloat16*>(p.act) + (size_t)g * p.act_group_stride + (size_t)row * p.ldact + n0 : nullptr;
    for (int c0 = 0; c0 < n_cnt; c0 += 16) {
      float v[16];
      tc::tmem_ld16(tl + c0, v);
      tc::tmem_ld_wait();
      if (row_ok) {
        if (p.epilogue == FEDB_GEMM_FWD_BIAS_RELU) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i] + (bias ? __ldg(bias + c0 + i) : 0.f), 0.f);
        } else {
          const uint4 a0 = *reinterpret_cast<const uint4*>(act + c0), a1 = *reinterpret_cast<const uint4*>(act + c0 + 8);
          const __nv_bfloat16* ab0 = reinterpret_cast<const __nv_bfloat16*>(&a0);
          const __nv_bfloat16* ab1 = reinterpret_cast<const __nv_bfloat16*>(&a1);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            v[i] = __bfloat162float(ab0[i]) > 0.f ? v[i] : 0.f;
            v[8 + i] = __bfloat162float(ab1[i]) > 0.f ? v[8 + i] : 0.f;
          }
        }
        *reinterpret_cast<uint4*>(C + c0) = make_uint4(tc::pack_bf16(v[0], v[1]), tc::pack_bf16(v[2], v[3]), tc::pack_bf16(v[4], v[5]), tc::pack_bf16(v[6], v[7]));
        *reinterpret_cast<uint4*>(C + c0 + 8) = make_uint4(tc::pack_bf16(v[8], v[9]), tc::pack_bf16(v[10], v[11]), tc::pack_bf16(v[12], v[13]), tc::pack_bf16(v[14], v[15]));
      }
    }
  } else if (p.epilogue == FEDB_GEMM_WGRAD_OPT) {
    const FedbGroup& G = p.groups[g];
    const StepScalars sc = sc_s;
    const size_t off = (size_t)g * p.w_group_stride + (size_t)row * p.ldw + n0;
    float* W = p.W + off;
    float* Mo = p.Mom ? p.Mom + off : nullptr;
    float* Ve = (p.Vel && G.opt != FEDB_OPT_SGD) ? p.Vel + off : nullptr;
    __nv_bfloat16* Ws = p.Wsh ? reinterpret_cast<__nv_bfloat16*>(p.Wsh) + off : nullptr;
    const float l2 = G.alpha * p.l2_scale;
    for (int c0 = 0; c0 < n_cnt; c0 += 16) {
      float v[16];
      tc::tmem_ld16(tl + c0, v);
      tc::tmem_ld_wait();
      if (row_ok) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float4 w = *reinterpret_cast<float4*>(W + c0 + 4 * q);
          float4 mm = Mo ? *reinterpret_cast<float4*>(Mo + c0 + 4 * q) : make_float4(0.f, 0.f, 0.f, 0.f);
          float4 vv = Ve ? *reinterpret_cast<float4*>(Ve + c0 + 4 * q) : make_float4(0.f, 0.f, 0.f, 0.f);
          w.x = opt_update(G, sc, w.x, fmaf(l2, w.x, v[4 * q]), mm.x, vv.x);
          w.y = opt_update(G, sc, w.y, fmaf(l2, w.y, v[4 * q + 1]), mm.y, vv.y);
          w.z = opt_update(G, sc, w.z, fmaf(l2, w.z, v[4 * q + 2]), mm.z, vv.z);
          w.w = opt_update(G, sc, w.w, fmaf(l2, w.w, v[4 * q + 3]), mm.w, vv.w);
          *reinterpret_cast<float4*>(W + c0 + 4 * q) = w;
          if (Mo) *reinterpret_cast<float4*>(Mo + c0 + 4 * q) = mm;
          if (Ve) *reinterpret_cast<float4*>(Ve + c0 + 4 * q) = vv;
          if (Ws) *reinterpret_cast<uint2*>(Ws + c0 + 4 * q) = make_uint2(tc::pack_bf16(w.x, w.y), tc::pack_bf16(w.z, w.w));
        }
      }
    }
  } else {  // FEDB_GEMM_STORE_F32: plain fp32 result (tests, gradients for inspection)
    float* C = reinterpret_cast<float*>(p.C) + (size_t)g * p.c_group_stride + (size_t)row * p.ldc + n0;
    for (int c0 = 0; c0 < n_cnt; c0 += 16) {
      float v[16];
      tc::tmem_ld16(tl + c0, v);
      tc::tmem_ld_wait();
      if (row_ok) {
#pragma unroll
        for (int q = 0; q < 4; ++q) *reinterpret_cast<float4*>(C + c0 + 4 * q) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<128>(tmem);
}

}  // namespace fedb

using namespace fedb;

extern "C" int32_t fedb_gemm_args_size(void) { return (int32_t)sizeof(FedbGemm); }

extern "C" int fedb_gemm_tc(const FedbGemm* args_host, int32_t n_groups, void* stream) {
  const FedbGemm& p = *args_host;
  FEDB_REQUIRE(n_groups > 0 && p.M > 0 && p.N > 0 && p.K > 0, "bad GEMM sizes");
  FEDB_REQUIRE(p.M % 128 == 0 && p.N % 16 == 0 && p.K % 16 == 0, "GEMM needs M % 128 == 0, N % 16 == 0, K % 16 == 0");
  FEDB_REQUIRE(p.lda % 8 == 0 && p.ldb % 8 == 0, "operand leading dimensions must be multiples of 8 elements (16-byte rows)");
  static bool attr_set = false;
  const int smem = STAGES * STAGE_BYTES + 1024;
  if (!attr_set) {
    FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_gemm_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  dim3 grid((p.N + BN - 1) / BN, p.M / BM, n_groups);
  k_gemm_tc<<<grid, G_THREADS, smem, (cudaStream_t)stream>>>(p);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}
