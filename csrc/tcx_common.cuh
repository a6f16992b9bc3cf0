// Helpers shared by the split-operand tcgen05 kernels (csrc/fedb_tcx.cu: the federated round; csrc/fedb_fitx.cu: the
// sklearn-style fit): an fp32 value is held as two fp16 numbers, hi = fp16(a) and lo = fp16(a - hi); range scales; distributed
// shared memory by 32-bit shared::cluster addresses; bulk shared-to-shared copies counted on the receiver's mbarrier.
#pragma once
#include <cuda_fp16.h>

#include "tc_common.cuh"

namespace fedb {
namespace {

constexpr float WS = 16.f, INV_WS = 1.f / 16.f;   // weight panels hold WS * w
constexpr float DZS = 32.f;                        // backward operands hold DZS * n * (true gradient)
constexpr float HMAX = 60000.f;                    // activation clamp (fp16 max 65504)

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

// ---- distributed shared memory by 32-bit shared::cluster addresses --------------------------------------------
__device__ __forceinline__ uint32_t mapa(uint32_t saddr, int rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
// bytes [src, src + n) of THIS CTA's shared memory -> shared::cluster address dst; completion (n bytes) is counted on the
// mbarrier at shared::cluster address bar (the receiver's).  n: multiple of 16, addresses 16-byte aligned.
__device__ __forceinline__ void bulk_s2c(uint32_t dst_cluster, uint32_t src_cta, uint32_t n, uint32_t bar_cluster) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_cluster),
               "r"(src_cta), "r"(n), "r"(bar_cluster)
               : "memory");
}
__device__ __forceinline__ void mbar_expect(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc::smem_u32(bar)), "r"(bytes) : "memory");
}
// wait with cluster-scope acquire: the data counted on the barrier was written by other CTAs' bulk copies
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(done)
        : "r"(tc::smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

}  // namespace
}  // namespace fedb
