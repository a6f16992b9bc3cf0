// PTX wrappers shared by the CTA-pair (cta_group::2) GEMM kernels: fedb_gemm_tma2.cu, fedb_gemm_tma2p.cu.
#pragma once
#include <cuda.h>
#include <cstdint>

#include "tc_common.cuh"

namespace fedb {
namespace pair {

constexpr uint32_t PEER_MASK = 0xFEFFFFFFu;                // shared::cluster address of the same offset in the pair's even CTA
constexpr long long WAIT_LIMIT = 20000000ll;               // ~10 ms: a protocol bug must end the kernel, not hang the GPU

// TMA tile load whose completion is signalled on the barrier at `leader_bar` (a shared::cluster address, usually the
// leader CTA's copy of a barrier: own address & PEER_MASK).
__device__ __forceinline__ void tma2_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t leader_bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(dst),
      "l"(map), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2), "l"(0x1000000000000000ull)
      : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc::smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mma2_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate)
      : "memory");
}
__device__ __forceinline__ void commit2(uint64_t* bar) {   // arrive on the barrier at this offset in BOTH CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(tc::smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}
__device__ __forceinline__ void bulk_load_512(uint32_t dst, const float* src, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 512, [%2];" ::"r"(dst), "l"(src), "r"(bar) : "memory");
}
// arrive (count 1) on the barrier at the same offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}\n" ::"r"(tc::smem_u32(bar)),
      "r"(cta)
      : "memory");
}
// bounded wait: returns false on timeout
__device__ __forceinline__ bool mbar_wait_to(uint64_t* bar, uint32_t parity) {
  const long long t0 = clock64();
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"   // suspend-time hint: waiters should not poll
        "selp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(done)
        : "r"(tc::smem_u32(bar)), "r"(parity), "r"(20000u)
        : "memory");
    if (!done && clock64() - t0 > WAIT_LIMIT) return false;
  } while (!done);
  return true;
}
// acquire at cluster scope: pairs with remote arrivals
__device__ __forceinline__ bool mbar_wait_cluster_to(uint64_t* bar, uint32_t parity) {
  const long long t0 = clock64();
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(done)
        : "r"(tc::smem_u32(bar)), "r"(parity), "r"(20000u)
        : "memory");
    if (!done && clock64() - t0 > WAIT_LIMIT) return false;
  } while (!done);
  return true;
}

}  // namespace pair
}  // namespace fedb
