// Single-CTA tcgen05 probe: runs a chain of tcgen05.mma instructions on operand images that the
// host laid out byte-for-byte, and returns the fp32 accumulator.  tests/test_gpu_tc.py uses it
// to pin down, on hardware, every descriptor convention the fused kernels rely on (K-major and
// MN-major 128B-swizzled panels, N = 208, multi-panel K, accumulation) before those kernels are
// trusted.  Not used on any hot path.
#include "common.cuh"
#include "tc_common.cuh"

namespace fedb {

struct ProbeArgs {
  int a_bytes, b_bytes;
  uint32_t a_lbo, a_sbo, b_lbo, b_sbo;
  uint32_t idesc;
  int n_k;             // MMA instructions in the chain
  int a_step[32];      // byte offset of A's start address for instruction k
  int b_step[32];
  int N;
  int repeat;          // issue the whole chain this many times (accumulating)
};

__global__ void __launch_bounds__(128) k_tc_probe(const uint8_t* __restrict__ a_img, const uint8_t* __restrict__ b_img,
                                                  ProbeArgs args, float* __restrict__ out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint8_t* sa = smem;
  uint8_t* sb = smem + ((args.a_bytes + 1023) & ~1023);
  for (int i = threadIdx.x; i < args.a_bytes / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(sa)[i] = reinterpret_cast<const uint4*>(a_img)[i];
  for (int i = threadIdx.x; i < args.b_bytes / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(sb)[i] = reinterpret_cast<const uint4*>(b_img)[i];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tc::tmem_alloc<256>(&tmem_slot);
  if (threadIdx.x == 0) {
    tc::mbar_init(&bar, 1);
    tc::fence_mbar_init();
  }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    for (int rep = 0; rep < args.repeat; ++rep)
      for (int k = 0; k < args.n_k; ++k) {
        const uint64_t ad = tc::make_desc(tc::smem_u32(sa) + args.a_step[k], args.a_lbo, args.a_sbo);
        const uint64_t bd = tc::make_desc(tc::smem_u32(sb) + args.b_step[k], args.b_lbo, args.b_sbo);
        tc::mma_bf16(tmem, ad, bd, args.idesc, (rep | k) != 0);
      }
    tc::mma_commit(&bar);
  }
  tc::mbar_wait(&bar, 0);
  tc::tc_fence_after();
  const int row = warp * 32 + lane;
  for (int c = 0; c < args.N; c += 8) {
    float v[8];
    tc::tmem_ld8(tmem + ((uint32_t)(warp * 32) << 16) + c, v);
    tc::tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 8; ++i) out[(size_t)row * args.N + c + i] = v[i];
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<256>(tmem);
}

}  // namespace fedb

using namespace fedb;

// args_host: a ProbeArgs laid out as int32[ 9 + 64 + ... ] — see ops/tc.py:probe().
extern "C" int fedb_tc_probe(const void* a_img, const void* b_img, const int32_t* args_host, float* out, void* stream) {
  ProbeArgs a;
  memcpy(&a, args_host, sizeof(ProbeArgs));
  FEDB_REQUIRE(a.n_k > 0 && a.n_k <= 32 && a.N > 0 && a.N <= 256 && a.N % 8 == 0, "bad probe arguments");
  const int smem = ((a.a_bytes + 1023) & ~1023) + ((a.b_bytes + 1023) & ~1023) + 1024;
  FEDB_REQUIRE(smem <= 200 * 1024, "probe operands exceed shared memory");
  FEDB_CUDA_OK(cudaFuncSetAttribute((const void*)k_tc_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  k_tc_probe<<<1, 128, smem, (cudaStream_t)stream>>>((const uint8_t*)a_img, (const uint8_t*)b_img, a, out);
  FEDB_LAUNCH_OK();
  return FEDB_OK;
}
extern "C" int fedb_tc_probe_args_size(void) { return (int)sizeof(ProbeArgs); }
