// How fast can a kernel pull 1000 rows x 64 B from pinned host memory, depending on the access pattern?
// (sizing the in-kernel host-to-device copy of fedb_round_tc_hostin)   nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// (a) two threads per row, each two 16-byte volatile loads (the first implementation)
__global__ void k_pair(const float* __restrict__ xh, const int* __restrict__ idx, float* __restrict__ out, int n) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x, r = t >> 1, h = t & 1;
  if (r < n) {
    const float4* s = reinterpret_cast<const float4*>(xh + (size_t)idx[r] * 16) + 2 * h;
    float4 a = __ldcv(s), b = __ldcv(s + 1);
    float4* d = reinterpret_cast<float4*>(out + (size_t)r * 16) + 2 * h;
    d[0] = a; d[1] = b;
  }
}
// (b) four lanes per row, one 16-byte load each: one 64-byte request per row
__global__ void k_quad(const float* __restrict__ xh, const int* __restrict__ idx, float* __restrict__ out, int n, int nc) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x, r = t >> 2, c = t & 3;
  if (r < n) {
    const float4* s = reinterpret_cast<const float4*>(xh + (size_t)idx[r] * 16) + c;
    float4 a = nc ? __ldg(s) : __ldcv(s);
    reinterpret_cast<float4*>(out + (size_t)r * 16)[c] = a;
  }
}
// (c) one cp.async.bulk of 64 B per row into shared memory, then copy out
__global__ void k_bulk_rows(const float* __restrict__ xh, const int* __restrict__ idx, float* __restrict__ out, int n) {
  extern __shared__ __align__(128) float sm[];
  __shared__ uint64_t bar;
  const int r0 = blockIdx.x * 128, cnt = min(128, n - r0);
  if (threadIdx.x == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;"); }
  __syncthreads();
  if (threadIdx.x == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar)), "r"(cnt * 64) : "memory");
  __syncthreads();
  if (threadIdx.x < cnt)
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 64, [%2];" ::"r"(s32(sm + threadIdx.x * 16)),
                 "l"(xh + (size_t)idx[r0 + threadIdx.x] * 16), "r"(s32(&bar)) : "memory");
  uint32_t done;
  do { asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(done) : "r"(s32(&bar)) : "memory"); } while (!done);
  for (int i = threadIdx.x; i < cnt * 4; i += blockDim.x) reinterpret_cast<float4*>(out + (size_t)r0 * 16)[i] = reinterpret_cast<float4*>(sm)[i];
}
// (d) contiguous rows: one bulk copy of the block's 8 KB
__global__ void k_bulk_block(const float* __restrict__ xh, float* __restrict__ out, int n) {
  extern __shared__ __align__(128) float sm[];
  __shared__ uint64_t bar;
  const int r0 = blockIdx.x * 128, cnt = min(128, n - r0);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar)), "r"(cnt * 64) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %3, [%2];" ::"r"(s32(sm)), "l"(xh + (size_t)r0 * 16), "r"(s32(&bar)), "r"(cnt * 64) : "memory");
  }
  __syncthreads();
  uint32_t done;
  do { asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(done) : "r"(s32(&bar)) : "memory"); } while (!done);
  for (int i = threadIdx.x; i < cnt * 4; i += blockDim.x) reinterpret_cast<float4*>(out + (size_t)r0 * 16)[i] = reinterpret_cast<float4*>(sm)[i];
}
__global__ void k_empty() {}

template <typename F> float timeit(F f, int n = 200) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 10; ++i) f();
  cudaDeviceSynchronize();
  cudaEventRecord(a); for (int i = 0; i < n; ++i) f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms / n * 1e3f;
}

int main() {
  const int n = 1000;
  float* xh; int* idx_h; float* xd; float* out; int *idx_seq, *idx_perm;
  cudaHostAlloc(&xh, n * 64, cudaHostAllocDefault);
  for (int i = 0; i < n * 16; ++i) xh[i] = (float)i;
  cudaMalloc(&xd, n * 64); cudaMalloc(&out, n * 64); cudaMalloc(&idx_seq, n * 4); cudaMalloc(&idx_perm, n * 4);
  idx_h = new int[n];
  for (int i = 0; i < n; ++i) idx_h[i] = i;
  cudaMemcpy(idx_seq, idx_h, n * 4, cudaMemcpyHostToDevice);
  for (int i = 0; i < n; ++i) idx_h[i] = (int)((i * 617ll) % n);
  cudaMemcpy(idx_perm, idx_h, n * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(xd, xh, n * 64, cudaMemcpyHostToDevice);
  printf("empty kernel launch-to-launch           %6.2f us\n", timeit([&] { k_empty<<<8, 256>>>(); }));
  printf("cudaMemcpyAsync H2D 64 KB               %6.2f us\n", timeit([&] { cudaMemcpyAsync(xd, xh, n * 64, cudaMemcpyHostToDevice, 0); }));
  for (int perm = 0; perm < 2; ++perm) {
    const int* idx = perm ? idx_perm : idx_seq;
    const char* tag = perm ? "permuted rows" : "contiguous rows";
    printf("device source, 2 thr/row (%s)     %6.2f us\n", tag, timeit([&] { k_pair<<<8, 256>>>(xd, idx, out, n); }));
    printf("host: 2 thr/row, 2 x ld.cv.16B (%s)   %6.2f us\n", tag, timeit([&] { k_pair<<<8, 256>>>(xh, idx, out, n); }));
    printf("host: 4 lanes/row ld.cv (%s)      %6.2f us\n", tag, timeit([&] { k_quad<<<16, 256>>>(xh, idx, out, n, 0); }));
    printf("host: 4 lanes/row ld.nc (%s)      %6.2f us\n", tag, timeit([&] { k_quad<<<16, 256>>>(xh, idx, out, n, 1); }));
    printf("host: 4 lanes/row ld.nc, 32 CTAs (%s) %6.2f us\n", tag, timeit([&] { k_quad<<<32, 128>>>(xh, idx, out, n, 1); }));
    printf("host: cp.async.bulk 64 B/row (%s) %6.2f us\n", tag, timeit([&] { k_bulk_rows<<<8, 128, 8192>>>(xh, idx, out, n); }));
  }
  printf("host: cp.async.bulk 8 KB/CTA (contiguous)  %6.2f us\n", timeit([&] { k_bulk_block<<<8, 128, 8192>>>(xh, out, n); }));
  float* chk = new float[n * 16];
  cudaMemcpy(chk, out, n * 64, cudaMemcpyDeviceToHost);
  printf("check %g %g  %s\n", chk[17], chk[n * 16 - 1], cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
