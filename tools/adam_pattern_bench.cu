// Micro-benchmark: how fast can an Adam-style update (read W, m, v; write W, m, v + bf16 shadow = 26 B/element)
// stream through HBM for different traversal patterns and occupancies?  Used to size the optimizer epilogue of
// the wgrad GEMM (csrc/fedb_gemm_tma*.cu).   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/apb tools/adam_pattern_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

__device__ __forceinline__ void upd(float4& w, float4& m, float4& v, float g) {
  m.x = 0.9f * m.x + 0.1f * g; m.y = 0.9f * m.y + 0.1f * g; m.z = 0.9f * m.z + 0.1f * g; m.w = 0.9f * m.w + 0.1f * g;
  v.x = 0.999f * v.x + 0.001f * g * g; v.y = 0.999f * v.y + 0.001f * g * g; v.z = 0.999f * v.z + 0.001f * g * g; v.w = 0.999f * v.w + 0.001f * g * g;
  w.x -= 1e-3f * m.x * rsqrtf(v.x + 1e-8f); w.y -= 1e-3f * m.y * rsqrtf(v.y + 1e-8f); w.z -= 1e-3f * m.z * rsqrtf(v.z + 1e-8f); w.w -= 1e-3f * m.w * rsqrtf(v.w + 1e-8f);
}
__device__ __forceinline__ void one(float* W, float* M, float* V, __nv_bfloat16* S, size_t off) {
  float4 w = *(float4*)(W + off), m = *(float4*)(M + off), v = *(float4*)(V + off);
  upd(w, m, v, 0.5f);
  *(float4*)(W + off) = w; *(float4*)(M + off) = m; *(float4*)(V + off) = v;
  __nv_bfloat162 a = __floats2bfloat162_rn(w.x, w.y), b = __floats2bfloat162_rn(w.z, w.w);
  *(uint2*)(S + off) = make_uint2(*(uint32_t*)&a, *(uint32_t*)&b);
}
// linear: consecutive warps take consecutive 512-byte chunks
__global__ void k_linear(float* W, float* M, float* V, __nv_bfloat16* S, size_t n4) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) one(W, M, V, S, 4 * i);
}
// tiled: a CTA owns TR rows x TC columns of a [rows][ld] matrix; warps take rows, UR rows in flight per warp
template <int TR, int TC, int UR>
__global__ void k_tiled(float* W, float* M, float* V, __nv_bfloat16* S, int ld, int tiles_n, int tiles_m, int xfast_n) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int t = blockIdx.x;
  const int tn = xfast_n ? t % tiles_n : (t / tiles_m) % tiles_n, tm = xfast_n ? (t / tiles_n) % tiles_m : t % tiles_m;
  const size_t base = (size_t)(t / (tiles_n * tiles_m)) * ld * (size_t)(tiles_m * TR) + (size_t)tm * TR * ld + (size_t)tn * TC;
  for (int c = 0; c < TC; c += 128)
    for (int r0 = warp; r0 < TR; r0 += nw * UR) {
      float4 w[UR], m[UR], v[UR];
#pragma unroll
      for (int u = 0; u < UR; ++u) {
        const int r = r0 + u * nw;
        if (r < TR) { size_t off = base + (size_t)r * ld + c + 4 * lane; w[u] = *(float4*)(W + off); m[u] = *(float4*)(M + off); v[u] = *(float4*)(V + off); }
      }
#pragma unroll
      for (int u = 0; u < UR; ++u) {
        const int r = r0 + u * nw;
        if (r < TR) {
          size_t off = base + (size_t)r * ld + c + 4 * lane;
          upd(w[u], m[u], v[u], 0.5f);
          *(float4*)(W + off) = w[u]; *(float4*)(M + off) = m[u]; *(float4*)(V + off) = v[u];
          __nv_bfloat162 a = __floats2bfloat162_rn(w[u].x, w[u].y), b = __floats2bfloat162_rn(w[u].z, w[u].w);
          *(uint2*)(S + off) = make_uint2(*(uint32_t*)&a, *(uint32_t*)&b);
        }
      }
    }
}

// ring: loads go global -> shared through cp.async.bulk (per-warp ring of D rows x 3 arrays x 512 B), so bytes in
// flight do not depend on registers, warps or L1 lines; stores stay ordinary coalesced st.global.
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int TR, int TC, int D>
__global__ void k_ring(float* W, float* M, float* V, __nv_bfloat16* S, int ld, int tiles_n, int tiles_m) {
  extern __shared__ __align__(128) uint8_t sm[];
  __shared__ uint64_t bars[8][D];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int t = blockIdx.x, tn = (t / tiles_m) % tiles_n, tm = t % tiles_m;
  const size_t base = (size_t)(t / (tiles_n * tiles_m)) * ld * (size_t)(tiles_m * TR) + (size_t)tm * TR * ld + (size_t)tn * TC;
  float* ring = (float*)(sm + 65536) + warp * D * 384;      // 64 KB stands in for the accumulator tile
  if (lane == 0) for (int d = 0; d < D; ++d) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bars[warp][d])));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();
  const int rows_per_warp = (TR - warp + nw - 1) / nw, total = rows_per_warp * (TC / 128);
  auto issue = [&](int i) {
    const int c = (i / rows_per_warp) * 128, r = warp + (i % rows_per_warp) * nw, d = i % D;
    const size_t off = base + (size_t)r * ld + c;
    const uint32_t bar = s32(&bars[warp][d]), dst = s32(ring + d * 384);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], 1536;" ::"r"(bar) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 512, [%2];" ::"r"(dst), "l"(W + off), "r"(bar) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 512, [%2];" ::"r"(dst + 512), "l"(M + off), "r"(bar) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 512, [%2];" ::"r"(dst + 1024), "l"(V + off), "r"(bar) : "memory");
  };
  if (lane == 0) for (int i = 0; i < D && i < total; ++i) issue(i);
  for (int i = 0; i < total; ++i) {
    const int c = (i / rows_per_warp) * 128, r = warp + (i % rows_per_warp) * nw, d = i % D;
    const uint32_t bar = s32(&bars[warp][d]), par = (i / D) & 1;
    uint32_t done;
    do {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(done) : "r"(bar), "r"(par) : "memory");
    } while (!done);
    const float* slot = ring + d * 384;
    float4 w = *(const float4*)(slot + 4 * lane), m = *(const float4*)(slot + 128 + 4 * lane), v = *(const float4*)(slot + 256 + 4 * lane);
    __syncwarp();
    if (lane == 0 && i + D < total) { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); issue(i + D); }
    upd(w, m, v, 0.5f);
    const size_t off = base + (size_t)r * ld + c + 4 * lane;
    *(float4*)(W + off) = w; *(float4*)(M + off) = m; *(float4*)(V + off) = v;
    __nv_bfloat162 a = __floats2bfloat162_rn(w.x, w.y), b = __floats2bfloat162_rn(w.z, w.w);
    *(uint2*)(S + off) = make_uint2(*(uint32_t*)&a, *(uint32_t*)&b);
  }
}

template <typename F>
float timeit(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); for (int i = 0; i < 5; ++i) f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms / 5;
}

int main() {
  const int G = 16, R = 4096, LD = 4096;
  const size_t n = (size_t)G * R * LD;
  float *W, *M, *V; __nv_bfloat16* S;
  cudaMalloc(&W, n * 4); cudaMalloc(&M, n * 4); cudaMalloc(&V, n * 4); cudaMalloc(&S, n * 2);
  cudaMemset(W, 0, n * 4); cudaMemset(M, 0, n * 4); cudaMemset(V, 0, n * 4);
  const double bytes = (double)n * 26;
  auto rep = [&](const char* name, float ms) { printf("%-44s %.3f ms  %.0f GB/s\n", name, ms, bytes / ms / 1e6); };
  rep("linear 1024thr x 148*2", timeit([&] { k_linear<<<148 * 2, 1024>>>(W, M, V, S, n / 4); }));
  rep("linear 192thr x 148*2 (12 warps/SM)", timeit([&] { k_linear<<<148 * 2, 192>>>(W, M, V, S, n / 4); }));
  rep("linear 256thr x 148*8", timeit([&] { k_linear<<<148 * 8, 256>>>(W, M, V, S, n / 4); }));
#define TILED(TR, TC, UR, THR, XN)                                                                                              \
  {                                                                                                                              \
    int tn = LD / TC, tm = R / TR;                                                                                               \
    char nm[96]; snprintf(nm, 96, "tile %dx%d UR=%d thr=%d %s-fast", TR, TC, UR, THR, XN ? "N" : "M");                           \
    rep(nm, timeit([&] { k_tiled<TR, TC, UR><<<G * tn * tm, THR>>>(W, M, V, S, LD, tn, tm, XN); }));                             \
  }
  TILED(128, 128, 4, 192, 1) TILED(128, 128, 4, 192, 0) TILED(128, 256, 4, 192, 0) TILED(128, 256, 4, 192, 1)
  TILED(128, 128, 8, 192, 1) TILED(128, 128, 4, 384, 1) TILED(128, 128, 4, 768, 1) TILED(128, 256, 8, 192, 0)
  TILED(32, 1024, 4, 192, 1) TILED(128, 128, 2, 192, 1) TILED(128, 128, 1, 192, 1) TILED(128, 256, 2, 384, 0)
  // same tiles but with the launch limited to 2 CTAs/SM by dynamic smem (as in the GEMM)
  {
    auto k = k_tiled<128, 256, 4>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    rep("tile 128x256 UR=4 thr=192 M-fast, 2 CTA/SM", timeit([&] { k<<<G * 16 * 32, 192, 100 * 1024>>>(W, M, V, S, LD, 16, 32, 0); }));
    auto k2 = k_tiled<128, 128, 4>;
    cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    rep("tile 128x128 UR=4 thr=192 N-fast, 2 CTA/SM", timeit([&] { k2<<<G * 32 * 32, 192, 100 * 1024>>>(W, M, V, S, LD, 32, 32, 1); }));
    auto k3 = k_tiled<128, 256, 8>;
    cudaFuncSetAttribute(k3, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    rep("tile 128x256 UR=8 thr=192 M-fast, 2 CTA/SM", timeit([&] { k3<<<G * 16 * 32, 192, 100 * 1024>>>(W, M, V, S, LD, 16, 32, 0); }));
    auto k4 = k_tiled<128, 256, 4>;
    rep("tile 128x256 UR=4 thr=384 M-fast, 2 CTA/SM", timeit([&] { k4<<<G * 16 * 32, 384, 100 * 1024>>>(W, M, V, S, LD, 16, 32, 0); }));
    rep("tile 128x256 UR=4 thr=768 M-fast, 2 CTA/SM", timeit([&] { k4<<<G * 16 * 32, 768, 100 * 1024>>>(W, M, V, S, LD, 16, 32, 0); }));
  }
#define RING(D)                                                                                                   \
  {                                                                                                                \
    auto k = k_ring<128, 256, D>;                                                                                  \
    const int smem = 65536 + 6 * D * 1536;                                                                         \
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024);                              \
    char nm[96]; snprintf(nm, 96, "ring D=%d 128x256 thr=192 M-fast, 2 CTA/SM", D);                                \
    rep(nm, timeit([&] { k<<<G * 16 * 32, 192, smem < 100 * 1024 ? 100 * 1024 : smem>>>(W, M, V, S, LD, 16, 32); })); \
  }
  RING(2) RING(3) RING(4) RING(5)
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
