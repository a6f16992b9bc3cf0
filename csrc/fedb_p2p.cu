// K1b: cross-GPU FedAvg inside ONE kernel per rank, over NVLink peer memory — no NCCL call and no
// separate elementwise kernel on the aggregation path (the reference's gather -> rank-0 mean ->
// bcast: C:101-120, S:108-122, H:24-46).
//
// Every rank owns a "symmetric" buffer (same layout on every GPU, mapped into every peer through
// CUDA IPC):   partial[2][P_total] fp32 | conf[2][slots][64] int32 | flags[8] u64
// One launch per round on every rank:
//   A  local reduction: partial[par] = sum over the rank's resident clients of (w_k * n_k) / N,
//      in client order (the reference's arithmetic, term by term); local confusion counts are
//      published next to it.
//   -  grid barrier, then flag words are stored (st.release.sys) into EVERY peer's buffer.
//   B  spin (ld.acquire.sys, bounded by a timeout) until every peer's flag for this round arrived.
//   C  every rank sums the R partials in rank order with peer loads (ld.relaxed.sys over NVLink;
//      identical order on every rank => identical global model everywhere, no broadcast needed),
//      writes the global slab and writes it back into each resident client's slab; block 0 also
//      gathers all clients' confusion counts from the peers and runs the metric + early-stop
//      bookkeeping redundantly-but-identically on every rank.
// Buffers are double-buffered by round parity, so one cross-GPU barrier per round suffices.
//
// The same kernel runs R emulated ranks on ONE GPU (gridDim.y = R, cooperative launch), which is
// how the single-GPU test-suite exercises the protocol (ranks must never be separate launches on
// one device: nothing guarantees they are co-resident).
#include <cooperative_groups.h>

#include <cuda_bf16.h>

#include "common.cuh"
#include "p2p_common.cuh"

namespace fedb {

constexpr int P2P_THREADS = 256;

__global__ void __launch_bounds__(P2P_THREADS, 4) k_fedavg_p2p(const FedbP2PRank* __restrict__ ranks) {
  const FedbP2PRank& R = ranks[blockIdx.y];
  const int tid = threadIdx.x;
  const uint64_t seq = *reinterpret_cast<volatile uint64_t*>(R.seq);
  const int par = (int)(seq & 1);
  const int64_t conf_stride = (int64_t)R.conf_slots * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES;
  const bool nvls = R.nvls_mc_partial != nullptr;       // partial sums live in multicast-bound memory: the switch adds them
  float* my_partial = (nvls ? R.nvls_uc_partial : reinterpret_cast<float*>(R.sym[R.rank] + R.sym_partial_off)) + (int64_t)par * R.global_total;
  int32_t* my_confx = reinterpret_cast<int32_t*>(R.sym[R.rank] + R.sym_conf_off) + (par * 8 + R.rank) * conf_stride;   // area (parity, source rank)
  uint64_t* my_flags = reinterpret_cast<uint64_t*>(R.sym[R.rank] + R.sym_flag_off);
  const int gthreads = gridDim.x * blockDim.x, gtid = blockIdx.x * blockDim.x + tid;
  // Stop flags as they stand for THIS round, taken before the grid barrier: block 0 raises them for
  // the next round at the end of this very kernel (begin_next), which must not affect slower blocks.
  __shared__ unsigned char s_stop[256];
  for (int f = tid; f < R.n_feds; f += blockDim.x) s_stop[f] = R.fed_state[f].stop ? 1 : 0;
  __syncthreads();

  // ---- A: local weighted partial sums -----------------------------------------------------------
  for (int f = 0; f < R.n_feds; ++f) {
    const FedbFed& F = R.feds[f];
    if (s_stop[f]) continue;
    for (int e4 = gtid; e4 * 4 < F.slab_size; e4 += gthreads) {
      const int e = e4 * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k0 = 0; k0 < F.n_clients; k0 += 4) {   // 4 clients' loads in flight; accumulation stays in client order
        float4 w[4];
        float num[4], den[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (k0 + u < F.n_clients) {
            const FedbGroup& G = R.groups[F.first_group + k0 + u];
            w[u] = __ldcs(reinterpret_cast<const float4*>(R.params + G.slab_off + e));
            num[u] = G.agg_num; den[u] = G.agg_den;
          }
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (k0 + u < F.n_clients) {
            acc.x += __fdiv_rn(__fmul_rn(w[u].x, num[u]), den[u]);
            acc.y += __fdiv_rn(__fmul_rn(w[u].y, num[u]), den[u]);
            acc.z += __fdiv_rn(__fmul_rn(w[u].z, num[u]), den[u]);
            acc.w += __fdiv_rn(__fmul_rn(w[u].w, num[u]), den[u]);
          }
      }
      *reinterpret_cast<float4*>(my_partial + F.global_off + e) = acc;
    }
  }
  if (blockIdx.x == 0)
    for (int i = tid; i < R.n_groups_local * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES; i += blockDim.x) my_confx[i] = R.conf[i];
  __threadfence_system();

  // ---- grid barrier among this rank's blocks, then publish the arrival flag to every peer --------
  __syncthreads();
  bool timed_out = false;
  const long long t0 = clock64();
  if (tid == 0) {
    atomicAdd(R.bar, 1u);
    const uint32_t target = (uint32_t)(((nvls ? 2 * seq : seq) + 1) * gridDim.x);   // NVLS launches pass two grid barriers
    while ((int32_t)(ld_acquire_gpu_u32(R.bar) - target) < 0)
      if (clock64() - t0 > R.timeout_cycles) { timed_out = true; break; }
  }
  __syncthreads();
  if (blockIdx.x == 0 && tid < R.world) {
    uint64_t* peer_flags = reinterpret_cast<uint64_t*>(R.sym[tid] + R.sym_flag_off);
    st_release_sys_u64(peer_flags + R.rank, seq + 1);
  }
  // ---- B: wait for every peer ---------------------------------------------------------------------
  if (tid < R.world) {
    while (ld_acquire_sys_u64(my_flags + tid) < seq + 1)
      if (clock64() - t0 > R.timeout_cycles) { timed_out = true; break; }
  }
  if (timed_out) *R.err = FEDB_E_TIMEOUT;
  __syncthreads();

  if (nvls) {
    // ---- C (NVLS): reduce-scatter in the switch.  This rank owns a contiguous share of every slab: one multimem.ld_reduce
    // returns the sum of all ranks' partials, one multimem.st puts it into every rank's result buffer (all replicas receive
    // the SAME bits, whatever order the switch adds in).  Then a second cross-GPU barrier (flag C) before anyone reads.
    for (int f = 0; f < R.n_feds; ++f) {
      const FedbFed& F = R.feds[f];
      if (s_stop[f]) continue;
      const int n4 = F.slab_size >> 2, per = (n4 + R.world - 1) / R.world, lo = R.rank * per, hi = min(n4, lo + per);
      for (int e4 = lo + gtid; e4 < hi; e4 += gthreads) {
        const float4 acc = multimem_ld_reduce_add_f4(R.nvls_mc_partial + (int64_t)par * R.global_total + F.global_off + 4 * e4);
        multimem_st_f4(R.nvls_mc_result + F.global_off + 4 * e4, acc);
      }
    }
    __threadfence_system();
    __syncthreads();
    if (tid == 0) {
      atomicAdd(R.bar, 1u);
      const uint32_t target = (uint32_t)((2 * seq + 2) * gridDim.x);
      while ((int32_t)(ld_acquire_gpu_u32(R.bar) - target) < 0)
        if (clock64() - t0 > R.timeout_cycles) { timed_out = true; break; }
    }
    __syncthreads();
    if (blockIdx.x == 0 && tid < R.world) st_release_sys_u64(reinterpret_cast<uint64_t*>(R.sym[tid] + R.sym_flag_off) + 16 + R.rank, seq + 1);
    if (tid < R.world) {
      while (ld_acquire_sys_u64(my_flags + 16 + tid) < seq + 1)
        if (clock64() - t0 > R.timeout_cycles) { timed_out = true; break; }
    }
    if (timed_out) *R.err = FEDB_E_TIMEOUT;
    __syncthreads();
  }
  // ---- C: sum the partials of all ranks in rank order (or fetch the switch's sum); write global + every resident client ----
  for (int f = 0; f < R.n_feds; ++f) {
    const FedbFed& F = R.feds[f];
    if (s_stop[f]) continue;
    for (int e4 = gtid; e4 * 4 < F.slab_size; e4 += gthreads) {
      const int e = e4 * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      if (nvls) {
        acc = ld_relaxed_sys_f4(R.nvls_uc_result + F.global_off + e);   // written by its owner rank through the multicast mapping
      } else {
        for (int q = 0; q < R.world; ++q) {
          const float* pq = reinterpret_cast<const float*>(R.sym[q] + R.sym_partial_off) + (int64_t)par * R.global_total;
          const float4 p = ld_relaxed_sys_f4(pq + F.global_off + e);
          acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
        }
      }
      *reinterpret_cast<float4*>(R.global + F.global_off + e) = acc;
      if (R.write_back) {
        for (int k = 0; k < F.n_clients; ++k)
          __stcs(reinterpret_cast<float4*>(R.params + R.groups[F.first_group + k].slab_off + e), acc);
        if (R.shadow) {   // wide path: every client's bf16 GEMM operands follow the averaged weights in the same pass
          const __nv_bfloat162 lo2 = __floats2bfloat162_rn(acc.x, acc.y), hi2 = __floats2bfloat162_rn(acc.z, acc.w);
          const uint2 pk = make_uint2(*reinterpret_cast<const uint32_t*>(&lo2), *reinterpret_cast<const uint32_t*>(&hi2));
          for (int k = 0; k < F.n_clients; ++k)
            __stcs(reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(R.shadow) + R.groups[F.first_group + k].slab_off + e), pk);
        }
      }
    }
  }
  // ---- metrics + early stop, identically on every rank (C:165-192) --------------------------------
  if (blockIdx.x == 0 && R.do_metrics) {
    for (int i = tid; i < R.n_groups_total * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES; i += blockDim.x) {
      const int slot = i / (FEDB_MAX_CLASSES * FEDB_MAX_CLASSES), w = i - slot * (FEDB_MAX_CLASSES * FEDB_MAX_CLASSES);
      const int q = R.conf_src_rank[slot], ls = R.conf_src_slot[slot];
      const int32_t* cq = reinterpret_cast<const int32_t*>(R.sym[q] + R.sym_conf_off) + (par * 8 + q) * conf_stride;
      R.conf_all[i] = ld_relaxed_sys_i32(cq + ls * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES + w);
    }
    __syncthreads();
    __shared__ int pool[FEDB_MAX_CLASSES * FEDB_MAX_CLASSES];
    if (tid < 32)
      for (int f = 0; f < R.n_feds; ++f)
        if (!s_stop[f])
          round_metrics_warp(R.feds_global[f], R.fed_state[f], f, R.n_groups_total, R.conf_all, R.hist_local, R.hist_global, pool,
                             R.begin_next != 0);
    __syncthreads();
  }
  if (blockIdx.x == 0 && R.begin_next) {   // next round's fedb_round_begin, fused (the counts above were already published)
    if (tid == 0 && !R.do_metrics)
      for (int f = 0; f < R.n_feds; ++f) begin_next_round(R.fed_state[f]);
    int32_t* conf_rw = const_cast<int32_t*>(R.conf);
    for (int i = tid; i < R.n_groups_local * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES; i += blockDim.x) conf_rw[i] = 0;
  }
  if (blockIdx.x == 0 && tid == 0) *reinterpret_cast<volatile uint64_t*>(R.seq) = seq + 1;
}

}  // namespace fedb

using namespace fedb;

extern "C" int32_t fedb_p2p_rank_size(void) { return (int32_t)sizeof(FedbP2PRank); }

extern "C" int fedb_fedavg_p2p(const FedbP2PRank* ranks_dev, int32_t n_ranks_here, int32_t blocks_x, void* stream) {
  FEDB_REQUIRE(n_ranks_here >= 1 && blocks_x >= 1, "bad grid");  // (n_feds <= 256 is checked by the caller: parallel/p2p.py)
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(blocks_x, n_ranks_here);
  cfg.blockDim = dim3(P2P_THREADS);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;  // blocks spin on each other: they must be co-resident
  at[0].val.cooperative = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  FEDB_CUDA_OK(cudaLaunchKernelEx(&cfg, k_fedavg_p2p, ranks_dev));
  return FEDB_OK;
}

// ---- symmetric memory plumbing (CUDA IPC; handles travel through torch.distributed) --------------
extern "C" int fedb_symm_alloc(int64_t bytes, void** out) {
  FEDB_REQUIRE(bytes > 0 && out, "bad arguments");
  FEDB_CUDA_OK(cudaMalloc(out, (size_t)bytes));
  FEDB_CUDA_OK(cudaMemset(*out, 0, (size_t)bytes));
  FEDB_CUDA_OK(cudaDeviceSynchronize());
  return FEDB_OK;
}
extern "C" int fedb_symm_free(void* p) {
  FEDB_CUDA_OK(cudaFree(p));
  return FEDB_OK;
}
extern "C" int fedb_ipc_get_handle(void* p, uint8_t* handle64) {
  cudaIpcMemHandle_t h;
  FEDB_CUDA_OK(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return FEDB_OK;
}
extern "C" int fedb_ipc_open_handle(const uint8_t* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  FEDB_CUDA_OK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return FEDB_OK;
}
extern "C" int fedb_ipc_close_handle(void* p) {
  FEDB_CUDA_OK(cudaIpcCloseMemHandle(p));
  return FEDB_OK;
}
