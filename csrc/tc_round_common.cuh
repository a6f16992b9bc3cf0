// Shared by the two fused round kernels (csrc/fedb_tc.cu: bf16 operands, 128-row tiles; csrc/fedb_tcx.cu: split fp16
// operands, 64-row tiles): the argument block and the two aggregation tails that turn "every client has trained and
// evaluated" into "the round is over" — single-GPU FedAvg by the last cluster to finish (round_tail), or the cross-GPU
// exchange over NVLink peer memory by the client's own cluster (p2p_tail).
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"
#include "p2p_common.cuh"

namespace fedb {

constexpr int TC_THREADS = 256;   // threads per CTA of the bf16 round kernel (the tails below take the count from blockDim.x)

struct TcArgs {
  const FedbGroup* groups;
  const float* X;
  const int32_t* y;
  // optional: this step's inputs in (pinned, device-visible) host memory, same layout as X / y.  The kernel then performs the
  // host-to-device copy itself: rows are read over the host link and written through to X / y on their way to the X panel.
  const float* x_host;
  const int32_t* y_host;
  const int32_t* idx;
  float* params;
  float* m;
  float* v;
  int32_t* step;
  const int32_t* active;
  const FedbFedState* fed_state;
  float* loss_out;
  int32_t* conf;
  float* scratch;   // per-group gradient records (FedbGroup::scratch_off; >= cluster_size records each)
  int ldx, batch, n_steps, do_eval;
  long long* prof;  // optional: clock64 stamps of CTA 0 / thread 0 (debug, tools/tc_timeline.py)
  int prof_stride;  // 0: as above; n > 0: globaltimer nanoseconds, CTA r of the first cluster writes prof[r * n + i]
  // ---- fused round tail (tail_mode 1): the last cluster to finish aggregates every federation ----
  int tail_mode, n_groups, n_feds, n_conf;
  const FedbFed* feds;
  FedbFedState* fed_state_rw;
  float* global;
  float* hist_local;
  float* hist_global;
  unsigned int* tail_count;   // wraps to 0 by itself (atomicInc)
  // ---- fused cross-GPU aggregation (tail_mode 2): one federation, one resident client per rank ----
  const FedbP2PRank* p2p;
  FedbFedState* state_mirror;   // optional: mapped pinned host copy of the federation state, written by the tail
};

// What a round kernel can hand its tail in shared memory of the cluster's LAST CTA (the one that scores the round):
// the federation record and state fetched at kernel start (they change only when the tail itself writes them), and — when the
// grid is a single cluster — the confusion counts of the local evaluation, added up over distributed shared memory, so the
// tail starts computing without a chain of L2 round trips.
struct TailPre {
  FedbFed F;
  FedbFedState S;
  int32_t cnt[FEDB_MAX_CLASSES * FEDB_MAX_CLASSES];
  int32_t conf_all[8 * FEDB_MAX_CLASSES * FEDB_MAX_CLASSES];   // cross-GPU tail: every rank's counts (one client per rank, <= 8 ranks)
  int32_t have_fs, have_cnt;
};
// (whole warp) fetch F / S of the only federation into *pre
__device__ __forceinline__ void tail_prefetch(const TcArgs& a, TailPre* pre) {
  const int lane = threadIdx.x & 31;
  const FedbFed* F = nullptr;
  const FedbFedState* S = nullptr;
  if (a.tail_mode == 1 && a.n_feds == 1) { F = a.feds; S = a.fed_state_rw; }
  if (a.tail_mode == 2) { F = a.p2p->feds_global; S = a.p2p->fed_state; }
  if (!F) return;
  for (int i = lane; i < (int)(sizeof(FedbFed) / 4); i += 32) reinterpret_cast<int32_t*>(&pre->F)[i] = reinterpret_cast<const int32_t*>(F)[i];
  for (int i = lane; i < (int)(sizeof(FedbFedState) / 4); i += 32) reinterpret_cast<int32_t*>(&pre->S)[i] = reinterpret_cast<const int32_t*>(S)[i];
  __syncwarp();
  if (lane == 0) pre->have_fs = 1;
}
__device__ __forceinline__ void st_relaxed_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ void tc_stamp(const TcArgs& a, int crank, int i) {
  long long t;
  if (a.prof_stride) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  else t = clock64();
  a.prof[crank * a.prof_stride + i] = t;
}

// One copy of the metrics / early-stop code for both tails (it runs once per round on one warp; inlined twice it is ~4000
// instructions of mostly double-precision code that a cache-cold round has to fetch).
// S_out (optional): S is a shared-memory copy fetched early; the updated state is stored to *S_out as well.
static __device__ __noinline__ void round_metrics_ni(const FedbFed* F, FedbFedState* S, int f, int n_groups, const int32_t* conf,
                                                     float* hist_local, float* hist_global, int* pool, FedbFedState* host_mirror,
                                                     void* host_result, FedbFedState* S_out = nullptr, long long* stamps = nullptr) {
  round_metrics_warp(*F, *S, f, n_groups, conf, hist_local, hist_global, pool, true, host_mirror, host_result, stamps);
  if (S_out && (threadIdx.x & 31) == 0) *S_out = *S;
}

// K1 + K5 for small problems, executed by the last cluster to finish: FedAvg of every federation in
// client order with the reference's arithmetic (C:108-116), write-back (C:119-120), metrics and early
// stop (C:165-192), then the next round's begin bookkeeping — same results as k_fedavg.
constexpr int TAIL_MAX_FEDS = 256, TAIL_MAX_CLIENTS = 32;

// skip_avg: the caller's optimizer phase has already written the FedAvg result into the global arena (the single-client
// federation of the split-operand kernel); what is left is the write-back into the client's slab (C:119-120), metrics, early
// stop and bookkeeping.
static __device__ void round_tail(const TcArgs& a, int crank, int csize, const unsigned char* s_stop, bool skip_avg = false,
                                  TailPre* pre = nullptr) {
  const int tid = threadIdx.x, NT = (int)blockDim.x;
  __shared__ long long t_off[TAIL_MAX_CLIENTS];
  __shared__ float t_num[TAIL_MAX_CLIENTS], t_den[TAIL_MAX_CLIENTS];
  // the last CTA of the cluster scores the round (metrics need only the confusion counts) while the
  // others average; with a single CTA it does both
  const int mrank = csize - 1, workers = csize > 1 ? csize - 1 : 1;
  const int nthr = workers * NT, gt = crank * NT + tid;
  if (skip_avg && (csize == 1 || crank != mrank)) {
    for (int f = 0; f < a.n_feds; ++f) {
      if (s_stop[f]) continue;
      const FedbFed F = a.feds[f];
      const float4* src = reinterpret_cast<const float4*>(a.global + F.global_off);
      for (int k = 0; k < F.n_clients; ++k) {
        float4* dst = reinterpret_cast<float4*>(a.params + a.groups[F.first_group + k].slab_off);
        for (int e4 = gt; e4 < (F.slab_size >> 2); e4 += nthr) dst[e4] = __ldcg(src + e4);
      }
    }
  }
  for (int f = 0; f < a.n_feds && !skip_avg && (csize == 1 || crank != mrank); ++f) {
    if (s_stop[f]) continue;
    const FedbFed F = a.feds[f];
    __syncthreads();
    if (tid < F.n_clients) {
      const FedbGroup* G = a.groups + F.first_group + tid;
      t_off[tid] = G->slab_off; t_num[tid] = G->agg_num; t_den[tid] = G->agg_den;
    }
    __syncthreads();
    const int n4 = F.slab_size >> 2;
    for (int base = gt; base < n4; base += 4 * nthr) {
      float4 acc[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
      for (int k = 0; k < F.n_clients; ++k) {
        const float num = t_num[k], den = t_den[k];
        const float4* src = reinterpret_cast<const float4*>(a.params + t_off[k]);
        float4 w[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (base + u * nthr < n4) w[u] = __ldcg(src + base + u * nthr);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (base + u * nthr < n4) {
            acc[u].x += __fdiv_rn(__fmul_rn(w[u].x, num), den);
            acc[u].y += __fdiv_rn(__fmul_rn(w[u].y, num), den);
            acc[u].z += __fdiv_rn(__fmul_rn(w[u].z, num), den);
            acc[u].w += __fdiv_rn(__fmul_rn(w[u].w, num), den);
          }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int e4 = base + u * nthr;
        if (e4 < n4) {
          reinterpret_cast<float4*>(a.global + F.global_off)[e4] = acc[u];
          for (int k = 0; k < F.n_clients; ++k) reinterpret_cast<float4*>(a.params + t_off[k])[e4] = acc[u];
        }
      }
    }
  }
  if (a.prof && crank == 0 && tid == 0) tc_stamp(a, crank, 20);
  if (crank == mrank) {
    __shared__ int pool[FEDB_MAX_CLASSES * FEDB_MAX_CLASSES];
    if (tid < 32)
      for (int f = 0; f < a.n_feds; ++f) {
        if (s_stop[f]) continue;
        const bool fs = pre && pre->have_fs && a.n_feds == 1;
        round_metrics_ni(fs ? &pre->F : a.feds + f, fs ? &pre->S : a.fed_state_rw + f, f, a.n_groups,
                           (pre && pre->have_cnt) ? pre->cnt : a.conf, a.hist_local, a.hist_global, pool,
                           a.state_mirror ? a.state_mirror + f : nullptr,
                           a.state_mirror ? reinterpret_cast<char*>(a.state_mirror) + ((sizeof(FedbFedState) * a.n_feds + 15) & ~size_t(15)) + 64 * f : nullptr,
                           fs ? a.fed_state_rw + f : nullptr, (a.prof && a.prof_stride) ? a.prof + crank * a.prof_stride + 22 : nullptr);
      }
    __syncthreads();
    if (a.prof && tid == 0) tc_stamp(a, crank, 21);
    for (int i = tid; i < a.n_conf; i += NT) a.conf[i] = 0;
  }
}

// Cross-GPU FedAvg fused into the round kernel (one federation, one resident client per rank): the protocol of k_fedavg_p2p
// (csrc/fedb_p2p.cu) executed by the client's own cluster.  The partial (w * n_k) / N was published by the optimizer phase
// and flag A raised BEFORE the local evaluation, so that NVLink flight hides behind it.  After the evaluation the last CTA
// pushes this rank's confusion counts into every peer's buffer as SELF-VALIDATING words — the count in the low 24 bits, the
// low 8 bits of the round's sequence number above them — so the receiver polls the words themselves: one one-way NVLink
// flight on the critical path, no flag, no system-scope fence (the area is double-buffered by round parity; a shard has
// fewer than 2^24 rows).  Meanwhile every other CTA sums the R partials in rank order with peer loads and writes the global
// model + the client's slab; the last CTA scores the round from all ranks' counts and does the next round's bookkeeping.
static __device__ void p2p_tail(const TcArgs& a, TailPre* pre = nullptr) {
  cooperative_groups::cluster_group cluster = cooperative_groups::this_cluster();
  const int CS = (int)cluster.num_blocks(), crank = (int)cluster.block_rank(), tid = threadIdx.x, NT = (int)blockDim.x;
  __shared__ FedbP2PRank R;
  for (int i = tid; i < (int)(sizeof(FedbP2PRank) / 4); i += NT)
    reinterpret_cast<int32_t*>(&R)[i] = reinterpret_cast<const int32_t*>(a.p2p)[i];
  __shared__ int s_timeout;
  if (tid == 0) s_timeout = 0;
  __syncthreads();
  const uint64_t seq = *reinterpret_cast<volatile uint64_t*>(R.seq);
  const int par = (int)(seq & 1);
  const uint32_t tag = (uint32_t)((seq + 1) & 0xFF) << 24;
  const bool stopped = R.fed_state[0].stop != 0;   // as it stands for THIS round
  __threadfence();
  cluster.sync();                                   // all CTAs: evaluation counts are complete; seq / stop have been read
  if (stopped) return;
  const long long t0 = clock64();
  constexpr int C2 = FEDB_MAX_CLASSES * FEDB_MAX_CLASSES;
  const int64_t conf_stride = (int64_t)R.conf_slots * C2;
  uint64_t* my_flags = reinterpret_cast<uint64_t*>(R.sym[R.rank] + R.sym_flag_off);
  const int mrank = CS - 1, workers = CS > 1 ? CS - 1 : 1;
  const bool smem_cnt = pre && pre->have_cnt;
  if (crank == mrank) {   // PUSH this rank's tagged confusion counts into every peer's area (parity, my rank)
    const int nconf = R.n_groups_local * C2;
    for (int i = tid; i < nconf * R.world; i += NT) {
      const int qq = i / nconf, w = i - qq * nconf;
      const uint32_t c = (uint32_t)(smem_cnt ? pre->cnt[w] : __ldcg(a.conf + w));
      st_relaxed_sys_u32(reinterpret_cast<uint32_t*>(R.sym[qq] + R.sym_conf_off) + (par * 8 + R.rank) * conf_stride + w, (c & 0xFFFFFFu) | tag);
    }
  }
  if (CS == 1 || crank != mrank) {   // ---- aggregation: wait for every peer's partial (flag A), sum in rank order
    if (tid < R.world)
      while (ld_acquire_sys_u64(my_flags + tid) < seq + 1)
        if (clock64() - t0 > R.timeout_cycles) { s_timeout = 1; break; }
    __syncthreads();
    const FedbFed F = R.feds[0];
    const FedbGroup* G = R.groups + F.first_group;
    float4* slab = reinterpret_cast<float4*>(R.params + G->slab_off);
    float4* glob = reinterpret_cast<float4*>(R.global + F.global_off);
    const int n4 = F.slab_size >> 2;
    for (int e4 = crank * NT + tid; e4 < n4; e4 += workers * NT) {
      float4 pr[8];
#pragma unroll
      for (int qq = 0; qq < 8; ++qq)
        if (qq < R.world)
          pr[qq] = ld_relaxed_sys_f4(reinterpret_cast<const float*>(R.sym[qq] + R.sym_partial_off) + (int64_t)par * R.global_total + F.global_off + 4 * e4);
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int qq = 0; qq < 8; ++qq)
        if (qq < R.world) { acc.x += pr[qq].x; acc.y += pr[qq].y; acc.z += pr[qq].z; acc.w += pr[qq].w; }
      glob[e4] = acc;
      if (R.write_back) slab[e4] = acc;
    }
  }
  if (crank == mrank) {              // ---- metrics from every rank's counts, early stop, next round's begin
    __shared__ int pool[C2];
    int* conf_all = (pre && R.n_groups_total <= 8) ? pre->conf_all : R.conf_all;
    for (int i = tid; i < R.n_groups_total * C2; i += NT) {
      const int slot = i / C2, w = i - slot * C2;
      const int qq = R.conf_src_rank[slot], ls = R.conf_src_slot[slot];
      // pushed into THIS rank's buffer by their owners: poll the local word until it carries this round's tag
      const int32_t* cq = reinterpret_cast<const int32_t*>(R.sym[R.rank] + R.sym_conf_off) + (par * 8 + qq) * conf_stride + ls * C2 + w;
      uint32_t v = (uint32_t)ld_relaxed_sys_i32(cq);
      while ((v & 0xFF000000u) != tag) {
        if (clock64() - t0 > R.timeout_cycles) { s_timeout = 1; break; }
        v = (uint32_t)ld_relaxed_sys_i32(cq);
      }
      conf_all[i] = (int32_t)(v & 0xFFFFFFu);
    }
    __syncthreads();
    if (tid < 32 && R.do_metrics) {
      const bool fs = pre && pre->have_fs;
      round_metrics_ni(fs ? &pre->F : R.feds_global, fs ? &pre->S : R.fed_state, 0, R.n_groups_total, conf_all, R.hist_local, R.hist_global,
                       pool, a.state_mirror, a.state_mirror ? reinterpret_cast<char*>(a.state_mirror) + ((sizeof(FedbFedState) + 15) & ~size_t(15)) : nullptr,
                       fs ? R.fed_state : nullptr);
    }
    __syncthreads();
    if (!smem_cnt)
      for (int i = tid; i < R.n_groups_local * C2; i += NT) a.conf[i] = 0;
    if (tid == 0) *reinterpret_cast<volatile uint64_t*>(R.seq) = seq + 1;
  }
  __syncthreads();
  if (s_timeout && tid == 0) *R.err = FEDB_E_TIMEOUT;
}

// The same cross-GPU round for ANY number of resident clients per rank (tail_mode 3), executed by the last cluster of the
// rank to finish: it first folds the resident clients' weighted slabs into this rank's partial, sum_k (w_k n_k) / N in
// client order (C:110-115) — what k_fedavg_p2p's first phase does — publishes it (flag A; nothing is left to hide that
// flight behind), pushes the tagged confusion counts of all resident clients, then sums the ranks' partials in rank order
// over NVLink peer loads, writes the global model and every resident client's slab, and scores the round.  One launch per
// rank per round; bit-identical to the local-round + k_fedavg_p2p pair.
static __device__ void p2p_tail_many(const TcArgs& a, TailPre* pre = nullptr) {
  cooperative_groups::cluster_group cluster = cooperative_groups::this_cluster();
  const int CS = (int)cluster.num_blocks(), crank = (int)cluster.block_rank(), tid = threadIdx.x, NT = (int)blockDim.x;
  __shared__ FedbP2PRank R;
  for (int i = tid; i < (int)(sizeof(FedbP2PRank) / 4); i += NT)
    reinterpret_cast<int32_t*>(&R)[i] = reinterpret_cast<const int32_t*>(a.p2p)[i];
  __shared__ int s_timeout;
  __shared__ unsigned char m_stop[TAIL_MAX_FEDS];
  if (tid == 0) s_timeout = 0;
  __syncthreads();
  const uint64_t seq = *reinterpret_cast<volatile uint64_t*>(R.seq);
  const int par = (int)(seq & 1);
  const uint32_t tag = (uint32_t)((seq + 1) & 0xFF) << 24;
  for (int f = tid; f < R.n_feds; f += NT) m_stop[f] = R.fed_state[f].stop ? 1 : 0;   // as it stands for THIS round
  __syncthreads();
  constexpr int C2 = FEDB_MAX_CLASSES * FEDB_MAX_CLASSES;
  const int64_t conf_stride = (int64_t)R.conf_slots * C2;
  const int nthr = CS * NT, gt = crank * NT + tid;
  float* my_partial = reinterpret_cast<float*>(R.sym[R.rank] + R.sym_partial_off) + (int64_t)par * R.global_total;
  // ---- 1. this rank's partial: the resident clients' terms in client order
  for (int f = 0; f < R.n_feds; ++f) {
    if (m_stop[f]) continue;
    const FedbFed F = R.feds[f];
    const int n4 = F.slab_size >> 2;
    for (int e4 = gt; e4 < n4; e4 += nthr) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = 0; k < F.n_clients; ++k) {
        const FedbGroup* G = R.groups + F.first_group + k;
        const float num = G->agg_num, den = G->agg_den;
        const float4 w = __ldcg(reinterpret_cast<const float4*>(R.params + G->slab_off) + e4);
        acc.x += __fdiv_rn(__fmul_rn(w.x, num), den); acc.y += __fdiv_rn(__fmul_rn(w.y, num), den);
        acc.z += __fdiv_rn(__fmul_rn(w.z, num), den); acc.w += __fdiv_rn(__fmul_rn(w.w, num), den);
      }
      reinterpret_cast<float4*>(my_partial + F.global_off)[e4] = acc;
    }
  }
  __threadfence();
  cluster.sync();
  const long long t0 = clock64();
  uint64_t* my_flags = reinterpret_cast<uint64_t*>(R.sym[R.rank] + R.sym_flag_off);
  if (crank == 0 && tid < R.world) {      // flag A: the cluster barrier ordered every CTA's partial stores before this thread
    __threadfence_system();
    st_release_sys_u64(reinterpret_cast<uint64_t*>(R.sym[tid] + R.sym_flag_off) + R.rank, seq + 1);
  }
  const int mrank = CS - 1, workers = CS > 1 ? CS - 1 : 1;
  if (crank == mrank) {   // ---- 2. tagged confusion counts of every resident client into every peer's area
    const int nconf = R.n_groups_local * C2;
    for (int i = tid; i < nconf * R.world; i += NT) {
      const int qq = i / nconf, w = i - qq * nconf;
      st_relaxed_sys_u32(reinterpret_cast<uint32_t*>(R.sym[qq] + R.sym_conf_off) + (par * 8 + R.rank) * conf_stride + w,
                         ((uint32_t)__ldcg(a.conf + w) & 0xFFFFFFu) | tag);
    }
  }
  if (CS == 1 || crank != mrank) {   // ---- 3. sum the ranks' partials in rank order; global model + write-back
    if (tid < R.world)
      while (ld_acquire_sys_u64(my_flags + tid) < seq + 1)
        if (clock64() - t0 > R.timeout_cycles) { s_timeout = 1; break; }
    __syncthreads();
    const int wt = (CS == 1 ? 0 : crank) * NT + tid, wn = workers * NT;
    for (int f = 0; f < R.n_feds; ++f) {
      if (m_stop[f]) continue;
      const FedbFed F = R.feds[f];
      float4* glob = reinterpret_cast<float4*>(R.global + F.global_off);
      const int n4 = F.slab_size >> 2;
      for (int e4 = wt; e4 < n4; e4 += wn) {
        float4 pr[8];
#pragma unroll
        for (int qq = 0; qq < 8; ++qq)
          if (qq < R.world)
            pr[qq] = ld_relaxed_sys_f4(reinterpret_cast<const float*>(R.sym[qq] + R.sym_partial_off) + (int64_t)par * R.global_total + F.global_off + 4 * e4);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int qq = 0; qq < 8; ++qq)
          if (qq < R.world) { acc.x += pr[qq].x; acc.y += pr[qq].y; acc.z += pr[qq].z; acc.w += pr[qq].w; }
        glob[e4] = acc;
        if (R.write_back)
          for (int k = 0; k < F.n_clients; ++k) reinterpret_cast<float4*>(R.params + R.groups[F.first_group + k].slab_off)[e4] = acc;
      }
    }
  }
  if (crank == mrank) {              // ---- 4. metrics from every rank's counts, early stop, next round's begin
    __shared__ int pool[C2];
    int* conf_all = (pre && R.n_groups_total <= 8) ? pre->conf_all : R.conf_all;
    for (int i = tid; i < R.n_groups_total * C2; i += NT) {
      const int slot = i / C2, w = i - slot * C2;
      const int qq = R.conf_src_rank[slot], ls = R.conf_src_slot[slot];
      const int32_t* cq = reinterpret_cast<const int32_t*>(R.sym[R.rank] + R.sym_conf_off) + (par * 8 + qq) * conf_stride + ls * C2 + w;
      uint32_t v = (uint32_t)ld_relaxed_sys_i32(cq);
      while ((v & 0xFF000000u) != tag) {
        if (clock64() - t0 > R.timeout_cycles) { s_timeout = 1; break; }
        v = (uint32_t)ld_relaxed_sys_i32(cq);
      }
      conf_all[i] = (int32_t)(v & 0xFFFFFFu);
    }
    __syncthreads();
    if (tid < 32 && R.do_metrics)
      for (int f = 0; f < R.n_feds; ++f) {
        if (m_stop[f]) continue;
        round_metrics_ni(R.feds_global + f, R.fed_state + f, f, R.n_groups_total, conf_all, R.hist_local, R.hist_global, pool,
                         a.state_mirror ? a.state_mirror + f : nullptr,
                         a.state_mirror ? reinterpret_cast<char*>(a.state_mirror) + ((sizeof(FedbFedState) * R.n_feds + 15) & ~size_t(15)) + 64 * f : nullptr);
      }
    __syncthreads();
    for (int i = tid; i < R.n_groups_local * C2; i += NT) a.conf[i] = 0;
    if (tid == 0) *reinterpret_cast<volatile uint64_t*>(R.seq) = seq + 1;
  }
  __syncthreads();
  if (s_timeout && tid == 0) *R.err = FEDB_E_TIMEOUT;
}

}  // namespace fedb
