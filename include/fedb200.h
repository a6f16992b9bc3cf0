/*
 * fedb200 — C ABI of the native library (libfedb200.so), sm_100a only.
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on (as void*),
 * returns 0 on success or a negative FEDB_E_* code, and never synchronises the device unless
 * its comment says so.  fedb_last_error() returns a human-readable description of the last
 * failure on the calling thread.  Python binds these with ctypes
 * (federated-learning-with-mpi_b200/ops/_native.py); the same header serves C/C++ callers.
 *
 * Citations of the form "C:", "S:", "H:" refer to the reference scripts
 *   C: FL_CustomMLPCLassifierImplementation_Multiple_Rounds.py
 *   S: FL_SkLearn_MLPClassifier_Limitation.py
 *   H: hyperparameters_tuning.py
 * and "SK:" to sklearn/neural_network/_multilayer_perceptron.py, whose behaviour S:/H: rely on.
 */
#ifndef FEDB200_H
#define FEDB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FEDB_MAX_LAYERS 6  /* Linear layers per MLP */
#define FEDB_MAX_CLASSES 8 /* confusion matrix is FEDB_MAX_CLASSES^2 counters per slot */
#define FEDB_TILE_ROWS 32  /* rows per CTA in the SIMT kernels (one lane per row) */
#define FEDB_LOSS_SLOTS 32 /* floats appended to every partial-gradient record */

enum { FEDB_OK = 0, FEDB_E_CUDA = -1, FEDB_E_ARG = -2, FEDB_E_UNSUPPORTED = -3, FEDB_E_TIMEOUT = -4 };
enum { FEDB_HEAD_SOFTMAX_CE = 0, FEDB_HEAD_SIGMOID_LOGLOSS = 1 };
enum { FEDB_ACT_RELU = 0, FEDB_ACT_LOGISTIC = 1, FEDB_ACT_TANH = 2, FEDB_ACT_IDENTITY = 3 };
enum { FEDB_OPT_SGD = 0, FEDB_OPT_ADAM_TORCH = 1, FEDB_OPT_ADAM_SKLEARN = 2 };
enum { FEDB_METRIC_MEAN_OF_LOCAL = 0, FEDB_METRIC_POOLED = 1 };
enum { FEDB_INIT_ZEROS = 0, FEDB_INIT_TORCH_DEFAULT = 1, FEDB_INIT_SKLEARN_GLOROT = 2 };

/* One simulated client of one federation ("group").  The reference has one OS process per
 * client (C:30-46, S:11-15); here a client is a row of this table and all of them are
 * trained by the same grouped launches.  Groups of one federation are contiguous. */
typedef struct FedbGroup {
  int32_t n_layers;
  int32_t dims[FEDB_MAX_LAYERS + 1]; /* in, h1.., out (out = #classes, or 1 for sigmoid) */
  int32_t head;
  int32_t act;
  int32_t slab_size;   /* P: padded floats per client, multiple of 32 */
  int32_t fed;         /* federation index */
  int64_t slab_off;    /* float offset of this client's slab in the params/m/v arenas */
  int64_t idx_off;     /* offset of this client's shard in the row-index array */
  int32_t n_rows;      /* shard length */
  int32_t batch_size;  /* 0 = full batch (C:63-73); 200 = sklearn default (SK:691-692) */
  int32_t opt;
  int32_t nesterov;
  float lr, beta1, beta2, eps;
  float momentum;
  float alpha;         /* L2 penalty, sklearn semantics (SK:235, SK:353-361) */
  int32_t lr_step_size; /* StepLR (C:46); 0 = constant */
  float lr_gamma;
  float agg_num;       /* FedAvg weight numerator (n_k, C:110-115) or 1 (S:114-117) */
  float agg_den;       /* denominator (N, or number of clients) */
  int64_t scratch_off; /* float offset of this client's partial-gradient records */
  int32_t max_tiles;   /* records reserved at scratch_off, each slab_size+FEDB_LOSS_SLOTS floats */
  int32_t client;      /* client index inside the federation */
} FedbGroup;

/* One federation (one FedAvg problem; the grid search packs 90 of them, H:83-84). */
typedef struct FedbFed {
  int32_t first_group;
  int32_t n_clients;
  int32_t slab_size;
  int32_t n_classes;
  int64_t global_off;  /* float offset of the federation's global slab in the global arena */
  int32_t metric_mode; /* mean of local metrics (C:169) or pooled predictions (S:130-134) */
  int32_t patience;    /* C:122 termination_patience; 0 disables early stopping */
  float atol, rtol;    /* np.allclose(atol=tolerance) with its default rtol=1e-5 (C:181-184) */
  int32_t max_rounds;  /* capacity of the history buffers */
  int32_t pad_;
} FedbFed;

/* Device-resident, kernel-updated state of a federation. */
typedef struct FedbFedState {
  int32_t stop;      /* set at the top of the round after stop_pending was raised (C:132-136) */
  int32_t stop_pending;
  int32_t round;     /* rounds started so far */
  int32_t count;     /* C:124 termination_count */
  int32_t has_anchor;
  int32_t done;      /* rounds whose metrics have been recorded */
  double anchor[4];  /* C:125 prev_metric */
  double last[4];    /* last global metrics */
} FedbFedState;

/* Evaluate `n_rows` rows (indices at idx_off) with the parameters at slab_off, using the
 * shape of `group`; counts go to confusion slot `conf_slot`. */
typedef struct FedbEvalJob {
  int32_t group;
  int32_t n_rows;
  int64_t slab_off;
  int64_t idx_off;
  int32_t conf_slot;
  int32_t from_global; /* 1: slab_off indexes the global arena instead of the params arena */
} FedbEvalJob;

typedef struct FedbTile {
  int32_t owner; /* group (training) or job (evaluation) */
  int32_t tile;  /* tile index inside the owner's current batch */
} FedbTile;

/* Per-rank argument block of the cross-GPU aggregation kernel (fedb_fedavg_p2p).  All pointers are
 * device pointers valid on the launching GPU; sym[q] is rank q's symmetric buffer as mapped into
 * this process (CUDA IPC).  Symmetric buffer layout (bytes): partial[2][global_total] fp32 at
 * sym_partial_off, conf[2][8 source ranks][conf_slots][64] int32 at sym_conf_off, flags[24] u64 at sym_flag_off
 * (arrival A, counts B, NVLS result C; one word per source rank each). */
typedef struct FedbP2PRank {
  int32_t rank, world;
  int32_t n_feds, n_groups_local;
  int32_t n_groups_total, conf_slots;   /* conf_slots: slots per parity in the symmetric buffer */
  int64_t global_total;                 /* floats in the global arena */
  const FedbGroup* groups;              /* local tables (resident clients only) */
  const FedbFed* feds;                  /* first_group / n_clients index the local group table */
  const FedbFed* feds_global;           /* first_group / n_clients index all clients of all ranks */
  FedbFedState* fed_state;
  float* params;
  float* global;
  const int32_t* conf;                  /* local confusion slots [n_groups_local][64] */
  int32_t* conf_all;                    /* scratch [n_groups_total][64] */
  float* hist_local;                    /* [max_rounds][n_groups_total][4] */
  float* hist_global;                   /* [n_feds][max_rounds][4] */
  const int32_t* conf_src_rank;         /* [n_groups_total]: which rank hosts global client i */
  const int32_t* conf_src_slot;         /* [n_groups_total]: its local slot on that rank */
  uint8_t* sym[8];
  int64_t sym_partial_off, sym_conf_off, sym_flag_off;
  uint64_t* seq;                        /* local: launches completed (monotonic) */
  uint32_t* bar;                        /* local grid-barrier counter (monotonic) */
  int32_t* err;                         /* set to FEDB_E_TIMEOUT when a peer never arrived */
  int64_t timeout_cycles;
  int32_t write_back, do_metrics;
  int32_t begin_next, pad_;
  void* shadow;                         /* optional bf16 arena (params layout) that receives the averaged weights too */
  /* optional NVLS (NVLink-SHARP multicast) buffers, see fedb_nvls_*: when nvls_mc_partial is set, k_fedavg_p2p keeps the ranks'
   * partial sums in multicast-bound memory and lets the NVSwitch add them (multimem.ld_reduce) instead of loading every peer */
  float* nvls_uc_partial;               /* this GPU's copy: partial[2][global_total] */
  float* nvls_mc_partial;               /* multicast view of the same */
  float* nvls_uc_result;                /* this GPU's copy of the reduced global arena [global_total] */
  float* nvls_mc_result;                /* multicast view: a multimem.st reaches every GPU's copy */
} FedbP2PRank;

/* Grouped tensor-core GEMM of the wide-MLP path (csrc/fedb_gemm.cu).  blockIdx.z = group (client);
 * operand X of group g starts at X + g * x_group_stride (elements).  a_mn / b_mn: 0 = K-major source
 * ([M or N][K], K contiguous), 1 = MN-major source ([K][M or N], M or N contiguous). */
enum { FEDB_GEMM_FWD_BIAS_RELU = 0, FEDB_GEMM_DGRAD_MASK = 1, FEDB_GEMM_WGRAD_OPT = 2, FEDB_GEMM_STORE_F32 = 3 };
typedef struct FedbGemm {
  const void* A;                 /* bf16 */
  const void* B;                 /* bf16 */
  int64_t a_group_stride, b_group_stride;
  int32_t lda, ldb;
  int32_t M, N, K;
  int32_t a_mn, b_mn;
  int32_t epilogue;
  const float* bias;             /* FWD: fp32 [N] per group */
  int64_t bias_group_stride;
  void* C;                       /* FWD / DGRAD: bf16 [M][ldc]; STORE_F32: fp32 */
  int64_t c_group_stride;
  int32_t ldc, ldact;
  const void* act;               /* DGRAD: bf16 forward activation [M][ldact] whose sign gates the gradient */
  int64_t act_group_stride;
  float* W;                      /* WGRAD_OPT: fp32 master weights [M][ldw] (+ moments, bf16 shadow), updated in place */
  float* Mom;
  float* Vel;
  void* Wsh;
  int64_t w_group_stride;
  int32_t ldw;
  float l2_scale;                /* WGRAD_OPT: gradient += alpha * l2_scale * w (sklearn: 1 / n_batch) */
  const FedbGroup* groups;       /* WGRAD_OPT: optimizer hyper-parameters per group */
  const int32_t* step;           /* WGRAD_OPT: optimizer step count t per group (already incremented) */
} FedbGemm;

const char* fedb_last_error(void);
int fedb_version(void);
void fedb_struct_sizes(int32_t* out5); /* sizeof of the five structs above, for binding checks */
/* Dynamic shared memory (bytes) k_fwd_bwd / k_eval need for a model; <0 if it cannot run. */
int64_t fedb_simt_smem_bytes(const int32_t* dims, int32_t n_layers);
int32_t fedb_batch_stats_stride(int32_t max_slab);

/* K4 — parameter init (C:18-21 torch default; SK:445-460 Glorot).  Philox4x32-10, key =
 * seed, counter = (element/4, stream, 0, 0).  streams[g] < 0 leaves group g untouched. */
int fedb_init_params(const FedbGroup* groups, int32_t n_groups, int32_t max_slab, float* params,
                     const int32_t* streams, int32_t kind, uint64_t seed, void* stream);

/* Top of a round: promote stop_pending -> stop, count the round, clear the confusion
 * counters of the federation's slots (C:130-142). conf_slots_per_fed slots per federation. */
int fedb_round_begin(const FedbFed* feds, FedbFedState* state, int32_t n_feds, int32_t* conf,
                     int32_t conf_slots_total, void* stream);

/* Per-epoch reshuffle of every client's shard order (SK:704-718): order_out = shard indices
 * permuted by a keyed Feistel bijection of (seed, group, epoch[group]). */
int fedb_shuffle(const FedbGroup* groups, int32_t n_groups, const int32_t* idx_in, int32_t* idx_out,
                 const int32_t* epoch, const int32_t* active, uint64_t seed, void* stream);

/* K2 (SIMT fp32) part 1 — forward, loss, backward of one optimizer step for every group:
 * per 32-row tile, writes a partial-gradient record (+ loss sum) to `scratch`
 * (C:68-71; SK:301-393).  batch = minibatch index inside the epoch (0 for full batch). */
int fedb_fwd_bwd(const FedbGroup* groups, const FedbTile* tiles, int32_t n_tiles, const float* params,
                 const float* X, int32_t ldx, const int32_t* y, const int32_t* idx, int32_t batch,
                 float* scratch, int32_t* step, const int32_t* active, const FedbFedState* fed_state,
                 int64_t smem_bytes, void* stream);

/* K2 part 2 — fixed-order reduction of the tile records, L2, optimizer update (C:72-73 Adam +
 * StepLR; SKO AdamOptimizer / SGDOptimizer).  loss_out[g] = mean data loss of the step;
 * batch_stats (optional; fedb_batch_stats_stride(max_slab) doubles per (group, batch)) receives
 * {sum of row losses, n_b, per-block sum W^2...} for the sklearn stopping rule (SK:353-361). */
int fedb_reduce_opt(const FedbGroup* groups, int32_t n_groups, int32_t max_slab, float* params, float* m,
                    float* v, const float* scratch, int32_t batch, const int32_t* step,
                    const int32_t* active, const FedbFedState* fed_state, float* loss_out,
                    double* batch_stats, int32_t max_batches, void* stream);

/* End of an sklearn-style epoch: loss curve, tol / n_iter_no_change rule (SK:750-777). */
int fedb_epoch_end(const FedbGroup* groups, int32_t n_groups, const double* batch_stats,
                   int32_t max_batches, int32_t max_slab, double tol, int32_t n_iter_no_change, int32_t max_iter,
                   int32_t* active, int32_t* epoch, int32_t* no_improve, double* best_loss,
                   double* last_loss, void* stream);
/* K2b in persistent form: the whole local fit of every client (per-epoch shuffle, mini-batch steps, tol / n_iter_no_change
 * stopping: SK:703-777) in ONE launch.  A cluster of cluster_size CTAs (1..8) owns a client and loops on the device; arguments
 * as for fedb_shuffle / fedb_fwd_bwd / fedb_reduce_opt / fedb_epoch_end, which it replaces call for call with bit-identical
 * results.  Runs at most max_epochs further epochs; clients with active == 0 or a stopped federation return at once. */
int fedb_fit_cluster(const FedbGroup* groups, int32_t n_groups, int32_t max_slab, int32_t cluster_size, float* params, float* m,
                     float* v, const float* X, int32_t ldx, const int32_t* y, const int32_t* idx_in, int32_t* idx_epoch,
                     float* scratch, int32_t* step, int32_t* active, const FedbFedState* fed_state, int32_t* epoch,
                     int32_t* no_improve, double* best_loss, double* last_loss, float* loss_out, double* batch_stats,
                     int32_t max_batches, int32_t n_batches, double tol, int32_t n_iter_no_change, int32_t max_iter,
                     int32_t max_epochs, int32_t shuffle, uint64_t seed, int64_t smem_bytes, void* stream);

/* K3 — forward + argmax / threshold + confusion counts (C:75-83, S:104; SK:1283-1290). */
int fedb_eval(const FedbGroup* groups, const FedbEvalJob* jobs, const FedbTile* tiles, int32_t n_tiles,
              const float* params, const float* global, const float* X, int32_t ldx, const int32_t* y,
              const int32_t* idx, int32_t* conf, const FedbFedState* fed_state, int64_t smem_bytes,
              void* stream);

/* K1 (single GPU) + K5 — FedAvg: global = sum_k (w_k * n_k) / N in client order (C:108-116;
 * unweighted: S:114-117, H:37), written to the global arena and back into every client slab
 * (C:119-120); block 0 of each federation also turns the round's confusion counts into
 * accuracy / weighted precision / recall / F1 (C:85-90), appends them to the history and
 * runs the early-stop rule (C:181-192).  hist_local: [max_rounds][n_groups][4] floats,
 * hist_global: [n_feds][max_rounds][4].  write_back=0 leaves client slabs untouched.
 * begin_next=1 additionally does the next round's fedb_round_begin work for the federation
 * (stop promotion, round count, clearing its clients' confusion slots); only valid when one block
 * covers a federation's slab (max_slab <= 1024 floats) — larger problems call fedb_round_begin.
 * shadow_bf16 (optional): bf16 arena with the params layout that receives the averaged weights too. */
int fedb_fedavg(const FedbGroup* groups, const FedbFed* feds, FedbFedState* state, int32_t n_feds,
                int32_t n_groups, int32_t max_slab, float* params, float* global, const int32_t* conf,
                float* hist_local, float* hist_global, int32_t write_back, int32_t do_metrics,
                int32_t begin_next, void* shadow_bf16, void* stream);

/* K2a/K3 on tcgen05: n_steps local optimizer steps (+ local evaluation into conf slot g) for every
 * group in one launch, one thread-block cluster per client (csrc/fedb_tc.cu; C:63-91).  Groups
 * must have 2 hidden ReLU layers with in <= 15, h1 <= 63, h2 <= 256, out <= 8.  scratch: the
 * gradient-record arena of fedb_fwd_bwd; every group needs max_tiles >= cluster_size records. */
int64_t fedb_tc_smem_bytes(void);
int fedb_local_round_tc(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X,
                        int32_t ldx, const int32_t* y, const int32_t* idx, float* params, float* m, float* v,
                        int32_t* step, const int32_t* active, const FedbFedState* fed_state, float* loss_out,
                        int32_t* conf, float* scratch, int32_t batch, int32_t n_steps, int32_t do_eval, void* stream);

/* The whole round in ONE launch for small single-GPU problems: fedb_local_round_tc's work, then the
 * last cluster to finish performs fedb_fedavg(write_back=1, do_metrics=1, begin_next=1) for every
 * federation.  conf_ints counters are cleared for the next round; tail_count: one zeroed uint32.
 * state_mirror (optional): mapped pinned HOST memory, n_feds records; the tail stores the updated
 * federation state there as well, so a host that syncs on the stream reads the round's result
 * without a device-to-host memcpy. */
/* state_mirror (fedb_round_tc*, optional): mapped pinned host memory of n_feds FedbFedState records FOLLOWED, at the next multiple
 * of 16 bytes, by n_feds x 4 chunks of 16 bytes {int32 done, int32 stop, double metric_i}.  The records are complete once the stream
 * has synchronised; each chunk is a single store that carries the round counter, so a host may poll the chunks instead. */
int fedb_round_tc(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X, int32_t ldx,
                  const int32_t* y, const int32_t* idx, float* params, float* m, float* v, int32_t* step,
                  const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf, float* scratch,
                  int32_t n_steps, const FedbFed* feds, int32_t n_feds, float* global, float* hist_local,
                  float* hist_global, int32_t conf_ints, uint32_t* tail_count, FedbFedState* state_mirror, void* stream);

/* The same launches with an optional trace buffer: prof (may be null) = 32 int64 on the device that receive clock64 stamps of
 * CTA 0 / thread 0 at the phase boundaries (tools/tc_timeline.py lists them).  The bindings call these forms. */
int fedb_local_round_tc_prof(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X, int32_t ldx,
                             const int32_t* y, const int32_t* idx, float* params, float* m, float* v, int32_t* step,
                             const int32_t* active, const FedbFedState* fed_state, float* loss_out, int32_t* conf,
                             float* scratch, int32_t batch, int32_t n_steps, int32_t do_eval, long long* prof, void* stream);
int fedb_round_tc_prof(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X, int32_t ldx,
                       const int32_t* y, const int32_t* idx, float* params, float* m, float* v, int32_t* step,
                       const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf, float* scratch,
                       int32_t n_steps, const FedbFed* feds, int32_t n_feds, float* global, float* hist_local,
                       float* hist_global, int32_t conf_ints, uint32_t* tail_count, FedbFedState* state_mirror,
                       long long* prof, void* stream);
/* Largest slab (floats) the bf16 round kernel accepts (gradient slices of a cluster must fit one CTA's exchange area). */
int32_t fedb_tc_max_slab(void);
/* Hardware probe behind tests/test_gpu_tc.py: a chain of M = 128 tcgen05.mma on operand images laid out by the host; args_host
 * is the int32 block ops/tc.py:probe() fills (fedb_tc_probe_args_size() bytes).  Synchronises nothing; not a hot path. */
int fedb_tc_probe(const void* a_img, const void* b_img, const int32_t* args_host, float* out, void* stream);
int fedb_tc_probe_args_size(void);

/* K2a/K3 at fp32-grade accuracy on tcgen05 (csrc/fedb_tcx.cu): every operand is held as two fp16 numbers (hi + lo, 22
 * significant bits) and every GEMM is three MMAs into the same fp32 accumulator; the reference trains in fp32 (C:65, C:69-72).
 * Tiles are 64 rows, cluster_size 1..16 (fedb_tcx_max_cluster(): 16 where the device can co-schedule a 16-CTA cluster, else
 * 8).  Same model limits as the bf16 kernel; arguments as fedb_local_round_tc_prof / fedb_round_tc_hostin /
 * fedb_round_tc_p2p_hostin respectively. */
int64_t fedb_tcx_smem_bytes(void);
int32_t fedb_tcx_rows(void);
int32_t fedb_tcx_max_cluster(void);
/* 1 if a model in-h1-h2-out runs with `cluster` CTAs per client (the per-owner gradient blocks of all CTAs must fit the
 * receive and staging areas in shared memory); the kernel itself trusts the device table, so check before launching. */
int32_t fedb_tcx_fits(int32_t d0, int32_t h1, int32_t h2, int32_t n_out, int32_t cluster);
int fedb_local_round_tcx(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, const float* X, int32_t ldx,
                         const int32_t* y, const int32_t* idx, float* params, float* m, float* v, int32_t* step,
                         const int32_t* active, const FedbFedState* fed_state, float* loss_out, int32_t* conf,
                         float* scratch, int32_t batch, int32_t n_steps, int32_t do_eval, long long* prof, void* stream);
int fedb_round_tcx(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, float* X, int32_t ldx, int32_t* y,
                   const float* x_host, const int32_t* y_host, const int32_t* idx, float* params, float* m, float* v,
                   int32_t* step, const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf,
                   float* scratch, int32_t n_steps, const FedbFed* feds, int32_t n_feds, float* global, float* hist_local,
                   float* hist_global, int32_t conf_ints, uint32_t* tail_count, FedbFedState* state_mirror, long long* prof,
                   void* stream);
int fedb_round_tcx_p2p(const FedbGroup* groups, int32_t cluster_size, float* X, int32_t ldx, int32_t* y, const float* x_host,
                       const int32_t* y_host, const int32_t* idx, float* params, float* m, float* v, int32_t* step,
                       const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf, float* scratch,
                       int32_t n_steps, const FedbP2PRank* rank_dev, FedbFedState* state_mirror, void* stream);

/* The multi-GPU round in one launch per rank for ANY number of resident clients: the rank's last cluster to finish folds its
 * clients' weighted slabs into the rank partial (C:110-115), then runs the cross-GPU exchange, write-back, metrics and early
 * stop; tail_count = one zeroed uint32.  Bit-identical to fedb_local_round_tcx + fedb_fedavg_p2p. */
int fedb_round_tcx_p2p_many(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, float* X, int32_t ldx, int32_t* y,
                            const float* x_host, const int32_t* y_host, const int32_t* idx, float* params, float* m, float* v,
                            int32_t* step, const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf,
                            float* scratch, int32_t n_steps, const FedbP2PRank* rank_dev, uint32_t* tail_count,
                            FedbFedState* state_mirror, void* stream);

/* K2b on the tensor cores: the sklearn-style local fit (fedb_fit_cluster's contract: mini-batch Adam / SGD with the L2 term,
 * loss curve, tol / n_iter_no_change rule, per-epoch reshuffle, up to max_epochs further epochs) of every client in ONE
 * launch, epoch loop on the device.  A cluster of 4 CTAs per client splits hidden layer 2 (tcgen05 on split fp16 operands,
 * fp32 accumulators in tensor memory).  fedb_fitx_fits: 2 hidden ReLU layers, in <= 15, h1 <= 63, h2 <= 512, out <= 8.
 * prof (optional, device): per-CTA globaltimer stamps for tools/fitx_timeline.py; pass NULL otherwise. */
int64_t fedb_fitx_smem_bytes(void);
int32_t fedb_fitx_fits(int32_t d0, int32_t h1, int32_t h2, int32_t n_out);
int fedb_fit_tcx(const FedbGroup* groups, int32_t n_groups, float* params, float* m, float* v, const float* X, int32_t ldx,
                 const int32_t* y, const int32_t* idx_in, int32_t* idx_epoch, int32_t* step, int32_t* active,
                 const FedbFedState* fed_state, int32_t* epoch, int32_t* no_improve, double* best_loss, double* last_loss,
                 float* loss_out, double tol, int32_t n_iter_no_change, int32_t max_iter, int32_t max_epochs, int32_t shuffle,
                 uint64_t seed, long long* prof, void* stream);

/* K2d: grouped tcgen05 GEMM with fused epilogues (forward bias+ReLU, dgrad ReLU mask, weight gradient
 * + optimizer); args_host is read on the host.  M % 128 == 0, N % 16 == 0, K % 16 == 0. */
int32_t fedb_gemm_args_size(void);
int fedb_gemm_tc(const FedbGemm* args_host, int32_t n_groups, void* stream);   /* cp.async-fed variant (fallback) */
/* Main variant: TMA-fed (3-D tensor maps encoded per call), warp-specialised, coalesced epilogues. */
int fedb_gemm_tma(const FedbGemm* args_host, int32_t n_groups, void* stream);
/* CTA-pair variant (cluster of 2, tcgen05 cta_group::2, 256 x 256 tile per pair).  Needs M % 256 == 0 and
 * N % 256 == 0.  err_dev (optional device int32) receives FEDB_E_TIMEOUT if a pair's pipeline stalls. */
int fedb_gemm_tma2(const FedbGemm* args_host, int32_t n_groups, int32_t* err_dev, void* stream);
/* Persistent CTA-pair variant: one cluster per SM pair walks a static tile list, two accumulators in tensor memory, so
 * the epilogue of a tile overlaps the main loop of the next.  Same requirements and err_dev as fedb_gemm_tma2. */
int fedb_gemm_tma2p(const FedbGemm* args_host, int32_t n_groups, int32_t* err_dev, void* stream);
/* Same launch with per-CTA phase stamps: prof_dev = 8 uint64 per CTA in grid order (x, y, group): globaltimer ns at
 * start, main loop done, pass 0 staged, pass 0 done, pass 1 staged, end; then the SM id (tools/gemm_timeline.py). */
int fedb_gemm_tma2_prof(const FedbGemm* args_host, int32_t n_groups, int32_t* err_dev, uint64_t* prof_dev, void* stream);

/* Wide-MLP path, the kernels around fedb_gemm_tc (csrc/fedb_wide.cu).  A* / dZ are bf16 [G][B][d]
 * activation buffers (B rows per client, padded shards must be equal-sized); dz is fp32 [G][B][8]. */
int fedb_wide_gather_x(const FedbGroup* groups, int32_t n_groups, const float* X, int32_t ldx, const int32_t* idx,
                       void* A0, int32_t B, int32_t ld0, void* stream);
int fedb_wide_shadow(const float* src, void* dst, int64_t n_floats, void* stream);
int fedb_wide_head_fwd(const FedbGroup* groups, int32_t n_groups, const float* params, const void* A, int32_t B, int32_t d,
                       const int32_t* y, const int32_t* idx, float* dz, float* loss_rows, int32_t* conf, int32_t* step,
                       const FedbFedState* fed_state, int32_t eval_mode, void* stream);
int fedb_wide_head_bwd(const FedbGroup* groups, int32_t n_groups, float* params, float* m, float* v, void* shadow,
                       const void* A, int32_t B, int32_t d, const float* dz, const float* loss_rows, void* dZ,
                       const int32_t* step, const FedbFedState* fed_state, float* loss_out, void* stream);
int fedb_wide_bias_opt(const FedbGroup* groups, int32_t n_groups, float* params, float* m, float* v, const void* dZ,
                       int32_t B, int32_t d, int32_t layer, const int32_t* step, const FedbFedState* fed_state, void* stream);

/* K2a + K3 + K1b in ONE launch per rank: a multi-GPU federation with one resident client per rank (the
 * reference's own topology, C:36) — local steps, local evaluation, cross-GPU FedAvg over NVLink peer
 * memory, write-back, metrics, early stop.  rank_dev: this rank's FedbP2PRank (same record and symmetric
 * buffer as fedb_fedavg_p2p; flags[8..15] of the buffer carry the second arrival flag). */
int fedb_round_tc_p2p(const FedbGroup* groups, int32_t cluster_size, const float* X, int32_t ldx, const int32_t* y,
                      const int32_t* idx, float* params, float* m, float* v, int32_t* step, const int32_t* active,
                      FedbFedState* fed_state, float* loss_out, int32_t* conf, float* scratch, int32_t n_steps,
                      const FedbP2PRank* rank_dev, FedbFedState* state_mirror, void* stream);
/* The two single-launch rounds with the step's inputs pulled from pinned host memory by the kernel itself: x_host / y_host are
 * device-visible addresses of host arrays laid out like X / y (cudaHostAlloc / torch pin_memory under UVA); every row the round
 * uses is read over the host link, written through to X / y, and fed to the first layer.  This replaces two copy-engine nodes in
 * front of the round (13 us for 64 KB inside a CUDA graph) by ~3 us of loads that overlap the weight staging.  Both null: the
 * plain entry points.  prof as in fedb_round_tc_prof (may be null). */
int fedb_round_tc_hostin(const FedbGroup* groups, int32_t n_groups, int32_t cluster_size, float* X, int32_t ldx, int32_t* y,
                         const float* x_host, const int32_t* y_host, const int32_t* idx, float* params, float* m, float* v,
                         int32_t* step, const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf,
                         float* scratch, int32_t n_steps, const FedbFed* feds, int32_t n_feds, float* global, float* hist_local,
                         float* hist_global, int32_t conf_ints, uint32_t* tail_count, FedbFedState* state_mirror, long long* prof,
                         void* stream);
int fedb_round_tc_p2p_hostin(const FedbGroup* groups, int32_t cluster_size, float* X, int32_t ldx, int32_t* y, const float* x_host,
                             const int32_t* y_host, const int32_t* idx, float* params, float* m, float* v, int32_t* step,
                             const int32_t* active, FedbFedState* fed_state, float* loss_out, int32_t* conf, float* scratch,
                             int32_t n_steps, const FedbP2PRank* rank_dev, FedbFedState* state_mirror, void* stream);

/* K1b: one launch per rank per round — local weighted partial sums, cross-GPU arrival flags,
 * peer loads over NVLink of every rank's partial, write-back, metrics and early stop
 * (csrc/fedb_p2p.cu; C:101-120, C:165-192).  ranks_dev[0..n_ranks_here): one entry in real
 * multi-GPU use; R entries emulate R ranks on one GPU (tests).  Cooperative launch. */
int32_t fedb_p2p_rank_size(void);
int fedb_fedavg_p2p(const FedbP2PRank* ranks_dev, int32_t n_ranks_here, int32_t blocks_x, void* stream);
/* NVLS plumbing (csrc/fedb_nvls.cu).  Rank 0 creates a multicast object for n_devices and exports it as a POSIX file descriptor
 * (pass it to the other ranks with SCM_RIGHTS); they import it; every rank adds its device; after ALL have, every rank binds an
 * allocation of *bytes_out bytes and gets two views of it: uc_ptr (own copy) and mc_ptr (multicast, multimem.* only). */
int fedb_nvls_supported(int32_t device);   /* 1 / 0, negative on error */
int fedb_nvls_create(int32_t n_devices, int64_t bytes, uint64_t* mc_handle_out, int32_t* fd_out, int64_t* bytes_out);
int fedb_nvls_import(int32_t fd, uint64_t* mc_handle_out);
int fedb_nvls_add_device(uint64_t mc_handle, int32_t device);
int fedb_nvls_bind_map(uint64_t mc_handle, int32_t device, int64_t bytes, uint64_t* mem_handle_out, void** uc_ptr, void** mc_ptr);
int fedb_nvls_release(uint64_t mc_handle, uint64_t mem_handle, void* uc_ptr, void* mc_ptr, int64_t bytes);
/* Symmetric-memory plumbing: cudaMalloc'ed, zeroed buffer; 64-byte CUDA IPC handle out / in.
 * These four allocate / synchronise and are for set-up only. */
int fedb_symm_alloc(int64_t bytes, void** out);
int fedb_symm_free(void* p);
int fedb_ipc_get_handle(void* p, uint8_t* handle64);
int fedb_ipc_open_handle(const uint8_t* handle64, void** out);
int fedb_ipc_close_handle(void* p);

/* One optimizer step of every group (the work of fedb_fwd_bwd followed by fedb_reduce_opt) for
 * callers that must not allocate device memory.  No counterpart in the reference.
 *
 * Who needs it: a host that has handed all device memory to its own pool up front (an inference
 * server, a simulator with a fixed budget), code that runs where cudaMalloc is off limits because
 * it synchronises the whole device (a latency-critical loop, a callback on a stream), and
 * embedders that may not link an allocator at all.  Such callers own params/m/v/X/... already, but
 * the step also needs room for per-tile gradient records; here that room is a fixed 16 MiB array
 * inside the library (fedb_noalloc_scratch_floats() floats), resident on every GPU that loads the
 * library until the process exits.  The call performs no cudaMalloc/cudaFree and uses no allocating
 * library routine.
 *
 * groups_host is a host copy of the table at groups_dev and must match it byte for byte (the
 * kernels index the array from the device table); it is used to check that every group's records
 * (scratch_off + max_tiles * (slab_size + FEDB_LOSS_SLOTS)) fit in the array, that max_tiles covers
 * the tiles of one step (ceil(min(batch_size or n_rows, n_rows) / FEDB_TILE_ROWS)) and that n_tiles
 * does not exceed the records reserved.  A step that needs more room returns FEDB_E_UNSUPPORTED, a
 * malformed table FEDB_E_ARG, each with an explanatory fedb_last_error(); nothing is launched.
 *
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 * that device current) and remembers it; make that first call outside any stream capture.
 * All calls on a device share the one array and it keeps whatever the last call left in it.  The
 * step does not rely on its initial contents (every record it reads is written first), but two
 * calls whose kernels can overlap on the GPU — different streams, or different host threads — race
 * on it: serialise them on one stream.  (Host threads may enter concurrently; only the GPU work
 * must not overlap.)  Results are bit-identical to fedb_fwd_bwd + fedb_reduce_opt with a
 * caller-provided buffer (batch_stats is not produced). */
int64_t fedb_noalloc_scratch_floats(void);
int fedb_local_step_noalloc(const FedbGroup* groups_host, const FedbGroup* groups_dev, int32_t n_groups,
                            const FedbTile* tiles, int32_t n_tiles, int32_t max_slab, float* params, float* m,
                            float* v, const float* X, int32_t ldx, const int32_t* y, const int32_t* idx,
                            int32_t batch, int32_t* step, const int32_t* active, const FedbFedState* fed_state,
                            float* loss_out, int64_t smem_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FEDB200_H */
