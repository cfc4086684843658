/*
 * qpolicy.h — C ABI of the B200-native Q-learning decision policy.
 *
 * Drop-in boundary for ONE hot path of cosmir17/Scala-akka-tensorflow-sharetrade-helper: the body
 * of QDecisionPolicyActor (the tensorflow_scala session) and the per-step arithmetic of
 * TrainerChildActor.computeWithFolding.  The Scala/Akka layer keeps its messages; a JNI/Panama
 * shim forwards them to these entry points.  All `file:line` citations are
 * relative to the reference's src/main/scala/.
 *
 * Conventions
 *   - plain C, no exceptions; every function returns a qp_status unless stated; the text of the
 *     last failure on the calling thread is available from qp_last_error().
 *   - host pointers unless the name says otherwise; the caller owns host buffers, the library
 *     copies or consumes them before returning.  Device memory is owned by the handle.
 *   - a handle is thread-compatible (one call at a time), not thread-bound: each entry point
 *     selects the handle's device itself, so any dispatcher thread may call it.
 *   - calls are synchronous on return (the actor replies Action / Updated right after), except
 *     the *_async ones, which only enqueue on the handle's stream.
 *   - parameter layout: params = [ W1 (D x H, row-major) | W2 (H x 3, row-major) ], D = window + 2;
 *     the AdaGrad accumulators use the same layout.  Biases are the constants 0.1f
 *     (QDecisionPolicyActor.scala:42, :46) and are not parameters.
 *   - actions: 0 = Buy, 1 = Sell, 2 = Hold (QDecisionPolicyActor.scala:17).
 */
#ifndef QPOLICY_H
#define QPOLICY_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    QP_OK = 0,
    QP_EINVAL = 1,       /* IllegalArgumentException in the reference (size guards) */
    QP_ECUDA = 2,        /* CUDA runtime failure */
    QP_ENCCL = 3,        /* NCCL failure */
    QP_ESTATE = 4,       /* call made in the wrong state (e.g. no episodes created) */
    QP_EUNSUPPORTED = 5, /* shape outside what the kernels cover */
    QP_EIO = 6           /* checkpoint / file errors */
} qp_status;

enum { QP_BUY = 0, QP_SELL = 1, QP_HOLD = 2, QP_ACTIONS = 3 };
enum { QP_BOOK_REFERENCE = 0, QP_BOOK_RLPY = 1 };

/* Hyper-parameters: compile-time constants in the reference
 * (QDecisionPolicyActor.scala:17-22, :50), gathered into one POD. */
typedef struct {
    int32_t window;          /* price-window length; reference 201 (TrainerChildActor.scala:90) */
    int32_t hidden;          /* h1Dim = 200; must be a multiple of 4 */
    float epsilon;           /* 0.9f */
    float gamma;             /* 0.001f */
    float lr;                /* AdaGrad(0.01f) */
    float bias;              /* 0.1f, both layers */
    float step_div;          /* 1000f in `step / 1000f` (QDecisionPolicyActor.scala:58) */
    float adagrad_init_acc;  /* initial accumulator value, 1e-8f */
    int32_t bookkeeping;     /* QP_BOOK_REFERENCE: Buy/Sell decided from the constructor's position
                                (TrainerChildActor.scala:118-123); QP_BOOK_RLPY: from the carry */
    int32_t device;          /* CUDA device ordinal */
    uint64_t seed;           /* Philox key for weights (stream 2) and action draws (stream 1) */
} qp_config;

typedef struct qp_handle qp_handle;

/* Fills *cfg with the reference's constants (QDecisionPolicyActor.scala:17-22, :50). */
int qp_config_default(qp_config* cfg);

/* QDecisionPolicyActor constructor + initialiseSession (QDecisionPolicyActor.scala:38-52, :79-83):
 * allocates the policy on cfg->device and draws W1, W2 ~ N(0,1) from the seed. */
int qp_create(const qp_config* cfg, qp_handle** out);
int qp_destroy(qp_handle* h);

/* D = window + 2 (inputDim, QDecisionPolicyActor.scala:18); P = D*H + 3*H.  Return the count. */
int64_t qp_input_dim(const qp_handle* h);
int64_t qp_param_count(const qp_handle* h);

/* No counterpart in the reference (its weights live inside the Session): inject
 * or export weights and accumulators for parity runs and checkpoints.  acc may be NULL. */
int qp_set_weights(qp_handle* h, const float* params, const float* acc);
int qp_get_weights(qp_handle* h, float* params, float* acc);
/* The summed gradient of the most recent update (parity aid; no reference counterpart). */
int qp_get_last_grad(qp_handle* h, float* grad);

/* SelectionAction(currentState, step) -> Action  (QDecisionPolicyActor.scala:32, :56-62), batched.
 *   states [B x state_len] row-major; state_len != D -> QP_EINVAL (the guard at :56-57).
 *   steps [B]; env_ids [B] or NULL (then env id = row index) pick the Philox counter
 *   (env id, (uint32)step, stream 1): one 24-bit uniform per row, one bounded draw if exploring.
 *   q_out [B x 3] may be NULL. */
int qp_select_action(qp_handle* h, const float* states, int64_t state_len, const float* steps,
                     const int64_t* env_ids, int32_t* actions, float* q_out, int64_t B);

/* UpdateQ(state, reward, nextState) -> Updated  (QDecisionPolicyActor.scala:33, :64-76), batched:
 * TD target on the argmax of q(nextState), gradient SUM over the B rows, one AdaGrad step.
 * B = 1 is exactly the reference's message.  A length other than D -> QP_EINVAL (:64-65). */
int qp_update_q(qp_handle* h, const float* states, int64_t state_len, const float* rewards,
                const float* next_states, int64_t next_state_len, int64_t B);

/* ---- episode level: TrainerChildActor.train / computeWithFolding (TrainerChildActor.scala:64-103)
 * for n_envs independent episodes that share this policy (TrainerRouterActor.scala:44, :60-66). */

/* Uploads the price series ([n_series x series_len] fp32, already date-sorted as at
 * TrainerChildActor.scala:66) and creates n_envs episodes; series_of_env[e] names the series of
 * episode e.  series_len <= window -> QP_EINVAL (the guard at :65, :70).  env_id_base is added
 * to the episode index to form the RNG counter, so a shard draws its global episodes' numbers.
 * Every episode starts from (budget0, shares0, shareValue 0.0) (:84-86). */
int qp_env_create(qp_handle* h, const float* prices, int64_t series_len, int64_t n_series,
                  const int32_t* series_of_env, int64_t n_envs, int64_t env_id_base,
                  double budget0, int32_t shares0);
/* Initialise (TrainerChildActor.scala:58-60): back to step 0 and the initial position. */
int qp_env_reset(qp_handle* h);
/* nPrices - window (TrainerChildActor.scala:67), and how many env-steps have run.  Counts. */
int64_t qp_env_steps_total(const qp_handle* h);
int64_t qp_env_steps_done(const qp_handle* h);

/* Runs n_steps iterations of the fold at TrainerChildActor.scala:87-102 for every episode:
 * state build, select (QDecisionPolicyActor.scala:56-62), decision (:118-123), reward and next
 * state (:136-146), TD update (QDecisionPolicyActor.scala:64-76, gradient SUM over episodes). */
int qp_env_step(qp_handle* h, int32_t n_steps);
int qp_env_step_async(qp_handle* h, int32_t n_steps);
int qp_sync(qp_handle* h);

/* Execution options (no reference counterpart); all default to 1.
 *   "fused"   qp_env_step* run as ONE persistent cooperative kernel for all requested steps (price windows
 *             stay in shared memory); 0 forces the three-kernels-per-step path.
 *   "ep_tc"   the persistent kernel does its two contractions on the tcgen05 tensor cores (hi/lo TF32 pairs,
 *             fp32-grade).  Taken when the prices are tf32-exact (checked at qp_env_create) and the shape fits
 *             (window 193..203 at hidden 200); any number of episodes (waves).  Otherwise, and
 *             with 0, the FFMA form of the kernel runs, which needs the episodes to fit one 32-episode tile per SM.
 *   "ep_c2"   the tensor-core kernel as CTA PAIRS (clusters of two) that split the hidden dimension: one 64-episode
 *             tile per pair, each CTA holds half of W1^T's columns, Q-value halves meet over distributed shared
 *             memory.  Taken when "ep_tc" is and the shape fits (hidden a multiple of 8, <= 256; window + 50 <= 256
 *             rows); more episodes than one tile per pair (64 x SMs/2; 4736 on a B200) run in waves.  Prices that are
 *             not tf32-exact: 32 episodes per pair, one tile per pair only.  Otherwise, and with 0, the
 *             one-CTA-per-tile tensor-core kernel runs (tf32-exact prices, any number of episodes, in waves).
 *   "ep_c2_waves"  the pair kernel also where a pair has several tiles per env-step (tf32-exact prices); 0: the
 *             one-tile-per-CTA kernel takes those batches.
 *   "p2p"     several GPUs: gradient exchange inside the persistent kernel over NVLink peer memory
 *             (qp_dist_p2p_export / qp_dist_p2p_import, or qp_dist_p2p_attach); 0: one fused step kernel per
 *             env-step + NCCL all-reduce.
 *   "nvls"    several GPUs, pair kernel, blocks attached with a multicast address: each share leaves as ONE
 *             multimem store that the NVSwitch delivers to every rank; 0: world - 1 unicast stores.
 *   "rsag"    several GPUs, pair kernel: 1 = the share exchange as reduce-scatter + all-gather (each share goes to ONE
 *             owner rank, which adds the ranks' values in rank order and sends the sum to all), 0 = all-to-all (each
 *             share to every rank, every rank adds), -1 (default) = reduce-scatter + all-gather from 8 ranks on.
 *             Same sums bit for bit either way.
 *   "fanout"  the router fan-out path for episodes grouped by series (series_of_env non-decreasing): the window part of
 *             the hidden layer once per (series, env-step), the backward of the window rows as an outer product per
 *             series.  1 (default): taken where it pays (one GPU, at least 8 episodes per series on average, more
 *             episodes than the pair kernel takes); 2: wherever the grouping allows (several GPUs: set it on every rank;
 *             NCCL all-reduce of the gradient per step); 0: never.  Sums over episodes are re-associated: same tolerance.
 *   "wide", "tc", "tc_passes"   the GEMM-organised path for policies whose W1 does not fit in shared memory.
 *   "phase_clocks"  tuning aid, see qp_debug_phase_clocks (1: CTA 0's phase stamps; 2: per-CTA wall-clock stamps;
 *             3: fine stamps of the pair kernel).  "c2_dbg": reserved for timing experiments (no effect in this build).
 * Read-only: "fanout_active", "fused_active", "ep_tc_active" (a tensor-core persistent kernel: either form),
 * "ep_c2_active", "ep_c2_split_active" (the pair kernel on prices that are not tf32-exact), "p2p_active", "nvls_active",
 * "rsag_active", "fused_step_active", "tc_active", "tc_a_lo_zero": what the next qp_env_step would take.  With several GPUs all of
 * them are facts the ranks have AGREED on at qp_env_create (the step protocols do not interoperate). */
int qp_set_option(qp_handle* h, const char* key, int64_t value);
int qp_get_option(qp_handle* h, const char* key, int64_t* value);
/* Test hook for the tcgen05 tensor-core GEMM of the GEMM-organised path, on its own:
 * C[M,N] = A[M,K] · B[N,K]^T with host buffers; passes = 3 (3xTF32, fp32-grade) or 1 (plain TF32);
 * K must be a multiple of 32*splits.  ms_out (may be NULL) receives the device time of one launch. */
int qp_debug_tc_gemm(int32_t device, int64_t M, int64_t N, int64_t K, const float* A, const float* B, float* C,
                     int32_t passes, int32_t splits, float* ms_out);
/* Test hook for single tcgen05.mma instructions (kind::tf32) on operand images a test lays out byte by byte:
 * a_img / b_img are copied to shared memory (A at offset 0, B at the next 1024-byte boundary), desc_a / desc_b
 * are the 64-bit shared-memory matrix descriptors with their start-address field relative to the image, n_mma
 * instructions are issued (descriptor start addresses advanced by a_step / b_step 16-byte units each), and all
 * 128 lanes x n_cols columns of tensor memory (preset to `fill`) come back in out.  a_tmem_k != 0: the A operand comes
 * from tensor memory instead (a_img = [128 rows][a_tmem_k] fp32, written to columns a_col.. of lane = row; a_step then
 * counts columns).  Pins the operand layouts the
 * persistent kernels rely on (tests/test_gpu_umma_layouts.py).  No counterpart in the reference. */
int qp_debug_umma(int32_t device, const void* a_img, int64_t a_bytes, const void* b_img, int64_t b_bytes,
                  uint64_t desc_a, uint64_t desc_b, uint32_t idesc, int32_t n_mma, uint32_t a_step, uint32_t b_step,
                  uint32_t d_col, uint32_t d_lane, int32_t n_cols, float fill, float* out, int32_t a_tmem_k, uint32_t a_col);
/* Tuning aid: with option "phase_clocks" = 1 the persistent kernel's CTA 0 stamps clock64() at up
 * to 16 phase boundaries of every env-step of a launch (unused slots stay 0); this copies the
 * [n_steps x 16] stamps of the last launch to out. */
int qp_debug_phase_clocks(qp_handle* h, int32_t n_steps, int64_t* out);

/* Per-step record for parity checks; host arrays of n_steps x n_envs entries (x3 for q_*),
 * step-major.  Any member may be NULL.  No counterpart in the reference. */
typedef struct {
    int32_t* action_sel;  /* index the policy returned */
    int32_t* action_eff;  /* after demotion to Hold (TrainerChildActor.scala:118-123) */
    int32_t* explored;    /* 1 if the random branch was taken */
    int32_t* k_next;      /* argmax of q(nextState) (QDecisionPolicyActor.scala:69) */
    float* reward;        /* fp64 reward cast to fp32 (TrainerChildActor.scala:158) */
    float* q_s;           /* q(state) */
    float* q_n;           /* q(nextState) */
    float* target;        /* reward + gamma * q_n[k_next] (QDecisionPolicyActor.scala:70) */
    double* budget;       /* carry after the step */
    int32_t* shares;
} qp_trace;
int qp_env_step_trace(qp_handle* h, int32_t n_steps, const qp_trace* out);

/* budget + shares * shareValue per episode (TrainerChildActor.scala:68) -> TrainedData. */
int qp_env_get_portfolio(qp_handle* h, double* out);
/* The fold's carry (budget, shares, shareValue) (TrainerChildActor.scala:99-102); get / set.
 * Set also places every episode at env-step `step`.  Any pointer may be NULL. */
int qp_env_get_state(qp_handle* h, double* budget, int32_t* shares, double* share_value);
int qp_env_set_state(qp_handle* h, const double* budget, const int32_t* shares,
                     const double* share_value, int64_t step);
/* GetAvg / GetStd: mean and POPULATION standard deviation of the portfolios, reduced on the
 * device (TrainerRouterActor.scala:89-94, :149-151). */
int qp_env_portfolio_stats(qp_handle* h, double* mean, double* stddev);

/* TrainerChildActor.train (:64-71) in one call with host buffers: qp_env_create, every step of
 * the episode, qp_env_get_portfolio. */
int qp_train(qp_handle* h, const float* prices, int64_t series_len, int64_t n_series,
             const int32_t* series_of_env, int64_t n_envs, int64_t env_id_base, double budget0,
             int32_t shares0, double* portfolios_out);

/* ---- batching front-end for the router fan-out (TrainerRouterActor.scala:60-66, :88): many child
 * actors, on any threads, queue their SelectionAction / UpdateQ messages; one qp_flush answers all
 * queued selects with ONE launch (using the weights as they are at the flush) and then turns all
 * queued updates into ONE gradient-SUM step.  Enqueue calls are thread-safe and never launch GPU work: a message is one atomic slot
 * reservation plus a copy straight into the pinned staging buffer the flush uploads from (a queue that
 * has to grow allocates pinned host memory);
 * the size guards of QDecisionPolicyActor.scala:56-57, :64-65 apply at enqueue time (QP_EINVAL).
 * qp_take_action returns the action of a ticket from the most recent flush that held selects. */
int qp_enqueue_select(qp_handle* h, const float* state, int64_t state_len, float step, int64_t env_id,
                      int64_t* ticket);
int qp_enqueue_update(qp_handle* h, const float* state, int64_t state_len, float reward,
                      const float* next_state, int64_t next_state_len);
int qp_flush(qp_handle* h, int64_t* n_selects, int64_t* n_updates);
int qp_take_action(qp_handle* h, int64_t ticket, int32_t* action);
int64_t qp_pending(qp_handle* h);   /* queued messages; returns the count */

/* ---- policy checkpoint: weights, AdaGrad accumulators, UpdateQ counter and (if episodes exist) the
 * episodes' carry and step.  Fills the reference's empty saveSnapshot hook, called every 500
 * updates (QDecisionPolicyActor.scala:74, :91-93): qp_set_snapshot(path, 500) arms exactly that.
 * The RNG is counter-based, so (env id, step) in the checkpoint is all the RNG state there is. */
int qp_save_checkpoint(qp_handle* h, const char* path);
int qp_load_checkpoint(qp_handle* h, const char* path);
int qp_set_snapshot(qp_handle* h, const char* path, int64_t every);
int64_t qp_iteration(const qp_handle* h);   /* `iteration` of QDecisionPolicyActor.scala:54, :76 */

/* ---- price ingest: the text format of MSFT-stock-prices-revised.txt, `price, yyyy-MM-dd` per line,
 * parsed as SharePriceGetter.queryData does (SharePriceGetter.scala:89-101, :104-108): fields split on
 * ',' and trimmed, rows with fewer than two fields or an unparsable price/date dropped, a repeated
 * date keeps its last row, result sorted by date and narrowed to fp32 (TrainerChildActor.scala:66).
 * Host-only (needs no GPU).  out may be NULL to query the count. */
int qp_parse_prices(const char* text, size_t len, float* out, int64_t capacity, int64_t* n_out);

/* ---- several GPUs: episodes are sharded over ranks (one process per GPU, env_id_base = first
 * global episode of the shard), weights are replicated, and each env-step exchanges ONE thing: a SUM
 * all-reduce of the P-float gradient (NCCL over NVLink), after which every rank applies the same
 * AdaGrad step.  This is the synchronous form of the reference's children sharing one policy actor
 * (TrainerRouterActor.scala:44, :60-66); the reference itself has no collective.
 * qp_dist_unique_id: rank 0 obtains a 128-byte rendezvous id and hands it to the other ranks by
 * whatever channel the host has (torch.distributed, Akka remoting, a file).  qp_dist_init: every
 * rank joins; subsequent qp_env_step* calls all-reduce.  NCCL is resolved with dlopen at first use. */
#define QP_DIST_ID_BYTES 128
int qp_dist_unique_id(void* out, size_t out_bytes);
int qp_dist_init(qp_handle* h, const void* unique_id, int32_t rank, int32_t world);
int qp_dist_nccl_version(int32_t* version);
/* Optional, after qp_dist_init, ranks on ONE node (<= 8): gradient all-reduce INSIDE the persistent
 * episode kernel over NVLink peer memory instead of NCCL launches between kernels.  Each rank exports a
 * 64-byte CUDA IPC handle of its exchange buffer (qp_dist_p2p_export), the host gathers the world's
 * handles in rank order and gives them to every rank (qp_dist_p2p_import).  Per env-step every rank
 * then sends its share of the gradient sum into every peer's buffer and adds all ranks' shares in rank
 * order before AdaGrad — bit-identical on every rank.  Tensor-core kernel (the default): each value
 * travels with its step tag as ONE 64-bit store, the receiver polls its local slots, no fence and no
 * flag; FFMA kernel: whole-gradient buffers, a system fence and one flag per rank.  All ranks must step
 * in lock-step with the same n_steps.  If either call fails (no peer access) the NCCL path stays in use.
 * Call order with several GPUs: qp_dist_init, the p2p calls and every qp_set_option BEFORE qp_env_create —
 * qp_env_create is collective (one small NCCL MIN all-reduce) and fixes the execution path from what ALL
 * ranks can do (tf32-exact prices, tiles per SM or per CTA pair, peer access, multicast mapping, exchange
 * scheme), because the multi-GPU step protocols do not interoperate. */
#define QP_DIST_IPC_BYTES 64
int qp_dist_p2p_export(qp_handle* h, void* out, size_t out_bytes);
int qp_dist_p2p_import(qp_handle* h, const void* handles, int32_t world);
/* The same exchange over blocks the HOST runtime has set up (instead of export / import): block_ptrs[r] is rank r's block
 * of qp_dist_p2p_bytes(h) bytes as mapped in THIS process (zero-filled, 256-byte aligned, the caller keeps them alive
 * until qp_destroy), multicast_ptr the address of a multicast (NVLS) mapping of the same blocks, or 0.  With a
 * multicast address the pair kernel sends each share as ONE store that the NVSwitch delivers to every rank, instead of
 * world - 1 stores (option "nvls" = 0 keeps the unicast stores; "nvls_active" reads what the next step takes).  No
 * counterpart in the reference. */
int64_t qp_dist_p2p_bytes(const qp_handle* h);
int qp_dist_p2p_attach(qp_handle* h, const uint64_t* block_ptrs, int32_t world, uint64_t multicast_ptr);

/* ---- allocation-free variant.  No counterpart in the reference.
 *
 * qp_td_grad_noalloc computes what qp_update_q computes up to, but not including, the AdaGrad
 * step: the TD gradient of B (state, reward, nextState) rows, summed over the rows, written to
 * d_grad_out [P floats].  It exists for callers that must not touch the CUDA allocator while they
 * are in here: an actor on a dispatcher shared with other GPU work (cudaMalloc and cudaFree
 * serialise the whole device, and the handle-based calls above grow their scratch on demand), or a
 * process that has given all device memory to another runtime's pool and can only pass pointers
 * taken from that pool.  Every pointer argument is a DEVICE pointer owned by the caller; d_params
 * has the usual layout [W1 | W2].  The call makes no cudaMalloc / cudaFree and uses no library that
 * allocates.
 *
 * Its working memory is a 4 MiB array built into the library (each GPU that loads libqpolicy.so
 * holds those 4 MiB until the process exits).  Sizes that need more are refused with QP_EINVAL and
 * a message that states both numbers: at the reference shape (window 201, hidden 200) batches
 * of up to 1792 rows fit.  All calls on a GPU share that one array, so two calls whose kernels overlap
 * on the same GPU (different streams or threads) race: serialise them.  Contents left by one call
 * are never read by the next — every word consumed is rewritten first — so nothing needs clearing.
 * The first call on each device also does that device's one-time set-up (it resolves the array's
 * device address and raises the kernels' shared-memory limit).  `stream` is a cudaStream_t (NULL =
 * default stream); the call synchronises that stream before it returns, so it cannot be part of a
 * stream capture.  Alignment: d_params 16 bytes (the kernels fetch W1 in 16-byte granules), every
 * other pointer 4 bytes; anything else is refused with QP_EINVAL.  The calling thread's current
 * device is the same after the call as before it. */
int qp_td_grad_noalloc(int32_t device, const qp_config* cfg, const float* d_params,
                       const float* d_states, const float* d_rewards, const float* d_next_states,
                       int64_t B, float* d_grad_out, void* stream);

/* ---- measurement aids (no reference counterpart) */
enum { QP_KERNEL_STEP = 0, QP_KERNEL_DW1 = 1, QP_KERNEL_APPLY = 2, QP_N_KERNELS = 3 };
/* The CUDA stream (cudaStream_t) every kernel of this handle is launched on. */
void* qp_stream(qp_handle* h);
/* Runs n_steps with CUDA events around each kernel; ms_sum[QP_N_KERNELS] receives the summed
 * device time per kernel in milliseconds. */
int qp_env_step_timed(qp_handle* h, int32_t n_steps, float* ms_sum);
/* Measured fp32-FMA throughput of the handle's GPU in TFLOP/s (dependency-free FFMA streams):
 * the roof that binds the fp32 path; MEASURED_PEAKS.json only carries HBM and bf16 tensor peaks. */
int qp_measure_fp32_peak(qp_handle* h, double* tflops);
/* The same for the FFMA pattern a register-tiled contraction issues (8 x 8 outer product per step,
 * three register sources per FFMA): the attainable roof of the fp32 kernels of this library. */
int qp_measure_fp32_outer_peak(qp_handle* h, double* tflops);
/* Kernels launched by this handle so far.  Returns the count. */
int64_t qp_launch_count(const qp_handle* h);

/* Text of the last failure on this thread ("" if none).  Never NULL. */
const char* qp_last_error(void);
const char* qp_version(void);

#ifdef __cplusplus
}
#endif
#endif
