/*
 * apexb200.h — C ABI of libapexb200.so: B200 (sm_100a) prioritized experience
 * replay and learner data path for Ape-X DQN.
 *
 * The reference (lefarov/Ape-X-DQN) is pure Python and has no FFI; the boundary
 * this library sits behind is the Python class surface of replay.py / learner.py
 * / actor.py.  Each entry point below names the reference code it replaces
 * (file:line in the reference repository).  The host-side mirror classes in
 * ape-x-dqn_b200/ bind these symbols with ctypes.
 *
 * Conventions
 *   - every function returns 0 (APEXB200_OK) or a negative APEXB200_E* code and
 *     never throws; apexb200_last_error() gives a per-thread message;
 *   - pointers are DEVICE pointers unless the name ends in _host; the caller
 *     (PyTorch) owns every input/output buffer and keeps it alive until the
 *     stream has run past the call;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *     the per-step entry points (sample*, gather, td_*, tree_update*, shard_exchange,
 *     sample_gather_sharded, nstep, initial_priority, is_weights, params_publish/load,
 *     actor_policy) only enqueue work: they never synchronise and never allocate.
 *     Allocation or synchronisation happens only in replay_create/destroy,
 *     params_create/destroy, replay_status, peer_connect (maps peer memory) and in the
 *     first call of host_step (copy streams + staging);
 *   - one replay handle is driven from one host thread at a time; calls on one
 *     stream are totally ordered, which is what gives add/sample/set_priorities
 *     their reference (sequential) semantics.
 *
 * Fixed-point priorities: a leaf of the sum tree is the unsigned 64-bit integer
 *     leaf = min(floor(double(float(priority^alpha)) * 2^24), 2^36 - 1)
 * so every tree sum is exact and independent of reduction order (deterministic),
 * and the proportional search is bit-exact against an integer
 * searchsorted(cumsum(leaves), target, side='right').
 */
#ifndef APEXB200_H_
#define APEXB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define APEXB200_VERSION 100            /* 0.1.0 */

#define APEXB200_OK        0
#define APEXB200_EINVAL   -1            /* bad argument (null pointer, size, alignment) */
#define APEXB200_ECUDA    -2            /* CUDA runtime error (see last_error) */
#define APEXB200_ENOMEM   -3            /* device allocation failed */
#define APEXB200_ERANGE   -4            /* size beyond a documented limit */

#define APEXB200_FRAC_BITS   24
#define APEXB200_LEAF_MAX    ((1ull << 36) - 1)
#define APEXB200_MAX_CAPACITY (1ll << 24)   /* records per handle (shard) */
#define APEXB200_MAX_ACTIONS 64

#define APEXB200_HOST_STEP_SLOTS 4           /* buffer sets of apexb200_host_step */
#define APEXB200_TD_NOALLOC_MAX_B 262144    /* rows per call of apexb200_td_fwd_noalloc */

#define APEXB200_FMT_U8  0              /* gathered batches as stored: uint8 [rows, C, F] */
#define APEXB200_FMT_F32 1              /* gathered batches widened to float32 [rows, C, F] (learner.py:37-38) */

#define APEXB200_LOSS_HALF_SQ 0         /* 1/2 delta^2, learner.py:48 */
#define APEXB200_LOSS_HUBER   1         /* Huber(kappa) */

/* device-side status bits, accumulated in the handle (replay_status) */
#define APEXB200_ST_EMPTY_TREE   1      /* sample on a tree whose total is zero */
#define APEXB200_ST_BAD_INDEX    2      /* record index outside [0, capacity) */
#define APEXB200_ST_BAD_TARGET   4      /* target >= total in sample_targets */
#define APEXB200_ST_BAD_ACTION   8      /* action outside [0, num_actions) */
#define APEXB200_ST_PEER_TIMEOUT 16     /* shard_exchange: a peer never delivered its total */
#define APEXB200_ST_QUOTA        32     /* sharded sample: the caller's output arrays (max_rows < global_count) were too small */

#define APEXB200_MAX_SHARDS       16
#define APEXB200_IPC_HANDLE_BYTES 64

typedef struct apexb200_replay *apexb200_replay_t;

/* Layout of one replay shard in HBM.  capacity records are split into `lanes`
 * equal rings (one per actor group); each record owns `frames_per_record`
 * consecutive frame slots and may also reference frames of up to `lookback`
 * earlier records of its lane (stream layout: frames_per_record = 1, lookback =
 * frames_per_state - 1 + n_step; materialised layout used by the reference-shaped
 * add(): frames_per_record = 2 * frames_per_state, lookback = 0).
 * Replaces the python list/dicts of ReplayMemory.__init__ (replay.py:9-16) and
 * the N_Step_Transition record (replay.py:5). */
typedef struct apexb200_replay_config {
    int64_t capacity;            /* records; multiple of lanes; <= APEXB200_MAX_CAPACITY */
    int32_t lanes;
    int32_t frames_per_record;
    int32_t lookback;            /* records */
    int32_t frames_per_state;    /* C */
    int32_t frame_bytes;         /* 7056 for 84x84 uint8; multiple of 16 */
    int32_t num_actions;         /* >0: also store the actor's qS_t/qS_tpn rows */
    float   alpha;               /* priority exponent, params['priority_exponent'] replay.py:14 */
    int32_t device;              /* CUDA device ordinal */
} apexb200_replay_config;

int         apexb200_version(void);
const char *apexb200_strerror(int code);
const char *apexb200_last_error(void);

/* ---- lifetime: ReplayMemory.__init__ (replay.py:9-16) ------------------- */
int apexb200_replay_create(apexb200_replay_t *out, const apexb200_replay_config *cfg);
int apexb200_replay_destroy(apexb200_replay_t h);
/* ReplayMemory.size (replay.py:82-83): records held, min(inserted, capacity) per lane. */
int apexb200_replay_size(apexb200_replay_t h, int64_t *size_host);
int apexb200_replay_device_bytes(apexb200_replay_t h, int64_t *bytes_host);
/* Reads and clears the device status word (synchronises `stream`). */
int apexb200_replay_status(apexb200_replay_t h, int32_t *status_host, void *stream);

/* ---- K6 insert: ReplayMemory.add (replay.py:59-69) + Actor.compute_priorities
 * (actor.py:127-143) + FIFO eviction (replay.py:71-80, as ring overwrite) ------
 * Appends `count` records to ring `lane`.  frames [count*frames_per_record,
 * frame_bytes] u8; rel [count, 2C] int32 offsets (<= 0, frame units, relative to
 * the record's own last frame, wrapping inside the lane) naming the C frames of
 * S_t then the C frames of S_tpn; A/R/Gamma [count].  prio = raw |TD| priorities
 * [count] or NULL, in which case |R + Gamma*max_a q_tpn - q_t[A]| is computed from
 * q_t/q_tpn [count, num_actions] (which are also stored when the handle was
 * created with num_actions > 0).  The `lookback` records after the new head are
 * invalidated (leaf = 0) because frames they reference were overwritten.
 * first_slot_host (nullable) receives the record index of the first new record. */
int apexb200_replay_insert(apexb200_replay_t h, int32_t lane, int32_t count,
                           const uint8_t *frames, const int32_t *rel,
                           const int32_t *A, const float *R, const float *Gamma,
                           const float *prio, const float *q_t, const float *q_tpn,
                           int64_t *first_slot_host, void *stream);

/* The same for `n_lanes` distinct lanes at once, `count` records each (E lock-stepped actors: one
 * launch pair instead of E).  lanes_host [n_lanes] (host); every device input carries a leading
 * lane dimension: frames [n_lanes, count*frames_per_record, frame_bytes], rel [n_lanes, count, 2C],
 * A/R/Gamma/prio [n_lanes, count], q_t/q_tpn [n_lanes, count, num_actions]. */
int apexb200_replay_insert_lanes(apexb200_replay_t h, const int32_t *lanes_host, int32_t n_lanes, int32_t count,
                                 const uint8_t *frames, const int32_t *rel,
                                 const int32_t *A, const float *R, const float *Gamma,
                                 const float *prio, const float *q_t, const float *q_tpn, void *stream);

/* ---- K1 build: ReplayMemory.update_sample_probabilities (replay.py:18-30) --
 * Recomputes every inner level from the leaves, level by level. */
int apexb200_tree_build(apexb200_replay_t h, void *stream);

/* ---- K2 point update: ReplayMemory.set_priorities (replay.py:32-42) -------
 * idx [count] record indices; duplicates resolve as "last in batch order wins"
 * (dict.update, replay.py:40).  tree_update takes raw priorities |delta| and
 * stores max(|delta|, eps)^alpha; tree_set_leaves stores exact u64 leaves. */
int apexb200_tree_update(apexb200_replay_t h, const int32_t *idx, const float *prio,
                         int32_t count, float eps, void *stream);
int apexb200_tree_set_leaves(apexb200_replay_t h, const int32_t *idx, const uint64_t *leaves,
                             int32_t count, void *stream);
/* Sum of all leaves (the normaliser of replay.py:24) -> *total (device u64). */
int apexb200_tree_total(apexb200_replay_t h, uint64_t *total, void *stream);
/* Tree geometry: number of levels (leaves = level 0) and entries per level. */
int apexb200_tree_geometry(apexb200_replay_t h, int32_t *n_levels_host, int64_t *sizes_host /*[8]*/);
/* Copy `count` nodes of `level` starting at `offset` to out (device or host). */
int apexb200_tree_read(apexb200_replay_t h, int32_t level, int64_t offset, int64_t count,
                       uint64_t *out, void *stream);
/* Bulk restore: overwrite leaves [offset, offset+count) (then call tree_build). */
int apexb200_tree_load_leaves(apexb200_replay_t h, const uint64_t *leaves, int64_t offset,
                              int64_t count, void *stream);

/* ---- K3 proportional search: the np.random.choice call of ReplayMemory.sample
 * (replay.py:52-53) ---------------------------------------------------------
 * u [count] fp64 uniform draws in [0,1) (supplied by the caller so that both
 * sides consume the same stream); one warp per draw descends the 32-ary tree.
 * idx_out[j] = #{i : cumsum(leaves)[i] <= target_j}, target_j =
 * min(trunc(u_j * double(total)), total-1).  leaf_out (nullable) = the selected
 * leaf, total_out (nullable, 1 element) = tree total (for importance weights).
 * fsel_out (nullable) [count, 2C] = the frame slots of the selected records. */
int apexb200_sample(apexb200_replay_t h, const double *u, int32_t count, int32_t *idx_out,
                    uint64_t *leaf_out, uint64_t *total_out, int32_t *fsel_out, void *stream);
/* K3 with the PER stratification done in the kernel (the reference draws i.i.d. from
 * numpy's global RNG, replay.py:52; stratified draws are the Ape-X/PER paper's variant
 * of the same law): u_j = (j + v_j)/count, v_j = splitmix64(seed, epoch, call, j) * 2^-53.
 * `epoch` is a device-resident counter that every tree-changing kernel of this handle
 * (insert, tree_update, tree_set_leaves, td_fwd_update) advances by one, so a CUDA-graph
 * replay of a sample -> ... -> update step draws fresh numbers without host state;
 * `call` counts this handle's stratified calls on the host and separates several samples
 * taken between two updates (it is baked into a captured launch).  u_out (nullable)
 * [count] receives the draws so that a checker can replay them. */
int apexb200_sample_stratified(apexb200_replay_t h, uint64_t seed, int32_t count, int32_t *idx_out,
                               uint64_t *leaf_out, uint64_t *total_out, int32_t *fsel_out, double *u_out,
                               void *stream);
/* Same search from explicit integer targets (sharded sampling); count_dev
 * (nullable) is a device int32 holding the live count <= max_count. */
int apexb200_sample_targets(apexb200_replay_t h, const uint64_t *targets, const int32_t *count_dev,
                            int32_t max_count, int32_t *idx_out, uint64_t *leaf_out,
                            int32_t *fsel_out, void *stream);

/* ---- K4 gather: the batch assembly of ReplayMemory.sample (replay.py:54-56)
 * and the np.array stacking of learner.py:35-41 -----------------------------
 * Copies, for each of `count` record indices (row order preserved), the C frames
 * of S_t and of S_tpn into S_t/S_tpn [count, C, frame_bytes] u8 and A/R/Gamma
 * (and q_t/q_tpn if stored and non-NULL).  fsel (nullable) = frame slots from
 * apexb200_sample, which saves one dependent lookup.  count_dev as above. */
int apexb200_gather(apexb200_replay_t h, const int32_t *idx, const int32_t *fsel, int32_t count,
                    const int32_t *count_dev, uint8_t *S_t, uint8_t *S_tpn, int32_t *A, float *R,
                    float *Gamma, float *q_t, float *q_tpn, void *stream);
/* 0 = per-warp 16-byte vector copy, 1 (default) = TMA bulk copy staged through shared memory.
 * Measurement aid for the fused kernel only: 3 = the default kernel with per-warp timestamps, read back
 * (synchronously) with apexb200_debug_timeline: out_host [units][20] u64 = global timer at entry, SM-clock
 * offsets of {dependency wait done, search done, loads issued, first / last frame landed, stores issued,
 * staging free, writes complete}, global timer at exit, SM id, CTA id, then SM-clock offsets of {draw formed,
 * top level(s) searched, each further tree level searched (up to 5, 0 = unused)} (units = half draws, <= 2048). */
int apexb200_set_gather_variant(int32_t variant);
int apexb200_debug_timeline(uint64_t *out_host, int32_t units);
/* out_host [8] u64: global-timer stamps of the last single-CTA TD/update launch — [0] entry, [1] dependency wait
 * done, [3] rows evaluated, [4] leaves written, [5] ancestor updates issued, [2] end of thread 0, [6] end of the
 * last thread; all zero unless the library was built with -DAPEX_TD_TIMELINE (tools/timeline.py). */
int apexb200_debug_td_timeline(uint64_t *out_host);
/* K3 + K4 in one launch: ReplayMemory.sample (replay.py:44-57) as a whole.  u [count]
 * as in apexb200_sample, or NULL for the in-kernel stratified draws of
 * apexb200_sample_stratified (seed, u_out).  One warp per half draw (S_t or S_tpn): the
 * warp descends the tree, then moves the C frames of its half with the TMA engine.  Needs
 * frames_per_state <= 8 and C frames fitting in shared memory (otherwise the two kernels
 * run back to back).  Outputs as in
 * apexb200_sample + apexb200_gather (fsel_out, leaf_out, total_out, u_out, A, R, Gamma,
 * q_t, q_tpn nullable).  out_format: APEXB200_FMT_U8 — S_t/S_tpn are uint8 [count, C, frame_bytes]
 * as stored; APEXB200_FMT_F32 — float32 arrays of the same shape, every byte widened exactly
 * (the torch.from_numpy(...).float() of learner.py:37-38 done by the gather itself: frames land in
 * shared memory by bulk copy, the SM converts and writes 16-byte vectors; fused kernel only). */
int apexb200_sample_gather(apexb200_replay_t h, const double *u, uint64_t seed, int32_t count,
                           int32_t *idx_out, int32_t *fsel_out, uint64_t *leaf_out, uint64_t *total_out,
                           double *u_out, void *S_t, void *S_tpn, int32_t *A, float *R, float *Gamma,
                           float *q_t, float *q_tpn, int32_t out_format, void *stream);
/* Programmatic dependent launch for the per-step kernels (default 1): each kernel is
 * scheduled while its predecessor on the stream drains and blocks on
 * griddepcontrol.wait before touching memory.  0 = plain stream-ordered launches. */
int apexb200_set_pdl(int32_t enabled);
/* K5 / K5+K2 for B <= 1024 as ONE thread-block cluster of ceil(B/128) CTAs instead of one CTA:
 * the rows are spread over up to 8 SMs, the loss partials and the multi-GPU total delta meet in
 * the first CTA's shared memory (distributed shared memory), and cluster barriers order the
 * duplicate-resolution phases.  Default 0: measured on B200 at B=512 the two cluster barriers
 * cost more than the extra SMs save (4.7 vs 3.8 us per launch, 16.6 vs 16.2 us per step). */
int apexb200_set_td_cluster(int32_t enabled);

/* ---- K5 loss / TD / priority: Learner.compute_loss_and_priorities
 * (learner.py:43-52) ---------------------------------------------------------
 * Inputs are the three Q tensors the unchanged model produced, [B, num_actions]
 * fp32 row-major: q_on_t = Q(S_t), q_on_tpn = Q(S_tpn), q_tg_tpn = Q_target(S_tpn).
 * Outputs [B]: delta, prio = max(|delta|, eps)^alpha, loss; loss_mean [1] summed
 * in a fixed order.  workspace: apexb200_td_workspace_bytes(B) bytes of device
 * scratch owned by the caller (256-byte aligned), contents need not be preserved. */
int apexb200_td_workspace_bytes(int32_t B, int64_t *bytes_host);
int apexb200_td_fwd(const float *q_on_t, const float *q_on_tpn, const float *q_tg_tpn,
                    const int32_t *A, const float *R, const float *Gamma, int32_t B,
                    int32_t num_actions, float alpha, float eps, int32_t loss_kind, float kappa,
                    const float *weights, float *delta, float *prio, float *loss, float *loss_mean,
                    void *workspace, int64_t workspace_bytes, void *stream);
/* Importance-sampling weights of prioritized replay (parameters.json:30
 * importance_sampling_exponent, README.md:25 TODO: never read by the reference):
 * weights[j] = (n_records * leaf[j] / total)^(-beta) / max_j, from the leaf and total
 * outputs of the sample calls.  Passed as `weights` (nullable everywhere) they multiply the
 * per-record loss and its gradient in td_fwd / td_fwd_update / td_bwd. */
int apexb200_is_weights(const uint64_t *leaf, const uint64_t *total, int32_t B, int64_t n_records,
                        float beta, float *weights, void *stream);
/* apexb200_td_fwd without a workspace argument.  (No counterpart in the reference: it
 * exists for the way this library gets embedded.)  The loss reduction needs a few KB of
 * device scratch.  Some callers have nowhere to get it from: a C or ctypes binding that
 * is handed input and output pointers by its host framework but is not itself linked to
 * the CUDA runtime cannot call cudaMalloc; a learner that has already given the whole
 * device to its framework's allocator (or to a replay shard sized to fill HBM) may find
 * cudaMalloc failing; and a steady-state actor/learner loop that must not stall other
 * streams cannot afford cudaMalloc/cudaFree, which synchronise the device.  For them the
 * scratch is compiled into the library: APEXB200_TD_NOALLOC_MAX_B/256 floats + 256 bytes
 * (4352 bytes) of static device memory, held by every GPU that loads the library until
 * the process exits.  The call never calls cudaMalloc, cudaFree or any allocating
 * routine; inputs and outputs are still the caller's.
 *   - B > APEXB200_TD_NOALLOC_MAX_B returns APEXB200_ERANGE (use apexb200_td_fwd).
 *   - The device that owns the pointers and the stream must be current.  The first call on
 *     each device looks up the scratch address (cudaGetSymbolAddress) and caches it; make
 *     that first call before starting any stream capture.
 *   - All calls on a GPU share the one scratch, whatever stream they use: two calls that
 *     can overlap on the GPU race on it, so keep them on one stream or order them with
 *     events.  The scratch keeps the previous call's contents; the call itself clears the
 *     one word that has to start at zero.
 * Results equal apexb200_td_fwd (bit for bit when B > 1024, where both run the same kernel). */
int apexb200_td_fwd_noalloc(const float *q_on_t, const float *q_on_tpn, const float *q_tg_tpn,
                            const int32_t *A, const float *R, const float *Gamma, int32_t B,
                            int32_t num_actions, float alpha, float eps, int32_t loss_kind, float kappa,
                            float *delta, float *prio, float *loss, float *loss_mean, void *stream);
/* Gradient of loss_mean w.r.t. q_on_t (loss.backward() through learner.py:46-48):
 * grad_q [B, num_actions], zero except column A_b. grad_loss_mean: device scalar.
 * count_dev / mean_divisor as in td_fwd_update (NULL / 0 for a plain batch). */
int apexb200_td_bwd(const float *delta, const int32_t *A, const float *grad_loss_mean, int32_t B,
                    int32_t num_actions, int32_t loss_kind, float kappa, const int32_t *count_dev,
                    float mean_divisor, const float *weights, float *grad_q, void *stream);
/* K5 + K2 fused: td_fwd, then the new priorities are written into the leaves of
 * idx [B] and their ancestors refreshed (learner.py:70 + :76 in one call).
 * Sharded batches: count_dev (nullable) is a device int32 with the
 * number of live rows <= B (the rank's data-dependent quota; rows beyond it are
 * ignored) and mean_divisor (0 = B) is the GLOBAL batch size, so that summing
 * loss_mean / gradients over ranks gives the single-batch mean.  With count_dev and
 * mean_divisor the arrays may hold the whole global batch (B up to the global count):
 * the single-CTA kernel is sized for the rank's fair share (mean_divisor / world) and
 * takes further passes when the rank owns more, in batch order ("last row wins" holds
 * across passes).  An action outside [0, num_actions) is evaluated for the clamped column
 * (forward and backward alike) and raises APEXB200_ST_BAD_ACTION. */
int apexb200_td_fwd_update(apexb200_replay_t h, const int32_t *idx,
                           const float *q_on_t, const float *q_on_tpn, const float *q_tg_tpn,
                           const int32_t *A, const float *R, const float *Gamma, int32_t B,
                           int32_t num_actions, float eps, int32_t loss_kind, float kappa,
                           const float *weights, float *delta, float *prio, float *loss, float *loss_mean,
                           void *workspace, int64_t workspace_bytes,
                           const int32_t *count_dev, float mean_divisor, void *stream);

/* ---- host-facing step: learner.py:68 (sample) -> :70 (loss, priorities) -> :76
 * (set_priorities) for a caller that keeps its numbers on the host, as the reference does.
 * u_host [B] fp64 draws and q_host [3, B, num_actions] fp32 (Q(S_t), Q(S_tpn), Q_target(S_tpn))
 * are in pinned HOST memory (one H2D copy if q_host == u_host + B); S_t/S_tpn/A/R/Gamma are
 * the learner's device batch tensors; result_host [4B+4] fp32, pinned, receives
 * delta[B] | priority[B] | loss[B] | record index[B] (int32 bits) | loss_mean.  The kernels
 * run on `stream`; the two copies run on copy streams owned by the handle, so with slot cycling
 * through 0 .. APEXB200_HOST_STEP_SLOTS-1 (or a prefix of them) the H2D copies of later steps
 * and the D2H copies of earlier ones overlap the kernels.  The D2H copy is ordered after the
 * kernels by an event; the arrival of the inputs is signalled by a stream-ordered write of a
 * device word behind the H2D copy, which the sampling kernel polls (no cross-stream wait on
 * `stream`, so the kernels keep chaining by programmatic dependent launch across steps).  done_event (cudaEvent_t, nullable) is recorded
 * after the D2H copy; the call returns without synchronising.  A slot's host buffers may be
 * refilled once its done_event has completed.  B <= 1024.  The first call (and any call
 * with a larger B) creates the copy streams and allocates the device staging. */
int apexb200_host_step(apexb200_replay_t h, int32_t slot, const double *u_host, const float *q_host, int32_t B,
                       int32_t num_actions, float eps, int32_t loss_kind, float kappa,
                       uint8_t *S_t, uint8_t *S_tpn, int32_t *A, float *R, float *Gamma,
                       float *result_host, void *done_event, void *stream);

/* The same step for one rank of a sharded replay (learner.py:68 -> :70 -> :76 over N GPUs; the sampling side is
 * apexb200_sample_gather_sharded, so the draws are generated on the device, identically on every rank, and no
 * draws travel).  q_host [3, max_rows, num_actions] fp32, pinned: row r of each block belongs to the r-th row this
 * rank owns; S_t/S_tpn/A/R/Gamma are device batch arrays of max_rows rows.  result_host [5*max_rows + 4] fp32,
 * pinned, receives delta[max_rows] | priority | loss | record index (int32 bits) | global draw index (int32
 * bits) | loss_mean (this rank's part: sum of its rows' losses / global_count) | rows owned (int32 bits) | 2 pad
 * words; only the first `rows owned` entries of each array are meaningful.  max_rows is the caller's row
 * capacity: with max_rows >= global_count no draw can be lost whatever the split of priority mass; with a smaller
 * max_rows a rank that owns more keeps the first max_rows, sets APEXB200_ST_QUOTA and reports max_rows.  Needs
 * the peers connected (apexb200_peer_connect[_local]) and push-on-update or an exchange per step; the TD side
 * runs as the single-CTA kernel (max_rows <= 1024, or a fair share global_count/world of at most ~900 rows).
 * Streams, slots, done_event and ordering as for apexb200_host_step. */
int apexb200_host_step_sharded(apexb200_replay_t h, int32_t slot, uint64_t seed, int32_t global_count, int32_t max_rows,
                               const float *q_host, int32_t num_actions, float eps, int32_t loss_kind, float kappa,
                               uint8_t *S_t, uint8_t *S_tpn, int32_t *A, float *R, float *Gamma,
                               float *result_host, void *done_event, void *stream);

/* ---- K7 n-step return: ExperienceBuffer.update_buffer (actor.py:29-43),
 * intended semantics, vectorised over E actors x T steps --------
 * R[e,t] = sum_{k<m} gamma^k r[e,t+k], Gamma = gamma^m, or 0 when a terminal lies
 * in the window; span = m, or 0 while the window is still incomplete. */
int apexb200_nstep(const float *rewards, const uint8_t *dones, int32_t E, int32_t T, int32_t n,
                   float gamma, float *R, float *Gamma, int32_t *span, void *stream);
/* Actor.compute_priorities (actor.py:127-143): |R + Gamma*max_a q_tpn - q_t[A]|. */
int apexb200_initial_priority(const float *q_t, const float *q_tpn, const int32_t *A,
                              const float *R, const float *Gamma, int32_t count,
                              int32_t num_actions, float *abs_td, void *stream);

/* ---- device-side ingest of actor step streams: replaces, for E environments that each own
 * one lane of a frame-once ("stream") replay, ExperienceBuffer.add / update_buffer / construct_nstep_transition
 * (actor.py:29-70), Actor.compute_priorities (actor.py:127-143), the queue pump (main.py:21-25) and
 * ReplayMemory.add (replay.py:59-69).  The n-step carry (last n_step actions, rewards, dones, Q rows), the
 * episode ages and the ring heads of every environment live on the device, so a push is a fixed sequence of
 * launches (frame copy, record build + insert, state roll) with no host arithmetic between them.
 * The frame seen at env step tau is stored with the record of step t = tau - n_step: A_t, R = sum_k gamma^k
 * r_{t+k}, Gamma = gamma^n_step or 0 when the episode ended inside the window, initial priority
 * |R + Gamma max_a Q(s_tau) - Q(s_t)[A_t]|; S_t / S_tpn are the C frames ending at t / tau, repeating the
 * episode's first frame where the episode is younger.  The first n_step frames of a stream carry no record.
 * create: lanes_host [num_envs] distinct lanes (heads are taken over from the handle); max_chunk_steps > 0
 * also allocates device staging for push_blobs.  Allocates and synchronises; push / push_blobs only enqueue. */
typedef struct apexb200_ingest *apexb200_ingest_t;
int apexb200_ingest_create(apexb200_replay_t h, int32_t num_envs, const int32_t *lanes_host, int32_t n_step, float gamma,
                           int32_t num_actions, int32_t max_chunk_steps, apexb200_ingest_t *out);
int apexb200_ingest_destroy(apexb200_ingest_t g);
/* One chunk of T env steps for n_envs environments (envs_host [n_envs] distinct environment numbers, NULL = the
 * first n_envs environments in order).  Device inputs with a leading environment dimension: frames [n_envs, T,
 * frame_bytes] u8 (s_tau), actions [n_envs, T] (taken at s_tau), rewards [n_envs, T], dones [n_envs, T] u8
 * (1 = the episode ended with that step), q [n_envs, T, num_actions] = Q(s_tau).  T + lookback <= lane capacity. */
int apexb200_ingest_push(apexb200_ingest_t g, const int32_t *envs_host, int32_t n_envs, int32_t T, const uint8_t *frames,
                         const int32_t *actions, const float *rewards, const uint8_t *dones, const float *q, void *stream);
/* The same for chunks that sit in page-locked, device-mapped HOST memory (apexb200_host_register_mapped), one blob
 * per environment as actor processes write them: frames [T, frame_bytes] at offset 0, then actions i32 [T],
 * rewards f32 [T], dones u8 [T], q f32 [T, num_actions] at the given byte offsets.  blobs_host [n_envs] are the
 * DEVICE-visible addresses of the blobs.  One pack kernel reads all of them over PCIe into the ingest's staging,
 * then the three launches of ingest_push follow: no per-environment copy calls.  The blobs may be reused once
 * work enqueued on `stream` up to here has completed (record an event). */
int apexb200_ingest_push_blobs(apexb200_ingest_t g, const int32_t *envs_host, const uint64_t *blobs_host, int32_t n_envs,
                               int32_t T, int32_t off_actions, int32_t off_rewards, int32_t off_dones, int32_t off_q,
                               void *stream);
/* The handle's host-side ring heads follow every push call; after replaying a CUDA graph that contains pushes
 * (the host was not involved) copy the device-side heads back with this (synchronises `stream`). */
int apexb200_ingest_sync_heads(apexb200_ingest_t g, void *stream);

/* ---- K8 shard routing (no reference counterpart: the reference is single
 * process; this is the multi-GPU split of ReplayMemory.sample) ----
 * totals [G] = all-gathered shard totals.  For each draw u_j the global target is
 * min(trunc(u_j*double(T)), T-1); this rank keeps the draws whose target falls in
 * its prefix interval, in increasing j: local_targets/local_pos [count] (target
 * minus the rank's prefix, and j), local_count [1]. */
int apexb200_shard_targets(const uint64_t *totals, int32_t G, int32_t rank, const double *u,
                           int32_t count, uint64_t *local_targets, int32_t *local_pos,
                           int32_t *local_count, void *stream);

/* ---- multi-GPU: exchange of shard totals over peer memory, fused sharded sample
 * (no reference counterpart).  One process per GPU, one handle per rank.
 * peer_export writes a CUDA-IPC handle (APEXB200_IPC_HANDLE_BYTES) of this shard's
 * exchange buffer; the host all-gathers the handles (torch.distributed) and gives every
 * rank the full list [world][64] through peer_connect, which maps the peers' buffers
 * (NVLink peer access).  world <= APEXB200_MAX_SHARDS. */
int apexb200_peer_export(apexb200_replay_t h, void *ipc_handle_host);
int apexb200_peer_connect(apexb200_replay_t h, int32_t rank, int32_t world, const void *ipc_handles_host);
/* The same for shards that live in ONE process (one handle per shard, on one GPU or on several GPUs
 * with direct peer access): shards_host [world] are the handles of all shards in rank order,
 * shards_host[rank] == h.  No IPC handles are involved.  Kernels of different shards must not be
 * made to wait for one another on one GPU: issue apexb200_shard_push (not shard_exchange) for every
 * shard before the sharded samples, and keep push-on-update steps in lock step (all samples, then
 * all updates), so that every word a kernel polls for has already been written. */
int apexb200_peer_connect_local(apexb200_replay_t h, int32_t rank, int32_t world, const apexb200_replay_t *shards_host);
/* One launch replacing tree_total + all-gather: every rank writes {total, round} into
 * every peer's buffer and waits for all entries of the round.  Collective: every rank
 * must call it the same number of times.  The totals are kept in the handle for
 * apexb200_sample_gather_sharded and copied to totals_out (device u64 [world], nullable). */
int apexb200_shard_exchange(apexb200_replay_t h, uint64_t *totals_out, void *stream);
/* The publishing half of shard_exchange alone: start a new exchange round and write this shard's total
 * into every peer's buffer without waiting for theirs (what td_fwd_update does with push-on-update). */
int apexb200_shard_push(apexb200_replay_t h, void *stream);
/* Attribution of the multi-GPU step: stats_host [2] u64 = {SM clocks the sharded sampling kernels of this handle
 * spent waiting for their peers' totals (slowest peer per launch, summed), number of launches}; reset != 0 clears
 * the counters.  Synchronises `stream`. */
int apexb200_peer_wait_stats(apexb200_replay_t h, uint64_t *stats_host, int32_t reset, void *stream);
/* Fused compute + collective: with this enabled every apexb200_td_fwd_update on the handle
 * also publishes the shard's new total to all peers from the kernel that produced it (as a
 * new exchange round), so the steady-state step is sample_gather_sharded -> model ->
 * td_fwd_update with no exchange launch in between.  Every rank must then issue the same
 * sequence of td_fwd_update / shard_exchange calls (collective discipline); after inserts,
 * which change totals at different rates per rank, call apexb200_shard_exchange on every
 * rank before the next sharded sample. */
int apexb200_peer_push_on_update(apexb200_replay_t h, int32_t enabled);
/* Fused K8 + K3 + K4 for one rank of a sharded replay.  The kernel itself waits for every
 * peer's total of the current exchange round — published by apexb200_shard_exchange or,
 * with push-on-update, by the peers' apexb200_td_fwd_update.
 * The global batch is global_count stratified draws u_j = (j + v_j)/global_count with
 * v_j = splitmix64(seed, exchange round, call index, j): identical on every rank.  This
 * rank owns the contiguous run of draws whose global target falls into its prefix
 * interval; each warp finds the run by a 32-ary search on the draw function, takes local
 * row r = j - j_first, descends the local tree with (target - prefix) and gathers the
 * record into row r of the outputs ([max_rows, ...]).  count_out [1] = rows owned (rows
 * beyond it are stale), pos_out [max_rows] = global draw index j of each row.
 * Every draw the rank owns is delivered whatever the split of priority mass between the shards
 * (replay.py:52-57: B requested, B returned): the launch covers the rank's fair share of the batch
 * (global_count / world, plus head-room) in one wave of CTAs and a rank that owns more runs further
 * waves inside the same launch.  max_rows is only the number of rows the output arrays hold: pass
 * max_rows >= global_count (one rank can own the whole batch) and nothing is ever dropped; a smaller
 * max_rows is a caller-side limit — a rank that owns more draws than that keeps the first max_rows
 * and sets APEXB200_ST_QUOTA.  out_format as for apexb200_sample_gather.  Needs frames_per_state <= 8
 * and half a record (C frames) fitting in shared memory. */
int apexb200_sample_gather_sharded(apexb200_replay_t h, uint64_t seed, int32_t global_count, int32_t max_rows,
                                   int32_t *idx_out, int32_t *fsel_out, int32_t *pos_out, double *u_out,
                                   int32_t *count_out, void *S_t, void *S_tpn, int32_t *A, float *R,
                                   float *Gamma, float *q_t, float *q_tpn, int32_t out_format, void *stream);

/* ---- K9 learner -> actor weight publication ----------------
 * Replaces shared_state["Q_state_dict"] = Q.state_dict() (learner.py:24, :74 — a 13.3 MB
 * pickle through a Manager proxy on every step) and Q.load_state_dict(shared_state[...])
 * (actor.py:106, :189-191).  A "params board" owns one flat device snapshot of a set of
 * tensors (span i = `bytes` bytes at device address `tensor`, stored at `flat_offset`,
 * a multiple of 256, inside the snapshot).  create/destroy allocate and synchronise;
 * publish/load only enqueue. */
typedef struct apexb200_param_span {
    uint64_t tensor;        /* device address of the tensor's storage */
    uint64_t flat_offset;   /* byte offset inside the flat snapshot, multiple of 256 */
    uint64_t bytes;
} apexb200_param_span;
typedef struct apexb200_params *apexb200_params_t;
int apexb200_params_create(int32_t device, const apexb200_param_span *spans_host, int32_t n_spans, uint64_t flat_bytes,
                           apexb200_params_t *out);
int apexb200_params_destroy(apexb200_params_t p);
/* Device address / size of the snapshot (peers may read it over NVLink or broadcast it). */
int apexb200_params_flat(apexb200_params_t p, void **flat_dev_host, uint64_t *bytes_host);
/* Learner side (learner.py:74).  One kernel on `stream` packs the tensors into the snapshot.
 * With slot_host (page-locked, flat_bytes long) and ctrl_host (int64 [>=2], host) the snapshot
 * is then copied out on the board's own copy stream — the learner stream does not wait for
 * PCIe — and when the copy has landed a host callback stores `version` into ctrl_host[0];
 * ctrl_host[1] = version is stored at once ("begun").  Versions of one board increase by 1 and
 * alternate between two slots, so a reader that copied slot v&1 after seeing ctrl[0] == v
 * has a consistent copy iff ctrl[1] <= v + 1 afterwards.  slot_host == ctrl_host == NULL: pack only. */
int apexb200_params_publish(apexb200_params_t p, void *slot_host, int64_t *ctrl_host, int64_t version, void *stream);
/* Actor side with a GPU model (actor.py:189-191): copy flat_src (a host slot when src_is_host,
 * else a device or peer snapshot; NULL = the board's own snapshot as it stands) into the
 * snapshot and scatter it into the tensors, on `stream`. */
int apexb200_params_load(apexb200_params_t p, const void *flat_src, int32_t src_is_host, void *stream);

/* ---- K10 batched actor policy (actor.py:121-125, :160-163; duelling_network.py:27) ----
 * For E environments at once: q[e,a] = value[e] + (adv[e,a] - (1/A) sum_a adv[e,a]) — the
 * duelling combine with the batch-1 semantics every reference actor has (the model's own
 * combine couples the rows of a batch) — then epsilon-greedy per environment:
 * explore iff u_explore[e] < eps[e] (random.random() >= eps is the greedy branch), random
 * action min(floor(u_action[e]*A), A-1), greedy action = first maximum (np.argmax; NaN counts
 * as largest).  value == NULL: advantage_or_q already holds Q rows.  u_explore == u_action ==
 * NULL: the uniforms are splitmix64(seed, step, e).  q_out [E, A] and explored [E] are nullable.
 * Enqueue only. */
int apexb200_actor_policy(const float *value, const float *advantage_or_q, const float *eps, const double *u_explore,
                          const double *u_action, uint64_t seed, uint64_t step, int32_t E, int32_t A, float *q_out,
                          int32_t *actions, uint8_t *explored, void *stream);

/* ---- page-locking of shared host memory (actor staging rings, weight boards: the transport that
 * replaces the Manager queue / dict of main.py:21-25, 40-47).  Registers [address, address+bytes)
 * with CUDA so that copies to and from it are asynchronous; a stale registration of the same
 * virtual range is replaced.  Unregister before unmapping the memory. */
int apexb200_host_register(void *address_host, uint64_t bytes);
/* The same, and the range becomes readable by kernels: *device_pointer_host receives its device-visible address
 * (what apexb200_ingest_push_blobs takes). */
int apexb200_host_register_mapped(void *address_host, uint64_t bytes, void **device_pointer_host);
int apexb200_host_unregister(void *address_host);

/* ---- raw access for checkpoint / tests (learner.py:18-23 is load-only; the
 * replay itself is never persisted by the reference) ---------------------- */
#define APEXB200_BUF_FRAMES 0
#define APEXB200_BUF_FIDX   1
#define APEXB200_BUF_ACTION 2
#define APEXB200_BUF_REWARD 3
#define APEXB200_BUF_GAMMA  4
#define APEXB200_BUF_QT     5
#define APEXB200_BUF_QTPN   6
#define APEXB200_BUF_LEAVES 7
int apexb200_replay_buffer(apexb200_replay_t h, int32_t which, void **ptr_host, int64_t *bytes_host);
/* ---- lossless packing of the frame store for checkpoints (README.md:24 of the reference lists
 * "observation compression" as a TODO, the reference never persists its replay).  Not on the sampling path: the
 * ring stays raw.  Frames [first_slot, first_slot + count) of the store, in slot order; every frame is XORed with
 * its predecessor (every 64th frame of the range is a key frame, kept as is) and only the 16-byte blocks with a
 * non-zero byte are stored behind a bitmap: 4*ceil(nb/32) bytes (rounded up to 16) + 16 bytes per stored block,
 * nb = frame_bytes/16.  pack_sizes writes the encoded size of every frame (multiples of 16); the caller forms the
 * exclusive prefix sum `offsets` [count] (u64) and allocates the output; pack encodes; unpack writes the frames
 * back into the store (one CTA per 64-frame segment).  All device pointers, enqueue only. */
int apexb200_frames_pack_sizes(apexb200_replay_t h, int64_t first_slot, int32_t count, uint32_t *sizes, void *stream);
int apexb200_frames_pack(apexb200_replay_t h, int64_t first_slot, int32_t count, const uint64_t *offsets, uint8_t *out,
                         void *stream);
int apexb200_frames_unpack(apexb200_replay_t h, int64_t first_slot, int32_t count, const uint64_t *offsets,
                           const uint8_t *packed, void *stream);
/* Host-side ring heads: total records ever inserted per lane ([lanes] int64). */
int apexb200_replay_get_heads(apexb200_replay_t h, int64_t *lane_count_host);
int apexb200_replay_set_heads(apexb200_replay_t h, const int64_t *lane_count_host);

#ifdef __cplusplus
}
#endif
#endif /* APEXB200_H_ */
