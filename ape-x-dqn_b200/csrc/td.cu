// K5: fused n-step double-DQN target / TD error / |delta|^alpha priority / loss
// (+ backward, + in-place priority write-back), K7 n-step returns, and the actor's
// initial priority.  All fp32, one thread per record; these kernels move a few
// hundred bytes per record and are launch/latency bound, not bandwidth bound.
#include <cooperative_groups.h>

#include <mutex>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace apexb200 {

int run_tree_update(apexb200_replay *h, const int32_t *idx, const void *values, int mode, int count,
                    const int32_t *count_dev, float eps, cudaStream_t st);
int run_tree_apply(apexb200_replay *h, const int32_t *idx, const void *values, int mode, int count,
                   const int32_t *count_dev, cudaStream_t st);
int run_peer_push(apexb200_replay *h, cudaStream_t st);

struct TdArgs {
    const float *q_on_t, *q_on_tpn, *q_tg_tpn;
    const int32_t *A;
    const float *R, *G;
    float *delta, *prio, *loss, *loss_mean;
    int B, nA, loss_kind;
    float alpha, eps, kappa;
    const int32_t *count_dev;       // nullable: live rows = min(B, *count_dev) (sharded batches)
    float divisor;                  // loss_mean = sum / divisor (0 -> B); global batch size when sharded
    const float *weights;           // nullable: per-record importance weights multiplying the loss
};

__device__ __forceinline__ int live_rows(const TdArgs &a) { return a.count_dev ? min(a.B, *a.count_dev) : a.B; }
__device__ __forceinline__ float mean_div(const TdArgs &a) { return a.divisor > 0.0f ? a.divisor : (float)a.B; }

// learner.py:43-50 for one record.  Explicit _rn intrinsics keep the compiler from
// contracting R + Gamma*q into an FMA, so delta/loss equal a plain fp32 evaluation.
__device__ __forceinline__ float td_one(const TdArgs &a, int b, float *loss_out, float *prio_out = nullptr) {
    const float *row_t = a.q_on_t + (long long)b * a.nA;
    const float *row_n = a.q_on_tpn + (long long)b * a.nA;
    const float *row_g = a.q_tg_tpn + (long long)b * a.nA;
    int arg = 0;
    float best = row_n[0];
    for (int k = 1; k < a.nA; ++k) {
        float q = row_n[k];
        if (q > best) { best = q; arg = k; }                 // first maximum on ties (argmax, learner.py:45)
    }
    int act = a.A[b];
    act = act < 0 ? 0 : (act >= a.nA ? a.nA - 1 : act);
    const float target = __fadd_rn(a.R[b], __fmul_rn(a.G[b], row_g[arg]));     // learner.py:44
    const float d = __fsub_rn(target, row_t[act]);                              // learner.py:47
    const float ad = fabsf(d);
    float l;
    if (a.loss_kind == APEXB200_LOSS_HALF_SQ) l = __fmul_rn(__fmul_rn(0.5f, d), d);   // learner.py:48
    else l = ad <= a.kappa ? __fmul_rn(__fmul_rn(0.5f, d), d) : __fmul_rn(a.kappa, __fsub_rn(ad, 0.5f * a.kappa));
    if (a.weights) l = __fmul_rn(l, a.weights[b]);                              // importance-sampling correction
    const float pr = prio_pow(ad, a.alpha, a.eps);                              // learner.py:50 + replay.py:24
    a.delta[b] = d;
    a.prio[b] = pr;
    a.loss[b] = l;
    *loss_out = l;
    if (prio_out) *prio_out = pr;
    return d;
}

// The inputs of one record fetched up front, before anything is used: a warp issues in order, so a dependent pair
// (row of Q(S_tpn) -> the selected entry of Q_target(S_tpn); record index -> its leaf) otherwise costs two memory
// round trips one after the other.  Rows of 4 actions (the Atari action set of parameters.json) are single 16-byte
// loads.  td_row_finish performs the same operations on the same values as td_one: delta / loss / priority are
// bit-identical.
struct TdRow4 {
    float4 qn, qg, qt;
    int act;
    float r, g, w;
};
__device__ __forceinline__ bool td_rows_are_vec4(const TdArgs &a) {
    return a.nA == 4 && ((((uintptr_t)a.q_on_t | (uintptr_t)a.q_on_tpn | (uintptr_t)a.q_tg_tpn) & 15) == 0);
}
__device__ __forceinline__ TdRow4 td_row_load4(const TdArgs &a, int b) {
    TdRow4 x;
    x.qn = ((const float4 *)a.q_on_tpn)[b];
    x.qg = ((const float4 *)a.q_tg_tpn)[b];
    x.qt = ((const float4 *)a.q_on_t)[b];
    x.act = a.A[b];
    x.r = a.R[b];
    x.g = a.G[b];
    x.w = a.weights ? a.weights[b] : 1.0f;
    return x;
}
__device__ __forceinline__ float td_row_finish4(const TdArgs &a, int b, const TdRow4 &x, float *loss_out) {
    const float qn[4] = {x.qn.x, x.qn.y, x.qn.z, x.qn.w}, qg[4] = {x.qg.x, x.qg.y, x.qg.z, x.qg.w},
                qt[4] = {x.qt.x, x.qt.y, x.qt.z, x.qt.w};
    int arg = 0;
    float best = qn[0];
    #pragma unroll
    for (int k = 1; k < 4; ++k)
        if (qn[k] > best) { best = qn[k]; arg = k; }             // first maximum on ties (argmax, learner.py:45)
    const int act = x.act < 0 ? 0 : (x.act > 3 ? 3 : x.act);
    float g_sel = qg[0], t_sel = qt[0];
    #pragma unroll
    for (int k = 1; k < 4; ++k) {
        if (k == arg) g_sel = qg[k];
        if (k == act) t_sel = qt[k];
    }
    const float target = __fadd_rn(x.r, __fmul_rn(x.g, g_sel));                  // learner.py:44
    const float d = __fsub_rn(target, t_sel);                                    // learner.py:47
    const float ad = fabsf(d);
    float l;
    if (a.loss_kind == APEXB200_LOSS_HALF_SQ) l = __fmul_rn(__fmul_rn(0.5f, d), d);   // learner.py:48
    else l = ad <= a.kappa ? __fmul_rn(__fmul_rn(0.5f, d), d) : __fmul_rn(a.kappa, __fsub_rn(ad, 0.5f * a.kappa));
    if (a.weights) l = __fmul_rn(l, x.w);
    const float pr = prio_pow(ad, a.alpha, a.eps);                               // learner.py:50 + replay.py:24
    a.delta[b] = d;
    a.prio[b] = pr;
    a.loss[b] = l;
    *loss_out = l;
    return pr;
}

// Fixed-shape block sum: xor-shuffle tree inside each warp, then warp partials
// added in warp order by thread 0 -> the same association for a given blockDim.
__device__ __forceinline__ float block_sum_fixed(float v, float *warp_part) {
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) warp_part[wid] = v;
    __syncthreads();
    float s = 0.0f;
    if (threadIdx.x == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        for (int k = 0; k < nw; ++k) s += warp_part[k];
    }
    return s;                                               // valid in thread 0
}

// ---- B <= 1024: one CTA, no scratch; optionally also the priority write-back.
// Duplicate record indices must resolve as "last row wins" (replay.py:40).  Stratified
// draws give non-decreasing indices, so duplicates are adjacent and the last of a run
// wins without any atomics; an unsorted batch falls back to the stamp protocol of K2.
// The old leaf is fetched up front, in parallel with the Q rows, to keep the chain of
// dependent memory round trips at two.
// Measurement aid (builds with -DAPEX_TD_TIMELINE only): global-timer stamps of the last td_small_kernel launch
// {entry, dependency wait done, thread 0 at the end}, read back with apexb200_debug_td_timeline.
__device__ unsigned long long g_dbg_td[8];
#ifdef APEX_TD_TIMELINE
#define TD_STAMP(k) do { if (threadIdx.x == 0) g_dbg_td[k] = td_gtimer(); } while (0)
#else
#define TD_STAMP(k) do { } while (0)
#endif
__device__ __forceinline__ unsigned long long td_gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__global__ void __launch_bounds__(1024)
td_small_kernel(TdArgs a, TreeView tv, const int32_t *__restrict__ idx, uint32_t *__restrict__ stamp,
                int32_t *status, PeerView pv, int push) {
    __shared__ float warp_part[32];
    __shared__ unsigned long long delta_part[32];
#ifdef APEX_TD_TIMELINE
    if (threadIdx.x == 0) g_dbg_td[0] = td_gtimer();
#endif
    pdl_prologue();
#ifdef APEX_TD_TIMELINE
    if (threadIdx.x == 0) g_dbg_td[1] = td_gtimer();
#endif
    if (idx != nullptr) bump_epoch(status);
    const int live = live_rows(a);
    // multi-GPU push (see below): the shard total before this update, read by warp 0 while the batch
    // rows are being fetched
    unsigned long long total_before = 0ull, cur_round = 0ull, peer_word = 0ull;
    if (push && threadIdx.x < 32) {
        const int top = tv.n_levels - 1;
        total_before = threadIdx.x < tv.size[top] ? tv.lvl[top][threadIdx.x] : 0ull;
        // the peers' words of the current round (normally consumed by the preceding sample already):
        // fetched now so that the push does not wait for this read
        cur_round = read_round(status);
        if (threadIdx.x < pv.world && cur_round > 0ull)
            peer_word = *(volatile unsigned long long *)(pv.xbuf[pv.rank] +
                                                         (size_t)((int)(cur_round & 1ull) * kMaxShards + threadIdx.x));
    }
    // One row per thread; a batch with more live rows than threads (a sharded batch whose rank owns far more
    // than its fair share of the draws) takes further passes, `passes` is uniform over the CTA.  Each pass is
    // complete in itself (vote, leaf writes, ancestor updates), passes run in batch order, so "last row wins"
    // holds across them as well.
    const int passes = live > (int)blockDim.x ? (live + (int)blockDim.x - 1) / (int)blockDim.x : 1;
    const bool vec4 = td_rows_are_vec4(a);
    float l_acc = 0.0f;
    unsigned long long d_acc = 0ull;
    for (int pass = 0; pass < passes; ++pass) {
        const int b = pass * (int)blockDim.x + (int)threadIdx.x;
        float l = 0.0f;
        long long i = -1;
        bool in_order = true, last_of_run = true;
        unsigned long long old = 0ull;
        float pr = 0.0f;                                         // this row's |delta|^alpha, kept in a register
        if (b < live) {
            long long nxt = 1ll << 40;
            TdRow4 row;
            if (vec4) row = td_row_load4(a, b);                   // all loads of the row in flight ...
            if (idx != nullptr) {
                i = idx[b];                                       // ... together with the record index
                if (b + 1 < live) nxt = (long long)idx[b + 1];
                in_order = i <= nxt;
                last_of_run = i != nxt;
                if (i < 0 || i >= tv.size[0]) { atomicOr(status, APEXB200_ST_BAD_INDEX); i = -1; }
                else old = tv.lvl[0][i];                          // second (and last) dependent round trip
            }
            if (vec4) pr = td_row_finish4(a, b, row, &l);
            else td_one(a, b, &l, &pr);
            if (status != nullptr) {                              // an action outside [0, nA) was clamped: say so
                const int act = vec4 ? row.act : a.A[b];
                if (act < 0 || act >= a.nA) atomicOr(status, APEXB200_ST_BAD_ACTION);
            }
        }
        l_acc += l;
        TD_STAMP(3);
        const int sorted = __syncthreads_and(in_order);
        bool winner = (i >= 0) && last_of_run;
        if (!sorted) {
            if (i >= 0) atomicMax(stamp + i, (uint32_t)b + 1u);
            __syncthreads();
            winner = false;
            if (i >= 0 && __ldcg(stamp + i) == (uint32_t)b + 1u) { winner = true; stamp[i] = 0u; }
        }
        unsigned long long d = 0ull;
        if (winner) {
            unsigned long long nl = leaf_from_prio(pr);
            tv.lvl[0][i] = nl;
            d = nl - old;
        }
        d_acc += d;
        TD_STAMP(4);
        if (push && pass == passes - 1) {
            // Fused collective (multi-GPU): publish this shard's new total to every peer from the
            // kernel that produced it — the next step's sharded sample on each rank waits for these
            // words instead of an all-gather.  new total = total before + sum of this batch's leaf
            // deltas (exact in u64 arithmetic; this CTA is the only writer of the tree while it runs).
            // It is sent BEFORE the (last pass's) ancestor updates and the loss reduction below, so that
            // the NVLink flight overlaps them; peers only route draws with it, this shard's own tree is
            // complete when the kernel ends.
            unsigned long long ds = d_acc;
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) ds += __shfl_xor_sync(0xffffffffu, ds, o);
            if ((threadIdx.x & 31) == 0) delta_part[threadIdx.x >> 5] = ds;
            __syncthreads();
            if (threadIdx.x < 32) {
                const int lane = threadIdx.x;
                unsigned long long v = total_before;
                #pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                const int nw = (blockDim.x + 31) >> 5;
                for (int k = 0; k < nw; ++k) v += delta_part[k];
                const unsigned long long cur = cur_round;
                if (lane < pv.world) {
                    if (cur > 0ull && (peer_word & 15ull) != (cur & 15ull))
                        peer_wait_round(pv, cur, lane, status);             // not there at kernel start: wait now
                    peer_push_round(pv, cur + 1ull, lane, v);
                }
                __syncwarp();
                if (lane == 0) *(volatile unsigned long long *)(status + 6) = cur + 1ull;
            }
        }
        if (!sorted) {
            if (d != 0ull) propagate_delta(tv, i, d);
        } else if (idx != nullptr) {
            // Sorted record indices (stratified draws): one atomic per distinct ancestor per warp
            warp_propagate_adjacent(tv, i, d);
        }
        TD_STAMP(5);
        if (pass + 1 < passes) __syncthreads();                // the next pass reads leaves this one wrote
    }
    float s = block_sum_fixed(l_acc, warp_part);
    if (threadIdx.x == 0) *a.loss_mean = s / mean_div(a);
#ifdef APEX_TD_TIMELINE
    if (threadIdx.x == 0) g_dbg_td[2] = td_gtimer();
    if (threadIdx.x == blockDim.x - 1) g_dbg_td[6] = td_gtimer();
#endif
}

// ---- B <= 1024 as one thread-block cluster: ceil(B/128) CTAs of 128 rows on as many SMs
// (opt-in, apexb200_set_td_cluster; measured slower than td_small_kernel on B200, see the header).
// The single-CTA kernel issues all of a batch's leaf/ancestor atomics and row loads from one
// SM (~1.6 us of a 5.5 us in-step K5+K2 at B=512); here they issue from up to 8 SMs.  The
// CTAs meet in CTA 0's shared memory through DSMEM: per-CTA loss partials (summed there in
// CTA order -> deterministic), the "is the batch sorted" vote and the multi-GPU total delta.
constexpr int kTdClusterRows = 128;
constexpr int kTdClusterMax = 8;

__global__ void __launch_bounds__(kTdClusterRows)
td_cluster_kernel(TdArgs a, TreeView tv, const int32_t *__restrict__ idx, uint32_t *__restrict__ stamp,
                  int32_t *status, PeerView pv, int push) {
    cg::cluster_group cluster = cg::this_cluster();
    const int nc = (int)cluster.num_blocks(), cr = (int)cluster.block_rank();
    __shared__ float warp_part[kTdClusterRows / 32];
    __shared__ float cta_loss[kTdClusterMax];                 // filled in CTA 0
    __shared__ unsigned long long cta_delta[kTdClusterMax];   // filled in CTA 0
    __shared__ int cta_unsorted[kTdClusterMax];               // filled in every CTA
    pdl_prologue();
    if (idx != nullptr) bump_epoch(status);
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int live = live_rows(a);
    unsigned long long total_before = 0ull;
    if (push && cr == 0 && threadIdx.x < 32) {
        const int top = tv.n_levels - 1;
        total_before = threadIdx.x < tv.size[top] ? tv.lvl[top][threadIdx.x] : 0ull;
    }
    float l = 0.0f;
    long long i = -1;
    bool in_order = true, last_of_run = true;
    unsigned long long old = 0ull;
    if (b < live) {
        if (idx != nullptr) {
            i = idx[b];
            const long long nxt = (b + 1 < live) ? (long long)idx[b + 1] : (1ll << 40);
            in_order = i <= nxt;
            last_of_run = i != nxt;
            if (i < 0 || i >= tv.size[0]) { atomicOr(status, APEXB200_ST_BAD_INDEX); i = -1; }
            else old = tv.lvl[0][i];
        }
        td_one(a, b, &l);
    }
    cluster.sync();                                            // every CTA of the cluster is running: DSMEM is usable
    // cluster-wide vote: is the whole batch in non-decreasing index order?
    const int cta_sorted = __syncthreads_and(in_order);
    if (threadIdx.x < nc) *cluster.map_shared_rank(&cta_unsorted[cr], threadIdx.x) = !cta_sorted;
    const float s = block_sum_fixed(l, warp_part);             // per-CTA loss partial (valid in thread 0)
    if (threadIdx.x == 0) *cluster.map_shared_rank(&cta_loss[cr], 0) = s;
    cluster.sync();
    bool sorted = true;
    for (int r = 0; r < nc; ++r) sorted = sorted && !cta_unsorted[r];
    bool winner = (i >= 0) && last_of_run;
    if (!sorted) {
        if (i >= 0) atomicMax(stamp + i, (uint32_t)b + 1u);
        cluster.sync();
        winner = false;
        if (i >= 0 && __ldcg(stamp + i) == (uint32_t)b + 1u) { winner = true; stamp[i] = 0u; }
    }
    unsigned long long d = 0ull;
    if (winner) {
        unsigned long long nl = leaf_from_prio(a.prio[b]);
        tv.lvl[0][i] = nl;
        d = nl - old;
    }
    if (!sorted) {
        if (d != 0ull) propagate_delta(tv, i, d);
    } else if (idx != nullptr) {
        warp_propagate_adjacent(tv, i, d);                      // one atomic per distinct ancestor per warp
    }
    if (push) {                                                 // this CTA's share of the total delta -> CTA 0
        __shared__ unsigned long long dpart[kTdClusterRows / 32];
        unsigned long long ds = d;
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) ds += __shfl_xor_sync(0xffffffffu, ds, o);
        if (lane == 0) dpart[threadIdx.x >> 5] = ds;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long t = 0ull;
            for (int k = 0; k < kTdClusterRows / 32; ++k) t += dpart[k];
            *cluster.map_shared_rank(&cta_delta[cr], 0) = t;
        }
        cluster.sync();
    }
    if (cr != 0) return;
    if (threadIdx.x == 0) {
        float tot = 0.0f;
        for (int r = 0; r < nc; ++r) tot += cta_loss[r];        // CTA order: same association every run
        *a.loss_mean = tot / mean_div(a);
    }
    if (push && threadIdx.x < 32) {                            // fused collective, see td_small_kernel
        unsigned long long v = total_before;
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        for (int r = 0; r < nc; ++r) v += cta_delta[r];
        const unsigned long long cur = read_round(status);
        if (lane < pv.world) {
            if (cur > 0ull) peer_wait_round(pv, cur, lane, status);
            peer_push_round(pv, cur + 1ull, lane, v);
        }
        __syncwarp();
        if (lane == 0) *(volatile unsigned long long *)(status + 6) = cur + 1ull;
    }
}

// td_one with every load issued before the first use (rows of up to kWideActions actions live in
// registers; the selected entries are picked with compare-selects): one memory round trip per
// record instead of two dependent ones.  Same operations on the same values as td_one.
constexpr int kWideActions = 8;
__device__ __forceinline__ void td_one_wide(const TdArgs &a, int b, float *loss_out) {
    const float *row_t = a.q_on_t + (long long)b * a.nA;
    const float *row_n = a.q_on_tpn + (long long)b * a.nA;
    const float *row_g = a.q_tg_tpn + (long long)b * a.nA;
    float qt[kWideActions], qn[kWideActions], qg[kWideActions];
    #pragma unroll
    for (int k = 0; k < kWideActions; ++k) {
        const bool in = k < a.nA;
        qn[k] = in ? __ldcs(row_n + k) : 0.0f;
        qg[k] = in ? __ldcs(row_g + k) : 0.0f;
        qt[k] = in ? __ldcs(row_t + k) : 0.0f;
    }
    int act = a.A[b];
    const float r = a.R[b], g = a.G[b];
    const float w = a.weights ? a.weights[b] : 1.0f;
    act = act < 0 ? 0 : (act >= a.nA ? a.nA - 1 : act);
    int arg = 0;
    float best = qn[0];
    #pragma unroll
    for (int k = 1; k < kWideActions; ++k)
        if (k < a.nA && qn[k] > best) { best = qn[k]; arg = k; }   // first maximum on ties (argmax, learner.py:45)
    float g_sel = qg[0], t_sel = qt[0];
    #pragma unroll
    for (int k = 1; k < kWideActions; ++k) {
        if (k == arg) g_sel = qg[k];
        if (k == act) t_sel = qt[k];
    }
    const float target = __fadd_rn(r, __fmul_rn(g, g_sel));                      // learner.py:44
    const float d = __fsub_rn(target, t_sel);                                    // learner.py:47
    const float ad = fabsf(d);
    float l;
    if (a.loss_kind == APEXB200_LOSS_HALF_SQ) l = __fmul_rn(__fmul_rn(0.5f, d), d);   // learner.py:48
    else l = ad <= a.kappa ? __fmul_rn(__fmul_rn(0.5f, d), d) : __fmul_rn(a.kappa, __fsub_rn(ad, 0.5f * a.kappa));
    if (a.weights) l = __fmul_rn(l, w);
    a.delta[b] = d;
    a.prio[b] = prio_pow(ad, a.alpha, a.eps);
    a.loss[b] = l;
    *loss_out = l;
}

// ---- any B: 256 records per CTA; per-CTA loss partials go to scratch and the last CTA to finish
// (ticket counter in scratch) adds them: thread t sums partials t, t+256, ... in that order, then
// the fixed-shape block sum -> the same association for a given grid, whichever CTA is last.
// scratch layout: [0] u32 ticket (zeroed by the host before launch), [64..] float partials.
__global__ void __launch_bounds__(256)
td_multi_kernel(TdArgs a, const int32_t *__restrict__ idx, uint32_t *__restrict__ stamp, long long N,
                int32_t *status, unsigned int *__restrict__ ticket, float *__restrict__ partials) {
    __shared__ float warp_part[8];
    __shared__ bool is_last;
    pdl_prologue();
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    float l = 0.0f;
    if (b < live_rows(a)) {
        long long i = -1;
        if (idx != nullptr) i = idx[b];                       // in flight with the rows
        if (a.nA <= kWideActions) td_one_wide(a, b, &l);
        else td_one(a, b, &l);
        if (idx != nullptr) {
            if (i < 0 || i >= N) atomicOr(status, APEXB200_ST_BAD_INDEX);
            else atomicMax(stamp + i, (uint32_t)b + 1u);
        }
    }
    float s = block_sum_fixed(l, warp_part);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = s;
        __threadfence();
        unsigned int t = atomicAdd(ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        float tot = 0.0f;
        for (unsigned k = threadIdx.x; k < gridDim.x; k += blockDim.x) tot += __ldcg(partials + k);
        __syncthreads();                                      // warp_part is reused
        tot = block_sum_fixed(tot, warp_part);
        if (threadIdx.x == 0) *a.loss_mean = tot / mean_div(a);
    }
}

__global__ void __launch_bounds__(256)
td_bwd_kernel(const float *__restrict__ delta, const int32_t *__restrict__ A, const float *__restrict__ grad_out,
              int B, int nA, int loss_kind, float kappa, const int32_t *__restrict__ count_dev, float divisor,
              const float *__restrict__ weights, float *__restrict__ grad_q) {
    pdl_prologue();
    long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)B * nA) return;
    int b = (int)(e / nA), k = (int)(e - (long long)b * nA);
    const int live = count_dev ? min(B, *count_dev) : B;
    float g = 0.0f;
    int act = b < live ? A[b] : -1;
    if (b < live) act = act < 0 ? 0 : (act >= nA ? nA - 1 : act);      // same clamp as the forward (td_one)
    if (b < live && k == act) {
        float d = delta[b];
        if (loss_kind == APEXB200_LOSS_HUBER) d = fminf(fmaxf(d, -kappa), kappa);
        g = -d * (*grad_out) / (divisor > 0.0f ? divisor : (float)B);   // d(mean loss)/dQ(S_t)[b, A_b]
        if (weights) g *= weights[b];
    }
    grad_q[e] = g;
}

// ---- K7
__global__ void __launch_bounds__(256)
nstep_kernel(const float *__restrict__ rewards, const uint8_t *__restrict__ dones, int E, int T, int n, float gamma,
             float *__restrict__ R, float *__restrict__ G, int32_t *__restrict__ span) {
    long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)E * T) return;
    const int t = (int)(e % T);
    float acc = 0.0f, disc = 1.0f;
    int m = 0;
    bool term = false;
    for (int k = 0; k < n && t + k < T; ++k) {
        acc = __fadd_rn(acc, __fmul_rn(disc, rewards[e + k]));
        disc = __fmul_rn(disc, gamma);
        ++m;
        if (dones[e + k]) { term = true; break; }
    }
    const bool valid = term || m == n;
    R[e] = valid ? acc : 0.0f;
    G[e] = valid ? (term ? 0.0f : disc) : 0.0f;
    span[e] = valid ? m : 0;
}

__global__ void __launch_bounds__(256)
initial_priority_kernel(const float *__restrict__ q_t, const float *__restrict__ q_tpn, const int32_t *__restrict__ A,
                        const float *__restrict__ R, const float *__restrict__ G, int count, int nA,
                        float *__restrict__ abs_td) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    const float *qa = q_t + (long long)j * nA, *qb = q_tpn + (long long)j * nA;
    float mx = qb[0];
    for (int k = 1; k < nA; ++k) mx = fmaxf(mx, qb[k]);
    int a = A[j];
    a = a < 0 ? 0 : (a >= nA ? nA - 1 : a);
    abs_td[j] = fabsf(__fsub_rn(__fadd_rn(R[j], __fmul_rn(G[j], mx)), qa[a]));   // actor.py:138-140
}

// Importance-sampling weights of PER (parameters.json:30, README.md:25 TODO — never read by
// the reference): w_j = (N * leaf_j / total)^(-beta), divided by the batch maximum.  Single CTA.
__global__ void __launch_bounds__(1024)
is_weights_kernel(const unsigned long long *__restrict__ leaf, const unsigned long long *__restrict__ total, int B,
                  float n_records, float beta, float *__restrict__ w) {
    __shared__ float wmax[32];
    pdl_prologue();
    const double tot = (double)(*total);
    float local = 0.0f;
    for (int j = threadIdx.x; j < B; j += blockDim.x) {
        const double p = (double)leaf[j] / tot;
        const float v = p > 0.0 ? powf((float)(n_records * p), -beta) : 0.0f;
        w[j] = v;
        local = fmaxf(local, v);
    }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) local = fmaxf(local, __shfl_xor_sync(0xffffffffu, local, o));
    if ((threadIdx.x & 31) == 0) wmax[threadIdx.x >> 5] = local;
    __syncthreads();
    float m = 0.0f;
    for (int k = 0; k < (int)((blockDim.x + 31) >> 5); ++k) m = fmaxf(m, wmax[k]);
    if (m > 0.0f)
        for (int j = threadIdx.x; j < B; j += blockDim.x) w[j] = w[j] / m;
}

static int td_check(const float *q1, const float *q2, const float *q3, const int32_t *A, const float *R,
                    const float *G, int32_t B, int32_t nA, float *d, float *p, float *l, float *lm) {
    APEX_REQUIRE(q1 && q2 && q3 && A && R && G && d && p && l && lm, APEXB200_EINVAL, "td_fwd: null argument");
    APEX_REQUIRE(B > 0, APEXB200_EINVAL, "td_fwd: B must be positive");
    APEX_REQUIRE(nA >= 1 && nA <= APEXB200_MAX_ACTIONS, APEXB200_ERANGE, "td_fwd: num_actions %d outside [1,%d]", nA,
                 APEXB200_MAX_ACTIONS);
    return APEXB200_OK;
}

static int64_t td_ws_bytes(int B) {
    int64_t blocks = (B + 255) / 256;
    return 256 + (blocks * 4 + 255) / 256 * 256;
}

static int td_launch(apexb200_replay *h, const int32_t *idx, const TdArgs &a, void *workspace, int64_t ws_bytes,
                     cudaStream_t st) {
    TreeView tv_none{};
    if (a.B <= 1024 && g_td_cluster) {
        PeerView pv_none{};
        const int nc = (a.B + kTdClusterRows - 1) / kTdClusterRows;
        return launch_step_cluster("td_cluster_kernel", td_cluster_kernel, dim3(nc), dim3(kTdClusterRows), 0, st, nc, a,
                                   h ? h->tree : tv_none, idx, h ? h->stamp : (uint32_t *)nullptr,
                                   h ? h->status : (int32_t *)nullptr, h ? h->peers : pv_none,
                                   (h && idx && h->push_on_update) ? 1 : 0);
    }
    // Sharded batches: the arrays hold up to a.B rows (the whole global batch, so that no draw is ever
    // dropped) but the live count is on the device and is normally the rank's fair share, divisor / world.
    // The single-CTA kernel is sized for that share and takes extra passes when the rank owns more.
    int small_threads = a.B <= 1024 ? ((a.B + 31) / 32) * 32 : 0;
    if (a.count_dev && h && a.divisor > 0.0f) {
        const int world = h->peers.world > 0 ? h->peers.world : 1;
        long long share = ((long long)a.divisor + world - 1) / world;
        share += (share + 7) / 8 + 2;
        if (share < a.B && share <= 1024) small_threads = (int)((share + 31) / 32) * 32;
    }
    if (small_threads > 0) {
        PeerView pv_none{};
        return launch_step("td_small_kernel", td_small_kernel, dim3(1), dim3(small_threads), 0, st, a,
                           h ? h->tree : tv_none, idx, h ? h->stamp : (uint32_t *)nullptr,
                           h ? h->status : (int32_t *)nullptr, h ? h->peers : pv_none,
                           (h && idx && h->push_on_update) ? 1 : 0);
    }
    APEX_REQUIRE(workspace && ws_bytes >= td_ws_bytes(a.B), APEXB200_EINVAL,
                 "td_fwd: B=%d needs %lld bytes of workspace (got %lld)", a.B, (long long)td_ws_bytes(a.B),
                 (long long)ws_bytes);
    APEX_REQUIRE(((uintptr_t)workspace & 255) == 0, APEXB200_EINVAL, "td_fwd: workspace must be 256-byte aligned");
    APEX_CUDA(cudaMemsetAsync(workspace, 0, 256, st));
    const int blocks = (a.B + 255) / 256;
    int rc = launch_step("td_multi_kernel", td_multi_kernel, dim3(blocks), dim3(256), 0, st, a, idx,
                         h ? h->stamp : (uint32_t *)nullptr, h ? (long long)h->ring.N : 0ll,
                         h ? h->status : (int32_t *)nullptr, (unsigned int *)workspace,
                         (float *)((char *)workspace + 256));
    if (rc || !h) return rc;
    rc = run_tree_apply(h, idx, a.prio, 2, a.B, a.count_dev, st);     // stamps were taken inside td_multi_kernel
    if (rc || !h->push_on_update) return rc;
    return run_peer_push(h, st);                                      // large batches: publish from a trailing launch
}

// ---------------------------------------------------------------------------
// Scratch for apexb200_td_fwd_noalloc.  It is part of the loaded module: every GPU that
// loads the library carries these bytes until the process exits, and every call on that
// GPU uses the same bytes.  Layout as in td_multi_kernel: [0,256) ticket, then one float
// partial per 256-row CTA.
constexpr size_t kNoAllocScratchBytes = 256 + (size_t)(APEXB200_TD_NOALLOC_MAX_B / 256) * sizeof(float);
static __device__ __align__(256) unsigned char g_td_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
static std::mutex g_noalloc_mutex;
static void *g_noalloc_ptr[kMaxDevices] = {};

// In host code `g_td_noalloc_scratch` is only a handle for the symbol, not an address a
// kernel can use; the device address has to be asked for, per device, with that device
// current.  Done once per device and cached.
static int noalloc_scratch_of_current_device(void **out) {
    int dev = -1;
    APEX_CUDA(cudaGetDevice(&dev));
    APEX_REQUIRE(dev >= 0 && dev < kMaxDevices, APEXB200_ERANGE, "td_fwd_noalloc: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[dev] == nullptr) {
        void *p = nullptr;
        APEX_CUDA(cudaGetSymbolAddress(&p, g_td_noalloc_scratch));
        APEX_REQUIRE(p != nullptr, APEXB200_ECUDA, "td_fwd_noalloc: no device address for the static scratch");
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return APEXB200_OK;
}

}  // namespace apexb200

using namespace apexb200;

extern "C" {

int apexb200_debug_td_timeline(uint64_t *out_host) {
    APEX_REQUIRE(out_host, APEXB200_EINVAL, "debug_td_timeline: null argument");
    APEX_CUDA(cudaMemcpyFromSymbol(out_host, g_dbg_td, sizeof(unsigned long long) * 8));
    return APEXB200_OK;
}

int apexb200_td_workspace_bytes(int32_t B, int64_t *bytes_host) {
    APEX_REQUIRE(bytes_host && B > 0, APEXB200_EINVAL, "td_workspace_bytes: bad argument");
    *bytes_host = td_ws_bytes(B);
    return APEXB200_OK;
}

int apexb200_td_fwd(const float *q_on_t, const float *q_on_tpn, const float *q_tg_tpn, const int32_t *A,
                    const float *R, const float *Gamma, int32_t B, int32_t num_actions, float alpha, float eps,
                    int32_t loss_kind, float kappa, const float *weights, float *delta, float *prio, float *loss,
                    float *loss_mean, void *workspace, int64_t workspace_bytes, void *stream) {
    int rc = td_check(q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, B, num_actions, delta, prio, loss, loss_mean);
    if (rc) return rc;
    TdArgs a{q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, delta, prio, loss, loss_mean, B, num_actions, loss_kind,
             alpha, eps, kappa, nullptr, 0.0f, weights};
    return td_launch(nullptr, nullptr, a, workspace, workspace_bytes, (cudaStream_t)stream);
}

int apexb200_td_fwd_noalloc(const float *q_on_t, const float *q_on_tpn, const float *q_tg_tpn, const int32_t *A,
                            const float *R, const float *Gamma, int32_t B, int32_t num_actions, float alpha,
                            float eps, int32_t loss_kind, float kappa, float *delta, float *prio, float *loss,
                            float *loss_mean, void *stream) {
    int rc = td_check(q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, B, num_actions, delta, prio, loss, loss_mean);
    if (rc) return rc;
    APEX_REQUIRE(B <= APEXB200_TD_NOALLOC_MAX_B, APEXB200_ERANGE,
                 "td_fwd_noalloc: B=%d exceeds the built-in scratch (at most %d rows per call); "
                 "use apexb200_td_fwd with a caller-provided workspace", B, APEXB200_TD_NOALLOC_MAX_B);
    void *scratch = nullptr;
    rc = noalloc_scratch_of_current_device(&scratch);        // first call on a device: symbol look-up
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // The bytes still hold whatever the previous call left; only the ticket has to start at
    // zero (each CTA overwrites its own partial before the last CTA reads them).
    APEX_CUDA(cudaMemsetAsync(scratch, 0, 256, st));
    TdArgs a{q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, delta, prio, loss, loss_mean, B, num_actions, loss_kind,
             alpha, eps, kappa, nullptr, 0.0f, nullptr};
    return launch_step("td_multi_kernel", td_multi_kernel, dim3((B + 255) / 256), dim3(256), 0, st, a,
                       (const int32_t *)nullptr, (uint32_t *)nullptr, 0ll, (int32_t *)nullptr,
                       (unsigned int *)scratch, (float *)((char *)scratch + 256));
}

int apexb200_is_weights(const uint64_t *leaf, const uint64_t *total, int32_t B, int64_t n_records, float beta,
                        float *weights, void *stream) {
    APEX_REQUIRE(leaf && total && weights && B > 0 && n_records > 0, APEXB200_EINVAL, "is_weights: bad argument");
    return launch_step("is_weights_kernel", is_weights_kernel, dim3(1), dim3(B < 1024 ? ((B + 31) / 32) * 32 : 1024), 0,
                       (cudaStream_t)stream, (const unsigned long long *)leaf, (const unsigned long long *)total, B,
                       (float)n_records, beta, weights);
}

int apexb200_td_fwd_update(apexb200_replay_t h, const int32_t *idx, const float *q_on_t, const float *q_on_tpn,
                           const float *q_tg_tpn, const int32_t *A, const float *R, const float *Gamma, int32_t B,
                           int32_t num_actions, float eps, int32_t loss_kind, float kappa, const float *weights,
                           float *delta, float *prio, float *loss, float *loss_mean, void *workspace,
                           int64_t workspace_bytes, const int32_t *count_dev, float mean_divisor, void *stream) {
    APEX_REQUIRE(h && idx, APEXB200_EINVAL, "td_fwd_update: null handle/idx");
    int rc = td_check(q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, B, num_actions, delta, prio, loss, loss_mean);
    if (rc) return rc;
    TdArgs a{q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, delta, prio, loss, loss_mean, B, num_actions, loss_kind,
             h->ring.alpha, eps, kappa, count_dev, mean_divisor, weights};
    return td_launch(h, idx, a, workspace, workspace_bytes, (cudaStream_t)stream);
}

int apexb200_td_bwd(const float *delta, const int32_t *A, const float *grad_loss_mean, int32_t B,
                    int32_t num_actions, int32_t loss_kind, float kappa, const int32_t *count_dev,
                    float mean_divisor, const float *weights, float *grad_q, void *stream) {
    APEX_REQUIRE(delta && A && grad_loss_mean && grad_q && B > 0 && num_actions >= 1, APEXB200_EINVAL,
                 "td_bwd: bad argument");
    long long n = (long long)B * num_actions;
    return launch_step("td_bwd_kernel", td_bwd_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0,
                       (cudaStream_t)stream, delta, A, grad_loss_mean, B, num_actions, loss_kind, kappa, count_dev,
                       mean_divisor, weights, grad_q);
}

int apexb200_nstep(const float *rewards, const uint8_t *dones, int32_t E, int32_t T, int32_t n, float gamma,
                   float *R, float *Gamma, int32_t *span, void *stream) {
    APEX_REQUIRE(rewards && dones && R && Gamma && span, APEXB200_EINVAL, "nstep: null argument");
    APEX_REQUIRE(E > 0 && T > 0 && n >= 1, APEXB200_EINVAL, "nstep: bad E/T/n");
    long long tot = (long long)E * T;
    nstep_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, (cudaStream_t)stream>>>(rewards, dones, E, T, n, gamma, R,
                                                                                  Gamma, span);
    return launch_check("nstep_kernel");
}

int apexb200_initial_priority(const float *q_t, const float *q_tpn, const int32_t *A, const float *R,
                              const float *Gamma, int32_t count, int32_t num_actions, float *abs_td, void *stream) {
    APEX_REQUIRE(q_t && q_tpn && A && R && Gamma && abs_td, APEXB200_EINVAL, "initial_priority: null argument");
    APEX_REQUIRE(count > 0 && num_actions >= 1, APEXB200_EINVAL, "initial_priority: bad count/num_actions");
    initial_priority_kernel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>(q_t, q_tpn, A, R, Gamma, count,
                                                                                   num_actions, abs_td);
    return launch_check("initial_priority_kernel");
}

}  // extern "C"
