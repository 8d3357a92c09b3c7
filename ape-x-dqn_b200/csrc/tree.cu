// 32-ary exact (u64 fixed-point) sum tree: build (K1), point update (K2),
// proportional search (K3), shard routing (K8).
#include "common.cuh"

namespace apexb200 {

// ---------------------------------------------------------------------------
// K1: a warp builds kBuildParents consecutive parents; the 32 children of each are one coalesced
// 256-byte load and all kBuildParents loads are issued before the first sum (8 x 256 B in flight
// per warp: with one load per warp the kernel ran at 0.36 of the HBM peak, latency-bound).
// Integer sums: any order gives the same bits.
constexpr int kBuildParents = 8;
__global__ void __launch_bounds__(256)
tree_build_level_kernel(const unsigned long long *__restrict__ child, unsigned long long *__restrict__ parent,
                        long long n_child, long long n_parent) {
    pdl_prologue();                                          // the level below was written by the previous launch
    const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const long long p0 = w * kBuildParents;
    if (p0 >= n_parent) return;
    unsigned long long v[kBuildParents];
    #pragma unroll
    for (int k = 0; k < kBuildParents; ++k) {
        const long long c = (p0 + k) * kFanout + lane;
        v[k] = (p0 + k < n_parent && c < n_child) ? __ldcs(child + c) : 0ull;
    }
    unsigned long long mine = 0ull;
    #pragma unroll
    for (int k = 0; k < kBuildParents; ++k) {
        unsigned long long x = v[k];
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == k) mine = x;
    }
    if (lane < kBuildParents && p0 + lane < n_parent) parent[p0 + lane] = mine;
}

// ---------------------------------------------------------------------------
// K2.  Duplicate indices in one batch must resolve as "last in batch order wins"
// (dict.update, replay.py:40).  Phase 1 stamps every touched leaf with the largest
// batch position (+1) that names it (atomicMax on a per-leaf u32); phase 2 lets only
// the stamp owner write, and the owner clears the stamp again, so the array is all
// zero between calls and no host-side epoch is needed (the launches are replayable
// from a CUDA graph).  A loser can only ever read the winner's stamp or zero.
__global__ void __launch_bounds__(256)
tree_stamp_kernel(const int32_t *__restrict__ idx, int count, const int32_t *__restrict__ count_dev,
                  uint32_t *__restrict__ stamp, long long N, int32_t *status) {
    pdl_prologue();
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (count_dev) count = min(count, *count_dev);
    if (j >= count) return;
    long long i = idx[j];
    if (i < 0 || i >= N) { atomicOr(status, APEXB200_ST_BAD_INDEX); return; }
    atomicMax(stamp + i, (uint32_t)j + 1u);
}

// mode 0: values are raw priorities (fp32 |delta|) -> max(.,eps)^alpha -> leaf; mode 1: exact u64 leaves;
// mode 2: fp32 priorities already raised to alpha.
__device__ __forceinline__ unsigned long long new_leaf_value(const void *values, int mode, int j, float alpha, float eps) {
    if (mode == 1) {
        unsigned long long nl = ((const unsigned long long *)values)[j];
        return nl > APEXB200_LEAF_MAX ? APEXB200_LEAF_MAX : nl;
    }
    float p = ((const float *)values)[j];
    return leaf_from_prio(mode == 0 ? prio_pow(fabsf(p), alpha, eps) : p);
}

__global__ void __launch_bounds__(256)
tree_apply_kernel(TreeView tv, const int32_t *__restrict__ idx, const void *__restrict__ values, int mode,
                  int count, const int32_t *__restrict__ count_dev, uint32_t *__restrict__ stamp,
                  float alpha, float eps, int32_t *ctrl) {
    pdl_prologue();
    bump_epoch(ctrl);
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (count_dev) count = min(count, *count_dev);
    if (j >= count) return;
    long long i = idx[j];
    if (i < 0 || i >= tv.size[0]) return;
    if (__ldcg(stamp + i) != (uint32_t)j + 1u) return;                 // a later duplicate wins
    stamp[i] = 0u;
    unsigned long long nl = new_leaf_value(values, mode, j, alpha, eps);
    unsigned long long old = tv.lvl[0][i];
    tv.lvl[0][i] = nl;
    unsigned long long d = nl - old;
    if (d != 0ull) propagate_delta(tv, i, d);
}

// Single-CTA fusion of the two phases for count <= 1024 (one launch per learner step).
__global__ void __launch_bounds__(1024)
tree_update_small_kernel(TreeView tv, const int32_t *__restrict__ idx, const void *__restrict__ values, int mode,
                         int count, uint32_t *__restrict__ stamp, float alpha, float eps, int32_t *status) {
    pdl_prologue();
    bump_epoch(status);
    int j = threadIdx.x;
    long long i = -1;
    if (j < count) {
        i = idx[j];
        if (i < 0 || i >= tv.size[0]) { atomicOr(status, APEXB200_ST_BAD_INDEX); i = -1; }
        else atomicMax(stamp + i, (uint32_t)j + 1u);
    }
    __syncthreads();
    if (i < 0) return;
    if (__ldcg(stamp + i) != (uint32_t)j + 1u) return;
    stamp[i] = 0u;
    unsigned long long nl = new_leaf_value(values, mode, j, alpha, eps);
    unsigned long long old = tv.lvl[0][i];
    tv.lvl[0][i] = nl;
    unsigned long long d = nl - old;
    if (d != 0ull) propagate_delta(tv, i, d);
}

__global__ void tree_total_kernel(TreeView tv, unsigned long long *out) {
    pdl_prologue();
    int lane = threadIdx.x;
    const int top = tv.n_levels - 1;
    unsigned long long v = lane < tv.size[top] ? tv.lvl[top][lane] : 0ull;
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) *out = v;
}

// ---------------------------------------------------------------------------
// K3 (the descent itself, warp_descend, lives in common.cuh and is shared with the fused kernel).
// Draw source: u != null -> caller's uniforms; targets != null -> explicit integer targets;
// neither -> stratified draws generated here from (seed, device call counter).
__global__ void __launch_bounds__(128)
tree_sample_kernel(TreeView tv, const double *__restrict__ u, const unsigned long long *__restrict__ targets,
                   int count, const int32_t *__restrict__ count_dev, int32_t *__restrict__ idx_out,
                   unsigned long long *__restrict__ leaf_out, unsigned long long *__restrict__ total_out,
                   const int32_t *__restrict__ fidx, int twoC, int32_t *__restrict__ fsel_out, int32_t *ctrl,
                   unsigned long long seed, unsigned long long call, double *__restrict__ u_out) {
    pdl_prologue();
    int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (count_dev) count = min(count, *count_dev);
    if (w >= count) return;
    unsigned long long leaf, total;
    const bool strat = (u == nullptr && targets == nullptr);
    const bool from_u = strat || (u != nullptr);
    const TopValues vtop = top_level_value(tv);     // in flight while the draw is formed
    double uw = 0.0;
    if (strat) {
        uw = strat_uniform(seed, read_epoch(ctrl), call, w, count);
        if (u_out && lane == 0) u_out[w] = uw;
    } else if (u != nullptr) {
        uw = u[w];
    }
    long long node = warp_descend(tv, vtop, from_u ? 0ull : targets[w], from_u, uw, &leaf, &total, ctrl);
    if (lane == 0) {
        idx_out[w] = (int32_t)node;
        if (leaf_out) leaf_out[w] = node >= 0 ? leaf : 0ull;
        if (total_out && w == 0) *total_out = total;
    }
    if (fsel_out && lane < twoC)
        fsel_out[(long long)w * twoC + lane] = node >= 0 ? fidx[node * twoC + lane] : -1;
}

// ---------------------------------------------------------------------------
// K8.  Single CTA: global targets from the all-gathered shard totals, ownership
// test against this rank's prefix interval, order-preserving compaction.
__global__ void __launch_bounds__(1024)
shard_targets_kernel(const unsigned long long *__restrict__ totals, int G, int rank,
                     const double *__restrict__ u, int count, unsigned long long *__restrict__ local_targets,
                     int32_t *__restrict__ local_pos, int32_t *__restrict__ local_count) {
    pdl_prologue();
    __shared__ int warp_sums[32];
    __shared__ int running;
    unsigned long long T = 0ull, lo = 0ull, hi = 0ull;
    for (int g = 0; g < G; ++g) {
        unsigned long long tg = totals[g];
        if (g == rank) { lo = T; hi = T + tg; }
        T += tg;
    }
    if (threadIdx.x == 0) running = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int basej = 0; basej < count; basej += blockDim.x) {
        int j = basej + threadIdx.x;
        bool mine = false;
        unsigned long long t = 0ull;
        if (j < count && T > 0ull) {
            t = (unsigned long long)(u[j] * (double)T);
            if (t >= T) t = T - 1ull;
            mine = (t >= lo) && (t < hi);
        }
        unsigned m = __ballot_sync(0xffffffffu, mine);
        int before = __popc(m & ((1u << lane) - 1u));
        if (lane == 0) warp_sums[wid] = __popc(m);
        __syncthreads();
        int off = running;
        for (int k = 0; k < wid; ++k) off += warp_sums[k];
        if (mine) {
            local_targets[off + before] = t - lo;
            local_pos[off + before] = j;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int s = 0;
            for (int k = 0; k < (int)(blockDim.x >> 5); ++k) s += warp_sums[k];
            running += s;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) *local_count = running;
}

// ---------------------------------------------------------------------------
// All-gather of the shard totals over peer memory (one warp, one launch; protocol in
// common.cuh).  The warp sums its own top level, lane g waits for rank g's word of the
// current round (so that the slot about to be overwritten is no longer needed), writes
// the new word straight into rank g's exchange buffer over NVLink, then polls the local
// buffer for rank g's word of the new round.  Replaces tree_total + ncclAllGather
// (8 bytes per rank: pure latency) on the per-step path.
__global__ void __launch_bounds__(32)
shard_exchange_kernel(TreeView tv, PeerView pv, int32_t *ctrl, unsigned long long *totals_out) {
    pdl_prologue();
    const int lane = threadIdx.x;
    const int top = tv.n_levels - 1;
    unsigned long long v = lane < tv.size[top] ? tv.lvl[top][lane] : 0ull;
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const unsigned long long cur = read_round(ctrl);
    if (lane < pv.world) {
        if (cur > 0ull) peer_wait_round(pv, cur, lane, ctrl);
        peer_push_round(pv, cur + 1ull, lane, v);
        const unsigned long long t = peer_wait_round(pv, cur + 1ull, lane, ctrl);
        ((unsigned long long *)(ctrl + 16))[lane] = t;
        if (totals_out) totals_out[lane] = t;
    }
    __syncwarp();
    if (lane == 0) *(volatile unsigned long long *)(ctrl + 6) = cur + 1ull;
}

// Push without the trailing wait (what td_small_kernel does at its tail), as its own launch.
__global__ void __launch_bounds__(32)
shard_push_kernel(TreeView tv, PeerView pv, int32_t *ctrl) {
    pdl_prologue();
    const int lane = threadIdx.x;
    const int top = tv.n_levels - 1;
    unsigned long long v = lane < tv.size[top] ? tv.lvl[top][lane] : 0ull;
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const unsigned long long cur = read_round(ctrl);
    if (lane < pv.world) {
        if (cur > 0ull) peer_wait_round(pv, cur, lane, ctrl);
        peer_push_round(pv, cur + 1ull, lane, v);
    }
    __syncwarp();
    if (lane == 0) *(volatile unsigned long long *)(ctrl + 6) = cur + 1ull;
}

int run_peer_push(apexb200_replay *h, cudaStream_t st) {
    return launch_step("shard_push_kernel", shard_push_kernel, dim3(1), dim3(32), 0, st, h->tree, h->peers, h->status);
}

// Phase 2 alone, for callers that stamped inside their own kernel (td.cu).
int run_tree_apply(apexb200_replay *h, const int32_t *idx, const void *values, int mode, int count,
                   const int32_t *count_dev, cudaStream_t st) {
    return launch_step("tree_apply_kernel", tree_apply_kernel, dim3((count + 255) / 256), dim3(256), 0, st, h->tree, idx,
                       values, mode, count, count_dev, h->stamp, h->ring.alpha, 0.0f, h->status);
}

// Shared by tree.cu and td.cu: run K2 over (idx, values).
int run_tree_update(apexb200_replay *h, const int32_t *idx, const void *values, int mode, int count,
                    const int32_t *count_dev, float eps, cudaStream_t st) {
    if (count <= 0) return APEXB200_OK;
    if (count <= 1024 && count_dev == nullptr) {
        return launch_step("tree_update_small_kernel", tree_update_small_kernel, dim3(1), dim3(((count + 31) / 32) * 32),
                           0, st, h->tree, idx, values, mode, count, h->stamp, h->ring.alpha, eps, h->status);
    }
    const int blocks = (count + 255) / 256;
    int rc = launch_step("tree_stamp_kernel", tree_stamp_kernel, dim3(blocks), dim3(256), 0, st, idx, count, count_dev,
                         h->stamp, (long long)h->ring.N, h->status);
    if (rc) return rc;
    return launch_step("tree_apply_kernel", tree_apply_kernel, dim3(blocks), dim3(256), 0, st, h->tree, idx, values, mode,
                       count, count_dev, h->stamp, h->ring.alpha, eps, h->status);
}

}  // namespace apexb200

using namespace apexb200;

extern "C" {

int apexb200_tree_build(apexb200_replay_t h, void *stream) {
    APEX_REQUIRE(h, APEXB200_EINVAL, "tree_build: null handle");
    cudaStream_t st = (cudaStream_t)stream;
    const TreeView &tv = h->tree;
    for (int l = 0; l + 1 < tv.n_levels; ++l) {
        long long np = tv.size[l + 1];
        long long warps = (np + kBuildParents - 1) / kBuildParents;
        long long blocks = (warps * 32 + 255) / 256;
        int rc = launch_step("tree_build_level_kernel", tree_build_level_kernel, dim3((unsigned)blocks), dim3(256), 0, st,
                             (const unsigned long long *)tv.lvl[l], tv.lvl[l + 1], tv.size[l], np);
        if (rc) return rc;
    }
    return APEXB200_OK;
}

int apexb200_tree_update(apexb200_replay_t h, const int32_t *idx, const float *prio, int32_t count, float eps,
                         void *stream) {
    APEX_REQUIRE(h && (count == 0 || (idx && prio)), APEXB200_EINVAL, "tree_update: null argument");
    APEX_REQUIRE(count >= 0, APEXB200_EINVAL, "tree_update: negative count");
    return run_tree_update(h, idx, prio, 0, count, nullptr, eps, (cudaStream_t)stream);
}

int apexb200_tree_set_leaves(apexb200_replay_t h, const int32_t *idx, const uint64_t *leaves, int32_t count,
                             void *stream) {
    APEX_REQUIRE(h && (count == 0 || (idx && leaves)), APEXB200_EINVAL, "tree_set_leaves: null argument");
    APEX_REQUIRE(count >= 0, APEXB200_EINVAL, "tree_set_leaves: negative count");
    return run_tree_update(h, idx, leaves, 1, count, nullptr, 0.0f, (cudaStream_t)stream);
}

int apexb200_tree_total(apexb200_replay_t h, uint64_t *total, void *stream) {
    APEX_REQUIRE(h && total, APEXB200_EINVAL, "tree_total: null argument");
    return launch_step("tree_total_kernel", tree_total_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, h->tree,
                       (unsigned long long *)total);
}

int apexb200_tree_geometry(apexb200_replay_t h, int32_t *n_levels_host, int64_t *sizes_host) {
    APEX_REQUIRE(h && n_levels_host && sizes_host, APEXB200_EINVAL, "tree_geometry: null argument");
    *n_levels_host = h->tree.n_levels;
    for (int l = 0; l < kMaxLevels; ++l) sizes_host[l] = l < h->tree.n_levels ? h->tree.size[l] : 0;
    return APEXB200_OK;
}

int apexb200_tree_read(apexb200_replay_t h, int32_t level, int64_t offset, int64_t count, uint64_t *out,
                       void *stream) {
    APEX_REQUIRE(h && out, APEXB200_EINVAL, "tree_read: null argument");
    APEX_REQUIRE(level >= 0 && level < h->tree.n_levels, APEXB200_EINVAL, "tree_read: level %d out of range", level);
    APEX_REQUIRE(offset >= 0 && count >= 0 && offset + count <= h->tree.size[level], APEXB200_EINVAL,
                 "tree_read: range outside level");
    APEX_CUDA(cudaMemcpyAsync(out, h->tree.lvl[level] + offset, (size_t)count * 8, cudaMemcpyDefault,
                              (cudaStream_t)stream));
    return APEXB200_OK;
}

int apexb200_tree_load_leaves(apexb200_replay_t h, const uint64_t *leaves, int64_t offset, int64_t count,
                              void *stream) {
    APEX_REQUIRE(h && leaves, APEXB200_EINVAL, "tree_load_leaves: null argument");
    APEX_REQUIRE(offset >= 0 && count >= 0 && offset + count <= h->tree.size[0], APEXB200_EINVAL,
                 "tree_load_leaves: range outside leaves");
    APEX_CUDA(cudaMemcpyAsync(h->tree.lvl[0] + offset, leaves, (size_t)count * 8, cudaMemcpyDefault,
                              (cudaStream_t)stream));
    return APEXB200_OK;
}

static int launch_sample(apexb200_replay_t h, const double *u, const uint64_t *targets, int32_t count,
                         const int32_t *count_dev, int32_t *idx_out, uint64_t *leaf_out, uint64_t *total_out,
                         int32_t *fsel_out, void *stream, uint64_t seed = 0, uint64_t call = 0,
                         double *u_out = nullptr) {
    if (count == 0) return APEXB200_OK;
    const int twoC = 2 * h->ring.C;
    APEX_REQUIRE(!fsel_out || twoC <= 32, APEXB200_ERANGE, "sample: fsel_out needs 2*frames_per_state <= 32");
    const int warps_per_block = 4;
    const int blocks = (count + warps_per_block - 1) / warps_per_block;
    return launch_step("tree_sample_kernel", tree_sample_kernel, dim3(blocks), dim3(warps_per_block * 32), 0,
                       (cudaStream_t)stream, h->tree, u, (const unsigned long long *)targets, count, count_dev, idx_out,
                       (unsigned long long *)leaf_out, (unsigned long long *)total_out, (const int32_t *)h->ring.fidx,
                       twoC, fsel_out, h->status, (unsigned long long)seed, (unsigned long long)call, u_out);
}

int apexb200_sample(apexb200_replay_t h, const double *u, int32_t count, int32_t *idx_out, uint64_t *leaf_out,
                    uint64_t *total_out, int32_t *fsel_out, void *stream) {
    APEX_REQUIRE(h && count >= 0 && (count == 0 || (u && idx_out)), APEXB200_EINVAL, "sample: bad argument");
    return launch_sample(h, u, nullptr, count, nullptr, idx_out, leaf_out, total_out, fsel_out, stream);
}

int apexb200_sample_stratified(apexb200_replay_t h, uint64_t seed, int32_t count, int32_t *idx_out,
                               uint64_t *leaf_out, uint64_t *total_out, int32_t *fsel_out, double *u_out,
                               void *stream) {
    APEX_REQUIRE(h && count >= 0 && (count == 0 || idx_out), APEXB200_EINVAL, "sample_stratified: bad argument");
    return launch_sample(h, nullptr, nullptr, count, nullptr, idx_out, leaf_out, total_out, fsel_out, stream, seed,
                         h->draw_calls++, u_out);
}

int apexb200_sample_targets(apexb200_replay_t h, const uint64_t *targets, const int32_t *count_dev,
                            int32_t max_count, int32_t *idx_out, uint64_t *leaf_out, int32_t *fsel_out,
                            void *stream) {
    APEX_REQUIRE(h && max_count >= 0 && (max_count == 0 || (targets && idx_out)), APEXB200_EINVAL,
                 "sample_targets: bad argument");
    return launch_sample(h, nullptr, targets, max_count, count_dev, idx_out, leaf_out, nullptr, fsel_out, stream);
}

int apexb200_shard_exchange(apexb200_replay_t h, uint64_t *totals_out, void *stream) {
    APEX_REQUIRE(h, APEXB200_EINVAL, "shard_exchange: null handle");
    return launch_step("shard_exchange_kernel", shard_exchange_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, h->tree,
                       h->peers, h->status, (unsigned long long *)totals_out);
}

int apexb200_shard_push(apexb200_replay_t h, void *stream) {
    APEX_REQUIRE(h, APEXB200_EINVAL, "shard_push: null handle");
    return run_peer_push(h, (cudaStream_t)stream);
}

int apexb200_peer_wait_stats(apexb200_replay_t h, uint64_t *stats_host, int32_t reset, void *stream) {
    APEX_REQUIRE(h && stats_host, APEXB200_EINVAL, "peer_wait_stats: null argument");
    cudaStream_t st = (cudaStream_t)stream;
    APEX_CUDA(cudaMemcpyAsync(stats_host, h->status + 52, 16, cudaMemcpyDeviceToHost, st));
    if (reset) APEX_CUDA(cudaMemsetAsync(h->status + 52, 0, 16, st));
    APEX_CUDA(cudaStreamSynchronize(st));
    return APEXB200_OK;
}

int apexb200_peer_push_on_update(apexb200_replay_t h, int32_t enabled) {
    APEX_REQUIRE(h, APEXB200_EINVAL, "peer_push_on_update: null handle");
    h->push_on_update = enabled ? 1 : 0;
    return APEXB200_OK;
}

int apexb200_shard_targets(const uint64_t *totals, int32_t G, int32_t rank, const double *u, int32_t count,
                           uint64_t *local_targets, int32_t *local_pos, int32_t *local_count, void *stream) {
    APEX_REQUIRE(totals && u && local_targets && local_pos && local_count, APEXB200_EINVAL,
                 "shard_targets: null argument");
    APEX_REQUIRE(G >= 1 && rank >= 0 && rank < G && count >= 0, APEXB200_EINVAL, "shard_targets: bad G/rank/count");
    return launch_step("shard_targets_kernel", shard_targets_kernel, dim3(1), dim3(1024), 0, (cudaStream_t)stream,
                       (const unsigned long long *)totals, G, rank, u, count, (unsigned long long *)local_targets,
                       local_pos, local_count);
}

}  // extern "C"
