// Shared internals of libapexb200.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>

#include "apexb200.h"

namespace apexb200 {

constexpr int kFanout = 32;
constexpr int kMaxLevels = 8;
constexpr int kNumSMs = 148;                 // B200

// POD view of the 32-ary sum tree, passed to kernels by value.
// lvl[0] = leaves (N entries), lvl[l+1][i] = sum lvl[l][32i .. 32i+31]; every level
// is zero-padded to a multiple of 32 entries; the top level has <= 32 entries and
// its sum is the total.
struct TreeView {
    unsigned long long *lvl[kMaxLevels];
    long long size[kMaxLevels];
    int n_levels;
};

// Peer-mapped exchange buffers of a group of shards (one per GPU).  xbuf[g] is rank g's
// buffer as seen from this GPU (own allocation for g == rank, a CUDA-IPC mapping over
// NVLink otherwise).  Layout: u64 [2 slots][kMaxShards], one word = total << 4 | round & 15.
constexpr int kMaxShards = 16;
struct PeerView {
    unsigned long long *xbuf[kMaxShards];
    int world, rank;
};

struct RingView {
    uint8_t *frames;          // [N * fpr, F]
    int32_t *fidx;            // [N, 2C]   absolute frame slots of S_t (C) then S_tpn (C)
    int32_t *A;               // [N]
    float *R, *G;             // [N]
    float *q_t, *q_tpn;       // [N, nA] or null
    long long N, lane_cap;
    int lanes, fpr, lookback, C, F, nA;
    float alpha;
};

void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);
extern int g_pdl;                            // programmatic dependent launch for the step kernels (default on)
extern int g_td_cluster;                     // K5+K2 for B <= 1024 as one thread-block cluster (opt-in: measured slower)

#define APEX_CUDA(call)                                                    \
    do {                                                                   \
        cudaError_t e__ = (call);                                          \
        if (e__ != cudaSuccess) return ::apexb200::cuda_fail(e__, #call);  \
    } while (0)

#define APEX_REQUIRE(cond, code, ...)                                      \
    do {                                                                   \
        if (!(cond)) { ::apexb200::set_error(__VA_ARGS__); return (code); }\
    } while (0)

inline int launch_check(const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, what);
    return APEXB200_OK;
}

// Programmatic dependent launch (sm_90+): a kernel launched through launch_step() may be
// scheduled while its predecessor in the stream is still running; it must execute
// pdl_prologue() as its first statement, which blocks until the predecessor grid has
// completed and its writes are visible, and lets the *next* kernel start its own ramp-up.
// At 10-20 us per step the ~2 us launch ramp of every kernel is otherwise exposed.
__device__ __forceinline__ void pdl_prologue() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// `cluster` > 1 launches the grid as thread-block clusters of that many CTAs (grid.x must be
// a multiple of it).
template <typename... KArgs, typename... Args>
inline int launch_step_cluster(const char *what, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                               cudaStream_t st, int cluster, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int n = 0;
    if (g_pdl) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    if (cluster > 1) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = (unsigned)cluster;
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
    if (e != cudaSuccess) return cuda_fail(e, what);
    return APEXB200_OK;
}

template <typename... KArgs, typename... Args>
inline int launch_step(const char *what, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                       cudaStream_t st, Args... args) {
    return launch_step_cluster(what, kernel, grid, block, smem, st, 1, args...);
}

// fp32 priority^alpha -> u64 fixed-point leaf (exact function of the fp32 value).
__device__ __forceinline__ unsigned long long leaf_from_prio(float p) {
    if (!(p > 0.0f) || isinf(p)) return (p > 0.0f) ? APEXB200_LEAF_MAX : 0ull;   // NaN/neg -> 0, +inf -> max
    double s = (double)p * 16777216.0;
    if (s >= (double)APEXB200_LEAF_MAX) return APEXB200_LEAF_MAX;
    return (unsigned long long)s;
}

__device__ __forceinline__ float prio_pow(float abs_td, float alpha, float eps) {
    float b = eps > 0.0f ? fmaxf(abs_td, eps) : abs_td;
    return powf(b, alpha);
}

// Device control block of a handle: int32 words [0] status, [4..5] u64 update epoch,
// [6..7] u64 exchange round, [16..47] u64 totals[kMaxShards] of the last exchange, [48..51] arrival words of the
// host-facing step, [52..53] u64 SM clocks the sharded sampling kernel spent waiting for its peers' totals (CTA 0,
// slowest peer, summed over launches), [54..55] u64 number of such launches (apexb200_peer_wait_stats).
__device__ __forceinline__ unsigned long long read_round(const int32_t *ctrl) {
    return *(const volatile unsigned long long *)(ctrl + 6);
}
__device__ __forceinline__ const unsigned long long *exchanged_totals(const int32_t *ctrl) {
    return (const unsigned long long *)(ctrl + 16);
}
__device__ __forceinline__ unsigned long long read_epoch(const int32_t *ctrl) {
    return __ldcg((const unsigned long long *)(ctrl + 4));     // L2 (never a stale L1 line); written by earlier kernels only
}
__device__ __forceinline__ void bump_epoch(int32_t *ctrl) {        // one thread of a tree-changing kernel
    // a reduction without return value: the issuing warp does not wait for the round trip a volatile read-modify-
    // write would cost it at the head of the kernel (writers of one handle are ordered by the stream)
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0)
        atomicAdd((unsigned long long *)(ctrl + 4), 1ull);
}

// The same for a full warp whose leaves are adjacent in index order (sorted batch rows, or the
// consecutive slots of an insert): rows that share an ancestor are neighbours, so the deltas are
// summed per ancestor with a segmented shuffle scan and one atomic per distinct ancestor is issued
// instead of one per row (hundreds of rows hit 2 nodes at the top level of a 2M-record tree).
// Must be called by all 32 lanes; lanes without an update pass i < 0 (they never merge) or d = 0.
// Levels below `first_merged` get one atomic per row (batch rows rarely share a level-1 parent;
// the consecutive slots of an insert always do: first_merged = 1 there).
__device__ __forceinline__ void warp_propagate_adjacent(const TreeView &tv, long long i, unsigned long long d,
                                                        int first_merged = 2) {
    const int lane = threadIdx.x & 31;
    for (int l = 1; l < first_merged && l < tv.n_levels; ++l)
        if (i >= 0 && d != 0ull) atomicAdd(tv.lvl[l] + (i >> (5 * l)), d);
    for (int l = first_merged; l < tv.n_levels; ++l) {
        const long long key = i >= 0 ? (i >> (5 * l)) : (long long)(-1 - lane);
        unsigned long long sum = i >= 0 ? d : 0ull;
        // Segmented inclusive scan with head flags: a segment is a run of neighbouring lanes with the same
        // ancestor.  Accumulation stops at a segment head, so a key that reappears later in the warp (ring
        // wrap inside the warp: A.. B.. A..) forms a second segment with its own atomic instead of pulling
        // in the first run's sum (comparing keys at distance o would add it twice).
        const long long kp = __shfl_up_sync(0xffffffffu, key, 1);
        unsigned flag = (lane == 0 || kp != key) ? 1u : 0u;
        const unsigned heads = __ballot_sync(0xffffffffu, flag != 0u);
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long vs = __shfl_up_sync(0xffffffffu, sum, o);
            const unsigned fs = __shfl_up_sync(0xffffffffu, flag, o);
            if (lane >= o && !flag) { sum += vs; flag |= fs; }
        }
        const bool tail = lane == 31 || ((heads >> (lane + 1)) & 1u);
        if (tail && sum != 0ull && key >= 0) atomicAdd(tv.lvl[l] + key, sum);
    }
}

// ---- exchange of shard totals over peer memory (multi-GPU) ---------------------------
// One 8-byte word per (slot, rank): total << 4 | round & 15.  Total and tag travel in the
// same naturally aligned word, so no fence is needed between them (a shard total < 2^60).
// Round r lives in slot r & 1.  A rank only pushes round r+1 after it has seen every
// peer's round-r word (peer_wait_round), which proves the peers are done with round r-1,
// whose slot the push overwrites; hence a slot sees tags r, r+2, r+4, ... and the 4-bit
// tag cannot alias.  Both helpers are called by the lanes g < world of one warp.
// `waited` (nullable): SM clocks from the first look at the word until it carried the round's tag.
__device__ __forceinline__ unsigned long long peer_wait_round(const PeerView &pv, unsigned long long round, int g,
                                                              int32_t *ctrl, long long *waited = nullptr) {
    volatile unsigned long long *in = pv.xbuf[pv.rank] + (size_t)((int)(round & 1ull) * kMaxShards + g);
    const unsigned long long tag = round & 15ull;
    const long long t0 = clock64();
    unsigned long long w = *in;
    while ((w & 15ull) != tag) {
        if (clock64() - t0 > (5ll << 31)) { atomicOr(ctrl, APEXB200_ST_PEER_TIMEOUT); return 0ull; }   // ~5 s: peer gone
        w = *in;
    }
    if (waited) *waited = clock64() - t0;
    return w >> 4;
}
__device__ __forceinline__ void peer_push_round(const PeerView &pv, unsigned long long round, int g,
                                                unsigned long long total) {
    volatile unsigned long long *out = pv.xbuf[g] + (size_t)((int)(round & 1ull) * kMaxShards + pv.rank);
    *out = (total << 4) | (round & 15ull);
}

// Add `d` (mod 2^64) to every ancestor of leaf i.  Integer adds commute, so the
// tree is exact whatever order concurrent updates land in.
__device__ __forceinline__ void propagate_delta(const TreeView &tv, long long i, unsigned long long d) {
    #pragma unroll 1
    for (int l = 1; l < tv.n_levels; ++l) {
        i >>= 5;
        atomicAdd(tv.lvl[l] + i, d);
    }
}

// Counter-based uniform for the in-kernel stratified draws: splitmix64 finaliser over
// (seed, update epoch, host call index, draw position) -> 53 random bits -> [0,1).
// The update epoch lives on the device and is advanced by every kernel that changes the
// tree (bump_epoch), so a CUDA-graph replay of a sample -> ... -> update step draws fresh
// numbers each time without any host-side state; the call index separates several
// sample calls between two updates.
__device__ __forceinline__ double strat_uniform(unsigned long long seed, unsigned long long epoch,
                                                unsigned long long call, int j, int count) {
    unsigned long long z = seed ^ (epoch * 0xD1342543DE82EF95ull) ^ (call * 0xA24BAED4963EE407ull);
    z += (unsigned long long)(j + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    const double v = (double)(z >> 11) * 0x1.0p-53;
    // u_j = (j + v_j) / B.  For a power-of-two B the quotient is an exact scaling by 2^-k: one multiply gives the
    // same bits as the fp64 division, which is a long dependent instruction sequence in front of the tree search
    // (0.3 us of every sampling launch, tools/timeline.py)
    if ((count & (count - 1)) == 0) {
        const int k = 31 - __clz(count);
        return ((double)j + v) * __longlong_as_double((long long)(1023 - k) << 52);
    }
    return ((double)j + v) / (double)count;
}

// K3 core.  One warp per draw.  At each level the warp loads the 32 children of the
// current node (one 256-byte line pair), forms their inclusive prefix sums with
// shuffles, and the number of prefixes <= target is the child to enter — exactly
// searchsorted(cumsum, target, side='right') restricted to that node.  The chosen
// child and the remaining target are broadcast by shuffle; no shared memory.
// The top-level load does not depend on the draw, so callers fetch it first
// (top_level_value) and compute the draw while it is in flight.
// When the level below the top has at most 64 entries (N in (32^k, 2*32^k]: 2M, 64k records ...) the top
// level is nearly empty; its children are then fetched together with it (two values per lane) and the two
// levels are searched as one 64-entry level: one dependent L2 round trip less per draw (0.33 us).  The flat
// right-sided search over the 64 prefix sums selects exactly the entry the two-level search selects.
struct TopValues {
    unsigned long long top;        // lvl[top][lane]
    unsigned long long a, b;       // merged mode: lvl[top-1][lane], lvl[top-1][lane + 32]
};
__device__ __forceinline__ bool top_is_merged(const TreeView &tv) {
    return tv.n_levels >= 2 && tv.size[tv.n_levels - 2] <= 2 * kFanout;
}
__device__ __forceinline__ TopValues top_level_value(const TreeView &tv) {
    const int lane = threadIdx.x & 31;
    TopValues tvv;
    tvv.top = tv.lvl[tv.n_levels - 1][lane];
    tvv.a = tvv.b = 0ull;
    if (top_is_merged(tv)) {                                   // padded to a multiple of 32: lane + 32 only if it exists
        const unsigned long long *l2 = tv.lvl[tv.n_levels - 2];
        tvv.a = l2[lane];
        if (tv.size[tv.n_levels - 2] > kFanout) tvv.b = l2[lane + kFanout];
    }
    return tvv;
}

__device__ __forceinline__ unsigned long long warp_inclusive_scan(unsigned long long v, int lane) {
    unsigned long long incl = v;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned long long n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    return incl;
}

// `stamps` (measurement aid, normally null and compiled away): clock64() when the search of the top level(s) is
// done and after every further level.
__device__ __forceinline__ long long warp_descend(const TreeView &tv, const TopValues &vtop, unsigned long long t,
                                                  bool from_u, double u, unsigned long long *leaf,
                                                  unsigned long long *total, int32_t *status,
                                                  long long *stamps = nullptr) {
    const int lane = threadIdx.x & 31;
    const int top = tv.n_levels - 1;
    const bool merged = top_is_merged(tv);
    long long node = 0;
    unsigned long long lf = 0ull;
    int first = top;                                           // first level the loop below has to visit
    {
        // total and target from the top level (or from its 64 children: the same sum)
        const unsigned long long v1 = merged ? vtop.a : vtop.top;
        const unsigned long long incl1 = warp_inclusive_scan(v1, lane);
        const unsigned long long sum1 = __shfl_sync(0xffffffffu, incl1, 31);
        unsigned long long incl2 = 0ull, tot = sum1;
        if (merged) {
            incl2 = sum1 + warp_inclusive_scan(vtop.b, lane);
            tot = __shfl_sync(0xffffffffu, incl2, 31);
        }
        *total = tot;
        if (tot == 0ull) { if (lane == 0) atomicOr(status, APEXB200_ST_EMPTY_TREE); return -1; }
        if (from_u) {
            t = (unsigned long long)(u * (double)tot);              // one fp64 multiply, truncation
            if (t >= tot) t = tot - 1ull;
        } else if (t >= tot) {
            if (lane == 0) atomicOr(status, APEXB200_ST_BAD_TARGET);
            t = tot - 1ull;
        }
        int c = __popc(__ballot_sync(0xffffffffu, incl1 <= t));
        unsigned long long excl, val;
        if (merged) {
            c += __popc(__ballot_sync(0xffffffffu, incl2 <= t));
            c = min(c, 2 * kFanout - 1);
            const bool hi = c >= kFanout;
            excl = __shfl_sync(0xffffffffu, hi ? incl2 - vtop.b : incl1 - v1, c & 31);
            val = __shfl_sync(0xffffffffu, hi ? vtop.b : v1, c & 31);
            first = top - 2;
        } else {
            c = min(c, 31);
            excl = __shfl_sync(0xffffffffu, incl1 - v1, c);
            val = __shfl_sync(0xffffffffu, v1, c);
            first = top - 1;
        }
        if (first < 0) lf = val;                                   // the searched level was the leaf level
        t -= excl;
        node = c;
    }
    int n_stamp = 0;
    if (stamps) stamps[n_stamp++] = clock64();
    for (int l = first; l >= 0; --l) {
        const unsigned long long v = tv.lvl[l][node * kFanout + lane];      // padded levels: in bounds
        const unsigned long long incl = warp_inclusive_scan(v, lane);
        int c = __popc(__ballot_sync(0xffffffffu, incl <= t));
        c = min(c, 31);
        const unsigned long long excl = __shfl_sync(0xffffffffu, incl - v, c);
        if (l == 0) lf = __shfl_sync(0xffffffffu, v, c);
        t -= excl;
        node = node * kFanout + c;
        if (stamps && n_stamp < 6) stamps[n_stamp++] = clock64();
    }
    *leaf = lf;
    return node;
}

}  // namespace apexb200

struct apexb200_replay {
    apexb200_replay_config cfg;
    apexb200::TreeView tree;
    apexb200::RingView ring;
    void *tree_block = nullptr;        // one allocation for all levels
    uint32_t *stamp = nullptr;         // [N] duplicate-resolution stamps (K2); all zero between calls
    int32_t *status = nullptr;         // device control block (256 B), see read_epoch / read_round
    unsigned long long *xbuf = nullptr; // this shard's exchange buffer (peers write into it)
    apexb200::PeerView peers;          // xbuf of every rank; world = 1, xbuf[0] = own until peer_connect
    void *opened[apexb200::kMaxShards] = {};   // IPC mappings to close
    int push_on_update = 0;            // td_fwd_update also publishes the new total to the peers
    // apexb200_host_step: two sets of device staging + copy streams, so that the H2D copy of step
    // k+1 and the D2H copy of step k overlap the kernels of the other step
    unsigned char *hs_in[APEXB200_HOST_STEP_SLOTS] = {};
    float *hs_out[APEXB200_HOST_STEP_SLOTS] = {};
    size_t hs_in_cap = 0, hs_out_cap = 0;
    cudaStream_t hs_copy_in = nullptr, hs_copy_out = nullptr;
    cudaEvent_t hs_ev_in[APEXB200_HOST_STEP_SLOTS] = {}, hs_ev_k[APEXB200_HOST_STEP_SLOTS] = {},
                hs_ev_out[APEXB200_HOST_STEP_SLOTS] = {};
    unsigned *hs_seq_host = nullptr;   // pinned [APEXB200_HOST_STEP_SLOTS]: source of the "inputs landed" words
    unsigned hs_seq[APEXB200_HOST_STEP_SLOTS] = {};
    unsigned long long draw_calls = 0; // host: stratified-draw calls issued (the `call` of strat_uniform)
    std::vector<long long> lane_count; // host ring heads
    int64_t device_bytes = 0;
};

namespace apexb200 {
// fused K3+K4 launcher shared with the host-facing step (gather.cu -> replay.cu)
bool sample_gather_is_fused(const apexb200_replay *h);
int run_sample_gather(apexb200_replay *h, const double *u, uint64_t seed, int32_t count, int32_t *idx_out,
                      int32_t *fsel_out, uint64_t *leaf_out, uint64_t *total_out, double *u_out, uint8_t *S_t,
                      uint8_t *S_tpn, int32_t *A, float *R, float *Gamma, float *q_t, float *q_tpn,
                      const unsigned *ready, unsigned ready_value, cudaStream_t stream,
                      int out_format = APEXB200_FMT_U8);
int sharded_wave_rows(const apexb200_replay *h, int global_count, int max_rows);
int run_sample_gather_sharded(apexb200_replay *h, uint64_t seed, int32_t global_count, int32_t max_rows,
                              int32_t *idx_out, int32_t *fsel_out, int32_t *pos_out, double *u_out,
                              int32_t *count_out, uint8_t *S_t, uint8_t *S_tpn, int32_t *A, float *R, float *Gamma,
                              float *q_t, float *q_tpn, const unsigned *ready, unsigned ready_value,
                              cudaStream_t stream, int out_format = APEXB200_FMT_U8);
}  // namespace apexb200
