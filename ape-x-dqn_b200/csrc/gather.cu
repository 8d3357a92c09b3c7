// K4: gather of the sampled stacked frames (+ n-step metadata) straight into the
// learner's batch tensors, and the fused K3+K4 kernel (proportional search + gather in
// one launch).  Byte movement only: 2C frames of F bytes per record are read from
// random ring slots and written to dense [B, C, F] outputs.  The frame reads come from
// HBM (the frame store is far larger than L2); the batch that is written stays in the
// 126 MB L2 for its consumer.
#include "common.cuh"

namespace apexb200 {

static int g_gather_variant = 1;

__device__ __forceinline__ uint4 ld_stream(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

struct GatherArgs {
    const int32_t *idx, *fsel, *count_dev;
    uint8_t *S_t, *S_tpn;            // [rows, C, F] uint8, or (elem_bytes = 4) float32 in the same shape
    int32_t *A;
    float *R, *G, *q_t, *q_tpn;
    int count;
    int elem_bytes;                  // bytes per output element: 1 (uint8 as stored) or 4 (float32, learner.py:37-38)
};

struct SampleArgs {
    const double *u;                 // caller's draws, or null -> stratified draws generated here
    unsigned long long seed, call;
    int32_t *idx_out, *fsel_out;
    unsigned long long *leaf_out, *total_out;
    double *u_out;
};

// Work item `it` in [0, count*2C): row b = it / 2C, k = it % 2C; k < C is frame k of S_t,
// else frame k-C of S_tpn.
__device__ __forceinline__ uint8_t *item_dst(const RingView &rv, const GatherArgs &ga, long long b, int k) {
    uint8_t *base = k < rv.C ? ga.S_t : ga.S_tpn;
    int kk = k < rv.C ? k : k - rv.C;
    return base + (b * rv.C + kk) * (long long)rv.F * ga.elem_bytes;
}

__device__ __forceinline__ bool resolve_item(const RingView &rv, const GatherArgs &ga, long long it, int count,
                                             const uint8_t **src, uint8_t **dst, int32_t *status) {
    const int twoC = 2 * rv.C;
    long long b = it / twoC;
    int k = (int)(it - b * twoC);
    if (b >= count) return false;
    long long f;
    if (ga.fsel) {
        f = ga.fsel[it];
    } else {
        long long slot = ga.idx[b];
        if (slot < 0 || slot >= rv.N) { atomicOr(status, APEXB200_ST_BAD_INDEX); return false; }
        f = rv.fidx[slot * twoC + k];
    }
    if (f < 0 || f >= rv.N * rv.fpr) { atomicOr(status, APEXB200_ST_BAD_INDEX); return false; }
    *src = rv.frames + f * rv.F;
    *dst = item_dst(rv, ga, b, k);
    return true;
}

__device__ __forceinline__ void copy_meta_row(const RingView &rv, const GatherArgs &ga, int b, long long slot) {
    if (ga.A) ga.A[b] = rv.A[slot];
    if (ga.R) ga.R[b] = rv.R[slot];
    if (ga.G) ga.G[b] = rv.G[slot];
    if (ga.q_t && rv.q_t)
        for (int a = 0; a < rv.nA; ++a) ga.q_t[(long long)b * rv.nA + a] = rv.q_t[slot * rv.nA + a];
    if (ga.q_tpn && rv.q_tpn)
        for (int a = 0; a < rv.nA; ++a) ga.q_tpn[(long long)b * rv.nA + a] = rv.q_tpn[slot * rv.nA + a];
}

__device__ __forceinline__ void gather_meta(const RingView &rv, const GatherArgs &ga, int b, int count, int32_t *status) {
    if (b >= count) return;
    long long slot = ga.idx[b];
    if (slot < 0 || slot >= rv.N) { atomicOr(status, APEXB200_ST_BAD_INDEX); return; }
    copy_meta_row(rv, ga, b, slot);
}

// One frame through shared memory with the TMA (bulk async-copy) engine: global -> shared
// completes on an mbarrier, shared -> global drains it; the bytes never enter registers.
// A frame can move as several chunks, each with its own mbarrier (chunk c stored while
// chunks c+1.. are landing); measured on B200 (profiles/r01_kernel_experiments.md) whole
// 7 KB frames are faster than 3 x 2.3 KB chunks, because the launch is bound by L2 slice
// throughput rather than by phase serialisation, so frames up to 16 KB move as one chunk.
// Frame reads are single-use (evict_first) so that they do not push the sum tree or the
// freshly written batch out of L2.
constexpr int kMaxChunks = 4;

__device__ __forceinline__ int frame_chunks(int frame_bytes) {
    if (frame_bytes <= 16384) return 1;
    const int v = frame_bytes >> 4;
    return (v % 4 == 0) ? 4 : (v % 3 == 0) ? 3 : (v % 2 == 0) ? 2 : 1;
}

__device__ __forceinline__ void tma_load_frame(const uint8_t *src, uint8_t *stage, unsigned long long *bars, unsigned bytes,
                                               int nch) {
    const unsigned buf_s = (unsigned)__cvta_generic_to_shared(stage);
    const unsigned chunk = bytes / nch;
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    for (int c = 0; c < nch; ++c) {
        const unsigned bar_s = (unsigned)__cvta_generic_to_shared(bars + c);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (int c = 0; c < nch; ++c) {
        const unsigned bar_s = (unsigned)__cvta_generic_to_shared(bars + c);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(chunk) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                     ::"r"(buf_s + c * chunk), "l"(src + c * chunk), "r"(chunk), "r"(bar_s), "l"(pol) : "memory");
    }
}

// The same in two steps for kernels that reuse a lane's staging buffer over several waves: the
// mbarriers are initialised once, every wave arms them again (phase parity = wave & 1).
__device__ __forceinline__ void tma_init_bars(unsigned long long *bars, int nch) {
    for (int c = 0; c < nch; ++c) {
        const unsigned bar_s = (unsigned)__cvta_generic_to_shared(bars + c);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void tma_issue_frame(const uint8_t *src, uint8_t *stage, unsigned long long *bars,
                                                unsigned bytes, int nch) {
    const unsigned buf_s = (unsigned)__cvta_generic_to_shared(stage);
    const unsigned chunk = bytes / nch;
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    for (int c = 0; c < nch; ++c) {
        const unsigned bar_s = (unsigned)__cvta_generic_to_shared(bars + c);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(chunk) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                     ::"r"(buf_s + c * chunk), "l"(src + c * chunk), "r"(chunk), "r"(bar_s), "l"(pol) : "memory");
    }
}

// Drain: every lane that moves a frame polls its own mbarriers without blocking and
// issues the shared -> global copy of a chunk the moment that chunk has landed, so the
// stores of early frames overlap the loads of late ones (a blocking wait would hold the
// whole warp until its slowest frame arrived).  Must be called by all 32 lanes.
__device__ __forceinline__ void tma_drain_frames(bool moving, uint8_t *dst, uint8_t *stage, unsigned long long *bars,
                                                 unsigned bytes, int nch, unsigned parity = 0u,
                                                 long long *landed = nullptr) {
    const unsigned buf_s = (unsigned)__cvta_generic_to_shared(stage);
    const unsigned chunk = bytes / nch;
    int next = moving ? 0 : nch;                     // next chunk this lane has to store
    while (__any_sync(0xffffffffu, next < nch)) {
        if (next < nch) {
            const unsigned bar_s = (unsigned)__cvta_generic_to_shared(bars + next);
            unsigned done;
            asm volatile("{\n .reg .pred p;\n mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
                         " selp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(bar_s), "r"(parity) : "memory");
            if (done) {
                if (landed) *landed = clock64();             // measurement aid (null and compiled away otherwise)
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             ::"l"(dst + next * chunk), "r"(buf_s + next * chunk), "r"(chunk) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                ++next;
            }
        }
    }
    if (landed == nullptr) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// Blocking wait for every chunk of this lane's frame (phase `parity`).
__device__ __forceinline__ void tma_wait_frame(unsigned long long *bars, int nch, unsigned parity) {
    for (int c = 0; c < nch; ++c) {
        const unsigned bar_s = (unsigned)__cvta_generic_to_shared(bars + c);
        unsigned done = 0u;
        while (!done)
            asm volatile("{\n .reg .pred p;\n mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
                         " selp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(bar_s), "r"(parity) : "memory");
    }
}

// The frame slot a record's frame-slot row holds in the common case, computed without touching memory: record
// `node` of a frame-once ("stream") ring looks back over fixed offsets (S_tpn = the C frames ending at its own,
// S_t = the C frames ending n steps earlier, n = lookback - (C - 1)) unless it lies within C - 1 + n steps of an
// episode start; a record that owns its frames (materialised layout) names them in order.  Always a valid slot of
// the record's own lane; the row itself decides (see the fused kernels).
__device__ __forceinline__ long long guess_frame_slot(const RingView &rv, long long node, int k) {
    const unsigned lane_cap = (unsigned)rv.lane_cap;
    const unsigned ring_lane = (unsigned)node / lane_cap, p = (unsigned)node - ring_lane * lane_cap;
    const long long lane_frames = rv.lane_cap * rv.fpr;
    int rel;
    if (rv.fpr == 1) {
        const int back = rv.C - 1 - (k < rv.C ? k : k - rv.C);
        rel = k < rv.C ? -(rv.lookback - (rv.C - 1) + back) : -back;
    } else {
        rel = k - (2 * rv.C - 1);
    }
    long long f = (long long)p * rv.fpr + rv.fpr - 1 + rel;
    if (f < 0) f += lane_frames;
    return (long long)ring_lane * lane_frames + f;
}

// float32 outputs (the dtype the learner's network consumes, learner.py:37-38): the bulk-copy engine cannot
// convert, so a frame that has landed in shared memory is widened by the whole warp — 4 bytes in, one 16-byte
// vector of 4 floats out per lane and iteration — and written with ordinary vector stores.  Frames are taken in
// lane order; frame c's mbarrier is polled by all lanes.  `dst` is the owning lane's output pointer.
// Must be called by all 32 lanes.
__device__ __forceinline__ void f32_drain_frames(bool moving, uint8_t *dst, const uint8_t *warp_stage,
                                                 unsigned long long *warp_bars, int C, unsigned bytes, int nch,
                                                 unsigned my_parity) {
    const int lane = threadIdx.x & 31;
    const unsigned mask = __ballot_sync(0xffffffffu, moving);
    for (int c = 0; c < C; ++c) {
        const unsigned parity = __shfl_sync(0xffffffffu, my_parity, c);    // phase of the owning lane's barrier
        if (!((mask >> c) & 1u)) continue;
        for (int ch = 0; ch < nch; ++ch) {
            const unsigned bar_s = (unsigned)__cvta_generic_to_shared(warp_bars + c * kMaxChunks + ch);
            unsigned done = 0u;
            while (!done)
                asm volatile("{\n .reg .pred p;\n mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
                             " selp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(bar_s), "r"(parity) : "memory");
        }
        float4 *out = (float4 *)__shfl_sync(0xffffffffu, (unsigned long long)dst, c);
        const uchar4 *in = (const uchar4 *)(warp_stage + (size_t)c * bytes);
        const int nv = (int)(bytes >> 2);
        #pragma unroll 4
        for (int v = lane; v < nv; v += 32) {
            const uchar4 x = in[v];
            __stcs(out + v, make_float4((float)x.x, (float)x.y, (float)x.z, (float)x.w));
        }
    }
    __syncwarp();
}

// ---- variant 0: one warp per frame, 16-byte vectors, all loads of a frame in flight
// before the first store.  Kept as the non-TMA comparison point.
template <int UNROLL>
__global__ void __launch_bounds__(256)
gather_vec_kernel(RingView rv, GatherArgs ga, int32_t *status) {
    pdl_prologue();
    const int lane = threadIdx.x & 31;
    const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int count = ga.count;
    if (ga.count_dev) count = min(count, *ga.count_dev);
    const long long host_items = (long long)ga.count * 2 * rv.C;
    if (w >= host_items) {                                 // trailing warps copy the metadata rows
        gather_meta(rv, ga, (int)((w - host_items) * 32 + lane), count, status);
        return;
    }
    const uint8_t *src; uint8_t *dst;
    if (!resolve_item(rv, ga, w, count, &src, &dst, status)) return;
    const uint4 *s = (const uint4 *)src;
    uint4 *d = (uint4 *)dst;
    const int nv = rv.F >> 4;
    for (int base = 0; base < nv; base += 32 * UNROLL) {
        uint4 r[UNROLL];
        #pragma unroll
        for (int k = 0; k < UNROLL; ++k) {
            int v = base + k * 32 + lane;
            if (v < nv) r[k] = ld_stream(s + v);
        }
        #pragma unroll
        for (int k = 0; k < UNROLL; ++k) {
            int v = base + k * 32 + lane;
            if (v < nv) d[v] = r[k];
        }
    }
}

// Shared-memory budget for frame staging: one fat CTA per SM instead of many thin ones.
// With ~7 KB frames a thin (1-2 frame) CTA spends longer being launched than moving
// data; up to 32 frames per CTA (220.5 KB of the 227 KB an SM offers), spread so that a
// whole B=512, C=4 batch (4096 frames) is in flight at once in one launch wave.
constexpr size_t kStageBudget = 226 * 1024;
constexpr size_t kBarBytes = 32 * kMaxChunks * 8;           // mbarriers sit in front of the staging area

// ---- variant 1: TMA.  One lane owns one frame; up to 32 frames per 32-thread CTA.
__global__ void __launch_bounds__(32)
gather_tma_kernel(RingView rv, GatherArgs ga, int per_cta, int32_t *status) {
    extern __shared__ __align__(128) uint8_t smem[];
    unsigned long long *bars = (unsigned long long *)smem;
    uint8_t *stage = smem + kBarBytes;
    pdl_prologue();
    const int lane = threadIdx.x;
    const int nch = frame_chunks(rv.F);
    int count = ga.count;
    if (ga.count_dev) count = min(count, *ga.count_dev);
    const long long n_items = (long long)count * 2 * rv.C;
    const long long host_items = (long long)ga.count * 2 * rv.C;
    const long long first = (long long)blockIdx.x * per_cta;
    if (first >= host_items) {                              // trailing CTAs: metadata
        long long mb = blockIdx.x - (host_items + per_cta - 1) / per_cta;
        gather_meta(rv, ga, (int)(mb * 32 + lane), count, status);
        return;
    }
    const long long it = first + lane;
    const uint8_t *src = nullptr; uint8_t *dst = nullptr;
    const bool moving = lane < per_cta && it < n_items && resolve_item(rv, ga, it, count, &src, &dst, status);
    if (moving) tma_load_frame(src, stage + (size_t)lane * rv.F, bars + lane * kMaxChunks, (unsigned)rv.F, nch);
    tma_drain_frames(moving, dst, stage + (size_t)lane * rv.F, bars + lane * kMaxChunks, (unsigned)rv.F, nch);
}

// ---- fused K3 + K4: one warp per HALF draw (the C frames of S_t, or those of S_tpn),
// `wpc` warps per CTA.  Each warp descends the tree for its draw (the two halves of a draw
// repeat the descent: a few hundred bytes of L2 reads), lanes 0..C-1 read the frame slots
// of their half and each moves one frame with the TMA engine; the S_t half also writes the
// index and metadata outputs.  Half draws as the unit of work balance the frames over the
// SMs: a B=512, C=4 batch is 1024 units = 147 CTAs of 7 warps (28 frames, 193 KB of
// staging each), one CTA per SM, every frame of the batch in flight at once.
constexpr int kDbgWords = 20;
__device__ unsigned long long g_dbg_timeline[2048 * kDbgWords];      // written by the instrumented kernel only

__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// kTimed: the same code plus per-warp timestamps in g_dbg_timeline (apexb200_set_gather_variant(3), measurement aid).
template <bool kPolled, bool kF32, bool kTimed>
__device__ __forceinline__ void sample_gather_body(const TreeView &tv, const RingView &rv, const SampleArgs &sa,
                                                   const GatherArgs &ga, int32_t *ctrl, const unsigned *ready,
                                                   unsigned ready_value) {
    extern __shared__ __align__(128) uint8_t smem[];
    unsigned long long *bars = (unsigned long long *)smem;
    const unsigned long long g0 = kTimed ? gtimer() : 0ull;
    const long long c0 = kTimed ? clock64() : 0ll;
    pdl_prologue();
    const long long c1 = kTimed ? clock64() : 0ll;
    const int lane = threadIdx.x & 31;
    const int wl = threadIdx.x >> 5;
    const int wpc = blockDim.x >> 5;
    const int nch = frame_chunks(rv.F);
    const int unit = blockIdx.x * wpc + wl;
    const int w = unit >> 1, half = unit & 1;
    const int count = ga.count;
    const int C = rv.C, twoC = 2 * rv.C;
    if (w >= count) return;
    uint8_t *stage = smem + (size_t)wpc * 8 * kMaxChunks * 8 + (size_t)(wl * C + lane) * rv.F;
    unsigned long long *my_bars = bars + (wl * 8 + lane) * kMaxChunks;
    const bool strat = (sa.u == nullptr);
    const TopValues vtop = top_level_value(tv);     // in flight while the draw is formed
    double uw;
    if (strat) {
        uw = strat_uniform(sa.seed, read_epoch(ctrl), sa.call, w, count);
        if (sa.u_out && lane == 0 && half == 0) sa.u_out[w] = uw;
    } else if (!kPolled) {
        uw = sa.u[w];
    } else {
        // apexb200_host_step: the draws arrive by an H2D copy on another stream, followed by a stream-ordered
        // write of `ready_value` to *ready.  Polling it here keeps every cross-stream wait off the compute
        // stream, so this kernel still chains to its predecessor by programmatic dependent launch.  The copy
        // was issued a step or more ahead: normally the first load already sees the value.
        if (lane == 0) {
            const volatile unsigned *flag = ready;
            const long long t0 = clock64();
            while ((int)(*flag - ready_value) < 0) {
                if (clock64() - t0 > (5ll << 31)) { atomicOr(ctrl, APEXB200_ST_PEER_TIMEOUT); break; }   // ~5 s
            }
            __threadfence();
        }
        __syncwarp();
        uw = __ldcg(sa.u + w);                               // bypass L1: written by DMA while this CTA was resident
    }
    unsigned long long leaf, total;
    long long lv[6] = {0, 0, 0, 0, 0, 0};
    const long long c1b = kTimed ? clock64() : 0ll;
    const long long node = warp_descend(tv, vtop, 0ull, true, uw, &leaf, &total, ctrl, kTimed ? lv : nullptr);
    const long long c2 = kTimed ? clock64() : 0ll;
    const int k = half * C + lane;                           // this lane's frame among the record's 2C
    // The record's frame-slot row sits in HBM: one more dependent round trip before the first frame byte could
    // move.  The frame copy is therefore started on the slot the row holds in the common case (guess_frame_slot)
    // while the row itself is in flight, and the row decides afterwards: a lane whose guess was wrong (a record
    // next to an episode start) waits for the stray copy to land and fetches the right frame into the same
    // staging buffer (second phase of its mbarriers).
    long long f = -1, fg = -1, c3 = 0;
    const bool mine = node >= 0 && lane < C;
    if (mine) {
        f = rv.fidx[node * twoC + k];                        // issued now, used after the speculative copy is on its way
        fg = guess_frame_slot(rv, node, k);
        if (kTimed) c3 = clock64();
        tma_load_frame(rv.frames + fg * rv.F, stage, my_bars, (unsigned)rv.F, nch);
    }
    // bookkeeping overlaps the frame loads
    if (half == 0) {
        if (lane == 0) {
            sa.idx_out[w] = (int32_t)node;
            if (sa.leaf_out) sa.leaf_out[w] = node >= 0 ? leaf : 0ull;
            if (sa.total_out && w == 0) *sa.total_out = total;
        }
        if (node >= 0 && lane == 16) copy_meta_row(rv, ga, w, node);
    }
    if (lane < C && sa.fsel_out) sa.fsel_out[(long long)w * twoC + k] = (int32_t)f;
    bool moving = mine;
    unsigned parity = 0u;
    if (mine && f != fg) {
        const bool valid = f >= 0 && f < rv.N * rv.fpr;
        if (!valid) atomicOr(ctrl, APEXB200_ST_BAD_INDEX);
        tma_wait_frame(my_bars, nch, 0u);                    // the stray copy has landed: the buffer is free again
        moving = valid;
        if (valid) { tma_issue_frame(rv.frames + f * rv.F, stage, my_bars, (unsigned)rv.F, nch); parity = 1u; }
    }
    uint8_t *dst = moving ? item_dst(rv, ga, w, k) : nullptr;
    long long c_land = 0;
    if (kF32)
        f32_drain_frames(moving, dst, stage - (size_t)lane * rv.F, my_bars - lane * kMaxChunks, C, (unsigned)rv.F, nch, parity);
    else
        tma_drain_frames(moving, dst, stage, my_bars, (unsigned)rv.F, nch, parity, kTimed ? &c_land : nullptr);
    if (kTimed) {
        const long long c5 = clock64();
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        const long long c6 = clock64();
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");        // writes complete (measurement only)
        const long long c7 = clock64();
        long long lmin = moving ? c_land - c0 : (1ll << 60), lmax = moving ? c_land - c0 : 0, i3 = mine ? c3 - c0 : (1ll << 60);
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lmin = min(lmin, __shfl_xor_sync(0xffffffffu, lmin, o));
            lmax = max(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
            i3 = min(i3, __shfl_xor_sync(0xffffffffu, i3, o));
        }
        const unsigned wrong = __popc(__ballot_sync(0xffffffffu, mine && f != fg));
        if (lane == 0 && unit < 2048) {
            unsigned long long *d = g_dbg_timeline + (size_t)unit * kDbgWords;
            d[0] = g0; d[1] = (unsigned long long)(c1 - c0); d[2] = (unsigned long long)(c2 - c0); d[3] = (unsigned long long)i3;
            d[4] = (unsigned long long)lmin; d[5] = (unsigned long long)lmax; d[6] = (unsigned long long)(c5 - c0);
            d[7] = (unsigned long long)(c6 - c0); d[8] = (unsigned long long)(c7 - c0); d[9] = gtimer();
            unsigned smid;
            asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
            d[10] = smid; d[11] = (unsigned long long)blockIdx.x | ((unsigned long long)wrong << 32);
            d[12] = (unsigned long long)(c1b - c0);
            for (int q = 0; q < 6; ++q) d[13 + q] = lv[q] ? (unsigned long long)(lv[q] - c0) : 0ull;
        }
    }
}

__global__ void __launch_bounds__(512)
sample_gather_kernel(TreeView tv, RingView rv, SampleArgs sa, GatherArgs ga, int32_t *ctrl) {
    sample_gather_body<false, false, false>(tv, rv, sa, ga, ctrl, nullptr, 0u);
}

// Instrumented instantiation of the same body (apexb200_set_gather_variant(3)); read back with apexb200_debug_timeline.
__global__ void __launch_bounds__(512)
sample_gather_timed_kernel(TreeView tv, RingView rv, SampleArgs sa, GatherArgs ga, int32_t *ctrl) {
    sample_gather_body<false, false, true>(tv, rv, sa, ga, ctrl, nullptr, 0u);
}

// The same search + gather writing float32 batches (the network's input dtype) instead of uint8.
__global__ void __launch_bounds__(512)
sample_gather_f32_kernel(TreeView tv, RingView rv, SampleArgs sa, GatherArgs ga, int32_t *ctrl) {
    sample_gather_body<false, true, false>(tv, rv, sa, ga, ctrl, nullptr, 0u);
}

// Same kernel for the host-facing step: waits in-kernel for the arrival word of its draws.
__global__ void __launch_bounds__(512)
sample_gather_polled_kernel(TreeView tv, RingView rv, SampleArgs sa, GatherArgs ga, int32_t *ctrl,
                            const unsigned *ready, unsigned ready_value) {
    sample_gather_body<true, false, false>(tv, rv, sa, ga, ctrl, ready, ready_value);
}

// ---- fused K8 + K3 + K4 for one rank of a sharded replay (after shard_exchange_kernel).
// Rows are dealt to CTAs round-robin (row = cta + k * gridDim.x) so that the live rows —
// the first `count` of max_rows — spread over every CTA.
struct ShardArgs {
    unsigned long long seed, call;
    int global_count, max_rows;
    int32_t *pos_out, *count_out;
};

// target of global draw j: min(trunc(u_j * double(T)), T - 1); non-decreasing in j.
__device__ __forceinline__ unsigned long long draw_target(unsigned long long seed, unsigned long long round,
                                                          unsigned long long call, int j, int Bg,
                                                          unsigned long long T, double *u_out) {
    const double u = strat_uniform(seed, round, call, j, Bg);
    if (u_out) *u_out = u;
    unsigned long long t = (unsigned long long)(u * (double)T);
    return t >= T ? T - 1ull : t;
}

// Smallest j in [0, Bg] whose target is >= bound (Bg if none).  Stratified draws put
// target_j near j * T / Bg, so the answer lies within one draw of bound * Bg / T: the 32
// lanes first test a window around that estimate in one round; a 32-ary search over the
// whole range is the fallback.
__device__ __forceinline__ int first_draw_at_or_above(unsigned long long bound, unsigned long long seed,
                                                      unsigned long long round, unsigned long long call, int Bg,
                                                      unsigned long long T) {
    const int lane = threadIdx.x & 31;
    if (bound == 0ull) return 0;
    if (bound >= T) return Bg;
    {
        const long long est = (long long)((double)bound / (double)T * (double)Bg);
        const long long j = est - 15 + lane;                 // window [est-15, est+16]
        const bool ge = (j >= Bg) || (j >= 0 && draw_target(seed, round, call, (int)j, Bg, T, nullptr) >= bound);
        const unsigned m = __ballot_sync(0xffffffffu, ge);
        // monotone predicate: the answer is the first voting lane, provided the lane before it
        // exists in the window (did not vote) or the window starts at draw 0
        if (m != 0u) {
            const int f = __ffs(m) - 1;
            const long long ans = est - 15 + f;
            if (f > 0 || ans <= 0) return (int)(ans < 0 ? 0 : (ans > Bg ? Bg : ans));
        }
    }
    int lo = 0, hi = Bg;                       // answer in [lo, hi]
    while (lo < hi) {
        const int span = hi - lo;
        const int step = (span + 31) >> 5;     // candidates lo + (i+1)*step - 1, i = 0..31
        const int j = lo + (lane + 1) * step - 1;
        const bool ge = (j >= hi) || draw_target(seed, round, call, j, Bg, T, nullptr) >= bound;
        const unsigned m = __ballot_sync(0xffffffffu, ge);
        const int f = m ? __ffs(m) - 1 : 32;   // no vote: every candidate is below the bound and below hi
        const int new_hi = min(hi, lo + (f + 1) * step - 1);
        lo = lo + f * step;
        hi = new_hi;
    }
    return lo;
}

// Every draw the rank owns is delivered, whatever the split of priority mass between the shards
// (replay.py:52-57: B requested, B returned): the grid covers `wave_rows` rows at a time — sized by the
// launcher for the rank's expected share plus head-room — and a warp whose row is live goes on to row
// + wave_rows, + 2 wave_rows, ... until it passes the end of the rank's run of draws.  Targets are
// non-decreasing in the draw index, so the first dead row of a warp ends its loop.  The staging buffer
// and mbarriers of a lane are reused from wave to wave (phase parity = wave & 1).
template <bool kPolled, bool kF32>
__device__ __forceinline__ void sample_gather_sharded_body(const TreeView &tv, const RingView &rv, const SampleArgs &sa,
                                                           const GatherArgs &ga, const ShardArgs &sh, const PeerView &pv,
                                                           int32_t *ctrl, const unsigned *ready, unsigned ready_value) {
    extern __shared__ __align__(128) uint8_t smem[];
    unsigned long long *bars = (unsigned long long *)smem;
    pdl_prologue();
    const int lane = threadIdx.x & 31;
    const int wl = threadIdx.x >> 5;
    const int wpc = blockDim.x >> 5;
    const int nch = frame_chunks(rv.F);
    const int C = rv.C, twoC = 2 * rv.C;
    const TopValues vtop = top_level_value(tv);
    const int world = pv.world, rank = pv.rank;
    // the shard totals of the current round (pushed by shard_exchange_kernel or by the peers'
    // td_small_kernel) -> global total T and this rank's prefix interval [lo, hi).  One warp per
    // CTA polls the exchange buffer; the others take the words from shared memory.
    __shared__ unsigned long long s_totals[kMaxShards];
    // Both exchange slots are fetched together with the round word (which slot is current depends on the round):
    // one round trip instead of two dependent ones; only a word that has not arrived yet is polled for.
    unsigned long long w0 = 0ull, w1 = 0ull;
    if (wl == 0 && lane < world) {
        const unsigned long long *xb = pv.xbuf[rank];
        w0 = __ldcg(xb + lane);
        w1 = __ldcg(xb + kMaxShards + lane);
    }
    const unsigned long long round = __ldcg((const unsigned long long *)(ctrl + 6));
    if (wl == 0) {
        long long waited = 0;
        if (lane < world) {
            const unsigned long long w = (round & 1ull) ? w1 : w0;
            s_totals[lane] = (w & 15ull) == (round & 15ull) ? (w >> 4) : peer_wait_round(pv, round, lane, ctrl, &waited);
        }
        if (blockIdx.x == 0) {                               // attribution of the multi-GPU step: how long the peers' words took
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) waited = max(waited, __shfl_xor_sync(0xffffffffu, waited, o));
            if (lane == 0) {
                atomicAdd((unsigned long long *)(ctrl + 52), (unsigned long long)waited);
                atomicAdd((unsigned long long *)(ctrl + 54), 1ull);
            }
        }
    }
    if (kPolled && wl == 0 && lane == 0) {
        // apexb200_host_step_sharded: the Q rows of the following TD kernel arrive by an H2D copy on another
        // stream, followed by a stream-ordered write of `ready_value`; waiting for it here (the copy was issued
        // a step or more ahead) keeps every cross-stream wait off the compute stream.
        const volatile unsigned *flag = ready;
        const long long t0 = clock64();
        while ((int)(*flag - ready_value) < 0) {
            if (clock64() - t0 > (5ll << 31)) { atomicOr(ctrl, APEXB200_ST_PEER_TIMEOUT); break; }   // ~5 s
        }
        __threadfence();
    }
    __syncthreads();
    unsigned long long tg = lane < world ? s_totals[lane] : 0ull, incl = tg;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned long long n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    const unsigned long long T = __shfl_sync(0xffffffffu, incl, 31);
    const unsigned long long hi = __shfl_sync(0xffffffffu, incl, rank);
    const unsigned long long lo = hi - __shfl_sync(0xffffffffu, tg, rank);
    const int unit0 = blockIdx.x + wl * gridDim.x;           // half rows dealt round-robin over the CTAs
    if (T == 0ull) {
        if (lane == 0) { atomicOr(ctrl, APEXB200_ST_EMPTY_TREE); if (unit0 == 0) *sh.count_out = 0; }
        return;
    }
    const int Bg = sh.global_count;
    const int j_lo = first_draw_at_or_above(lo, sa.seed, round, sh.call, Bg, T);
    uint8_t *stage = smem + (size_t)wpc * 8 * kMaxChunks * 8 + (size_t)(wl * C + lane) * rv.F;
    unsigned long long *my_bars = bars + (wl * 8 + lane) * kMaxChunks;
    if (lane < C) tma_init_bars(my_bars, nch);
    const int wave_units = gridDim.x * wpc;
    unsigned phase = 0u;                                     // completed phases of this lane's mbarriers
    for (int unit = unit0;; unit += wave_units) {
        const int row = unit >> 1, half = unit & 1;
        const int j = j_lo + row;
        double uw = 0.0;
        const unsigned long long t = (row < sh.max_rows && j < Bg) ? draw_target(sa.seed, round, sh.call, j, Bg, T, &uw) : hi;
        if (t >= hi) {                                       // beyond this rank's run of draws (or the caller's rows)
            if (unit == 0 && lane == 0) *sh.count_out = 0;   // (unit 0 dead: the rank owns no draw)
            break;
        }
        unsigned long long leaf, total;
        const long long node = warp_descend(tv, vtop, t - lo, false, 0.0, &leaf, &total, ctrl);
        const int k = half * C + lane;
        // frame copy on the guessed slot while the frame-slot row is in flight; the row decides (see sample_gather_body)
        long long f = -1, fg = -1;
        const bool mine = node >= 0 && lane < C;
        if (mine) {
            f = rv.fidx[node * twoC + k];
            fg = guess_frame_slot(rv, node, k);
            tma_issue_frame(rv.frames + fg * rv.F, stage, my_bars, (unsigned)rv.F, nch);
        }
        if (half == 0) {
            if (lane == 0) {
                sa.idx_out[row] = (int32_t)node;
                if (sh.pos_out) sh.pos_out[row] = j;
                if (sa.u_out) sa.u_out[row] = uw;
            }
            if (node >= 0 && lane == 16) copy_meta_row(rv, ga, row, node);
        }
        if (unit == 0) {                                     // row count, computed in the shadow of the frame loads
            const int j_hi = first_draw_at_or_above(hi, sa.seed, round, sh.call, Bg, T);
            int cnt = j_hi - j_lo;
            // only a caller that passed fewer rows than the global batch can lose draws
            if (cnt > sh.max_rows) { cnt = sh.max_rows; if (lane == 0) atomicOr(ctrl, APEXB200_ST_QUOTA); }
            if (lane == 0) *sh.count_out = cnt;
        }
        if (lane < C && sa.fsel_out) sa.fsel_out[(long long)row * twoC + k] = (int32_t)f;
        bool moving = mine;
        if (mine && f != fg) {                               // wrong guess: let the stray copy land, fetch the right frame
            const bool valid = f >= 0 && f < rv.N * rv.fpr;
            if (!valid) atomicOr(ctrl, APEXB200_ST_BAD_INDEX);
            tma_wait_frame(my_bars, nch, phase & 1u);
            ++phase;                                         // the stray copy's phase is complete
            moving = valid;
            if (valid) tma_issue_frame(rv.frames + f * rv.F, stage, my_bars, (unsigned)rv.F, nch);
        }
        if (kF32)
            f32_drain_frames(moving, moving ? item_dst(rv, ga, row, k) : nullptr, stage - (size_t)lane * rv.F,
                             my_bars - lane * kMaxChunks, C, (unsigned)rv.F, nch, phase & 1u);
        else
            tma_drain_frames(moving, moving ? item_dst(rv, ga, row, k) : nullptr, stage, my_bars, (unsigned)rv.F, nch,
                             phase & 1u);
        if (moving) ++phase;                                 // the drained copy completed one more phase of this lane's mbarriers
    }
}

__global__ void __launch_bounds__(512)
sample_gather_sharded_kernel(TreeView tv, RingView rv, SampleArgs sa, GatherArgs ga, ShardArgs sh, PeerView pv,
                             int32_t *ctrl) {
    sample_gather_sharded_body<false, false>(tv, rv, sa, ga, sh, pv, ctrl, nullptr, 0u);
}

__global__ void __launch_bounds__(512)
sample_gather_sharded_f32_kernel(TreeView tv, RingView rv, SampleArgs sa, GatherArgs ga, ShardArgs sh, PeerView pv,
                                 int32_t *ctrl) {
    sample_gather_sharded_body<false, true>(tv, rv, sa, ga, sh, pv, ctrl, nullptr, 0u);
}

// Same kernel for the host-facing sharded step: also waits for the arrival word of the step's Q rows.
__global__ void __launch_bounds__(512)
sample_gather_sharded_polled_kernel(TreeView tv, RingView rv, SampleArgs sa, GatherArgs ga, ShardArgs sh, PeerView pv,
                                    int32_t *ctrl, const unsigned *ready, unsigned ready_value) {
    sample_gather_sharded_body<true, false>(tv, rv, sa, ga, sh, pv, ctrl, ready, ready_value);
}

static int ensure_smem(const void *fn, size_t smem, int device) {
    constexpr int kKernels = 8;
    static const void *known[kKernels] = {};
    static size_t configured[kKernels][64] = {};
    int which = 0;
    while (which < kKernels && known[which] != nullptr && known[which] != fn) ++which;
    APEX_REQUIRE(which < kKernels, APEXB200_ECUDA, "ensure_smem: kernel table full");
    known[which] = fn;
    if (smem > 48 * 1024 && device >= 0 && device < 64 && smem > configured[which][device]) {
        APEX_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[which][device] = smem;
    }
    return APEXB200_OK;
}

}  // namespace apexb200

using namespace apexb200;

extern "C" {

int apexb200_set_gather_variant(int32_t variant) {
    APEX_REQUIRE(variant == 0 || variant == 1 || variant == 3, APEXB200_EINVAL, "set_gather_variant: unknown variant %d", variant);
    g_gather_variant = variant;
    return APEXB200_OK;
}

int apexb200_debug_timeline(uint64_t *out_host, int32_t units) {
    APEX_REQUIRE(out_host && units > 0 && units <= 2048, APEXB200_EINVAL, "debug_timeline: bad argument");
    APEX_CUDA(cudaMemcpyFromSymbol(out_host, g_dbg_timeline, (size_t)units * kDbgWords * 8));
    return APEXB200_OK;
}

int apexb200_gather(apexb200_replay_t h, const int32_t *idx, const int32_t *fsel, int32_t count,
                    const int32_t *count_dev, uint8_t *S_t, uint8_t *S_tpn, int32_t *A, float *R, float *Gamma,
                    float *q_t, float *q_tpn, void *stream) {
    APEX_REQUIRE(h && count >= 0, APEXB200_EINVAL, "gather: bad argument");
    if (count == 0) return APEXB200_OK;
    APEX_REQUIRE(idx && S_t && S_tpn, APEXB200_EINVAL, "gather: null idx/S_t/S_tpn");
    APEX_REQUIRE((((uintptr_t)S_t | (uintptr_t)S_tpn) & 15) == 0, APEXB200_EINVAL,
                 "gather: output batches must be 16-byte aligned");
    const RingView &rv = h->ring;
    GatherArgs ga{idx, fsel, count_dev, S_t, S_tpn, A, R, Gamma, q_t, q_tpn, count, 1};
    cudaStream_t st = (cudaStream_t)stream;
    const long long items = (long long)count * 2 * rv.C;
    const long long meta_warps = (count + 31) / 32;
    if (g_gather_variant >= 1 && kBarBytes + (size_t)rv.F <= kStageBudget) {
        long long per_cta = (long long)((kStageBudget - kBarBytes) / rv.F);
        if (per_cta > 32) per_cta = 32;
        const long long spread = (items + kNumSMs - 1) / kNumSMs;      // small batches: use every SM
        if (spread < per_cta) per_cta = spread;
        const size_t smem = kBarBytes + (size_t)per_cta * rv.F;
        int rc = ensure_smem((const void *)gather_tma_kernel, smem, h->cfg.device);
        if (rc) return rc;
        const long long blocks = (items + per_cta - 1) / per_cta + meta_warps;
        return launch_step("gather_tma_kernel", gather_tma_kernel, dim3((unsigned)blocks), dim3(32), smem, st, rv, ga,
                           (int)per_cta, h->status);
    }
    const long long warps = items + meta_warps;
    return launch_step("gather_vec_kernel", gather_vec_kernel<14>, dim3((unsigned)((warps + 7) / 8)), dim3(256), 0, st,
                       rv, ga, h->status);
}

int apexb200_sample_gather(apexb200_replay_t h, const double *u, uint64_t seed, int32_t count, int32_t *idx_out,
                           int32_t *fsel_out, uint64_t *leaf_out, uint64_t *total_out, double *u_out, void *S_t,
                           void *S_tpn, int32_t *A, float *R, float *Gamma, float *q_t, float *q_tpn,
                           int32_t out_format, void *stream) {
    APEX_REQUIRE(out_format == APEXB200_FMT_U8 || out_format == APEXB200_FMT_F32, APEXB200_EINVAL,
                 "sample_gather: unknown output format %d", out_format);
    return apexb200::run_sample_gather(h, u, seed, count, idx_out, fsel_out, leaf_out, total_out, u_out, (uint8_t *)S_t,
                                       (uint8_t *)S_tpn, A, R, Gamma, q_t, q_tpn, nullptr, 0u, (cudaStream_t)stream,
                                       out_format);
}

}  // extern "C"

namespace apexb200 {
bool sample_gather_is_fused(const apexb200_replay *h) {
    const RingView &rv = h->ring;
    return rv.C <= 8 && (size_t)rv.C * rv.F + 8 * kMaxChunks * 8 <= kStageBudget;
}

// ready != null: the draws `u` land by a copy on another stream; the kernel polls *ready for ready_value.
int run_sample_gather(apexb200_replay *h, const double *u, uint64_t seed, int32_t count, int32_t *idx_out,
                      int32_t *fsel_out, uint64_t *leaf_out, uint64_t *total_out, double *u_out, uint8_t *S_t,
                      uint8_t *S_tpn, int32_t *A, float *R, float *Gamma, float *q_t, float *q_tpn,
                      const unsigned *ready, unsigned ready_value, cudaStream_t stream, int out_format) {
    APEX_REQUIRE(h && count >= 0, APEXB200_EINVAL, "sample_gather: bad argument");
    if (count == 0) return APEXB200_OK;
    APEX_REQUIRE(idx_out && S_t && S_tpn, APEXB200_EINVAL, "sample_gather: null idx_out/S_t/S_tpn");
    APEX_REQUIRE((((uintptr_t)S_t | (uintptr_t)S_tpn) & 15) == 0, APEXB200_EINVAL,
                 "sample_gather: output batches must be 16-byte aligned");
    const RingView &rv = h->ring;
    const size_t per_unit = (size_t)rv.C * rv.F + 8 * kMaxChunks * 8;       // one half draw: C frames + their mbarriers
    cudaStream_t st = stream;
    const bool f32 = out_format == APEXB200_FMT_F32;
    if (rv.C > 8 || per_unit > kStageBudget) {
        APEX_REQUIRE(!ready, APEXB200_ERANGE, "sample_gather: the polled variant needs the fused kernel");
        APEX_REQUIRE(!f32, APEXB200_ERANGE, "sample_gather: float32 output needs the fused kernel (half a record of "
                     "%d frames x %d bytes does not fit one CTA's staging)", rv.C, rv.F);
        // staging for half a record does not fit one CTA: run the two kernels back to back
        int rc = u ? apexb200_sample(h, u, count, idx_out, leaf_out, total_out, fsel_out, stream)
                   : apexb200_sample_stratified(h, seed, count, idx_out, leaf_out, total_out, fsel_out, u_out, stream);
        if (rc) return rc;
        return apexb200_gather(h, idx_out, fsel_out, count, nullptr, S_t, S_tpn, A, R, Gamma, q_t, q_tpn, stream);
    }
    const int units = 2 * count;
    int wpc = (int)(kStageBudget / per_unit);                              // units (warps) per CTA that fit
    if (wpc > 16) wpc = 16;
    // one wave: spread evenly over the SMs, but leave one SM free — the next kernel of the step
    // (single-CTA K5+K2, launched early by PDL) then has somewhere to wait without queueing behind
    // a 200 KB CTA (measured: 148 resident CTAs cost the step ~1.3 us)
    const int spread = (units + kNumSMs - 2) / (kNumSMs - 1);
    if (spread < wpc) wpc = spread;
    // A batch that needs several waves runs faster with thin CTAs (several CTAs per SM: one CTA's
    // tree descent overlaps another's frame movement; 78 vs 88 us at B=4096), a single-wave batch
    // with fat CTAs (profiles/r01_kernel_experiments.md).
    if (units > wpc * kNumSMs) wpc = 2;
    const size_t smem = (size_t)wpc * per_unit;
    int rc = ensure_smem((const void *)sample_gather_kernel, smem, h->cfg.device);
    if (rc) return rc;
    SampleArgs sa{u, (unsigned long long)seed, u ? 0ull : h->draw_calls++, idx_out, fsel_out, (unsigned long long *)leaf_out,
                  (unsigned long long *)total_out, u_out};
    GatherArgs ga{idx_out, fsel_out, nullptr, S_t, S_tpn, A, R, Gamma, q_t, q_tpn, count, f32 ? 4 : 1};
    if (f32) {
        APEX_REQUIRE(!ready, APEXB200_ERANGE, "sample_gather: the polled variant writes uint8 batches");
        rc = ensure_smem((const void *)sample_gather_f32_kernel, smem, h->cfg.device);
        if (rc) return rc;
        return launch_step("sample_gather_f32_kernel", sample_gather_f32_kernel, dim3((unsigned)((units + wpc - 1) / wpc)),
                           dim3(wpc * 32), smem, st, h->tree, rv, sa, ga, h->status);
    }
    if (ready) {
        APEX_REQUIRE(u, APEXB200_EINVAL, "sample_gather: the polled variant takes caller draws");
        rc = ensure_smem((const void *)sample_gather_polled_kernel, smem, h->cfg.device);
        if (rc) return rc;
        return launch_step("sample_gather_polled_kernel", sample_gather_polled_kernel,
                           dim3((unsigned)((units + wpc - 1) / wpc)), dim3(wpc * 32), smem, st, h->tree, rv, sa, ga,
                           h->status, ready, ready_value);
    }
    if (g_gather_variant == 3) {
        rc = ensure_smem((const void *)sample_gather_timed_kernel, smem, h->cfg.device);
        if (rc) return rc;
        return launch_step("sample_gather_timed_kernel", sample_gather_timed_kernel, dim3((unsigned)((units + wpc - 1) / wpc)),
                           dim3(wpc * 32), smem, st, h->tree, rv, sa, ga, h->status);
    }
    return launch_step("sample_gather_kernel", sample_gather_kernel, dim3((unsigned)((units + wpc - 1) / wpc)),
                       dim3(wpc * 32), smem, st, h->tree, rv, sa, ga, h->status);
}

// Rows one wave of the sharded kernel covers: the rank's fair share of the global batch plus ~15 % head-room,
// as whole rounds of the (kNumSMs - 1)-CTA grid.  A rank that owns more draws runs further waves in the
// same launch; max_rows only bounds what the caller's output arrays can hold.
int sharded_wave_rows(const apexb200_replay *h, int global_count, int max_rows) {
    const int world = h->peers.world > 0 ? h->peers.world : 1;
    long long share = ((long long)global_count + world - 1) / world;
    share += (share + 7) / 8 + 2;
    if (share > max_rows) share = max_rows;
    return (int)share;
}

int run_sample_gather_sharded(apexb200_replay *h, uint64_t seed, int32_t global_count, int32_t max_rows,
                              int32_t *idx_out, int32_t *fsel_out, int32_t *pos_out, double *u_out,
                              int32_t *count_out, uint8_t *S_t, uint8_t *S_tpn, int32_t *A, float *R, float *Gamma,
                              float *q_t, float *q_tpn, const unsigned *ready, unsigned ready_value,
                              cudaStream_t stream, int out_format) {
    APEX_REQUIRE(h && global_count > 0 && max_rows > 0, APEXB200_EINVAL, "sample_gather_sharded: bad argument");
    const bool f32 = out_format == APEXB200_FMT_F32;
    APEX_REQUIRE(idx_out && count_out && S_t && S_tpn, APEXB200_EINVAL, "sample_gather_sharded: null output");
    APEX_REQUIRE((((uintptr_t)S_t | (uintptr_t)S_tpn) & 15) == 0, APEXB200_EINVAL,
                 "sample_gather_sharded: output batches must be 16-byte aligned");
    const RingView &rv = h->ring;
    const size_t per_unit = (size_t)rv.C * rv.F + 8 * kMaxChunks * 8;       // one half row: C frames + their mbarriers
    APEX_REQUIRE(rv.C <= 8 && per_unit <= kStageBudget, APEXB200_ERANGE,
                 "sample_gather_sharded: half a record (%d frames of %d bytes) does not fit one CTA's staging", rv.C, rv.F);
    const int units = 2 * sharded_wave_rows(h, global_count, max_rows);
    // one resident wave over kNumSMs - 1 CTAs (one SM stays free, as in sample_gather); half rows are dealt
    // round-robin, so the live ones spread evenly whatever the wave size; further rows = further waves
    const int ctas = units < kNumSMs - 1 ? units : kNumSMs - 1;
    int wpc = (units + ctas - 1) / ctas;
    const int fit = (int)(kStageBudget / per_unit);
    if (wpc > fit) wpc = fit;
    if (wpc > 16) wpc = 16;
    const size_t smem = (size_t)wpc * per_unit;
    SampleArgs sa{nullptr, (unsigned long long)seed, 0ull, idx_out, fsel_out, nullptr, nullptr, u_out};
    GatherArgs ga{idx_out, fsel_out, nullptr, S_t, S_tpn, A, R, Gamma, q_t, q_tpn, max_rows, f32 ? 4 : 1};
    ShardArgs sh{(unsigned long long)seed, h->draw_calls++, global_count, max_rows, pos_out, count_out};
    if (f32) {
        APEX_REQUIRE(!ready, APEXB200_ERANGE, "sample_gather_sharded: the polled variant writes uint8 batches");
        int rc = ensure_smem((const void *)sample_gather_sharded_f32_kernel, smem, h->cfg.device);
        if (rc) return rc;
        return launch_step("sample_gather_sharded_f32_kernel", sample_gather_sharded_f32_kernel, dim3((unsigned)ctas),
                           dim3(wpc * 32), smem, stream, h->tree, rv, sa, ga, sh, h->peers, h->status);
    }
    if (ready) {
        int rc = ensure_smem((const void *)sample_gather_sharded_polled_kernel, smem, h->cfg.device);
        if (rc) return rc;
        return launch_step("sample_gather_sharded_polled_kernel", sample_gather_sharded_polled_kernel, dim3((unsigned)ctas),
                           dim3(wpc * 32), smem, stream, h->tree, rv, sa, ga, sh, h->peers, h->status, ready, ready_value);
    }
    int rc = ensure_smem((const void *)sample_gather_sharded_kernel, smem, h->cfg.device);
    if (rc) return rc;
    return launch_step("sample_gather_sharded_kernel", sample_gather_sharded_kernel, dim3((unsigned)ctas), dim3(wpc * 32),
                       smem, stream, h->tree, rv, sa, ga, sh, h->peers, h->status);
}
}  // namespace apexb200

extern "C" {

int apexb200_sample_gather_sharded(apexb200_replay_t h, uint64_t seed, int32_t global_count, int32_t max_rows,
                                   int32_t *idx_out, int32_t *fsel_out, int32_t *pos_out, double *u_out,
                                   int32_t *count_out, void *S_t, void *S_tpn, int32_t *A, float *R,
                                   float *Gamma, float *q_t, float *q_tpn, int32_t out_format, void *stream) {
    APEX_REQUIRE(out_format == APEXB200_FMT_U8 || out_format == APEXB200_FMT_F32, APEXB200_EINVAL,
                 "sample_gather_sharded: unknown output format %d", out_format);
    return apexb200::run_sample_gather_sharded(h, seed, global_count, max_rows, idx_out, fsel_out, pos_out, u_out,
                                               count_out, (uint8_t *)S_t, (uint8_t *)S_tpn, A, R, Gamma, q_t, q_tpn,
                                               nullptr, 0u, (cudaStream_t)stream, out_format);
}

}  // extern "C"
