// Replay shard lifetime, ring insert (K6) and raw buffer access.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>

#include "common.cuh"

namespace apexb200 {

static thread_local char g_err[512] = "";
int g_pdl = 1;
int g_td_cluster = 0;

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what) {
    set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return e == cudaErrorMemoryAllocation ? APEXB200_ENOMEM : APEXB200_ECUDA;
}

// ---------------------------------------------------------------------------
// K6a: stream the new frames into their ring slots (16-byte vectors, coalesced; source is
// contiguous, destination is contiguous up to the ring wrap).  One CTA row (blockIdx.y, strided)
// per record, two vectors per thread in flight: no index division, ~2 loads outstanding per thread
// and tens of thousands of CTAs keep HBM busy in both directions.
// Up to kInsertLanes lanes per launch: lane ids and ring heads travel as a kernel parameter (blockIdx.z / .y
// selects the entry), so E lock-stepped actors insert with one launch pair instead of E.
constexpr int kInsertLanes = 64;
struct LaneTable {
    long long head[kInsertLanes];
    int lane[kInsertLanes];
};

__global__ void __launch_bounds__(256)
insert_frames_kernel(uint4 *__restrict__ ring_frames, const uint4 *__restrict__ src, int count,
                     int vecs_per_record, LaneTable lt, long long lane_cap) {
    const int v0 = blockIdx.x * 512 + threadIdx.x;
    const int v1 = v0 + 256;
    const int e = blockIdx.z;
    const long long lane_base = (long long)lt.lane[e] * lane_cap, head = lt.head[e];
    src += (long long)e * count * vecs_per_record;
    for (int j = blockIdx.y; j < count; j += gridDim.y) {
        const long long slot = lane_base + (head + j) % lane_cap;
        const uint4 *s = src + (long long)j * vecs_per_record;
        uint4 *d = ring_frames + slot * vecs_per_record;
        uint4 a, b;
        if (v0 < vecs_per_record) a = __ldcs(s + v0);                // staging data is read once
        if (v1 < vecs_per_record) b = __ldcs(s + v1);
        if (v0 < vecs_per_record) d[v0] = a;
        if (v1 < vecs_per_record) d[v1] = b;
    }
}

// K6b: metadata, frame-index rows, initial priority (actor.py:138-140), leaf write
// and ancestor update for the `count` new records, plus invalidation of the
// `lookback` records behind the new head.
__global__ void __launch_bounds__(128)
insert_meta_kernel(RingView rv, TreeView tv, LaneTable lt, int count,
                   const int32_t *__restrict__ rel, const int32_t *__restrict__ A,
                   const float *__restrict__ R, const float *__restrict__ G,
                   const float *__restrict__ prio, const float *__restrict__ q_t,
                   const float *__restrict__ q_tpn, int32_t *status) {
    bump_epoch(status);
    const int e = blockIdx.y;                                   // inputs of entry e start at record e * count
    const int lane = lt.lane[e];
    const long long head = lt.head[e];
    {
        const long long o = (long long)e * count;
        rel += o * 2 * rv.C; A += o; R += o; G += o;
        if (prio) prio += o;
        if (q_t) { q_t += o * rv.nA; q_tpn += o * rv.nA; }
    }
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = j < count + rv.lookback;
    long long p = (head + (active ? j : 0)) % rv.lane_cap;
    long long s = (long long)lane * rv.lane_cap + p;
    unsigned long long new_leaf = 0ull;
    if (active && j < count) {
        const int twoC = 2 * rv.C;
        const long long lane_frames = rv.lane_cap * rv.fpr;
        const long long lane_frame_base = (long long)lane * lane_frames;
        const long long last = p * rv.fpr + rv.fpr - 1;
        const long long reach = (long long)rv.lookback * rv.fpr + rv.fpr - 1;
        for (int k = 0; k < twoC; ++k) {
            long long r = rel[(long long)j * twoC + k];
            if (r > 0 || -r > reach) { atomicOr(status, APEXB200_ST_BAD_INDEX); r = 0; }
            long long f = (last + r) % lane_frames;
            if (f < 0) f += lane_frames;
            rv.fidx[s * twoC + k] = (int32_t)(lane_frame_base + f);
        }
        int a = A[j];
        float r = R[j], g = G[j];
        rv.A[s] = a; rv.R[s] = r; rv.G[s] = g;
        float p0;
        if (prio != nullptr) {
            p0 = fabsf(prio[j]);
        } else {
            const float *qa = q_t + (long long)j * rv.nA, *qb = q_tpn + (long long)j * rv.nA;
            float mx = qb[0];
            for (int k = 1; k < rv.nA; ++k) mx = fmaxf(mx, qb[k]);
            if (a < 0 || a >= rv.nA) { atomicOr(status, APEXB200_ST_BAD_ACTION); a = 0; }
            p0 = fabsf(__fadd_rn(__fadd_rn(r, __fmul_rn(g, mx)), -qa[a]));
        }
        if (rv.q_t != nullptr && q_t != nullptr) {
            for (int k = 0; k < rv.nA; ++k) {
                rv.q_t[s * rv.nA + k] = q_t[(long long)j * rv.nA + k];
                rv.q_tpn[s * rv.nA + k] = q_tpn[(long long)j * rv.nA + k];
            }
        }
        new_leaf = leaf_from_prio(prio_pow(p0, rv.alpha, 0.0f));
    }
    unsigned long long d = 0ull;
    if (active) {
        unsigned long long old = tv.lvl[0][s];
        tv.lvl[0][s] = new_leaf;
        d = new_leaf - old;
    }
    // consecutive slots: neighbours share ancestors -> one atomic per distinct ancestor per warp
    warp_propagate_adjacent(tv, active ? s : -1, d, 1);
}

}  // namespace apexb200

using namespace apexb200;

extern "C" {

int apexb200_version(void) { return APEXB200_VERSION; }

const char *apexb200_strerror(int code) {
    switch (code) {
        case APEXB200_OK: return "ok";
        case APEXB200_EINVAL: return "invalid argument";
        case APEXB200_ECUDA: return "CUDA runtime error";
        case APEXB200_ENOMEM: return "out of device memory";
        case APEXB200_ERANGE: return "size beyond supported limit";
        default: return "unknown error";
    }
}

const char *apexb200_last_error(void) { return g_err; }

int apexb200_set_td_cluster(int32_t enabled) {
    g_td_cluster = enabled ? 1 : 0;
    return APEXB200_OK;
}

int apexb200_set_pdl(int32_t enabled) {
    g_pdl = enabled ? 1 : 0;
    return APEXB200_OK;
}

int apexb200_replay_create(apexb200_replay_t *out, const apexb200_replay_config *cfg) {
    APEX_REQUIRE(out && cfg, APEXB200_EINVAL, "replay_create: null argument");
    *out = nullptr;
    APEX_REQUIRE(cfg->capacity > 0 && cfg->lanes > 0 && cfg->capacity % cfg->lanes == 0, APEXB200_EINVAL,
                 "replay_create: capacity %lld must be a positive multiple of lanes %d",
                 (long long)cfg->capacity, cfg->lanes);
    APEX_REQUIRE(cfg->capacity <= APEXB200_MAX_CAPACITY, APEXB200_ERANGE,
                 "replay_create: capacity %lld > %lld records per shard", (long long)cfg->capacity,
                 (long long)APEXB200_MAX_CAPACITY);
    APEX_REQUIRE(cfg->frames_per_record >= 1 && cfg->frames_per_state >= 1 && cfg->lookback >= 0,
                 APEXB200_EINVAL, "replay_create: bad frames_per_record/frames_per_state/lookback");
    APEX_REQUIRE(cfg->frame_bytes > 0 && cfg->frame_bytes % 16 == 0, APEXB200_EINVAL,
                 "replay_create: frame_bytes %d must be a positive multiple of 16", cfg->frame_bytes);
    APEX_REQUIRE(cfg->num_actions >= 0 && cfg->num_actions <= APEXB200_MAX_ACTIONS, APEXB200_ERANGE,
                 "replay_create: num_actions %d outside [0, %d]", cfg->num_actions, APEXB200_MAX_ACTIONS);
    APEX_REQUIRE(cfg->alpha >= 0.0f, APEXB200_EINVAL, "replay_create: alpha must be >= 0");
    const long long lane_cap = cfg->capacity / cfg->lanes;
    APEX_REQUIRE(cfg->lookback < lane_cap, APEXB200_EINVAL, "replay_create: lookback %d >= lane capacity %lld",
                 cfg->lookback, lane_cap);
    APEX_REQUIRE((long long)cfg->capacity * cfg->frames_per_record < (1ll << 31), APEXB200_ERANGE,
                 "replay_create: frame slot count overflows int32");
    APEX_CUDA(cudaSetDevice(cfg->device));

    apexb200_replay *h = new (std::nothrow) apexb200_replay();
    APEX_REQUIRE(h, APEXB200_ENOMEM, "replay_create: host allocation failed");
    h->cfg = *cfg;
    h->lane_count.assign(cfg->lanes, 0);
    const long long N = cfg->capacity;
    const int C = cfg->frames_per_state, nA = cfg->num_actions;

    // tree geometry
    TreeView &tv = h->tree;
    memset(&tv, 0, sizeof(tv));
    long long sz = N, padded_total = 0;
    int nl = 0;
    long long offs[kMaxLevels];
    for (;;) {
        tv.size[nl] = sz;
        offs[nl] = padded_total;
        padded_total += (sz + kFanout - 1) / kFanout * kFanout;
        ++nl;
        if (sz <= kFanout) break;
        sz = (sz + kFanout - 1) / kFanout;
    }
    tv.n_levels = nl;

    int rc = APEXB200_OK;
    auto alloc = [&](void **p, size_t bytes, bool zero) -> bool {
        cudaError_t e = cudaMalloc(p, bytes);
        if (e != cudaSuccess) { rc = cuda_fail(e, "cudaMalloc"); return false; }
        h->device_bytes += (int64_t)bytes;
        if (zero) {
            e = cudaMemset(*p, 0, bytes);
            if (e != cudaSuccess) { rc = cuda_fail(e, "cudaMemset"); return false; }
        }
        return true;
    };
    RingView &rv = h->ring;
    memset(&rv, 0, sizeof(rv));
    rv.N = N; rv.lane_cap = lane_cap; rv.lanes = cfg->lanes; rv.fpr = cfg->frames_per_record;
    rv.lookback = cfg->lookback; rv.C = C; rv.F = cfg->frame_bytes; rv.nA = nA; rv.alpha = cfg->alpha;
    bool ok = alloc(&h->tree_block, (size_t)padded_total * 8, true) &&
              alloc((void **)&h->stamp, (size_t)N * 4, true) &&
              alloc((void **)&h->status, 256, true) &&
              alloc((void **)&h->xbuf, 2 * kMaxShards * 2 * 8, true) &&
              alloc((void **)&rv.frames, (size_t)N * rv.fpr * rv.F, false) &&
              alloc((void **)&rv.fidx, (size_t)N * 2 * C * 4, true) &&
              alloc((void **)&rv.A, (size_t)N * 4, true) &&
              alloc((void **)&rv.R, (size_t)N * 4, true) &&
              alloc((void **)&rv.G, (size_t)N * 4, true);
    if (ok && nA > 0)
        ok = alloc((void **)&rv.q_t, (size_t)N * nA * 4, true) && alloc((void **)&rv.q_tpn, (size_t)N * nA * 4, true);
    if (!ok) { apexb200_replay_destroy(h); return rc; }
    for (int l = 0; l < nl; ++l) tv.lvl[l] = (unsigned long long *)h->tree_block + offs[l];
    memset(&h->peers, 0, sizeof(h->peers));
    h->peers.world = 1;
    h->peers.rank = 0;
    h->peers.xbuf[0] = h->xbuf;
    *out = h;
    return APEXB200_OK;
}

int apexb200_replay_destroy(apexb200_replay_t h) {
    if (!h) return APEXB200_OK;
    cudaSetDevice(h->cfg.device);
    for (void *m : h->opened) if (m) cudaIpcCloseMemHandle(m);
    if (h->hs_copy_in) cudaStreamDestroy(h->hs_copy_in);
    if (h->hs_seq_host) cudaFreeHost(h->hs_seq_host);
    if (h->hs_copy_out) cudaStreamDestroy(h->hs_copy_out);
    for (int k = 0; k < APEXB200_HOST_STEP_SLOTS; ++k) {
        if (h->hs_ev_in[k]) cudaEventDestroy(h->hs_ev_in[k]);
        if (h->hs_ev_k[k]) cudaEventDestroy(h->hs_ev_k[k]);
        if (h->hs_ev_out[k]) cudaEventDestroy(h->hs_ev_out[k]);
        if (h->hs_in[k]) cudaFree(h->hs_in[k]);
        if (h->hs_out[k]) cudaFree(h->hs_out[k]);
    }
    void *ptrs[] = {h->xbuf, h->tree_block, h->stamp, h->status, h->ring.frames, h->ring.fidx, h->ring.A,
                    h->ring.R, h->ring.G, h->ring.q_t, h->ring.q_tpn};
    for (void *p : ptrs) if (p) cudaFree(p);
    delete h;
    return APEXB200_OK;
}

int apexb200_replay_size(apexb200_replay_t h, int64_t *size_host) {
    APEX_REQUIRE(h && size_host, APEXB200_EINVAL, "replay_size: null argument");
    long long s = 0;
    for (long long c : h->lane_count) s += c < h->ring.lane_cap ? c : h->ring.lane_cap;
    *size_host = s;
    return APEXB200_OK;
}

int apexb200_replay_device_bytes(apexb200_replay_t h, int64_t *bytes_host) {
    APEX_REQUIRE(h && bytes_host, APEXB200_EINVAL, "replay_device_bytes: null argument");
    *bytes_host = h->device_bytes;
    return APEXB200_OK;
}

int apexb200_replay_status(apexb200_replay_t h, int32_t *status_host, void *stream) {
    APEX_REQUIRE(h && status_host, APEXB200_EINVAL, "replay_status: null argument");
    cudaStream_t st = (cudaStream_t)stream;
    APEX_CUDA(cudaMemcpyAsync(status_host, h->status, 4, cudaMemcpyDeviceToHost, st));
    APEX_CUDA(cudaMemsetAsync(h->status, 0, 4, st));
    APEX_CUDA(cudaStreamSynchronize(st));
    return APEXB200_OK;
}

}  // extern "C"

namespace apexb200 {
// `n_lanes` lanes, `count` records each; inputs are [n_lanes, count, ...] contiguous.
static int insert_lanes(apexb200_replay *h, const int32_t *lanes, int n_lanes, int32_t count, const uint8_t *frames,
                        const int32_t *rel, const int32_t *A, const float *R, const float *Gamma, const float *prio,
                        const float *q_t, const float *q_tpn, cudaStream_t st) {
    const RingView &rv = h->ring;
    APEX_REQUIRE((long long)count + rv.lookback <= rv.lane_cap, APEXB200_ERANGE,
                 "replay_insert: count %d + lookback %d exceeds lane capacity %lld", count, rv.lookback, rv.lane_cap);
    APEX_REQUIRE(frames && rel && A && R && Gamma, APEXB200_EINVAL, "replay_insert: null input");
    APEX_REQUIRE(prio || (q_t && q_tpn && rv.nA > 0), APEXB200_EINVAL,
                 "replay_insert: need prio, or q_t/q_tpn on a handle created with num_actions > 0");
    APEX_REQUIRE(((uintptr_t)frames & 15) == 0, APEXB200_EINVAL, "replay_insert: frames must be 16-byte aligned");
    const int vpr = rv.fpr * rv.F / 16;
    const int twoC = 2 * rv.C;
    for (int first = 0; first < n_lanes; first += kInsertLanes) {
        const int m = n_lanes - first < kInsertLanes ? n_lanes - first : kInsertLanes;
        LaneTable lt;
        for (int e = 0; e < m; ++e) {
            lt.lane[e] = lanes[first + e];
            lt.head[e] = h->lane_count[lanes[first + e]] % rv.lane_cap;
        }
        const long long o = (long long)first * count;                // records before this group of lanes
        int max_rows = kNumSMs * 8 / m;                              // ~8 resident CTAs per SM over all lanes
        if (max_rows < 1) max_rows = 1;
        const dim3 grid((unsigned)((vpr + 511) / 512), (unsigned)(count < max_rows ? count : max_rows), (unsigned)m);
        insert_frames_kernel<<<grid, 256, 0, st>>>((uint4 *)rv.frames, (const uint4 *)frames + o * vpr, count, vpr, lt,
                                                   rv.lane_cap);
        int rc = launch_check("insert_frames_kernel");
        if (rc) return rc;
        const int nmeta = count + rv.lookback;
        insert_meta_kernel<<<dim3((unsigned)((nmeta + 127) / 128), (unsigned)m), 128, 0, st>>>(
            rv, h->tree, lt, count, rel + o * twoC, A + o, R + o, Gamma + o, prio ? prio + o : nullptr,
            q_t ? q_t + o * rv.nA : nullptr, q_tpn ? q_tpn + o * rv.nA : nullptr, h->status);
        rc = launch_check("insert_meta_kernel");
        if (rc) return rc;
        for (int e = 0; e < m; ++e) h->lane_count[lanes[first + e]] += count;
    }
    return APEXB200_OK;
}
}  // namespace apexb200

extern "C" {

int apexb200_replay_insert(apexb200_replay_t h, int32_t lane, int32_t count, const uint8_t *frames,
                           const int32_t *rel, const int32_t *A, const float *R, const float *Gamma,
                           const float *prio, const float *q_t, const float *q_tpn,
                           int64_t *first_slot_host, void *stream) {
    APEX_REQUIRE(h, APEXB200_EINVAL, "replay_insert: null handle");
    const RingView &rv = h->ring;
    APEX_REQUIRE(lane >= 0 && lane < rv.lanes, APEXB200_EINVAL, "replay_insert: lane %d outside [0,%d)", lane, rv.lanes);
    APEX_REQUIRE(count >= 0, APEXB200_EINVAL, "replay_insert: negative count");
    const long long head = h->lane_count[lane] % rv.lane_cap;
    if (first_slot_host) *first_slot_host = (long long)lane * rv.lane_cap + head;
    if (count == 0) return APEXB200_OK;
    return insert_lanes(h, &lane, 1, count, frames, rel, A, R, Gamma, prio, q_t, q_tpn, (cudaStream_t)stream);
}

int apexb200_replay_insert_lanes(apexb200_replay_t h, const int32_t *lanes_host, int32_t n_lanes, int32_t count,
                                 const uint8_t *frames, const int32_t *rel, const int32_t *A, const float *R,
                                 const float *Gamma, const float *prio, const float *q_t, const float *q_tpn,
                                 void *stream) {
    APEX_REQUIRE(h && lanes_host, APEXB200_EINVAL, "replay_insert_lanes: null argument");
    const RingView &rv = h->ring;
    APEX_REQUIRE(n_lanes >= 0 && n_lanes <= rv.lanes && count >= 0, APEXB200_EINVAL,
                 "replay_insert_lanes: n_lanes %d / count %d out of range", n_lanes, count);
    std::vector<char> seen((size_t)rv.lanes, 0);
    for (int e = 0; e < n_lanes; ++e) {
        const int l = lanes_host[e];
        APEX_REQUIRE(l >= 0 && l < rv.lanes && !seen[l], APEXB200_EINVAL,
                     "replay_insert_lanes: lane %d outside [0,%d) or listed twice", l, rv.lanes);
        seen[l] = 1;
    }
    if (n_lanes == 0 || count == 0) return APEXB200_OK;
    return insert_lanes(h, lanes_host, n_lanes, count, frames, rel, A, R, Gamma, prio, q_t, q_tpn, (cudaStream_t)stream);
}

int apexb200_peer_export(apexb200_replay_t h, void *ipc_handle_host) {
    APEX_REQUIRE(h && ipc_handle_host, APEXB200_EINVAL, "peer_export: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == APEXB200_IPC_HANDLE_BYTES, "IPC handle size");
    APEX_CUDA(cudaSetDevice(h->cfg.device));
    cudaIpcMemHandle_t hd;
    APEX_CUDA(cudaIpcGetMemHandle(&hd, h->xbuf));
    memcpy(ipc_handle_host, &hd, sizeof(hd));
    return APEXB200_OK;
}

int apexb200_peer_connect(apexb200_replay_t h, int32_t rank, int32_t world, const void *ipc_handles_host) {
    APEX_REQUIRE(h && ipc_handles_host, APEXB200_EINVAL, "peer_connect: null argument");
    APEX_REQUIRE(world >= 1 && world <= kMaxShards && rank >= 0 && rank < world, APEXB200_ERANGE,
                 "peer_connect: rank %d / world %d outside [1, %d]", rank, world, kMaxShards);
    APEX_CUDA(cudaSetDevice(h->cfg.device));
    for (void *&m : h->opened) if (m) { cudaIpcCloseMemHandle(m); m = nullptr; }
    PeerView pv;
    memset(&pv, 0, sizeof(pv));
    pv.world = world;
    pv.rank = rank;
    for (int g = 0; g < world; ++g) {
        if (g == rank) { pv.xbuf[g] = h->xbuf; continue; }
        cudaIpcMemHandle_t hd;
        memcpy(&hd, (const char *)ipc_handles_host + (size_t)g * sizeof(hd), sizeof(hd));
        void *p = nullptr;
        APEX_CUDA(cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess));
        h->opened[g] = p;
        pv.xbuf[g] = (unsigned long long *)p;
    }
    h->peers = pv;
    return APEXB200_OK;
}

int apexb200_peer_connect_local(apexb200_replay_t h, int32_t rank, int32_t world, const apexb200_replay_t *shards_host) {
    APEX_REQUIRE(h && shards_host, APEXB200_EINVAL, "peer_connect_local: null argument");
    APEX_REQUIRE(world >= 1 && world <= kMaxShards && rank >= 0 && rank < world, APEXB200_ERANGE,
                 "peer_connect_local: rank %d / world %d outside [1, %d]", rank, world, kMaxShards);
    APEX_REQUIRE(shards_host[rank] == h, APEXB200_EINVAL, "peer_connect_local: shards[rank] must be the handle itself");
    APEX_CUDA(cudaSetDevice(h->cfg.device));
    for (void *&m : h->opened) if (m) { cudaIpcCloseMemHandle(m); m = nullptr; }
    PeerView pv;
    memset(&pv, 0, sizeof(pv));
    pv.world = world;
    pv.rank = rank;
    for (int g = 0; g < world; ++g) {
        const apexb200_replay *peer = shards_host[g];
        APEX_REQUIRE(peer, APEXB200_EINVAL, "peer_connect_local: null shard %d", g);
        if (peer->cfg.device != h->cfg.device) {             // another GPU of this process: direct peer access
            int can = 0;
            APEX_CUDA(cudaDeviceCanAccessPeer(&can, h->cfg.device, peer->cfg.device));
            APEX_REQUIRE(can, APEXB200_EINVAL, "peer_connect_local: device %d cannot access device %d", h->cfg.device,
                         peer->cfg.device);
            cudaError_t e = cudaDeviceEnablePeerAccess(peer->cfg.device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceEnablePeerAccess");
        }
        pv.xbuf[g] = peer->xbuf;
    }
    h->peers = pv;
    return APEXB200_OK;
}

}  // extern "C"

namespace apexb200 {
// Copy streams, events and device staging of the host-facing step (first call / larger batch).
static int host_step_prepare(apexb200_replay *h, size_t in_bytes, size_t out_bytes) {
    if (!h->hs_copy_in) {
        APEX_CUDA(cudaSetDevice(h->cfg.device));
        APEX_CUDA(cudaStreamCreateWithFlags(&h->hs_copy_in, cudaStreamNonBlocking));
        APEX_CUDA(cudaStreamCreateWithFlags(&h->hs_copy_out, cudaStreamNonBlocking));
        for (int k = 0; k < APEXB200_HOST_STEP_SLOTS; ++k) {
            APEX_CUDA(cudaEventCreateWithFlags(&h->hs_ev_in[k], cudaEventDisableTiming));
            APEX_CUDA(cudaEventCreateWithFlags(&h->hs_ev_k[k], cudaEventDisableTiming));
            APEX_CUDA(cudaEventCreateWithFlags(&h->hs_ev_out[k], cudaEventDisableTiming));
        }
        APEX_CUDA(cudaHostAlloc((void **)&h->hs_seq_host, APEXB200_HOST_STEP_SLOTS * sizeof(unsigned), cudaHostAllocDefault));
        for (int k = 0; k < APEXB200_HOST_STEP_SLOTS; ++k) h->hs_seq_host[k] = 0;
    }
    if (in_bytes > h->hs_in_cap || out_bytes > h->hs_out_cap) {
        APEX_CUDA(cudaDeviceSynchronize());
        for (int k = 0; k < APEXB200_HOST_STEP_SLOTS; ++k) {
            if (h->hs_in[k]) cudaFree(h->hs_in[k]);
            if (h->hs_out[k]) cudaFree(h->hs_out[k]);
            h->hs_in[k] = nullptr; h->hs_out[k] = nullptr;
        }
        h->hs_in_cap = h->hs_out_cap = 0;
        for (int k = 0; k < APEXB200_HOST_STEP_SLOTS; ++k) {
            APEX_CUDA(cudaMalloc((void **)&h->hs_in[k], in_bytes));
            APEX_CUDA(cudaMalloc((void **)&h->hs_out[k], out_bytes));
        }
        h->hs_in_cap = in_bytes;
        h->hs_out_cap = out_bytes;
    }
    return APEXB200_OK;
}
}  // namespace apexb200

extern "C" {

int apexb200_host_step(apexb200_replay_t h, int32_t slot, const double *u_host, const float *q_host, int32_t B,
                       int32_t num_actions, float eps, int32_t loss_kind, float kappa, uint8_t *S_t, uint8_t *S_tpn,
                       int32_t *A, float *R, float *Gamma, float *result_host, void *done_event, void *stream) {
    APEX_REQUIRE(h && u_host && q_host && result_host && S_t && S_tpn && A && R && Gamma, APEXB200_EINVAL,
                 "host_step: null argument");
    APEX_REQUIRE(slot >= 0 && slot < APEXB200_HOST_STEP_SLOTS, APEXB200_EINVAL, "host_step: slot must be in [0, %d)",
                 APEXB200_HOST_STEP_SLOTS);
    APEX_REQUIRE(B > 0 && B <= 1024 && num_actions >= 1 && num_actions <= APEXB200_MAX_ACTIONS, APEXB200_ERANGE,
                 "host_step: B=%d (max 1024) / num_actions=%d out of range", B, num_actions);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t in_bytes = (size_t)B * 8 + (size_t)3 * B * num_actions * 4;
    const size_t out_words = (size_t)4 * B + 4;
    int prc = host_step_prepare(h, in_bytes, out_words * 4);
    if (prc) return prc;
    unsigned char *d_in = h->hs_in[slot];
    float *d_res = h->hs_out[slot];
    // copy-in stream: the staging of this slot is free once the kernels of its previous use are done
    APEX_CUDA(cudaStreamWaitEvent(h->hs_copy_in, h->hs_ev_k[slot], 0));
    double *d_u = (double *)d_in;
    float *d_q = (float *)(d_in + (size_t)B * 8);
    if ((const void *)q_host == (const void *)(u_host + B)) {    // adjacent in the caller's pinned block: one copy
        APEX_CUDA(cudaMemcpyAsync(d_in, u_host, in_bytes, cudaMemcpyHostToDevice, h->hs_copy_in));
    } else {
        APEX_CUDA(cudaMemcpyAsync(d_u, u_host, (size_t)B * 8, cudaMemcpyHostToDevice, h->hs_copy_in));
        APEX_CUDA(cudaMemcpyAsync(d_q, q_host, (size_t)3 * B * num_actions * 4, cudaMemcpyHostToDevice, h->hs_copy_in));
    }
    // ... followed by the word the sampling kernel polls: no cross-stream wait sits on the compute stream, so
    // K3+K4 of this step still chains to K5+K2 of the previous one by programmatic dependent launch
    // (an event wait here cost 1.85 us of an 18.5 us step).
    const bool polled = sample_gather_is_fused(h);               // otherwise two plain kernels behind an event wait
    unsigned *ready = (unsigned *)(h->status + 48 + slot);
    const unsigned tick = ++h->hs_seq[slot];
    if (polled) {
        h->hs_seq_host[slot] = tick;
        APEX_CUDA(cudaMemcpyAsync(ready, h->hs_seq_host + slot, sizeof(unsigned), cudaMemcpyHostToDevice, h->hs_copy_in));
    } else {
        APEX_CUDA(cudaEventRecord(h->hs_ev_in[slot], h->hs_copy_in));
        APEX_CUDA(cudaStreamWaitEvent(st, h->hs_ev_in[slot], 0));
    }
    // compute stream: this slot's previous results have left the device
    APEX_CUDA(cudaStreamWaitEvent(st, h->hs_ev_out[slot], 0));
    float *d_delta = d_res, *d_prio = d_res + B, *d_loss = d_res + 2 * B, *d_mean = d_res + 4 * B;
    int32_t *d_idx = (int32_t *)(d_res + 3 * B);
    int rc = run_sample_gather(h, d_u, 0, B, d_idx, nullptr, nullptr, nullptr, nullptr, S_t, S_tpn, A, R, Gamma, nullptr,
                               nullptr, polled ? ready : nullptr, tick, st);
    if (rc) return rc;
    const size_t qs = (size_t)B * num_actions;
    rc = apexb200_td_fwd_update(h, d_idx, d_q, d_q + qs, d_q + 2 * qs, A, R, Gamma, B, num_actions, eps, loss_kind, kappa,
                                nullptr, d_delta, d_prio, d_loss, d_mean, nullptr, 0, nullptr, 0.0f, stream);
    if (rc) return rc;
    APEX_CUDA(cudaEventRecord(h->hs_ev_k[slot], st));
    // copy-out stream
    APEX_CUDA(cudaStreamWaitEvent(h->hs_copy_out, h->hs_ev_k[slot], 0));
    APEX_CUDA(cudaMemcpyAsync(result_host, d_res, out_words * 4, cudaMemcpyDeviceToHost, h->hs_copy_out));
    APEX_CUDA(cudaEventRecord(h->hs_ev_out[slot], h->hs_copy_out));
    if (done_event) APEX_CUDA(cudaEventRecord((cudaEvent_t)done_event, h->hs_copy_out));
    return APEXB200_OK;
}

int apexb200_host_step_sharded(apexb200_replay_t h, int32_t slot, uint64_t seed, int32_t global_count, int32_t max_rows,
                               const float *q_host, int32_t num_actions, float eps, int32_t loss_kind, float kappa,
                               uint8_t *S_t, uint8_t *S_tpn, int32_t *A, float *R, float *Gamma, float *result_host,
                               void *done_event, void *stream) {
    APEX_REQUIRE(h && q_host && result_host && S_t && S_tpn && A && R && Gamma, APEXB200_EINVAL,
                 "host_step_sharded: null argument");
    APEX_REQUIRE(slot >= 0 && slot < APEXB200_HOST_STEP_SLOTS, APEXB200_EINVAL,
                 "host_step_sharded: slot must be in [0, %d)", APEXB200_HOST_STEP_SLOTS);
    APEX_REQUIRE(global_count > 0 && max_rows > 0 && num_actions >= 1 && num_actions <= APEXB200_MAX_ACTIONS,
                 APEXB200_ERANGE, "host_step_sharded: global_count=%d / max_rows=%d / num_actions=%d out of range",
                 global_count, max_rows, num_actions);
    APEX_REQUIRE(sample_gather_is_fused(h), APEXB200_ERANGE,
                 "host_step_sharded: half a record does not fit one CTA's staging (no fused sharded kernel)");
    {
        const int world = h->peers.world > 0 ? h->peers.world : 1;
        long long share = ((long long)global_count + world - 1) / world;
        share += (share + 7) / 8 + 2;
        APEX_REQUIRE(max_rows <= 1024 || share <= 1024, APEXB200_ERANGE,
                     "host_step_sharded: the single-CTA TD kernel needs max_rows <= 1024 or a fair share <= 1024 rows");
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t M = (size_t)max_rows;
    const size_t in_bytes = (size_t)3 * M * num_actions * 4;
    const size_t out_words = 5 * M + 4;
    int prc = host_step_prepare(h, in_bytes, out_words * 4);
    if (prc) return prc;
    float *d_q = (float *)h->hs_in[slot];
    float *d_res = h->hs_out[slot];
    // copy-in stream: the staging of this slot is free once the kernels of its previous use are done
    APEX_CUDA(cudaStreamWaitEvent(h->hs_copy_in, h->hs_ev_k[slot], 0));
    APEX_CUDA(cudaMemcpyAsync(d_q, q_host, in_bytes, cudaMemcpyHostToDevice, h->hs_copy_in));
    unsigned *ready = (unsigned *)(h->status + 48 + slot);
    const unsigned tick = ++h->hs_seq[slot];
    h->hs_seq_host[slot] = tick;
    APEX_CUDA(cudaMemcpyAsync(ready, h->hs_seq_host + slot, sizeof(unsigned), cudaMemcpyHostToDevice, h->hs_copy_in));
    // compute stream: this slot's previous results have left the device
    APEX_CUDA(cudaStreamWaitEvent(st, h->hs_ev_out[slot], 0));
    float *d_delta = d_res, *d_prio = d_res + M, *d_loss = d_res + 2 * M, *d_mean = d_res + 5 * M;
    int32_t *d_idx = (int32_t *)(d_res + 3 * M), *d_pos = (int32_t *)(d_res + 4 * M), *d_count = (int32_t *)(d_res + 5 * M + 1);
    int rc = run_sample_gather_sharded(h, seed, global_count, max_rows, d_idx, nullptr, d_pos, nullptr, d_count, S_t, S_tpn,
                                       A, R, Gamma, nullptr, nullptr, ready, tick, st);
    if (rc) return rc;
    const size_t qs = M * num_actions;
    rc = apexb200_td_fwd_update(h, d_idx, d_q, d_q + qs, d_q + 2 * qs, A, R, Gamma, max_rows, num_actions, eps, loss_kind,
                                kappa, nullptr, d_delta, d_prio, d_loss, d_mean, nullptr, 0, d_count, (float)global_count,
                                stream);
    if (rc) return rc;
    APEX_CUDA(cudaEventRecord(h->hs_ev_k[slot], st));
    APEX_CUDA(cudaStreamWaitEvent(h->hs_copy_out, h->hs_ev_k[slot], 0));
    APEX_CUDA(cudaMemcpyAsync(result_host, d_res, out_words * 4, cudaMemcpyDeviceToHost, h->hs_copy_out));
    APEX_CUDA(cudaEventRecord(h->hs_ev_out[slot], h->hs_copy_out));
    if (done_event) APEX_CUDA(cudaEventRecord((cudaEvent_t)done_event, h->hs_copy_out));
    return APEXB200_OK;
}

int apexb200_host_register(void *address_host, uint64_t bytes) {
    APEX_REQUIRE(address_host && bytes > 0, APEXB200_EINVAL, "host_register: null or empty range");
    cudaError_t e = cudaHostRegister(address_host, bytes, cudaHostRegisterDefault);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        // a registration left behind by an earlier mapping of this virtual range (unmapped without unregister)
        cudaGetLastError();
        cudaHostUnregister(address_host);
        cudaGetLastError();
        e = cudaHostRegister(address_host, bytes, cudaHostRegisterDefault);
    }
    if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(e, "cudaHostRegister"); }
    return APEXB200_OK;
}

int apexb200_host_unregister(void *address_host) {
    APEX_REQUIRE(address_host, APEXB200_EINVAL, "host_unregister: null argument");
    cudaError_t e = cudaHostUnregister(address_host);
    if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(e, "cudaHostUnregister"); }
    return APEXB200_OK;
}

int apexb200_replay_buffer(apexb200_replay_t h, int32_t which, void **ptr_host, int64_t *bytes_host) {
    APEX_REQUIRE(h && ptr_host && bytes_host, APEXB200_EINVAL, "replay_buffer: null argument");
    const RingView &rv = h->ring;
    void *p = nullptr;
    int64_t b = 0;
    switch (which) {
        case APEXB200_BUF_FRAMES: p = rv.frames; b = rv.N * rv.fpr * rv.F; break;
        case APEXB200_BUF_FIDX: p = rv.fidx; b = rv.N * 2 * rv.C * 4; break;
        case APEXB200_BUF_ACTION: p = rv.A; b = rv.N * 4; break;
        case APEXB200_BUF_REWARD: p = rv.R; b = rv.N * 4; break;
        case APEXB200_BUF_GAMMA: p = rv.G; b = rv.N * 4; break;
        case APEXB200_BUF_QT: p = rv.q_t; b = rv.q_t ? rv.N * rv.nA * 4 : 0; break;
        case APEXB200_BUF_QTPN: p = rv.q_tpn; b = rv.q_tpn ? rv.N * rv.nA * 4 : 0; break;
        case APEXB200_BUF_LEAVES: p = h->tree.lvl[0]; b = rv.N * 8; break;
        default: set_error("replay_buffer: unknown buffer id %d", which); return APEXB200_EINVAL;
    }
    *ptr_host = p;
    *bytes_host = b;
    return APEXB200_OK;
}

int apexb200_replay_get_heads(apexb200_replay_t h, int64_t *lane_count_host) {
    APEX_REQUIRE(h && lane_count_host, APEXB200_EINVAL, "replay_get_heads: null argument");
    for (int i = 0; i < h->ring.lanes; ++i) lane_count_host[i] = h->lane_count[i];
    return APEXB200_OK;
}

int apexb200_replay_set_heads(apexb200_replay_t h, const int64_t *lane_count_host) {
    APEX_REQUIRE(h && lane_count_host, APEXB200_EINVAL, "replay_set_heads: null argument");
    for (int i = 0; i < h->ring.lanes; ++i) {
        APEX_REQUIRE(lane_count_host[i] >= 0, APEXB200_EINVAL, "replay_set_heads: negative head");
        h->lane_count[i] = lane_count_host[i];
    }
    return APEXB200_OK;
}

}  // extern "C"
