// Device-side ingest of actor step streams: the chain
//   ExperienceBuffer.add / update_buffer / construct_nstep_transition (actor.py:29-70)
//   -> Actor.compute_priorities (actor.py:127-143) -> queue pump (main.py:21-25) -> ReplayMemory.add (replay.py:59-69)
// for E environments, as a fixed sequence of launches with no host arithmetic in between: the n-step carry, the
// episode ages and the ring heads of every environment live on the device.
//
// One environment owns one lane of a stream-layout replay.  The frame observed at env step tau is stored at lane
// position tau together with the n-step record of step t = tau - n (its S_tpn ends at that frame):
//   A_t, R = sum_{k<m} gamma^k r_{t+k}, Gamma = gamma^m or 0 when a terminal lies inside the window (actor.py:35-43,
//   intended semantics), initial priority |R + Gamma max_a q_tau - q_t[A_t]| (actor.py:138-140).
// Everything a record needs lies in the past of its frame, so a chunk of T steps emits T records at once; only the
// last n steps are carried to the next chunk.  The first n frames of a stream carry no record (priority 0).
#include <new>

#include "common.cuh"

namespace apexb200 {

constexpr int kIngestMaxN = 16;              // n-step window
constexpr int kIngestEnvsPerLaunch = 256;    // environments named by one kernel-parameter table

struct IngestState {                          // device arrays, one row per environment
    int32_t *c_act;                           // [E, n]      actions of the last n steps
    float *c_rew;                             // [E, n]
    uint8_t *c_done;                          // [E, n]      (initially 1: no record before the stream starts)
    float *c_q;                               // [E, n, nA]
    int32_t *age;                             // [E]         frames since the episode start, at the NEXT frame
    int32_t *age_hist;                        // [E, n]      ages of the last n frames
    long long *seen;                          // [E]         frames pushed so far
    long long *head;                          // [E]         records ever inserted into the env's lane (ring head)
    int32_t *lane;                            // [E]         lane of each environment
    int n, nA;
    float gamma;
};

struct EnvTable {
    int count;
    int env[kIngestEnvsPerLaunch];
};

struct ChunkView {                            // one chunk of T steps for the table's environments, [k, T, ...] per field
    const uint8_t *frames;                    // [k, T, F]
    const int32_t *actions;                   // [k, T]
    const float *rewards;                     // [k, T]
    const uint8_t *dones;                     // [k, T]
    const float *q;                           // [k, T, nA]
    int T;
};

// K6a for device-side heads: stream the chunk's frames into the ring slots of every listed environment.
__global__ void __launch_bounds__(256)
ingest_frames_kernel(uint4 *__restrict__ ring_frames, ChunkView ch, int vecs_per_record, EnvTable et, IngestState stt,
                     long long lane_cap) {
    const int v0 = blockIdx.x * 512 + threadIdx.x;
    const int v1 = v0 + 256;
    const int k = blockIdx.z;
    const int env = et.env[k];
    const long long lane_base = (long long)stt.lane[env] * lane_cap, head = stt.head[env] % lane_cap;
    const uint4 *src = (const uint4 *)ch.frames + (long long)k * ch.T * vecs_per_record;
    for (int j = blockIdx.y; j < ch.T; j += gridDim.y) {
        const long long slot = lane_base + (head + j) % lane_cap;
        const uint4 *s = src + (long long)j * vecs_per_record;
        uint4 *d = ring_frames + slot * vecs_per_record;
        uint4 a, b;
        if (v0 < vecs_per_record) a = __ldcs(s + v0);
        if (v1 < vecs_per_record) b = __ldcs(s + v1);
        if (v0 < vecs_per_record) d[v0] = a;
        if (v1 < vecs_per_record) d[v1] = b;
    }
}

// Episode age of the chunk's frame f: frames since the last episode end before it (age0 = age at the chunk's
// first frame, carried from the previous chunk).
__device__ __forceinline__ int chunk_frame_age(const uint8_t *dones, int f, int age0) {
    for (int i = f - 1; i >= 0; --i)
        if (dones[i]) return f - 1 - i;
    return age0 + f;
}

// K7 + a11 + K6b in one kernel: thread j of environment k builds the record stored with the chunk's frame j.
// "cat" index i addresses the carried steps (i < n) followed by the chunk's steps (i - n).
__global__ void __launch_bounds__(128)
ingest_records_kernel(RingView rv, TreeView tv, IngestState stt, EnvTable et, ChunkView ch, int32_t *status) {
    bump_epoch(status);
    const int k = blockIdx.y;
    const int env = et.env[k];
    const int n = stt.n, nA = stt.nA, T = ch.T;
    const int lane = stt.lane[env];
    const long long head = stt.head[env] % rv.lane_cap;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = j < T + rv.lookback;
    const long long p = (head + (active ? j : 0)) % rv.lane_cap;
    const long long s = (long long)lane * rv.lane_cap + p;
    const int32_t *act_c = stt.c_act + (long long)env * n, *act_x = ch.actions + (long long)k * T;
    const float *rew_c = stt.c_rew + (long long)env * n, *rew_x = ch.rewards + (long long)k * T;
    const uint8_t *don_c = stt.c_done + (long long)env * n, *don_x = ch.dones + (long long)k * T;
    const float *q_c = stt.c_q + (long long)env * n * nA, *q_x = ch.q + (long long)k * T * nA;
    unsigned long long new_leaf = 0ull;
    if (active && j < T) {
        // n-step return over cat steps j .. j+n-1 (K7 arithmetic: explicit fp32 roundings, gamma^k by repeated product)
        float acc = 0.0f, disc = 1.0f;
        bool term = false;
        for (int w = 0; w < n; ++w) {
            const int i = j + w;
            const float r = i < n ? rew_c[i] : rew_x[i - n];
            acc = __fadd_rn(acc, __fmul_rn(disc, r));
            disc = __fmul_rn(disc, stt.gamma);
            if (i < n ? don_c[i] : don_x[i - n]) { term = true; break; }
        }
        const float R = acc, G = term ? 0.0f : disc;
        const int a = j < n ? act_c[j] : act_x[j - n];
        const int age0 = stt.age[env];
        const int age_tau = chunk_frame_age(don_x, j, age0);
        const int age_t = j < n ? stt.age_hist[(long long)env * n + j] : chunk_frame_age(don_x, j - n, age0);   // frame tau - n
        // frame-slot row: S_t = the C frames ending n frames back, S_tpn = the C frames ending here; offsets older
        // than the episode start repeat its first frame (stream_offsets)
        const int C = rv.C, twoC = 2 * rv.C;
        const long long lane_frames = rv.lane_cap * rv.fpr;
        const long long lane_frame_base = (long long)lane * lane_frames;
        const long long last = p * rv.fpr + rv.fpr - 1;
        for (int c = 0; c < twoC; ++c) {
            const int back = C - 1 - (c < C ? c : c - C);
            const long long rel = c < C ? -(long long)(n + min(back, age_t)) : -(long long)min(back, age_tau);
            long long f = (last + rel) % lane_frames;
            if (f < 0) f += lane_frames;
            rv.fidx[s * twoC + c] = (int32_t)(lane_frame_base + f);
        }
        // initial priority from the actor's own Q rows (actor.py:138-140): q_t = Q(s_t), q_tpn = Q(s_tau)
        const float *qa = j < n ? q_c + (long long)j * nA : q_x + (long long)(j - n) * nA;
        const float *qb = q_x + (long long)j * nA;
        float mx = qb[0];
        for (int c = 1; c < nA; ++c) mx = fmaxf(mx, qb[c]);
        const int a_sel = a < 0 ? 0 : (a >= nA ? nA - 1 : a);
        if (a != a_sel) atomicOr(status, APEXB200_ST_BAD_ACTION);
        float p0 = fabsf(__fsub_rn(__fadd_rn(R, __fmul_rn(G, mx)), qa[a_sel]));
        if (stt.seen[env] + j < n) p0 = 0.0f;                     // the stream's first n frames: no record yet
        rv.A[s] = a; rv.R[s] = R; rv.G[s] = G;
        if (rv.q_t != nullptr && rv.nA == nA) {
            for (int c = 0; c < nA; ++c) { rv.q_t[s * nA + c] = qa[c]; rv.q_tpn[s * nA + c] = qb[c]; }
        }
        new_leaf = leaf_from_prio(prio_pow(p0, rv.alpha, 0.0f));
    }
    unsigned long long d = 0ull;
    if (active) {                                                   // j >= T: records whose frames are overwritten next
        const unsigned long long old = tv.lvl[0][s];
        tv.lvl[0][s] = new_leaf;
        d = new_leaf - old;
    }
    warp_propagate_adjacent(tv, active ? s : -1, d, 1);
}

// After the two kernels above: roll the per-environment state forward by T steps (one warp per environment).
__global__ void __launch_bounds__(32)
ingest_commit_kernel(IngestState stt, EnvTable et, ChunkView ch) {
    const int k = blockIdx.x, env = et.env[k], n = stt.n, nA = stt.nA, T = ch.T, lane = threadIdx.x;
    int32_t *act_c = stt.c_act + (long long)env * n;
    float *rew_c = stt.c_rew + (long long)env * n;
    uint8_t *don_c = stt.c_done + (long long)env * n;
    float *q_c = stt.c_q + (long long)env * n * nA;
    int32_t *hist = stt.age_hist + (long long)env * n;
    const int32_t *act_x = ch.actions + (long long)k * T;
    const float *rew_x = ch.rewards + (long long)k * T;
    const uint8_t *don_x = ch.dones + (long long)k * T;
    const float *q_x = ch.q + (long long)k * T * nA;
    const int age0 = stt.age[env];
    // new carry entry i = cat step T + i; read everything before the first write (T < n shifts inside the carry)
    int a_new = 0, h_new = 0;
    float r_new = 0.0f;
    uint8_t d_new = 0;
    if (lane < n) {
        const int i = T + lane;
        a_new = i < n ? act_c[i] : act_x[i - n];
        r_new = i < n ? rew_c[i] : rew_x[i - n];
        d_new = i < n ? don_c[i] : don_x[i - n];
        h_new = i < n ? hist[i] : chunk_frame_age(don_x, i - n, age0);
    }
    float q_new[4];                                                  // entries lane, lane+32, ... of the n*nA carry
    const int nq = n * nA;
    #pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int e = lane + 32 * r;
        q_new[r] = 0.0f;
        if (e < nq) {
            const int i = T + e / nA, c = e % nA;
            q_new[r] = i < n ? q_c[(long long)i * nA + c] : q_x[(long long)(i - n) * nA + c];
        }
    }
    const int age_next = don_x[T - 1] ? 0 : chunk_frame_age(don_x, T - 1, age0) + 1;
    __syncwarp();
    if (lane < n) { act_c[lane] = a_new; rew_c[lane] = r_new; don_c[lane] = d_new; hist[lane] = h_new; }
    #pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int e = lane + 32 * r;
        if (e < nq) q_c[e] = q_new[r];
    }
    if (lane == 0) {
        stt.age[env] = age_next;
        stt.seen[env] += T;
        stt.head[env] += T;
    }
}

// Pack kernel of the shared-memory transport: chunk blobs written by actor processes (frames | actions | rewards |
// dones | q at fixed offsets, ActorStagingRing layout) are read straight out of page-locked, device-mapped HOST
// memory and laid out as the [k, T, ...] arrays the kernels above take.  One launch moves every ready chunk; no
// per-ring copy calls.
struct BlobTable {
    int count;
    unsigned long long blob[kIngestEnvsPerLaunch];                  // device-visible address of each chunk
};

__global__ void __launch_bounds__(256)
ingest_pack_kernel(BlobTable bt, int T, int F, int nA, int off_act, int off_rew, int off_done, int off_q,
                   uint8_t *frames, int32_t *actions, float *rewards, uint8_t *dones, float *q) {
    const int k = blockIdx.y;
    const uint8_t *blob = (const uint8_t *)bt.blob[k];
    const long long fv = (long long)T * F / 16;                      // 16-byte vectors of frame data
    const uint4 *src = (const uint4 *)blob;
    uint4 *dst = (uint4 *)(frames + (long long)k * T * F);
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < fv; v += (long long)gridDim.x * blockDim.x)
        dst[v] = src[v];
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i < T; i += blockDim.x) {
            actions[(long long)k * T + i] = ((const int32_t *)(blob + off_act))[i];
            rewards[(long long)k * T + i] = ((const float *)(blob + off_rew))[i];
            dones[(long long)k * T + i] = (blob + off_done)[i];
        }
        for (int i = threadIdx.x; i < T * nA; i += blockDim.x) q[(long long)k * T * nA + i] = ((const float *)(blob + off_q))[i];
    }
}

}  // namespace apexb200

using namespace apexb200;

struct apexb200_ingest {
    apexb200_replay *h = nullptr;
    IngestState st{};
    int E = 0, t_max = 0, device = 0;        // device ordinal kept here: destroy must not touch the replay handle
    void *block = nullptr;                    // one allocation for the whole state
    std::vector<int> lane_of_env;
    // staging of the shared-memory transport: [E, t_max, ...] per field
    uint8_t *sg_frames = nullptr, *sg_dones = nullptr;
    int32_t *sg_actions = nullptr;
    float *sg_rewards = nullptr, *sg_q = nullptr;
};

extern "C" {

int apexb200_ingest_create(apexb200_replay_t h, int32_t num_envs, const int32_t *lanes_host, int32_t n_step, float gamma,
                           int32_t num_actions, int32_t max_chunk_steps, apexb200_ingest_t *out) {
    APEX_REQUIRE(h && lanes_host && out, APEXB200_EINVAL, "ingest_create: null argument");
    *out = nullptr;
    const RingView &rv = h->ring;
    APEX_REQUIRE(rv.fpr == 1, APEXB200_EINVAL, "ingest_create: needs a frame-once (stream) replay, frames_per_record = 1");
    APEX_REQUIRE(num_envs >= 1 && num_envs <= rv.lanes, APEXB200_EINVAL, "ingest_create: %d environments for %d lanes",
                 num_envs, rv.lanes);
    APEX_REQUIRE(n_step >= 1 && n_step <= kIngestMaxN, APEXB200_ERANGE, "ingest_create: n_step %d outside [1, %d]", n_step,
                 kIngestMaxN);
    APEX_REQUIRE(num_actions >= 1 && num_actions <= APEXB200_MAX_ACTIONS && n_step * num_actions <= 128, APEXB200_ERANGE,
                 "ingest_create: num_actions %d out of range (n_step * num_actions <= 128)", num_actions);
    APEX_REQUIRE(rv.lookback >= rv.C - 1 + n_step, APEXB200_EINVAL,
                 "ingest_create: replay lookback %d < frames_per_state - 1 + n_step = %d", rv.lookback, rv.C - 1 + n_step);
    APEX_REQUIRE(max_chunk_steps >= 0 && (long long)max_chunk_steps + rv.lookback <= rv.lane_cap, APEXB200_ERANGE,
                 "ingest_create: max_chunk_steps %d + lookback %d exceeds the lane capacity %lld", max_chunk_steps,
                 rv.lookback, rv.lane_cap);
    std::vector<char> seen((size_t)rv.lanes, 0);
    for (int e = 0; e < num_envs; ++e) {
        const int l = lanes_host[e];
        APEX_REQUIRE(l >= 0 && l < rv.lanes && !seen[l], APEXB200_EINVAL,
                     "ingest_create: lane %d outside [0,%d) or given to two environments", l, rv.lanes);
        seen[l] = 1;
    }
    APEX_CUDA(cudaSetDevice(h->cfg.device));
    apexb200_ingest *g = new (std::nothrow) apexb200_ingest();
    APEX_REQUIRE(g, APEXB200_ENOMEM, "ingest_create: host allocation failed");
    g->h = h;
    g->device = h->cfg.device;
    g->E = num_envs;
    g->t_max = max_chunk_steps;
    g->lane_of_env.assign(lanes_host, lanes_host + num_envs);
    const size_t E = (size_t)num_envs, n = (size_t)n_step, nA = (size_t)num_actions;
    auto up = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t b_act = up(E * n * 4), b_rew = up(E * n * 4), b_done = up(E * n), b_q = up(E * n * nA * 4), b_age = up(E * 4),
                 b_hist = up(E * n * 4), b_seen = up(E * 8), b_head = up(E * 8), b_lane = up(E * 4);
    const size_t total = b_act + b_rew + b_done + b_q + b_age + b_hist + b_seen + b_head + b_lane;
    cudaError_t e = cudaMalloc(&g->block, total);
    if (e != cudaSuccess) { delete g; return cuda_fail(e, "cudaMalloc"); }
    e = cudaMemset(g->block, 0, total);
    if (e != cudaSuccess) { cudaFree(g->block); delete g; return cuda_fail(e, "cudaMemset"); }
    char *p = (char *)g->block;
    IngestState &st = g->st;
    st.c_act = (int32_t *)p; p += b_act;
    st.c_rew = (float *)p; p += b_rew;
    st.c_done = (uint8_t *)p; p += b_done;
    st.c_q = (float *)p; p += b_q;
    st.age = (int32_t *)p; p += b_age;
    st.age_hist = (int32_t *)p; p += b_hist;
    st.seen = (long long *)p; p += b_seen;
    st.head = (long long *)p; p += b_head;
    st.lane = (int32_t *)p;
    st.n = n_step; st.nA = num_actions; st.gamma = gamma;
    std::vector<long long> heads(E);
    for (size_t k = 0; k < E; ++k) heads[k] = h->lane_count[lanes_host[k]];
    bool ok = cudaMemset(st.c_done, 1, E * n) == cudaSuccess &&            // "terminal": nothing precedes the stream
              cudaMemcpy(st.head, heads.data(), E * 8, cudaMemcpyHostToDevice) == cudaSuccess &&
              cudaMemcpy(st.lane, lanes_host, E * 4, cudaMemcpyHostToDevice) == cudaSuccess;
    if (ok && max_chunk_steps > 0) {
        const size_t T = (size_t)max_chunk_steps;
        ok = cudaMalloc((void **)&g->sg_frames, E * T * rv.F) == cudaSuccess &&
             cudaMalloc((void **)&g->sg_actions, E * T * 4) == cudaSuccess &&
             cudaMalloc((void **)&g->sg_rewards, E * T * 4) == cudaSuccess &&
             cudaMalloc((void **)&g->sg_dones, up(E * T)) == cudaSuccess &&
             cudaMalloc((void **)&g->sg_q, E * T * nA * 4) == cudaSuccess;
    }
    if (!ok) {
        int rc = cuda_fail(cudaGetLastError(), "ingest_create: device set-up");
        apexb200_ingest_destroy(g);
        return rc;
    }
    *out = g;
    return APEXB200_OK;
}

int apexb200_ingest_destroy(apexb200_ingest_t g) {
    if (!g) return APEXB200_OK;
    cudaSetDevice(g->device);                 // (the replay handle may already be gone)
    void *ptrs[] = {g->block, g->sg_frames, g->sg_actions, g->sg_rewards, g->sg_dones, g->sg_q};
    for (void *p : ptrs) if (p) cudaFree(p);
    delete g;
    return APEXB200_OK;
}

}  // extern "C"

namespace apexb200 {
static int ingest_push_group(apexb200_ingest *g, const EnvTable &et, const ChunkView &ch, cudaStream_t st) {
    apexb200_replay *h = g->h;
    const RingView &rv = h->ring;
    const int vpr = rv.F / 16;
    int max_rows = kNumSMs * 8 / et.count;                                // ~8 resident CTAs per SM over all environments
    if (max_rows < 1) max_rows = 1;
    const dim3 grid((unsigned)((vpr + 511) / 512), (unsigned)(ch.T < max_rows ? ch.T : max_rows), (unsigned)et.count);
    ingest_frames_kernel<<<grid, 256, 0, st>>>((uint4 *)rv.frames, ch, vpr, et, g->st, rv.lane_cap);
    int rc = launch_check("ingest_frames_kernel");
    if (rc) return rc;
    const int nmeta = ch.T + rv.lookback;
    ingest_records_kernel<<<dim3((unsigned)((nmeta + 127) / 128), (unsigned)et.count), 128, 0, st>>>(rv, h->tree, g->st, et, ch,
                                                                                                 h->status);
    rc = launch_check("ingest_records_kernel");
    if (rc) return rc;
    ingest_commit_kernel<<<(unsigned)et.count, 32, 0, st>>>(g->st, et, ch);
    return launch_check("ingest_commit_kernel");
}

static int ingest_check_push(apexb200_ingest *g, const int32_t *envs_host, int32_t n_envs, int32_t T, const char *who) {
    APEX_REQUIRE(g && g->h, APEXB200_EINVAL, "%s: null handle", who);
    APEX_REQUIRE(T >= 1 && (long long)T + g->h->ring.lookback <= g->h->ring.lane_cap, APEXB200_ERANGE,
                 "%s: %d steps + lookback %d exceed the lane capacity %lld", who, T, g->h->ring.lookback, g->h->ring.lane_cap);
    APEX_REQUIRE(n_envs >= 1 && n_envs <= g->E, APEXB200_EINVAL, "%s: %d environments of %d", who, n_envs, g->E);
    if (envs_host) {
        std::vector<char> seen((size_t)g->E, 0);
        for (int k = 0; k < n_envs; ++k) {
            const int e = envs_host[k];
            APEX_REQUIRE(e >= 0 && e < g->E && !seen[e], APEXB200_EINVAL, "%s: environment %d outside [0,%d) or listed twice", who,
                         e, g->E);
            seen[e] = 1;
        }
    }
    return APEXB200_OK;
}
}  // namespace apexb200

extern "C" {

int apexb200_ingest_push(apexb200_ingest_t g, const int32_t *envs_host, int32_t n_envs, int32_t T, const uint8_t *frames,
                         const int32_t *actions, const float *rewards, const uint8_t *dones, const float *q, void *stream) {
    int rc = ingest_check_push(g, envs_host, n_envs, T, "ingest_push");
    if (rc) return rc;
    APEX_REQUIRE(frames && actions && rewards && dones && q, APEXB200_EINVAL, "ingest_push: null input");
    APEX_REQUIRE(((uintptr_t)frames & 15) == 0, APEXB200_EINVAL, "ingest_push: frames must be 16-byte aligned");
    const RingView &rv = g->h->ring;
    const int nA = g->st.nA;
    for (int first = 0; first < n_envs; first += kIngestEnvsPerLaunch) {
        EnvTable et;
        et.count = n_envs - first < kIngestEnvsPerLaunch ? n_envs - first : kIngestEnvsPerLaunch;
        for (int k = 0; k < et.count; ++k) et.env[k] = envs_host ? envs_host[first + k] : first + k;
        const long long o = (long long)first * T;
        ChunkView ch{frames + o * rv.F, actions + o, rewards + o, dones + o, q + o * nA, T};
        rc = ingest_push_group(g, et, ch, (cudaStream_t)stream);
        if (rc) return rc;
        for (int k = 0; k < et.count; ++k) g->h->lane_count[g->lane_of_env[et.env[k]]] += T;     // host mirror of the heads
    }
    return APEXB200_OK;
}

int apexb200_ingest_push_blobs(apexb200_ingest_t g, const int32_t *envs_host, const uint64_t *blobs_host, int32_t n_envs,
                               int32_t T, int32_t off_actions, int32_t off_rewards, int32_t off_dones, int32_t off_q,
                               void *stream) {
    int rc = ingest_check_push(g, envs_host, n_envs, T, "ingest_push_blobs");
    if (rc) return rc;
    APEX_REQUIRE(envs_host && blobs_host, APEXB200_EINVAL, "ingest_push_blobs: null argument");
    APEX_REQUIRE(T <= g->t_max, APEXB200_ERANGE, "ingest_push_blobs: %d steps per chunk, staging was sized for %d", T, g->t_max);
    APEX_REQUIRE(off_actions % 4 == 0 && off_rewards % 4 == 0 && off_q % 4 == 0 && off_actions >= T * g->h->ring.F,
                 APEXB200_EINVAL, "ingest_push_blobs: field offsets must be 4-byte aligned and follow the frames");
    const RingView &rv = g->h->ring;
    const int nA = g->st.nA;
    cudaStream_t st = (cudaStream_t)stream;
    for (int first = 0; first < n_envs; first += kIngestEnvsPerLaunch) {
        EnvTable et;
        BlobTable bt;
        et.count = bt.count = n_envs - first < kIngestEnvsPerLaunch ? n_envs - first : kIngestEnvsPerLaunch;
        for (int k = 0; k < et.count; ++k) {
            et.env[k] = envs_host[first + k];
            bt.blob[k] = blobs_host[first + k];
            APEX_REQUIRE((bt.blob[k] & 15) == 0, APEXB200_EINVAL, "ingest_push_blobs: chunk %d is not 16-byte aligned", first + k);
        }
        // staging rows first .. first+count of the [E, t_max, ...] arrays, used as [count, T, ...]
        const long long o = (long long)first * g->t_max;
        ChunkView ch{g->sg_frames + o * rv.F, g->sg_actions + o, g->sg_rewards + o, g->sg_dones + o, g->sg_q + o * nA, T};
        const long long fv = (long long)T * rv.F / 16;
        int bx = (int)((fv + 255) / 256);
        const int cap = kNumSMs * 8 / et.count > 0 ? kNumSMs * 8 / et.count : 1;
        if (bx > cap) bx = cap;
        ingest_pack_kernel<<<dim3((unsigned)bx, (unsigned)et.count), 256, 0, st>>>(
            bt, T, rv.F, nA, off_actions, off_rewards, off_dones, off_q, (uint8_t *)ch.frames, (int32_t *)ch.actions,
            (float *)ch.rewards, (uint8_t *)ch.dones, (float *)ch.q);
        rc = launch_check("ingest_pack_kernel");
        if (rc) return rc;
        rc = ingest_push_group(g, et, ch, st);
        if (rc) return rc;
        for (int k = 0; k < et.count; ++k) g->h->lane_count[g->lane_of_env[et.env[k]]] += T;
    }
    return APEXB200_OK;
}

int apexb200_ingest_sync_heads(apexb200_ingest_t g, void *stream) {
    APEX_REQUIRE(g && g->h, APEXB200_EINVAL, "ingest_sync_heads: null handle");
    std::vector<long long> heads((size_t)g->E);
    APEX_CUDA(cudaMemcpyAsync(heads.data(), g->st.head, (size_t)g->E * 8, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    APEX_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    for (int e = 0; e < g->E; ++e) g->h->lane_count[g->lane_of_env[e]] = heads[e];
    return APEXB200_OK;
}

int apexb200_host_register_mapped(void *address_host, uint64_t bytes, void **device_pointer_host) {
    APEX_REQUIRE(address_host && bytes > 0 && device_pointer_host, APEXB200_EINVAL, "host_register_mapped: null or empty range");
    cudaError_t e = cudaHostRegister(address_host, bytes, cudaHostRegisterMapped | cudaHostRegisterPortable);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        cudaGetLastError();
        cudaHostUnregister(address_host);
        cudaGetLastError();
        e = cudaHostRegister(address_host, bytes, cudaHostRegisterMapped | cudaHostRegisterPortable);
    }
    if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(e, "cudaHostRegister(mapped)"); }
    APEX_CUDA(cudaHostGetDevicePointer(device_pointer_host, address_host, 0));
    return APEXB200_OK;
}

}  // extern "C"
