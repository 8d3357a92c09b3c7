// Learner -> actor weight publication and the batched actor policy.
//
//   K9  params_pack / params_unpack: the tensors of a state_dict <-> one flat, 256-byte aligned
//       device snapshot (multi-tensor copy over a chunk table; HBM-bound, 2 x bytes of traffic);
//   K10 actor_policy: duelling combine with batch-1 semantics + epsilon-greedy for E environments.
//
// Replaces shared_state["Q_state_dict"] = Q.state_dict() (learner.py:74, a 13.3 MB pickle per
// step through a Manager proxy), Q.load_state_dict(shared_state[...]) (actor.py:106, 189-191) and
// the batch-1 CPU policy of actor.py:121-125, 160-163.
#include <cstring>
#include <new>

#include "common.cuh"

namespace apexb200 {

constexpr uint32_t kChunkBytes = 32768;       // one CTA per chunk: 256 threads x 8 x 16 B
constexpr int kPublishNodes = 8;              // host-callback arguments kept alive per board

struct ParamChunk {
    unsigned long long tensor;                // device address inside one tensor
    unsigned long long flat_off;              // byte offset inside the flat snapshot
    uint32_t bytes;
    uint32_t pad;
};

struct PublishNode {
    volatile int64_t *ctrl;
    int64_t version;
    volatile int pending;                     // set at enqueue, cleared by the host callback
};

}  // namespace apexb200

struct apexb200_params {
    int device = 0;
    uint64_t flat_bytes = 0;
    uint8_t *flat = nullptr;                  // device snapshot
    apexb200::ParamChunk *chunks = nullptr;   // device chunk table
    uint32_t n_chunks = 0;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t packed = nullptr, copied = nullptr;
    bool copy_pending = false;
    apexb200::PublishNode nodes[apexb200::kPublishNodes] = {};
    unsigned node_next = 0;
};

namespace apexb200 {

// K9.  to_flat != 0: tensors -> flat; else flat -> tensors.  Every flat offset is 256-byte aligned
// (checked at create) and torch allocations are at least 16-byte aligned, so the vector path is
// the normal one; a tensor at an odd address (a view) takes the byte path.
__global__ void __launch_bounds__(256)
params_copy_kernel(const ParamChunk *__restrict__ chunks, uint8_t *__restrict__ flat, int to_flat) {
    const ParamChunk c = chunks[blockIdx.x];
    uint8_t *t = reinterpret_cast<uint8_t *>(c.tensor);
    uint8_t *f = flat + c.flat_off;
    const uint8_t *src = to_flat ? t : f;
    uint8_t *dst = to_flat ? f : t;
    const uint32_t n = c.bytes;
    if ((c.tensor & 15ull) == 0) {
        const uint32_t nv = n >> 4;
        const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
        uint4 *d4 = reinterpret_cast<uint4 *>(dst);
        uint4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint32_t i = threadIdx.x + k * 256;
            if (i < nv) v[k] = __ldcs(s4 + i);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint32_t i = threadIdx.x + k * 256;
            if (i < nv) d4[i] = v[k];
        }
        for (uint32_t i = (nv << 4) + threadIdx.x; i < n; i += 256) dst[i] = src[i];
    } else {
        for (uint32_t i = threadIdx.x; i < n; i += 256) dst[i] = src[i];
    }
}

__device__ __forceinline__ double mix_uniform(unsigned long long seed, unsigned long long step, unsigned long long e,
                                              unsigned long long which) {
    unsigned long long z = seed ^ (step * 0xD1342543DE82EF95ull) ^ (which * 0xA24BAED4963EE407ull);
    z += (e + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (double)(z >> 11) * 0x1.0p-53;
}

// K10.  One thread per environment (A <= 64 floats per row).
//   q[e, a]  = value[e] + (adv[e, a] - (1/A) * sum_a adv[e, a])      duelling_network.py:27 at batch 1
//   greedy   = first a with the largest q (np.argmax: NaN counts as largest)   actor.py:122-123
//   explore  iff u_explore[e] < eps[e]  (random.random() >= eps -> greedy)     actor.py:122
//   random a = min(floor(u_action[e] * A), A - 1)                               actor.py:125
// With value == nullptr `adv` already holds Q rows and is used as is.
__global__ void __launch_bounds__(128)
actor_policy_kernel(const float *__restrict__ value, const float *__restrict__ adv, const float *__restrict__ eps,
                    const double *__restrict__ u_explore, const double *__restrict__ u_action,
                    unsigned long long seed, unsigned long long step, int E, int A, float inv_A,
                    float *__restrict__ q_out, int32_t *__restrict__ actions, uint8_t *__restrict__ explored) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    const float *row = adv + (size_t)e * A;
    float centre = 0.f, v = 0.f;
    if (value) {
        float s = 0.f;
        for (int a = 0; a < A; ++a) s += row[a];
        centre = inv_A * s;
        v = value[e];
    }
    int best = 0;
    float best_q = 0.f;
    for (int a = 0; a < A; ++a) {
        const float q = value ? v + (row[a] - centre) : row[a];
        if (q_out) q_out[(size_t)e * A + a] = q;
        if (a == 0 || (q > best_q) || (q != q && best_q == best_q)) { best = a; best_q = q; }
    }
    const double ue = u_explore ? u_explore[e] : mix_uniform(seed, step, (unsigned long long)e, 1);
    int act = best;
    const bool ex = ue < (double)eps[e];
    if (ex) {
        const double ua = u_action ? u_action[e] : mix_uniform(seed, step, (unsigned long long)e, 2);
        act = min((int)(ua * (double)A), A - 1);
    }
    actions[e] = act;
    if (explored) explored[e] = ex ? 1 : 0;
}

static void CUDART_CB publish_done(void *arg) {
    PublishNode *n = static_cast<PublishNode *>(arg);
    *n->ctrl = n->version;                    // "completed" word of the shared-memory board
    n->pending = 0;
}

}  // namespace apexb200

using namespace apexb200;

extern "C" {

int apexb200_params_create(int32_t device, const apexb200_param_span *spans_host, int32_t n_spans, uint64_t flat_bytes,
                           apexb200_params_t *out) {
    APEX_REQUIRE(out && spans_host && n_spans > 0 && flat_bytes > 0, APEXB200_EINVAL, "params_create: null or empty argument");
    *out = nullptr;
    std::vector<ParamChunk> chunks;
    for (int i = 0; i < n_spans; ++i) {
        const apexb200_param_span &s = spans_host[i];
        APEX_REQUIRE(s.tensor != 0 && (s.flat_offset & 255ull) == 0 && s.flat_offset + s.bytes <= flat_bytes, APEXB200_EINVAL,
                     "params_create: span %d is null, not 256-byte aligned in the flat buffer or beyond flat_bytes", i);
        for (uint64_t off = 0; off < s.bytes; off += kChunkBytes) {
            ParamChunk c;
            c.tensor = s.tensor + off;
            c.flat_off = s.flat_offset + off;
            c.bytes = (uint32_t)((s.bytes - off < kChunkBytes) ? (s.bytes - off) : kChunkBytes);
            c.pad = 0;
            chunks.push_back(c);
        }
    }
    APEX_REQUIRE(!chunks.empty(), APEXB200_EINVAL, "params_create: every span is empty");
    APEX_CUDA(cudaSetDevice(device));
    apexb200_params *p = new (std::nothrow) apexb200_params();
    APEX_REQUIRE(p, APEXB200_ENOMEM, "params_create: host allocation failed");
    p->device = device;
    p->flat_bytes = flat_bytes;
    p->n_chunks = (uint32_t)chunks.size();
    cudaError_t e = cudaMalloc((void **)&p->flat, flat_bytes);
    if (e == cudaSuccess) e = cudaMemset(p->flat, 0, flat_bytes);
    if (e == cudaSuccess) e = cudaMalloc((void **)&p->chunks, chunks.size() * sizeof(ParamChunk));
    if (e == cudaSuccess) e = cudaMemcpy(p->chunks, chunks.data(), chunks.size() * sizeof(ParamChunk), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&p->copy_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->packed, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->copied, cudaEventDisableTiming);
    if (e != cudaSuccess) {
        const int rc = cuda_fail(e, "params_create");
        apexb200_params_destroy(p);
        return rc;
    }
    *out = p;
    return APEXB200_OK;
}

int apexb200_params_destroy(apexb200_params_t p) {
    if (!p) return APEXB200_OK;
    cudaSetDevice(p->device);
    if (p->copy_stream) { cudaStreamSynchronize(p->copy_stream); cudaStreamDestroy(p->copy_stream); }
    if (p->packed) cudaEventDestroy(p->packed);
    if (p->copied) cudaEventDestroy(p->copied);
    if (p->flat) cudaFree(p->flat);
    if (p->chunks) cudaFree(p->chunks);
    delete p;
    return APEXB200_OK;
}

int apexb200_params_flat(apexb200_params_t p, void **flat_dev_host, uint64_t *bytes_host) {
    APEX_REQUIRE(p && flat_dev_host, APEXB200_EINVAL, "params_flat: null argument");
    *flat_dev_host = p->flat;
    if (bytes_host) *bytes_host = p->flat_bytes;
    return APEXB200_OK;
}

int apexb200_params_publish(apexb200_params_t p, void *slot_host, int64_t *ctrl_host, int64_t version, void *stream) {
    APEX_REQUIRE(p, APEXB200_EINVAL, "params_publish: null handle");
    APEX_REQUIRE((slot_host == nullptr) == (ctrl_host == nullptr), APEXB200_EINVAL,
                 "params_publish: slot_host and ctrl_host go together");
    cudaStream_t st = (cudaStream_t)stream;
    if (p->copy_pending) APEX_CUDA(cudaStreamWaitEvent(st, p->copied, 0));   // the snapshot is still being read out
    params_copy_kernel<<<p->n_chunks, 256, 0, st>>>(p->chunks, p->flat, 1);
    const int rc = launch_check("params_copy_kernel");
    if (rc != APEXB200_OK || !slot_host) return rc;
    APEX_CUDA(cudaEventRecord(p->packed, st));
    APEX_CUDA(cudaStreamWaitEvent(p->copy_stream, p->packed, 0));
    ctrl_host[1] = version;                                                   // "begun": readers of this slot must re-check
    APEX_CUDA(cudaMemcpyAsync(slot_host, p->flat, p->flat_bytes, cudaMemcpyDeviceToHost, p->copy_stream));
    PublishNode *node = &p->nodes[p->node_next++ % kPublishNodes];
    if (node->pending) APEX_CUDA(cudaStreamSynchronize(p->copy_stream));      // > kPublishNodes publications in flight
    node->ctrl = ctrl_host;
    node->version = version;
    node->pending = 1;
    APEX_CUDA(cudaLaunchHostFunc(p->copy_stream, publish_done, node));
    APEX_CUDA(cudaEventRecord(p->copied, p->copy_stream));
    p->copy_pending = true;
    return APEXB200_OK;
}

int apexb200_params_load(apexb200_params_t p, const void *flat_src, int32_t src_is_host, void *stream) {
    APEX_REQUIRE(p, APEXB200_EINVAL, "params_load: null handle");
    cudaStream_t st = (cudaStream_t)stream;
    if (flat_src && flat_src != p->flat) {
        if (p->copy_pending) APEX_CUDA(cudaStreamWaitEvent(st, p->copied, 0));
        APEX_CUDA(cudaMemcpyAsync(p->flat, flat_src, p->flat_bytes, src_is_host ? cudaMemcpyHostToDevice : cudaMemcpyDefault, st));
    }
    params_copy_kernel<<<p->n_chunks, 256, 0, st>>>(p->chunks, p->flat, 0);
    return launch_check("params_copy_kernel");
}

int apexb200_actor_policy(const float *value, const float *advantage_or_q, const float *eps, const double *u_explore,
                          const double *u_action, uint64_t seed, uint64_t step, int32_t E, int32_t A, float *q_out,
                          int32_t *actions, uint8_t *explored, void *stream) {
    APEX_REQUIRE(advantage_or_q && eps && actions, APEXB200_EINVAL, "actor_policy: null argument");
    APEX_REQUIRE(E >= 0 && A >= 1 && A <= APEXB200_MAX_ACTIONS, APEXB200_ERANGE, "actor_policy: E=%d A=%d out of range", E, A);
    APEX_REQUIRE((u_explore == nullptr) == (u_action == nullptr), APEXB200_EINVAL,
                 "actor_policy: give both uniform arrays or neither");
    if (E == 0) return APEXB200_OK;
    actor_policy_kernel<<<(E + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        value, advantage_or_q, eps, u_explore, u_action, seed, step, E, A, (float)(1.0 / (double)A), q_out, actions, explored);
    return launch_check("actor_policy_kernel");
}

}  // extern "C"
