// K4: gather of the sampled stacked frames (+ n-step metadata) straight into the
// learner's batch tensors.  HBM-bound byte movement: 2C frames of `F` bytes per
// record are read from random ring slots and written to dense [B, C, F] outputs.
#include "common.cuh"

namespace apexb200 {

static int g_gather_variant = 0;

__device__ __forceinline__ uint4 ld_stream(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// Resolve work item -> (source frame slot, destination pointer).  Item it in
// [0, count*2C): row b = it / 2C, k = it % 2C; k < C is frame k of S_t, else of S_tpn.
struct GatherArgs {
    const int32_t *idx, *fsel, *count_dev;
    uint8_t *S_t, *S_tpn;
    int32_t *A;
    float *R, *G, *q_t, *q_tpn;
    int count;
};

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
    uint8_t *base = k < rv.C ? ga.S_t : ga.S_tpn;
    int kk = k < rv.C ? k : k - rv.C;
    *dst = base + (b * rv.C + kk) * (long long)rv.F;
    return true;
}

__device__ __forceinline__ void gather_meta(const RingView &rv, const GatherArgs &ga, int b, int count, int32_t *status) {
    if (b >= count) return;
    long long slot = ga.idx[b];
    if (slot < 0 || slot >= rv.N) { atomicOr(status, APEXB200_ST_BAD_INDEX); return; }
    if (ga.A) ga.A[b] = rv.A[slot];
    if (ga.R) ga.R[b] = rv.R[slot];
    if (ga.G) ga.G[b] = rv.G[slot];
    if (ga.q_t && rv.q_t)
        for (int a = 0; a < rv.nA; ++a) ga.q_t[(long long)b * rv.nA + a] = rv.q_t[slot * rv.nA + a];
    if (ga.q_tpn && rv.q_tpn)
        for (int a = 0; a < rv.nA; ++a) ga.q_tpn[(long long)b * rv.nA + a] = rv.q_tpn[slot * rv.nA + a];
}

// ---- variant 0: one warp per frame, 16-byte vectors, all loads of a frame in
// flight before the first store (UNROLL x 512 B per warp per round).
template <int UNROLL>
__global__ void __launch_bounds__(256)
gather_vec_kernel(RingView rv, GatherArgs ga, int32_t *status) {
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

// ---- variant 1: TMA bulk copies staged through shared memory.  One lane owns one
// frame: cp.async.bulk global->shared completes on an mbarrier, then
// cp.async.bulk shared->global drains it; the bytes never enter registers.
constexpr int kTmaItemsPerCta = 4;

__global__ void __launch_bounds__(32)
gather_tma_kernel(RingView rv, GatherArgs ga, int32_t *status) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) unsigned long long bars[kTmaItemsPerCta];
    const int lane = threadIdx.x;
    int count = ga.count;
    if (ga.count_dev) count = min(count, *ga.count_dev);
    const long long n_items = (long long)count * 2 * rv.C;
    const long long first = (long long)blockIdx.x * kTmaItemsPerCta;
    if (first >= (long long)ga.count * 2 * rv.C) {          // trailing CTAs: metadata
        long long mb = blockIdx.x - ((long long)ga.count * 2 * rv.C + kTmaItemsPerCta - 1) / kTmaItemsPerCta;
        gather_meta(rv, ga, (int)(mb * 32 + lane), count, status);
        return;
    }
    if (lane < kTmaItemsPerCta) {
        const long long it = first + lane;
        const uint8_t *src; uint8_t *dst;
        if (it < n_items && resolve_item(rv, ga, it, count, &src, &dst, status)) {
            const unsigned bar = (unsigned)__cvta_generic_to_shared(&bars[lane]);
            const unsigned buf = (unsigned)__cvta_generic_to_shared(smem + (size_t)lane * rv.F);
            const unsigned bytes = (unsigned)rv.F;
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(buf), "l"(src), "r"(bytes), "r"(bar) : "memory");
            unsigned done = 0;
            while (!done) {
                asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n"
                             " selp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(bar) : "memory");
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(buf), "r"(bytes)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
    }
}

}  // namespace apexb200

using namespace apexb200;

extern "C" {

int apexb200_set_gather_variant(int32_t variant) {
    APEX_REQUIRE(variant == 0 || variant == 1, APEXB200_EINVAL, "set_gather_variant: unknown variant %d", variant);
    g_gather_variant = variant;
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
    GatherArgs ga{idx, fsel, count_dev, S_t, S_tpn, A, R, Gamma, q_t, q_tpn, count};
    cudaStream_t st = (cudaStream_t)stream;
    const long long items = (long long)count * 2 * rv.C;
    const long long meta_warps = (count + 31) / 32;
    if (g_gather_variant == 1) {
        const size_t smem = (size_t)kTmaItemsPerCta * rv.F;
        APEX_REQUIRE(smem <= 200 * 1024, APEXB200_ERANGE, "gather: frame too large for the TMA variant");
        static size_t configured = 0;
        if (smem > 48 * 1024 && smem > configured) {
            APEX_CUDA(cudaFuncSetAttribute(gather_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            configured = smem;
        }
        const long long blocks = (items + kTmaItemsPerCta - 1) / kTmaItemsPerCta + meta_warps;
        gather_tma_kernel<<<(unsigned)blocks, 32, smem, st>>>(rv, ga, h->status);
        return launch_check("gather_tma_kernel");
    }
    const long long warps = items + meta_warps;
    const long long blocks = (warps + 7) / 8;
    gather_vec_kernel<14><<<(unsigned)blocks, 256, 0, st>>>(rv, ga, h->status);
    return launch_check("gather_vec_kernel");
}

}  // extern "C"
