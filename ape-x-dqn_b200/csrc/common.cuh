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

// Add `d` (mod 2^64) to every ancestor of leaf i.  Integer adds commute, so the
// tree is exact whatever order concurrent updates land in.
__device__ __forceinline__ void propagate_delta(const TreeView &tv, long long i, unsigned long long d) {
    #pragma unroll 1
    for (int l = 1; l < tv.n_levels; ++l) {
        i >>= 5;
        atomicAdd(tv.lvl[l] + i, d);
    }
}

}  // namespace apexb200

struct apexb200_replay {
    apexb200_replay_config cfg;
    apexb200::TreeView tree;
    apexb200::RingView ring;
    void *tree_block = nullptr;        // one allocation for all levels
    uint32_t *stamp = nullptr;         // [N] duplicate-resolution stamps (K2); all zero between calls
    int32_t *status = nullptr;         // device control block: [0] status word, [2] u32 ticket, [4..5] u64 draw counter
    std::vector<long long> lane_count; // host ring heads
    int64_t device_bytes = 0;
};
