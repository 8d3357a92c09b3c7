/*
 * CPU restatement (plain C + OpenMP) of one learner-side replay step.
 * TEST INFRASTRUCTURE / timed CPU baseline only — see oracle/__init__.py.
 *
 * It follows the reference's algorithm, not a better one:
 *   - sampling: replay.py:52 calls np.random.choice(p=...), i.e. a full cumulative
 *     sum of the probabilities followed by a right-sided binary search per draw.
 *     Here the cumulative sum over all N leaves is rebuilt once per batch (the
 *     reference rebuilds it once per DRAW, replay.py:52-53) on exact u64 leaves;
 *   - batch assembly: replay.py:54-56 + learner.py:35-38 (copy S_t / S_tpn rows);
 *   - loss / TD / priority: learner.py:43-52;
 *   - priority write-back: replay.py:40 (last writer wins), leaf = |delta|^alpha.
 * Uses every thread OpenMP gives it (OMP_NUM_THREADS / omp_set_num_threads).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define LEAF_ONE 16777216.0
#define LEAF_MAX ((1ull << 36) - 1)

int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

static uint64_t leaf_from_prio(float p) {
    if (!(p > 0.0f)) return 0;
    double s = (double)p * LEAF_ONE;
    if (!(s < (double)LEAF_MAX)) return LEAF_MAX;
    return (uint64_t)s;
}

/* inclusive prefix sum of leaves into cum (two-pass block scan when threaded) */
void oracle_cumsum(const uint64_t *leaves, uint64_t *cum, int64_t n) {
    int nt = oracle_max_threads();
    if (nt <= 1 || n < (1 << 16)) {
        uint64_t acc = 0;
        for (int64_t i = 0; i < n; ++i) { acc += leaves[i]; cum[i] = acc; }
        return;
    }
    uint64_t *part = (uint64_t *)calloc((size_t)nt + 1, sizeof(uint64_t));
#pragma omp parallel num_threads(nt)
    {
        int t = omp_get_thread_num();
        int64_t lo = n * t / nt, hi = n * (t + 1) / nt;
        uint64_t acc = 0;
        for (int64_t i = lo; i < hi; ++i) { acc += leaves[i]; cum[i] = acc; }
        part[t + 1] = acc;
#pragma omp barrier
#pragma omp single
        for (int k = 1; k <= nt; ++k) part[k] += part[k - 1];
        uint64_t off = part[t];
        if (off) for (int64_t i = lo; i < hi; ++i) cum[i] += off;
    }
    free(part);
}

/* searchsorted(cum, target, side='right') */
static int64_t upper_bound(const uint64_t *cum, int64_t n, uint64_t t) {
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        int64_t mid = lo + ((hi - lo) >> 1);
        if (cum[mid] <= t) lo = mid + 1; else hi = mid;
    }
    return lo;
}

void oracle_sample(const uint64_t *cum, int64_t n, const double *u, int B, int32_t *idx) {
    uint64_t total = cum[n - 1];
#pragma omp parallel for schedule(static)
    for (int j = 0; j < B; ++j) {
        uint64_t t = (uint64_t)(u[j] * (double)total);
        if (t >= total) t = total - 1;
        idx[j] = (int32_t)upper_bound(cum, n, t);
    }
}

void oracle_gather(const uint8_t *frames, const int32_t *fidx, const int32_t *mA, const float *mR, const float *mG,
                   int C, int F, const int32_t *idx, int B, uint8_t *S_t, uint8_t *S_tpn, int32_t *A, float *R, float *G) {
    int twoC = 2 * C;
#pragma omp parallel for schedule(static)
    for (int it = 0; it < B * twoC; ++it) {
        int b = it / twoC, k = it % twoC;
        int64_t f = fidx[(int64_t)idx[b] * twoC + k];
        uint8_t *dst = (k < C ? S_t : S_tpn) + ((int64_t)b * C + (k < C ? k : k - C)) * F;
        memcpy(dst, frames + f * (int64_t)F, (size_t)F);
    }
    for (int b = 0; b < B; ++b) { A[b] = mA[idx[b]]; R[b] = mR[idx[b]]; G[b] = mG[idx[b]]; }
}

/* learner.py:43-52; returns loss mean */
float oracle_td(const float *q_on_t, const float *q_on_tpn, const float *q_tg_tpn, const int32_t *A, const float *R,
                const float *G, int B, int nA, float alpha, int huber, float kappa, float *delta, float *prio,
                float *loss) {
    double acc = 0.0;
    for (int b = 0; b < B; ++b) {
        const float *rn = q_on_tpn + (int64_t)b * nA;
        int arg = 0;
        for (int k = 1; k < nA; ++k) if (rn[k] > rn[arg]) arg = k;
        float prod = G[b] * q_tg_tpn[(int64_t)b * nA + arg];
        float target = R[b] + prod;
        float d = target - q_on_t[(int64_t)b * nA + A[b]];
        float ad = fabsf(d), l;
        if (!huber || ad <= kappa) { float h = 0.5f * d; l = h * d; }
        else l = kappa * (ad - 0.5f * kappa);
        delta[b] = d; prio[b] = powf(ad, alpha); loss[b] = l;
        acc += (double)l;
    }
    return (float)(acc / B);
}

void oracle_update(uint64_t *leaves, const int32_t *idx, const float *prio, int B) {
    for (int j = 0; j < B; ++j) leaves[idx[j]] = leaf_from_prio(prio[j]);     /* sequential: last writer wins */
}

/* One full step; cum is caller-provided scratch of n entries. */
float oracle_step(uint64_t *leaves, uint64_t *cum, int64_t n, const double *u, int B, const uint8_t *frames,
                  const int32_t *fidx, const int32_t *mA, const float *mR, const float *mG, int C, int F,
                  const float *q_on_t, const float *q_on_tpn, const float *q_tg_tpn, int nA, float alpha,
                  int32_t *idx, uint8_t *S_t, uint8_t *S_tpn, int32_t *A, float *R, float *G, float *delta,
                  float *prio, float *loss) {
    oracle_cumsum(leaves, cum, n);
    oracle_sample(cum, n, u, B, idx);
    oracle_gather(frames, fidx, mA, mR, mG, C, F, idx, B, S_t, S_tpn, A, R, G);
    float m = oracle_td(q_on_t, q_on_tpn, q_tg_tpn, A, R, G, B, nA, alpha, 0, 1.0f, delta, prio, loss);
    oracle_update(leaves, idx, prio, B);
    return m;
}
