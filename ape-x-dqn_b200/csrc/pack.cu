// Lossless packing of the frame store for checkpoints (the reference's README.md:24 TODO
// "observation compression"; the reference itself never persists its replay).  Not on the sampling path: the ring
// stays raw uint8 in HBM; this shrinks what state_dict() has to move to the host and to disk.
//
// Format, per frame of F bytes (F a multiple of 16, nb = F/16 blocks): the frame is XORed with its predecessor in
// slot order (consecutive env steps of one actor differ in a few sprites), except every kKeyEvery-th frame of the
// packed range, which is taken as is (a decoder can start at any key frame: segments decode in parallel).  Of the
// XOR image only the 16-byte blocks with a non-zero byte are kept:
//   [ bitmap: ceil(nb/32) little-endian u32 words, bit b set <=> block b is stored ][ the stored blocks, 16 B each ]
// Encoded size = 4*ceil(nb/32) + 16*popcount(bitmap).  Every frame starts at a caller-given 16-byte-aligned offset
// (exclusive prefix sum of the rounded sizes), so blocks are moved as 16-byte vectors.
#include "common.cuh"

namespace apexb200 {

constexpr int kKeyEvery = 64;

__device__ __forceinline__ bool block_nonzero(uint4 v) { return (v.x | v.y | v.z | v.w) != 0u; }

// One warp per frame.  sizes != null: only the encoded size (rounded up to 16) is written; else the frame is encoded
// at out + offsets[j].
__global__ void __launch_bounds__(128)
frames_pack_kernel(const uint8_t *__restrict__ frames, long long first_slot, int count, int F,
                   unsigned *__restrict__ sizes, const unsigned long long *__restrict__ offsets,
                   uint8_t *__restrict__ out) {
    const int j = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (j >= count) return;
    const int nb = F >> 4, nw = (nb + 31) >> 5;
    const uint4 *cur = (const uint4 *)(frames + (first_slot + j) * (long long)F);
    const uint4 *prev = (j % kKeyEvery) ? (const uint4 *)(frames + (first_slot + j - 1) * (long long)F) : nullptr;
    unsigned *bitmap = nullptr;
    uint4 *blocks = nullptr;
    if (sizes == nullptr) {
        uint8_t *dst = out + offsets[j];
        bitmap = (unsigned *)dst;
        const int head_words = (int)((((size_t)nw * 4 + 15) & ~(size_t)15) >> 2);
        blocks = (uint4 *)(dst + (size_t)head_words * 4);
        if (nw + lane < head_words) bitmap[nw + lane] = 0u;          // padding of the bitmap up to 16 bytes: defined bytes
    }
    int stored = 0;
    for (int w = 0; w < nw; ++w) {
        const int b = w * 32 + lane;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (b < nb) {
            v = cur[b];
            if (prev) { const uint4 p = prev[b]; v.x ^= p.x; v.y ^= p.y; v.z ^= p.z; v.w ^= p.w; }
        }
        const bool keep = b < nb && (prev == nullptr || block_nonzero(v));      // key frames keep every block
        const unsigned mask = __ballot_sync(0xffffffffu, keep);
        if (sizes == nullptr) {
            if (lane == 0) bitmap[w] = mask;
            if (keep) blocks[stored + __popc(mask & ((1u << lane) - 1u))] = v;
        }
        stored += __popc(mask);
    }
    if (sizes != nullptr && lane == 0)
        sizes[j] = (unsigned)((((size_t)nw * 4 + 15) & ~(size_t)15) + (size_t)stored * 16);
}

// One CTA per segment of kKeyEvery frames, one thread per 16-byte block column: the thread walks the segment's
// frames in order, XOR-accumulating its block (prefix XOR undoes the chain) and writing it into the ring.
__global__ void __launch_bounds__(512)
frames_unpack_kernel(uint8_t *__restrict__ frames, long long first_slot, int count, int F,
                     const unsigned long long *__restrict__ offsets, const uint8_t *__restrict__ packed) {
    const int nb = F >> 4, nw = (nb + 31) >> 5;
    const int seg = blockIdx.x;
    const int j0 = seg * kKeyEvery, j1 = min(count, j0 + kKeyEvery);
    for (int b = threadIdx.x; b < nb; b += blockDim.x) {
        uint4 acc = make_uint4(0u, 0u, 0u, 0u);
        for (int j = j0; j < j1; ++j) {
            const uint8_t *src = packed + offsets[j];
            const unsigned *bitmap = (const unsigned *)src;
            const uint4 *blocks = (const uint4 *)(src + (((size_t)nw * 4 + 15) & ~(size_t)15));
            const unsigned word = bitmap[b >> 5];
            if ((word >> (b & 31)) & 1u) {
                int rank = __popc(word & ((1u << (b & 31)) - 1u));
                for (int w = 0; w < (b >> 5); ++w) rank += __popc(bitmap[w]);
                const uint4 v = blocks[rank];
                if (j == j0) acc = v;                                            // key frame: the block itself
                else { acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w; }
            } else if (j == j0) {
                acc = make_uint4(0u, 0u, 0u, 0u);                                // (not produced by the encoder)
            }
            ((uint4 *)(frames + (first_slot + j) * (long long)F))[b] = acc;
        }
    }
}

}  // namespace apexb200

using namespace apexb200;

extern "C" {

static int pack_check(apexb200_replay_t h, int64_t first_slot, int32_t count, const char *who) {
    APEX_REQUIRE(h, APEXB200_EINVAL, "%s: null handle", who);
    const RingView &rv = h->ring;
    APEX_REQUIRE(count >= 0 && first_slot >= 0 && first_slot + count <= rv.N * rv.fpr, APEXB200_EINVAL,
                 "%s: frame slots [%lld, %lld) outside the store of %lld", who, (long long)first_slot,
                 (long long)(first_slot + count), (long long)(rv.N * rv.fpr));
    return APEXB200_OK;
}

int apexb200_frames_pack_sizes(apexb200_replay_t h, int64_t first_slot, int32_t count, uint32_t *sizes, void *stream) {
    int rc = pack_check(h, first_slot, count, "frames_pack_sizes");
    if (rc) return rc;
    APEX_REQUIRE(sizes || count == 0, APEXB200_EINVAL, "frames_pack_sizes: null output");
    if (count == 0) return APEXB200_OK;
    const long long threads = (long long)count * 32;
    frames_pack_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        h->ring.frames, first_slot, count, h->ring.F, sizes, nullptr, nullptr);
    return launch_check("frames_pack_kernel");
}

int apexb200_frames_pack(apexb200_replay_t h, int64_t first_slot, int32_t count, const uint64_t *offsets, uint8_t *out,
                         void *stream) {
    int rc = pack_check(h, first_slot, count, "frames_pack");
    if (rc) return rc;
    APEX_REQUIRE((offsets && out) || count == 0, APEXB200_EINVAL, "frames_pack: null argument");
    APEX_REQUIRE(((uintptr_t)out & 15) == 0, APEXB200_EINVAL, "frames_pack: output must be 16-byte aligned");
    if (count == 0) return APEXB200_OK;
    const long long threads = (long long)count * 32;
    frames_pack_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        h->ring.frames, first_slot, count, h->ring.F, nullptr, (const unsigned long long *)offsets, out);
    return launch_check("frames_pack_kernel");
}

int apexb200_frames_unpack(apexb200_replay_t h, int64_t first_slot, int32_t count, const uint64_t *offsets,
                           const uint8_t *packed, void *stream) {
    int rc = pack_check(h, first_slot, count, "frames_unpack");
    if (rc) return rc;
    APEX_REQUIRE((offsets && packed) || count == 0, APEXB200_EINVAL, "frames_unpack: null argument");
    APEX_REQUIRE(((uintptr_t)packed & 15) == 0, APEXB200_EINVAL, "frames_unpack: input must be 16-byte aligned");
    if (count == 0) return APEXB200_OK;
    const int segs = (count + kKeyEvery - 1) / kKeyEvery;
    const int nb = h->ring.F >> 4;
    frames_unpack_kernel<<<(unsigned)segs, nb < 512 ? ((nb + 31) / 32) * 32 : 512, 0, (cudaStream_t)stream>>>(
        h->ring.frames, first_slot, count, h->ring.F, (const unsigned long long *)offsets, packed);
    return launch_check("frames_unpack_kernel");
}

}  // extern "C"
