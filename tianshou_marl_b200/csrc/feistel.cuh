// feistel.cuh -- keyed pseudo-random permutation of [0, n) evaluated per element (no index array).
//
// Stands where the reference draws np.random.permutation(B) for the minibatch split
// (tianshou/data/batch.py:1209-1226).  MT19937 permutations cannot be reproduced on a GPU, so the
// throughput path uses an alternating unbalanced Feistel network over ceil(log2 n) bits with
// cycle walking (every round XORs one half with a hash of the other, hence a bijection; values
// >= n are walked until they land inside the domain).  oracle/philox.py restates it bit-for-bit.
// Parity mode passes explicit index arrays drawn by numpy instead.
#pragma once
#include <stdint.h>

namespace tsm {

__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x) {  // "lowbias32" integer hash
    x ^= x >> 16; x *= 0x7feb352dU;
    x ^= x >> 15; x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}

__host__ __device__ __forceinline__ int feistel_bits(uint64_t n) {
    int bits = 2;
    while (bits < 62 && (1ull << bits) < n) ++bits;
    return bits;
}

constexpr int FEISTEL_ROUNDS = 6;

__host__ __device__ __forceinline__ uint64_t feistel_permute(uint64_t x, uint64_t n, uint64_t key) {
    const int bits = feistel_bits(n);
    const int a = bits >> 1, b = bits - a;
    const uint32_t mask_a = (uint32_t)((1ull << a) - 1), mask_b = (uint32_t)((1ull << b) - 1);
    const uint32_t k0 = (uint32_t)key, k1 = (uint32_t)(key >> 32);
    do {
        uint32_t lo = (uint32_t)x & mask_a, hi = (uint32_t)(x >> a) & mask_b;
#pragma unroll
        for (int r = 0; r < FEISTEL_ROUNDS; ++r) {
            const uint32_t kr = mix32(k0 + 0x9E3779B9u * (uint32_t)(r + 1)) ^ k1;
            if (r & 1) lo ^= mix32(hi ^ kr) & mask_a;
            else hi ^= mix32(lo ^ kr) & mask_b;
        }
        x = ((uint64_t)hi << a) | lo;
    } while (x >= n);
    return x;
}

// The same permutation with everything that depends only on (n, key) prepared once on the host.
struct FeistelPre {
    uint32_t rk[FEISTEL_ROUNDS];
    uint32_t mask_a, mask_b;
    int a;
};
__host__ __device__ inline FeistelPre feistel_prepare(uint64_t n, uint64_t key) {
    FeistelPre p;
    const int bits = feistel_bits(n);
    p.a = bits >> 1;
    const int b = bits - p.a;
    p.mask_a = (uint32_t)((1ull << p.a) - 1);
    p.mask_b = (uint32_t)((1ull << b) - 1);
    const uint32_t k0 = (uint32_t)key, k1 = (uint32_t)(key >> 32);
    for (int r = 0; r < FEISTEL_ROUNDS; ++r) p.rk[r] = mix32(k0 + 0x9E3779B9u * (uint32_t)(r + 1)) ^ k1;
    return p;
}
__device__ __forceinline__ uint64_t feistel_permute(uint64_t x, uint64_t n, const FeistelPre& p) {
    do {
        uint32_t lo = (uint32_t)x & p.mask_a, hi = (uint32_t)(x >> p.a) & p.mask_b;
#pragma unroll
        for (int r = 0; r < FEISTEL_ROUNDS; ++r) {
            if (r & 1) lo ^= mix32(hi ^ p.rk[r]) & p.mask_a;
            else hi ^= mix32(lo ^ p.rk[r]) & p.mask_b;
        }
        x = ((uint64_t)hi << p.a) | lo;
    } while (x >= n);
    return x;
}

}  // namespace tsm
