/*
 * fair_baseline.c -- "B-fair": the strongest honest CPU competitor for the batched Filter, timed beside the GPU
 * path by bench.py.  TEST / MEASUREMENT INFRASTRUCTURE ONLY (same rules as gpushare_oracle.c: the product never
 * links or loads it).
 *
 * It takes the same decisions as Predicate.Handler over NodeInfo.Assume (pkg/scheduler/predicate.go:15-39,
 * pkg/cache/nodeinfo.go:113-137,254-284) but none of the reference's cost structure: flat arrays with integers
 * parsed once, no maps, no locks, no strings; the per-node maximum of (capacity - used) mod 2^64 is computed
 * once per batch (the rewrite `exists d: avail_d >= req  <=>  max_d avail_d >= req` that the GPU path uses too),
 * then every (pod, node) pair is one comparison, vectorised with AVX2 where the compiler targets it, rows spread
 * over all host threads with OpenMP.  Output: the same [P][row_stride_words] bitmap the GPU path writes.
 * It is checked against the numpy restatement in tests/test_fair_baseline.py.
 */
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#ifdef __AVX2__
#include <immintrin.h>
#endif

#define NODE_OK 0

int fair_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

int fair_has_avx2(void) {
#ifdef __AVX2__
    return 1;
#else
    return 0;
#endif
}

/* key[n] = max over the node's devices of (cap[n] - used[n][d]) mod 2^64; eligible[n] = class OK and >= 1 device */
static void node_keys(const uint64_t* cap, const uint64_t* used, const uint32_t* ndev, const uint8_t* cls, size_t n_nodes,
                      size_t dev_stride, uint64_t* key, uint32_t* key32, size_t padded) {
#pragma omp parallel for schedule(static)
    for (size_t n = 0; n < padded; ++n) {
        uint64_t k = 0;
        int ok = 0;
        if (n < n_nodes && cls[n] == NODE_OK && ndev[n] > 0) {
            ok = 1;
            const uint64_t c = cap[n];
            const uint64_t* row = used + n * dev_stride;
            for (uint32_t d = 0; d < ndev[n]; ++d) {
                const uint64_t a = c - row[d];
                if (a > k) k = a;
            }
        }
        /* an ineligible node must fail even a zero request: give it the key "-1" in the shifted domains below */
        key[n] = ok ? k : 0;
        key32[n] = ok ? (k > 0xfffffffeull ? 0xfffffffeu : (uint32_t)k) + 1u : 0u; /* 0 = never; k+1 saturating */
    }
}

/* one row: bit n = eligible[n] && key[n] >= req.  32-bit path: key32 holds min(key, 2^32-2) + 1 (0 = ineligible) and
 * the request is compared as req + 1, exact for every request below 2^32 - 1. */
static void row32(const uint32_t* key32, size_t words, uint32_t r1, uint32_t* out) {
#ifdef __AVX2__
    const __m256i bias = _mm256_set1_epi32((int)0x80000000u);
    const __m256i rv = _mm256_xor_si256(_mm256_set1_epi32((int)r1), bias);
    for (size_t w = 0; w < words; ++w) {
        uint32_t bits = 0;
        for (int q = 0; q < 4; ++q) {
            const __m256i k = _mm256_xor_si256(_mm256_loadu_si256((const __m256i*)(key32 + w * 32 + q * 8)), bias);
            /* key >= r  <=>  !(r > key) */
            const __m256i lt = _mm256_cmpgt_epi32(rv, k);
            bits |= ((uint32_t)(~_mm256_movemask_ps(_mm256_castsi256_ps(lt))) & 0xffu) << (q * 8);
        }
        out[w] = bits;
    }
#else
    for (size_t w = 0; w < words; ++w) {
        uint32_t bits = 0;
        for (int b = 0; b < 32; ++b) bits |= (uint32_t)(key32[w * 32 + b] >= r1) << b;
        out[w] = bits;
    }
#endif
}

static void row64(const uint64_t* key, const uint32_t* key32, size_t words, uint64_t r, uint32_t* out) {
    for (size_t w = 0; w < words; ++w) {
        uint32_t bits = 0;
        for (int b = 0; b < 32; ++b) bits |= (uint32_t)(key32[w * 32 + b] != 0 && key[w * 32 + b] >= r) << b;
        out[w] = bits;
    }
}

/* returns 0, or -1 when memory runs out.  Padding words [ceil(n/32), row_stride_words) are written as zero. */
int fair_filter(const uint64_t* cap, const uint64_t* used, const uint32_t* ndev, const uint8_t* cls, size_t n_nodes,
                size_t dev_stride, const uint64_t* req, size_t n_pods, uint32_t* out_bitmap, size_t row_stride_words) {
    const size_t words = (n_nodes + 31) / 32, padded = words * 32;
    uint64_t* key = (uint64_t*)malloc((padded ? padded : 1) * sizeof(uint64_t));
    uint32_t* key32 = (uint32_t*)malloc((padded ? padded : 1) * sizeof(uint32_t));
    if (!key || !key32) {
        free(key);
        free(key32);
        return -1;
    }
    node_keys(cap, used, ndev, cls, n_nodes, dev_stride, key, key32, padded);
#pragma omp parallel for schedule(dynamic, 16)
    for (size_t p = 0; p < n_pods; ++p) {
        uint32_t* out = out_bitmap + p * row_stride_words;
        const uint64_t r = req[p];
        if (r < 0xffffffffull)
            row32(key32, words, (uint32_t)r + 1u, out);
        else
            row64(key, key32, words, r, out);
        for (size_t w = words; w < row_stride_words; ++w) out[w] = 0;
    }
    free(key);
    free(key32);
    return 0;
}
