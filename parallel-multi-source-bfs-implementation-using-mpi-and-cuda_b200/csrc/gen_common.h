/*
 * gen_common.h — deterministic synthetic-input recipes shared by host (gcc) and
 * device (nvcc) code.  Everything is integer arithmetic on a counter-based hash,
 * so the GPU generators (gen.cu) and the CPU generators used by tests produce
 * byte-identical edge records and source lists for the same (kind, seed).
 *
 * The reference ships no generator and no sample data; these
 * recipes realise the five BASELINE.json configs.  Record layout is the
 * reference's graph.bin payload: (int32 u, int32 v) per undirected edge record,
 * duplicates and self-loops kept (/root/reference/main.cu:108-116).
 */
#ifndef MSBFS_GEN_COMMON_H
#define MSBFS_GEN_COMMON_H

#include <stdint.h>

#ifdef __CUDACC__
#define MSBFS_HD __host__ __device__ __forceinline__
#else
#define MSBFS_HD static inline
#endif

enum msbfs_graph_kind {
    MSBFS_GEN_RMAT    = 0, /* R-MAT (.57,.19,.19,.05), labels NOT permuted      */
    MSBFS_GEN_KRON    = 1, /* same recursion, vertex labels scrambled (Graph500) */
    MSBFS_GEN_LATTICE = 2, /* 2-D 4-neighbour grid, id = y*W + x                 */
    MSBFS_GEN_ER      = 3  /* Erdos-Renyi G(n, m): m uniform random pairs        */
};

/* splitmix64 finaliser */
MSBFS_HD uint64_t msbfs_mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ULL;
    x ^= x >> 27; x *= 0x94D049BB133111EBULL;
    x ^= x >> 31;
    return x;
}

/* counter-based random word: independent per (seed, stream, idx) */
MSBFS_HD uint64_t msbfs_rand64(uint64_t seed, uint64_t stream, uint64_t idx) {
    uint64_t k = msbfs_mix64(seed + 0x9E3779B97F4A7C15ULL * (stream + 1));
    return msbfs_mix64(k ^ (idx * 0xD1342543DE82EF95ULL + 0x2545F4914F6CDD1DULL));
}

/* bijection on [0, 2^scale) used to scramble Kronecker vertex labels */
MSBFS_HD uint32_t msbfs_scramble(uint32_t x, int scale, uint64_t seed) {
    const uint32_t mask = (scale >= 32) ? 0xFFFFFFFFu : ((1u << scale) - 1u);
    const uint32_t k0 = (uint32_t)msbfs_mix64(seed ^ 0xA5A5A5A5ULL) | 1u;
    const uint32_t k1 = (uint32_t)msbfs_mix64(seed ^ 0x5A5A5A5AULL) | 1u;
    const uint32_t c0 = (uint32_t)(msbfs_mix64(seed ^ 0x1234ULL) >> 17);
    const int sh = (scale + 1) / 2;
    x = (x * k0 + c0) & mask;   /* odd multiplier: invertible mod 2^scale */
    x ^= x >> sh;               /* xor-shift right: invertible            */
    x = (x * k1) & mask;
    x ^= x >> sh;
    return x & mask;
}

/* R-MAT quadrant thresholds in 2^32 units: a = .57, a+b = .76, a+b+c = .95 */
#define MSBFS_RMAT_T0 2448131358u /* floor(0.57 * 2^32) */
#define MSBFS_RMAT_T1 3264175144u /* floor(0.76 * 2^32) */
#define MSBFS_RMAT_T2 4080218931u /* floor(0.95 * 2^32) */

MSBFS_HD void msbfs_gen_rmat_edge(uint64_t seed, int scale, uint64_t i, int scrambled,
                                  int32_t* u_out, int32_t* v_out) {
    uint32_t u = 0, v = 0;
    for (int lvl = 0; lvl < scale; lvl += 2) {
        /* one hash word feeds two recursion levels */
        uint64_t h = msbfs_rand64(seed, (uint64_t)(16 + (lvl >> 1)), i);
        uint32_t r0 = (uint32_t)(h >> 32), r1 = (uint32_t)h;
        u = (u << 1) | (r0 >= MSBFS_RMAT_T1);
        v = (v << 1) | ((r0 >= MSBFS_RMAT_T0 && r0 < MSBFS_RMAT_T1) || r0 >= MSBFS_RMAT_T2);
        if (lvl + 1 < scale) {
            u = (u << 1) | (r1 >= MSBFS_RMAT_T1);
            v = (v << 1) | ((r1 >= MSBFS_RMAT_T0 && r1 < MSBFS_RMAT_T1) || r1 >= MSBFS_RMAT_T2);
        }
    }
    if (scrambled) {
        u = msbfs_scramble(u, scale, seed);
        v = msbfs_scramble(v, scale, seed);
    }
    *u_out = (int32_t)u;
    *v_out = (int32_t)v;
}

MSBFS_HD void msbfs_gen_er_edge(uint64_t seed, uint32_t n, uint64_t i,
                                int32_t* u_out, int32_t* v_out) {
    *u_out = (int32_t)(msbfs_rand64(seed, 3, i) % n);
    *v_out = (int32_t)(msbfs_rand64(seed, 4, i) % n);
}

/* W x H grid: first H*(W-1) horizontal records, then (H-1)*W vertical ones */
MSBFS_HD uint64_t msbfs_lattice_num_edges(uint32_t w, uint32_t h) {
    return (uint64_t)h * (w - 1) + (uint64_t)(h - 1) * w;
}
MSBFS_HD void msbfs_gen_lattice_edge(uint32_t w, uint32_t h, uint64_t i,
                                     int32_t* u_out, int32_t* v_out) {
    const uint64_t nh = (uint64_t)h * (w - 1);
    if (i < nh) {
        uint32_t y = (uint32_t)(i / (w - 1)), x = (uint32_t)(i % (w - 1));
        *u_out = (int32_t)(y * w + x);
        *v_out = (int32_t)(y * w + x + 1);
    } else {
        uint64_t j = i - nh;
        uint32_t y = (uint32_t)(j / w), x = (uint32_t)(j % w);
        *u_out = (int32_t)(y * w + x);
        *v_out = (int32_t)((y + 1) * w + x);
    }
}

/* k-th source of query group g: uniform over [0, n) (isolated vertices allowed) */
MSBFS_HD int32_t msbfs_gen_source(uint64_t seed, uint32_t n, uint64_t flat_index) {
    return (int32_t)(msbfs_rand64(seed, 9, flat_index) % n);
}

#endif /* MSBFS_GEN_COMMON_H */
