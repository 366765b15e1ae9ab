/*
 * host_util.cpp — host-only half of libmsbfs.so: error reporting, the query /
 * graph file formats of the reference (/root/reference/main.cu:101-116, 143-161),
 * the argmin rule (main.cu:379-397) and its packed form, and the CPU side of the
 * synthetic-input recipes.  Nothing here touches the GPU, so these entry points
 * work (and are tested) on a box without one.
 */
#include "internal.h"
#include "gen_common.h"

#include <cerrno>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

static thread_local std::string g_last_error;

int msbfs_fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}
void msbfs_clear_error() { g_last_error.clear(); }

extern "C" const char* msbfs_last_error(void) { return g_last_error.c_str(); }
extern "C" void msbfs_free(void* p) { std::free(p); }

/* ---------------------------------------------------------- query files --- */

static const char kExtMagic[7] = {'M', 'S', 'B', 'F', 'S', 'Q', '2'};

extern "C" int msbfs_load_query_file(const char* path, int32_t* num_groups_out,
                                     int64_t** group_offsets_out, int32_t** sources_out) {
    if (!path || !num_groups_out || !group_offsets_out || !sources_out)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_load_query_file: null argument");
    FILE* f = std::fopen(path, "rb");
    if (!f) return msbfs_fail(MSBFS_ERR_IO, "Could not open query file %s", path);
    std::fseek(f, 0, SEEK_END);
    long fsize = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<unsigned char> bytes((size_t)(fsize > 0 ? fsize : 0));
    if (fsize > 0 && std::fread(bytes.data(), 1, bytes.size(), f) != bytes.size()) {
        std::fclose(f);
        return msbfs_fail(MSBFS_ERR_IO, "short read on query file %s", path);
    }
    std::fclose(f);
    if (bytes.empty()) return msbfs_fail(MSBFS_ERR_IO, "query file %s is empty", path);

    size_t pos = 1;
    uint64_t K = bytes[0];
    bool extended = false;
    if (K == 0 && bytes.size() >= 12 && std::memcmp(&bytes[1], kExtMagic, 7) == 0) {
        uint32_t k32;
        std::memcpy(&k32, &bytes[8], 4);
        K = k32; pos = 12; extended = true;
    }
    int64_t* offs = (int64_t*)std::calloc((size_t)K + 1, sizeof(int64_t));
    if (!offs) return msbfs_fail(MSBFS_ERR_NOMEM, "out of host memory");
    /* first pass: sizes */
    size_t p = pos;
    int64_t total = 0;
    for (uint64_t i = 0; i < K; ++i) {
        uint64_t S = 0;
        if (extended) {
            if (p + 4 > bytes.size()) { std::free(offs); return msbfs_fail(MSBFS_ERR_IO, "query file %s is truncated", path); }
            uint32_t s32; std::memcpy(&s32, &bytes[p], 4); S = s32; p += 4;
        } else {
            if (p + 1 > bytes.size()) { std::free(offs); return msbfs_fail(MSBFS_ERR_IO, "query file %s is truncated", path); }
            S = bytes[p]; p += 1;
        }
        if (p + 4 * S > bytes.size()) { std::free(offs); return msbfs_fail(MSBFS_ERR_IO, "query file %s is truncated", path); }
        p += 4 * S;
        total += (int64_t)S;
        offs[i + 1] = total;
    }
    int32_t* src = (int32_t*)std::malloc(sizeof(int32_t) * (size_t)(total > 0 ? total : 1));
    if (!src) { std::free(offs); return msbfs_fail(MSBFS_ERR_NOMEM, "out of host memory"); }
    p = pos;
    for (uint64_t i = 0; i < K; ++i) {
        size_t S = (size_t)(offs[i + 1] - offs[i]);
        p += extended ? 4 : 1;
        std::memcpy(src + offs[i], &bytes[p], 4 * S);
        p += 4 * S;
    }
    *num_groups_out = (int32_t)K;
    *group_offsets_out = offs;
    *sources_out = src;
    return MSBFS_OK;
}

extern "C" int msbfs_write_query_file(const char* path, int32_t num_groups, const int64_t* group_offsets,
                                      const int32_t* sources, int force_extended) {
    if (!path || num_groups < 0 || (num_groups > 0 && (!group_offsets || !sources)))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_write_query_file: bad argument");
    bool extended = force_extended != 0 || num_groups > 255;
    for (int32_t i = 0; i < num_groups && !extended; ++i)
        if (group_offsets[i + 1] - group_offsets[i] > 255) extended = true;
    /* K = 0 in the reference format is the single byte 0x00 */
    FILE* f = std::fopen(path, "wb");
    if (!f) return msbfs_fail(MSBFS_ERR_IO, "Could not open query file %s for writing", path);
    if (extended) {
        unsigned char zero = 0; uint32_t k32 = (uint32_t)num_groups;
        std::fwrite(&zero, 1, 1, f); std::fwrite(kExtMagic, 1, 7, f); std::fwrite(&k32, 4, 1, f);
    } else {
        unsigned char k8 = (unsigned char)num_groups;
        std::fwrite(&k8, 1, 1, f);
    }
    for (int32_t i = 0; i < num_groups; ++i) {
        int64_t S = group_offsets[i + 1] - group_offsets[i];
        if (extended) { uint32_t s32 = (uint32_t)S; std::fwrite(&s32, 4, 1, f); }
        else { unsigned char s8 = (unsigned char)S; std::fwrite(&s8, 1, 1, f); }
        if (S > 0) std::fwrite(sources + group_offsets[i], 4, (size_t)S, f);
    }
    if (std::fclose(f) != 0) return msbfs_fail(MSBFS_ERR_IO, "write error on %s", path);
    return MSBFS_OK;
}

extern "C" int msbfs_write_graph_file(const char* path, int32_t n, int64_t m, const int32_t* uv_pairs) {
    if (!path || n < 0 || m < 0 || (m > 0 && !uv_pairs))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_write_graph_file: bad argument");
    FILE* f = std::fopen(path, "wb");
    if (!f) return msbfs_fail(MSBFS_ERR_IO, "Could not open graph file %s for writing", path);
    std::fwrite(&n, 4, 1, f);
    std::fwrite(&m, 8, 1, f);
    if (m > 0 && std::fwrite(uv_pairs, 8, (size_t)m, f) != (size_t)m) {
        std::fclose(f);
        return msbfs_fail(MSBFS_ERR_IO, "write error on %s", path);
    }
    if (std::fclose(f) != 0) return msbfs_fail(MSBFS_ERR_IO, "write error on %s", path);
    return MSBFS_OK;
}

/* ---------------------------------------------------------------- argmin --- */

extern "C" int msbfs_argmin(int32_t num_groups, const int64_t* sums, int32_t* min_k_out, int64_t* min_f_out) {
    if (num_groups < 0 || (num_groups > 0 && !sums) || !min_k_out || !min_f_out)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_argmin: bad argument");
    int64_t minF = -1; int32_t minK = -1;
    for (int32_t i = 0; i < num_groups; ++i)
        if (sums[i] >= 0 && (minK < 0 || sums[i] < minF)) { minF = sums[i]; minK = i; }
    *min_k_out = minK; *min_f_out = minF;
    return MSBFS_OK;
}

static const int kIdBits = 24;

extern "C" int msbfs_argmin_packed(const int64_t* sums, int32_t first_qid, int32_t count, uint64_t* packed_out) {
    if (count < 0 || first_qid < 0 || (count > 0 && !sums) || !packed_out)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_argmin_packed: bad argument");
    uint64_t best = UINT64_MAX;
    if ((int64_t)first_qid + count > ((int64_t)1 << kIdBits))
        return msbfs_fail(MSBFS_ERR_RANGE, "query id %lld does not fit %d bits",
                          (long long)first_qid + count - 1, kIdBits);
    for (int32_t i = 0; i < count; ++i) {
        if (sums[i] < 0) continue; /* never produced by the engine; skipped like main.cu:384 */
        if (sums[i] >= ((int64_t)1 << (63 - kIdBits)))
            return msbfs_fail(MSBFS_ERR_RANGE, "distance sum %lld does not fit the packed word", (long long)sums[i]);
        uint64_t w = ((uint64_t)sums[i] << kIdBits) | (uint64_t)(first_qid + i);
        if (w < best) best = w;
    }
    *packed_out = best;
    return MSBFS_OK;
}

extern "C" int msbfs_argmin_packed_ids(const int64_t* sums, const int32_t* query_ids, int32_t count, uint64_t* packed_out) {
    if (count < 0 || (count > 0 && (!sums || !query_ids)) || !packed_out)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_argmin_packed_ids: bad argument");
    uint64_t best = UINT64_MAX;
    for (int32_t i = 0; i < count; ++i) {
        if (sums[i] < 0) continue;
        if (query_ids[i] < 0 || query_ids[i] >= ((int32_t)1 << kIdBits))
            return msbfs_fail(MSBFS_ERR_RANGE, "query id %d does not fit %d bits", query_ids[i], kIdBits);
        if (sums[i] >= ((int64_t)1 << (63 - kIdBits)))
            return msbfs_fail(MSBFS_ERR_RANGE, "distance sum %lld does not fit the packed word", (long long)sums[i]);
        const uint64_t w = ((uint64_t)sums[i] << kIdBits) | (uint64_t)query_ids[i];
        if (w < best) best = w;
    }
    *packed_out = best;
    return MSBFS_OK;
}

extern "C" int msbfs_unpack_argmin(uint64_t packed, int32_t* min_k_out, int64_t* min_f_out) {
    if (!min_k_out || !min_f_out) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_unpack_argmin: null argument");
    if (packed == UINT64_MAX) { *min_k_out = -1; *min_f_out = -1; return MSBFS_OK; } /* K = 0, main.cu:379-380 */
    *min_k_out = (int32_t)(packed & (((uint64_t)1 << kIdBits) - 1));
    *min_f_out = (int64_t)(packed >> kIdBits);
    return MSBFS_OK;
}

/* ------------------------------------------------------------ generators --- */

extern "C" int msbfs_gen_num_vertices(int kind, int32_t p0, int32_t p1, int32_t* n_out) {
    if (!n_out) return msbfs_fail(MSBFS_ERR_INVALID, "null argument");
    switch (kind) {
        case MSBFS_GEN_RMAT: case MSBFS_GEN_KRON:
            if (p0 < 1 || p0 > 30) return msbfs_fail(MSBFS_ERR_INVALID, "scale %d out of range [1,30]", p0);
            *n_out = (int32_t)1 << p0; return MSBFS_OK;
        case MSBFS_GEN_LATTICE:
            if (p0 < 2 || p1 < 2 || (int64_t)p0 * p1 > INT32_MAX) return msbfs_fail(MSBFS_ERR_INVALID, "bad lattice size");
            *n_out = p0 * p1; return MSBFS_OK;
        case MSBFS_GEN_ER:
            if (p0 < 1) return msbfs_fail(MSBFS_ERR_INVALID, "bad vertex count");
            *n_out = p0; return MSBFS_OK;
        default: return msbfs_fail(MSBFS_ERR_INVALID, "unknown graph kind %d", kind);
    }
}

extern "C" int msbfs_gen_edges_host(int kind, int32_t p0, int32_t p1, int64_t m, uint64_t seed, int32_t* uv_out) {
    int32_t n;
    int rc = msbfs_gen_num_vertices(kind, p0, p1, &n);
    if (rc) return rc;
    if (m < 0 || (m > 0 && !uv_out)) return msbfs_fail(MSBFS_ERR_INVALID, "bad edge buffer");
    if (kind == MSBFS_GEN_LATTICE && (uint64_t)m != msbfs_lattice_num_edges((uint32_t)p0, (uint32_t)p1))
        return msbfs_fail(MSBFS_ERR_INVALID, "lattice %dx%d has %llu records, not %lld", p0, p1,
                          (unsigned long long)msbfs_lattice_num_edges((uint32_t)p0, (uint32_t)p1), (long long)m);
    for (int64_t i = 0; i < m; ++i) {
        int32_t u, v;
        switch (kind) {
            case MSBFS_GEN_RMAT: msbfs_gen_rmat_edge(seed, p0, (uint64_t)i, 0, &u, &v); break;
            case MSBFS_GEN_KRON: msbfs_gen_rmat_edge(seed, p0, (uint64_t)i, 1, &u, &v); break;
            case MSBFS_GEN_LATTICE: msbfs_gen_lattice_edge((uint32_t)p0, (uint32_t)p1, (uint64_t)i, &u, &v); break;
            default: msbfs_gen_er_edge(seed, (uint32_t)n, (uint64_t)i, &u, &v); break;
        }
        uv_out[2 * i] = u; uv_out[2 * i + 1] = v;
    }
    return MSBFS_OK;
}

extern "C" int msbfs_gen_queries_host(int32_t n, int32_t num_groups, int32_t sources_per_group, uint64_t seed,
                                      int64_t* group_offsets_out, int32_t* sources_out) {
    if (n < 1 || num_groups < 0 || sources_per_group < 0 || !group_offsets_out || (!sources_out && num_groups * (int64_t)sources_per_group > 0))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_gen_queries_host: bad argument");
    group_offsets_out[0] = 0;
    for (int32_t g = 0; g < num_groups; ++g) {
        for (int32_t k = 0; k < sources_per_group; ++k) {
            int64_t flat = (int64_t)g * sources_per_group + k;
            sources_out[flat] = msbfs_gen_source(seed, (uint32_t)n, (uint64_t)flat);
        }
        group_offsets_out[g + 1] = (int64_t)(g + 1) * sources_per_group;
    }
    return MSBFS_OK;
}
