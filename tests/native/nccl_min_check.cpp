// Test program (GPU box with >= 2 GPUs): msbfs_allreduce_min_u64 over a real NCCL
// communicator, one thread per device, against the expected packed minimum.
// Build: g++ -std=c++17 nccl_min_check.cpp -I<repo>/include -L<pkg> -lmsbfs -ldl -lpthread
// NCCL and the CUDA runtime are resolved with dlopen so that the test needs no headers.
#include "msbfs.h"

#include <dlfcn.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

typedef int (*comm_init_all_fn)(void** comms, int ndev, const int* devlist);
typedef int (*comm_destroy_fn)(void* comm);

int main(int argc, char** argv) {
    int ndev = argc > 1 ? atoi(argv[1]) : 2;
    void* nccl = dlopen("libnccl.so.2", RTLD_NOW);
    if (!nccl) { printf("SKIP: libnccl.so.2 not found (%s)\n", dlerror()); return 0; }
    auto init_all = (comm_init_all_fn)dlsym(nccl, "ncclCommInitAll");
    auto destroy = (comm_destroy_fn)dlsym(nccl, "ncclCommDestroy");
    if (!init_all || !destroy) { printf("FAIL: NCCL symbols missing\n"); return 1; }
    std::vector<msbfs_t*> eng(ndev, nullptr);
    for (int d = 0; d < ndev; ++d)
        if (msbfs_create(d, &eng[d]) != MSBFS_OK) { printf("SKIP: device %d: %s\n", d, msbfs_last_error()); return 0; }
    std::vector<void*> comms(ndev, nullptr);
    std::vector<int> devs(ndev);
    for (int d = 0; d < ndev; ++d) devs[d] = d;
    if (init_all(comms.data(), ndev, devs.data()) != 0) { printf("FAIL: ncclCommInitAll\n"); return 1; }

    // rank d owns sums {100 - d, 7, 7 + d} with ids starting at 10 * d: the global minimum is
    // F = 7, first reached at id 1 (rank 0) -> ties must resolve to the lowest id.
    std::vector<uint64_t> packed(ndev), result(ndev);
    for (int d = 0; d < ndev; ++d) {
        int64_t sums[3] = {100 - d, 7, 7 + d};
        if (msbfs_argmin_packed(sums, 10 * d, 3, &packed[d]) != MSBFS_OK) { printf("FAIL: pack\n"); return 1; }
        result[d] = packed[d];
    }
    std::vector<int> rc(ndev, 0);
    std::vector<std::thread> th;
    for (int d = 0; d < ndev; ++d)
        th.emplace_back([&, d] { rc[d] = msbfs_allreduce_min_u64(eng[d], comms[d], &result[d]); });
    for (auto& t : th) t.join();
    int bad = 0;
    for (int d = 0; d < ndev; ++d) {
        int32_t k; int64_t f;
        msbfs_unpack_argmin(result[d], &k, &f);
        if (rc[d] != MSBFS_OK || k != 1 || f != 7) { printf("FAIL: rank %d rc=%d k=%d f=%lld (%s)\n", d, rc[d], k, (long long)f, msbfs_last_error()); bad = 1; }
    }
    for (int d = 0; d < ndev; ++d) { destroy(comms[d]); msbfs_destroy(eng[d]); }
    if (!bad) printf("OK: %d ranks agree on k=1 F=7\n", ndev);
    return bad;
}
