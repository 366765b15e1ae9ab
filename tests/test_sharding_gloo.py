"""World-size-2/3 CPU test (gloo) of the multi-rank path: contiguous batch sharding plus
the packed-min all-reduce must give the reference's argmin, ties included.  Per-rank F
values come from the oracle here (no GPU in this test); on the GPU box bench.py runs the
same code with the CUDA engine and NCCL."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, sums, lanes, out_path):
    import torch.distributed as dist
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import importlib
    import __graft_entry__ as graft
    pkg = graft.load_package()
    sharding = importlib.import_module(pkg.__name__ + ".sharding")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    if lanes > 0:   # contiguous shares
        a, b = pkg.shard_range(len(sums), rank, world, lanes)
        k, f = sharding.distributed_argmin(np.asarray(sums[a:b], dtype=np.int64), a)
    else:           # groups dealt by cost: non-contiguous ids travel in the packed word
        costs = np.asarray([(7 * i) % 5 for i in range(len(sums))], dtype=np.int64)
        mine = pkg.deal_groups(costs, rank, world)
        packed = pkg.argmin_packed_ids(np.asarray(sums, dtype=np.int64)[mine], mine)
        k, f = pkg.unpack_argmin(sharding.allreduce_packed_min(packed))
    if rank == 0:
        with open(out_path, "w") as fh:
            fh.write(f"{k} {f}")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,case", [(2, "ties"), (2, "empty"), (3, "random"), (2, "single_batch"), (3, "too_big_to_pack"),
                                        (3, "dealt_ties")])
def test_distributed_argmin(pkg, ob, tmp_path, world, case):
    rng = np.random.default_rng(5)
    lanes = 4  # small batches so that several ranks own groups
    if case == "dealt_ties":
        lanes = 0
        sums = [9, 7, 8, 7, 7, 30, 7, 12, 7, 7, 11]
    elif case == "ties":
        sums = [9, 7, 8, 7, 7, 30, 7, 12, 7]
    elif case == "empty":
        sums = []
    elif case == "too_big_to_pack":
        sums = [(1 << 45) + 5, 1 << 41, (1 << 45) + 5, 1 << 41, 1 << 50, 7 << 40, 1 << 41, 9 << 40, 1 << 42]  # F >= 2^39
    elif case == "single_batch":
        sums = [5, 3, 3]  # one batch: rank 1 owns nothing and contributes UINT64_MAX
    else:
        sums = rng.integers(0, 1 << 38, size=37).tolist()
    out = str(tmp_path / "out.txt")
    mp.spawn(_worker, args=(world, _free_port(), sums, lanes, out), nprocs=world, join=True)
    k, f = (int(x) for x in open(out).read().split())
    assert (k, f) == ob.argmin(np.asarray(sums, dtype=np.int64))
