"""Multi-rank host logic on CPU (world_size 2 and 3, gloo): the ownership rules libb200svm uses on N GPUs
(b200svm_partition_rows / b200svm_partition_tiles, host-only entry points of the C ABI) cover the implicit matrix exactly
once, and the "partial K d per rank -> one allreduce(sum) -> replicated CG algebra" scheme of csrc/comm.cu + csrc/cg.cu
reproduces the single-rank operator.  The per-rank partial products are computed with the oracle (this is a test); the
GPU version of the same check is tests/multi_gpu_check.py (run under torchrun on >= 2 B200s).
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import port
from plssvm_sparse_fp22_b200 import capi, datagen


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port_no, kernel, out_q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n, F = 301, 23
        X, y = datagen.make_dense(n, F, seed=5, problem="planes")
        m = n - 1
        q = port.generate_q(X, kernel)
        qa = port.kernel_function(X[-1], X[-1], kernel, gamma=1.0 / F) + 1.0
        d = np.random.default_rng(0).uniform(-1, 2, m)
        lo, hi = capi.partition_rows(m, rank, world)
        # this rank's rows of  Abar d  (everything except the replicated O(m) diagonal term d_i / C, which cg.cu adds after the
        # allreduce): the oracle's full-row operator restricted to [lo, hi) with cost = inf
        part = np.zeros(m)
        part[lo:hi] = port.matvec(X, q, d, qa, 1.0, kernel, cost=1e300, rows=(lo, hi))[lo:hi]
        t = torch.from_numpy(part)
        dist.all_reduce(t)  # sum — what ncclAllReduce does on the GPUs
        full = t.numpy() + d / 1.0
        want = port.matvec(X, q, d, qa, 1.0, kernel, rows=(0, m))
        err = float(np.linalg.norm(full - want) / np.linalg.norm(want))
        # every rank ends with the same bits (replicated CG state relies on it)
        gathered = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(gathered, t)
        same = all(torch.equal(g, gathered[0]) for g in gathered)
        out_q.put((rank, lo, hi, err, same))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("kernel", ["linear", "rbf"])
def test_row_partition_allreduce_matches_single_rank(world, kernel):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port_no = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port_no, kernel, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # contiguous, disjoint, complete
    assert res[0][1] == 0 and res[-1][2] == 300
    for a, b in zip(res, res[1:]):
        assert a[2] == b[1]
    for _, _, _, err, same in res:
        assert err < 1e-13
        assert same


@pytest.mark.parametrize("m", [0, 1, 2, 7, 4999, 65535, 999_999])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_partition_rows_cover(m, world):
    edges = [capi.partition_rows(m, r, world) for r in range(world)]
    assert edges[0][0] == 0 and edges[-1][1] == m
    for (lo, hi), (lo2, _) in zip(edges, edges[1:]):
        assert lo <= hi == lo2
    sizes = [hi - lo for lo, hi in edges]
    assert max(sizes) - min(sizes) <= 1


@pytest.mark.parametrize("n", [2, 257, 5000, 65536])
@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_partition_tiles_cover(n, world):
    total = None
    nxt = 0
    counts = []
    for r in range(world):
        first, count, tot = capi.partition_tiles(n, r, world)
        total = tot if total is None else total
        assert tot == total and first == nxt
        nxt = first + count
        counts.append(count)
    assert nxt == total
    assert max(counts) - min(counts) <= 1
    # lower block triangle of 128 x 256 tiles over n rounded up to 256: T_c (2 T_c + ... ) tiles with i >= 2 j
    n_pad = (n + 255) // 256 * 256
    T_r, T_c = n_pad // 128, n_pad // 256
    assert total == sum(T_r - 2 * j for j in range(T_c))


def test_canonical_groups_are_nested_across_world_sizes():
    """The tile-partitioned paths return bit-identical results on 1, 2, 4 and 8 GPUs because every rank owns whole CANONICAL groups
    (DESIGN.md §6): the slice / row-block boundaries of 2, 4 and 8 ranks must be a subset of those of 8 ranks, for both partitions."""
    assert [capi.canonical_groups(w) for w in (1, 2, 4, 8)] == [8, 8, 8, 8]
    assert [capi.canonical_groups(w) for w in (3, 5, 6, 7, 16)] == [3, 5, 6, 7, 16]
    for n in (257, 5000, 65536, 262144):
        tile_edges, row_edges = {}, {}
        for w in (1, 2, 4, 8):
            te, re_ = {0}, {0}
            prev = 0
            for r in range(w):
                first, count, total = capi.partition_tiles(n, r, w)
                te.add(first + count)
                lo, hi = capi.partition_tile_rows(n - 1, r, w)
                assert lo == prev and hi >= lo and lo % 128 == 0
                prev = hi
                re_.add(hi)
            n_pad = (n - 1 + 255) // 256 * 256
            assert prev == n_pad and max(te) == total
            tile_edges[w], row_edges[w] = te, re_
        for w in (1, 2, 4):
            assert tile_edges[w] <= tile_edges[8]
            assert row_edges[w] <= row_edges[8]
    # equal tile counts: the row-block ranges of 8 ranks hold the same number of tiles to within one row block's worth
    n = 65536
    counts = []
    for r in range(8):
        lo, hi = capi.partition_tile_rows(n - 1, r, 8)
        counts.append(sum(I // 2 + 1 for I in range(lo // 128, hi // 128)))
    assert max(counts) - min(counts) <= 256 + 1
