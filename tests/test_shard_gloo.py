"""N>1 host logic on CPU: world_size-2 `gloo` processes shard a batch contiguously, map their slices
independently (the oracle stands in for the per-rank mapper -- there is no GPU here), reduce the shard counters
with the one collective of the design, and rank 0's concatenation in rank order equals the one-process run."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_partitions():
    from longreadmapper_b200.shard import shard_range
    for n in (0, 1, 7, 8, 1000, 1_000_003):
        for world in (1, 2, 3, 8):
            r = [shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [e - b for b, e in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    import longreadmapper_b200 as lm
    from longreadmapper_b200 import shard
    import oracle_lib
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = lm.sim_genome(41, 60_000)
    ix = oracle_lib.OracleIndex.build(["chr1"], [g.tobytes()])          # index replica of this rank
    reads, lens, _, _ = lm.sim_reads(42, g, 301, 100, 180, 0.0, 0.04, 0.7)
    b, e = shard.shard_range(len(lens), rank, world)
    mine = reads[b:e].copy()
    res, ops, cells = ix.map_batch(mine, mine.shape[1], lens[b:e], myers=False, ops_stride=2 * 180)
    st = {"total": e - b, "valid": int(res["valid"].sum()), "n_flag4": int((res["flag"] == 4).sum()),
          "n_flag16": int((res["flag"] == 16).sum()), "sum_ed": int(res["ed"][res["valid"] == 1].sum()), "dp_cells": cells}
    tot = shard.reduce_stats(st)
    tmax = shard.max_over_ranks(10.0 + rank)
    all_res = shard.gather_in_rank_order(res)
    all_ops = shard.gather_in_rank_order(ops)
    all_seq = shard.gather_in_rank_order(mine)
    if rank == 0:
        np.savez(os.path.join(out_dir, "gathered.npz"), res=all_res, ops=all_ops, seq=all_seq, tmax=tmax,
                 **{"st_" + k: v for k, v in tot.items()})
    else:
        assert all_res is None
    ix.close()
    dist.destroy_process_group()


def test_two_rank_sharding_equals_single_process(native, orc, tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    got = np.load(str(tmp_path / "gathered.npz"))
    g = native.sim_genome(41, 60_000)
    ix = orc.OracleIndex.build(["chr1"], [g.tobytes()])
    reads, lens, _, _ = native.sim_reads(42, g, 301, 100, 180, 0.0, 0.04, 0.7)
    res, ops, cells = ix.map_batch(reads, reads.shape[1], lens, myers=False, ops_stride=2 * 180)
    assert np.array_equal(got["res"], res) and np.array_equal(got["ops"], ops) and np.array_equal(got["seq"], reads)
    assert int(got["st_total"]) == 301 and int(got["st_valid"]) == int(res["valid"].sum())
    assert int(got["st_dp_cells"]) == cells and int(got["st_n_flag16"]) == int((res["flag"] == 16).sum())
    assert float(got["tmax"]) == 11.0
    ix.close()


def test_parse_cpulist():
    from longreadmapper_b200.shard import parse_cpulist
    assert parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert parse_cpulist("5") == {5}
    assert parse_cpulist("") == set()
