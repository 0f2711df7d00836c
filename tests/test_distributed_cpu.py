"""World-size-2 `gloo` tests of the multi-GPU protocol's host side, runnable without a GPU.

Each rank generates only the DatabasePartitions it owns (the same longest-processing-time assignment on
every rank), produces its LOCAL count table — here with the CPU oracle standing in for the device kernels,
which the -m gpu suite compares against the same oracle — the tables are summed onto rank 0 with the
collective bench.py uses (reduce, SUM on int32 = the uint32 bits), and rank 0 assembles the rows with the
product's host code. The result must be identical to the single-process oracle over the whole database.
"""
import json
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
N_ROWS, SEED, GENOME = 20000, 0x51100007, 4000
WORLD = 2


def _query(db):
    clade = db.clade_mutations()
    picks = [clade[(7 * k) % len(clade)] for k in range(10)]
    return {"action": {"type": "Mutations", "minProportion": 0.02},
            "filterExpression": {"type": "N-Of", "numberOfMatchers": 2, "matchExactly": False,
                                 "children": [{"type": "NucleotideEquals", "position": p + 1, "symbol": s} for p, s in picks]}}


def _worker(rank, port, out_dir):
    sys.path.insert(0, str(ROOT / "lapis-silo_b200"))
    sys.path.insert(0, str(ROOT / "tests"))
    import torch
    import torch.distributed as dist
    import pysilo
    import oracle_bind
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        db = pysilo.Database.synthetic(N_ROWS, SEED, genome_length=GENOME, world=WORLD, rank=rank)
        query = _query(db)
        table_len = 16 * GENOME
        # ownership: every partition on exactly one rank
        owned = torch.zeros(db.n_partitions, dtype=torch.int32)
        owned[db.owned] = 1
        dist.all_reduce(owned)
        assert bool((owned == 1).all()), "partition ownership is not a partition of the partitions"
        rows_local = torch.tensor([sum(db.partition_rows(p) for p in db.owned)], dtype=torch.int64)
        rows_all = [torch.zeros(1, dtype=torch.int64) for _ in range(WORLD)]
        dist.all_gather(rows_all, rows_local)
        # local table (oracle as the stand-in for sb_mutations_device), produced flag, reduce
        arr = (oracle_bind.ctypes.c_uint32 * max(len(db.owned), 1))(*db.owned)
        local = np.zeros(table_len, dtype=np.uint32)
        rc = oracle_bind.lib().oracle_mutation_counts_subset(db.handle, json.dumps(query).encode(), b"main", 1, 1, arr, len(db.owned), local.ctypes.data, table_len)
        assert rc in (0, 4), oracle_bind.lib().oracle_last_error()
        counts = torch.from_numpy(local.view(np.int32).copy())
        produced = torch.tensor([int(rc == 0)], dtype=torch.int32)
        dist.reduce(counts, dst=0, op=dist.ReduceOp.SUM)
        dist.all_reduce(produced, op=dist.ReduceOp.MAX)
        if rank == 0:
            reduced = counts.numpy().view(np.uint32)
            result = pysilo.rows_from_table(db, query, "main", reduced) if int(produced) else []
            Path(out_dir, "result.json").write_text(json.dumps({"rows": result, "owned_rows": [int(r) for r in rows_all], "table_sum": int(reduced.astype(np.uint64).sum())}))
        Path(out_dir, f"owned_{rank}.json").write_text(json.dumps(db.owned))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.fixture(scope="module")
def two_rank_run(libs, tmp_path_factory):
    import torch.multiprocessing as mp
    out_dir = tmp_path_factory.mktemp("dist")
    mp.spawn(_worker, args=(_free_port(), str(out_dir)), nprocs=WORLD, join=True)
    return out_dir


def test_ownership_is_disjoint_balanced_and_deterministic(libs, two_rank_run):
    pysilo, _ = libs
    owned = [json.loads((two_rank_run / f"owned_{r}.json").read_text()) for r in range(WORLD)]
    assert not set(owned[0]) & set(owned[1])
    single = pysilo.Database.synthetic(N_ROWS, SEED, genome_length=GENOME)
    assert sorted(owned[0] + owned[1]) == list(range(single.n_partitions))
    # the same assignment again in this process (what a restarted rank would compute)
    again = pysilo.Database.synthetic(N_ROWS, SEED, genome_length=GENOME, world=WORLD, rank=1)
    assert again.owned == owned[1]
    rows = json.loads((two_rank_run / "result.json").read_text())["owned_rows"]
    assert sum(rows) == N_ROWS
    largest = max(single.partition_rows(p) for p in range(single.n_partitions))
    assert abs(rows[0] - rows[1]) <= largest     # LPT bound: loads differ by at most one partition


def test_reduced_tables_give_the_single_process_result(libs, two_rank_run):
    pysilo, oracle = libs
    single = pysilo.Database.synthetic(N_ROWS, SEED, genome_length=GENOME)
    query = _query(single)
    expected = oracle.execute(single, query)
    got = json.loads((two_rank_run / "result.json").read_text())
    assert len(expected) > 0
    assert got["rows"] == expected
    table, produced = oracle.mutation_counts(single, query, "main", 16 * GENOME)
    assert produced and got["table_sum"] == int(table.astype(np.uint64).sum())


def test_uint32_wrap_survives_the_int32_sum():
    """The deleted-symbol arithmetic may wrap (mutations.cpp:127-133); an int32 SUM must keep the uint32 bits."""
    import torch
    a = np.array([0xFFFFFFF0, 5, 0x80000000], dtype=np.uint32)
    b = np.array([0x20, 0xFFFFFFFF, 0x80000000], dtype=np.uint32)
    s = (torch.from_numpy(a.view(np.int32).copy()) + torch.from_numpy(b.view(np.int32).copy())).numpy().view(np.uint32)
    assert s.tolist() == [0x10, 4, 0]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_row_sliced_shards_add_up_to_the_whole_database(libs, world):
    """bench.py --shard rows (the default for N > 1): every lineage partition is cut into `world` row slices, slice k on rank k.
    Each rank's database is built here in turn (host logic only), its local table and count come from the oracle over the partitions
    the rank owns; the sums over the ranks must be the whole database's table and count -- for the world sizes of the scaling run."""
    pysilo, oracle = libs
    n_rows, genome = 12000, 1500
    whole = pysilo.Database.synthetic(n_rows, SEED, genome_length=genome)
    clade = whole.clade_mutations()
    picks = [clade[(5 * k) % len(clade)] for k in range(8)]
    flt = {"type": "N-Of", "numberOfMatchers": 1, "matchExactly": False, "children": [{"type": "NucleotideEquals", "position": p + 1, "symbol": s} for p, s in picks]}
    query = {"action": {"type": "Mutations", "minProportion": 0.02}, "filterExpression": flt}
    want_table, produced = oracle.mutation_counts(whole, query, "main", 16 * genome)
    want_count = oracle.execute(whole, {"action": {"type": "Aggregated"}, "filterExpression": flt})[0]["count"]
    assert produced and want_count > 0
    grouped = {"action": {"type": "Aggregated", "groupByFields": ["region", "pangoLineage", "age"], "orderByFields": [{"field": "count", "order": "descending"}, "pangoLineage", "region", "age"], "limit": 25},
               "filterExpression": flt}
    partial_query = {"action": {"type": "Aggregated", "groupByFields": ["region", "pangoLineage", "age"]}, "filterExpression": flt}
    partials = []
    table, count, rows, loads = np.zeros(16 * genome, dtype=np.uint64), 0, 0, []
    for rank in range(world):
        shard = pysilo.Database.synthetic(n_rows, SEED, genome_length=genome, world=world, rank=rank, row_slices=world)
        assert shard.n_partitions == whole.n_partitions * world and len(shard.owned) == whole.n_partitions
        local, _ = oracle.mutation_counts_subset(shard, query, "main", 16 * genome, shard.owned)
        table += local
        count += sum(oracle.filter_cardinality(shard, flt, p) for p in shard.owned)
        partials.append(oracle.execute_subset_text(shard, partial_query, shard.owned))      # what Engine.execute_partial sends to rank 0
        loads.append(sum(shard.partition_rows(p) for p in shard.owned))
        rows += loads[-1]
        shard.close()
    assert rows == n_rows and max(loads) - min(loads) <= whole.n_partitions      # slices differ by at most one row per lineage
    assert count == want_count
    assert np.array_equal(table, want_table.astype(np.uint64))
    assert pysilo.merge_partials(whole, grouped, partials) == oracle.execute(whole, grouped)      # grouped Aggregated: partial tuples add up
