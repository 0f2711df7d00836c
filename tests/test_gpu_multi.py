"""Two GPUs, one process each (NCCL): a database sharded by DatabasePartition (row slices of every lineage partition), the count tables summed inside the query's
submission (sb_comm_init -> ncclReduce to rank 0), against the oracle over the WHOLE database in one process.
Also two host threads per rank issuing queries in an order both ranks agree on (collective calls: include/silo_b200.h).
Skipped with fewer than two devices."""
import json
import os
import socket
import sys
import threading
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parent.parent
N_ROWS, SEED, GENOME = 300_000, 0x51100009, 6000
WORLD = 2


def _queries(db):
    clade = db.clade_mutations()
    out = []
    for k in range(4):
        picks = [clade[(7 * j + 3 * k) % len(clade)] for j in range(8)]
        out.append({"action": {"type": "Mutations", "minProportion": 0.02 * (k + 1)},
                    "filterExpression": {"type": "N-Of", "numberOfMatchers": 2, "matchExactly": False,
                                         "children": [{"type": "NucleotideEquals", "position": p + 1, "symbol": s} for p, s in picks]}})
    out.append({"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": {"type": "True"}})
    out.append({"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": {"type": "Not", "child": {"type": "True"}}})
    return out


def _aggregated_queries(db):
    flt = _queries(db)[0]["filterExpression"]
    return [{"action": {"type": "Aggregated"}, "filterExpression": flt},
            {"action": {"type": "Aggregated", "groupByFields": ["region", "pangoLineage"], "orderByFields": ["pangoLineage", "region"]}, "filterExpression": flt},
            {"action": {"type": "Aggregated", "groupByFields": ["age", "date"], "orderByFields": [{"field": "count", "order": "descending"}, "age", "date"], "limit": 20}, "filterExpression": {"type": "True"}}]


def _worker(rank, port, out_dir):
    sys.path.insert(0, str(ROOT / "lapis-silo_b200"))
    sys.path.insert(0, str(ROOT / "tests"))
    import torch
    import torch.distributed as dist
    import pysilo
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    device = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=WORLD, device_id=device)
    # every lineage partition cut into WORLD row slices, one per rank (bench.py --shard rows); the oracle below sees whole partitions
    db = pysilo.Database.synthetic(N_ROWS, SEED, genome_length=GENOME, world=WORLD, rank=rank, row_slices=WORLD)
    engine = pysilo.Engine(db, rank, db.owned)
    id_t = torch.zeros(128, dtype=torch.uint8, device=device)
    if rank == 0:
        id_t.copy_(torch.frombuffer(bytearray(pysilo.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(id_t, 0)
    pysilo.comm_init(engine, bytes(id_t.cpu().numpy().tobytes()), rank, WORLD)
    queries = _queries(db)
    texts = [json.dumps(q) for q in queries]
    engine.set_plan_cache(False)
    # one thread, every query twice (one-shot path), then as prepared queries
    sequential = [engine.execute_sharded(t, rank == 0) for t in texts + texts]
    prepared = []
    for q in queries[:2]:
        p = pysilo.Prepared(engine, q, use_graph=False)
        prepared.append(p.run_text() if rank == 0 else (p.run_device(), "")[1])
        p.close()
    # two host threads per rank; the order of the collective calls is fixed by a turn counter both ranks walk identically
    schedule = [(i % 2, i % len(texts)) for i in range(12)]      # (thread, query) in issue order
    results, turn, cond = [None] * len(schedule), [0], threading.Condition()

    def run(thread_index):
        for slot, (who, q) in enumerate(schedule):
            if who != thread_index:
                continue
            with cond:
                cond.wait_for(lambda: turn[0] == slot)
            results[slot] = engine.execute_sharded(texts[q], rank == 0)      # collective: returns when the reduce is done
            with cond:
                turn[0] = slot + 1
                cond.notify_all()
    threads = [threading.Thread(target=run, args=(k,)) for k in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    torch.cuda.synchronize()
    dist.barrier()
    # Aggregated (count and groupByFields): every rank's partial answer travels to rank 0, which merges (host/group_by.hpp)
    aggregated = []
    for q in _aggregated_queries(db):
        gathered = [None] * WORLD if rank == 0 else None
        dist.gather_object(engine.execute_partial(q), gathered, dst=0)
        if rank == 0:
            aggregated.append(pysilo.merge_partials(db, q, gathered))
    if rank == 0:
        Path(out_dir, "rank0.json").write_text(json.dumps({"sequential": sequential, "prepared": prepared, "threaded": results, "schedule": schedule, "aggregated": aggregated}))
    pysilo.comm_free(engine)
    engine.close()
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_queries_equal_the_whole_database_oracle(libs, tmp_path):
    import torch
    if torch.cuda.device_count() < WORLD:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    pysilo, oracle = libs
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(port, str(tmp_path)), nprocs=WORLD, join=True)
    got = json.loads((tmp_path / "rank0.json").read_text())
    whole = pysilo.Database.synthetic(N_ROWS, SEED, genome_length=GENOME)
    want = [json.dumps({"queryResult": oracle.execute(whole, q)}, separators=(",", ":")) for q in _queries(whole)]
    expect = [json.loads(w) for w in want]
    assert [json.loads(t) for t in got["sequential"]] == expect + expect
    assert [json.loads(t) for t in got["prepared"]] == expect[:2]
    assert [json.loads(t) for t in got["threaded"]] == [expect[q] for _, q in got["schedule"]]
    assert got["aggregated"] == [oracle.execute(whole, q) for q in _aggregated_queries(whole)]
    assert len(expect[0]["queryResult"]) > 0 and len(expect[4]["queryResult"]) > 0 and expect[5]["queryResult"] == []
