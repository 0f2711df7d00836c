"""GPU parity tests: every call goes through the C ABI of libsilo_b200.so (via the C++ host engine)
and is compared bit-for-bit with the oracle / the reference's own golden vectors."""
import json

import numpy as np
import pytest

from conftest import GOLDEN, golden_queries, load_case

pytestmark = pytest.mark.gpu
KATS = json.loads((GOLDEN / "operator_kats.json").read_text())
REFS = json.loads((GOLDEN / "exampleDataset1000Sequences" / "reference_genomes.json").read_text())
GENE_LENGTH = {g["name"]: len(g["sequence"]) for g in REFS["genes"]}


@pytest.fixture(scope="module")
def example_engine(libs, example_db):
    pysilo, _ = libs
    e = pysilo.Engine(example_db, 0)
    yield e
    e.close()


@pytest.fixture(scope="module")
def thousand_engine(libs, thousand_db):
    pysilo, _ = libs
    e = pysilo.Engine(thousand_db, 0)
    yield e
    e.close()


@pytest.fixture(scope="module")
def small_engine(libs, synth_small):
    pysilo, _ = libs
    e = pysilo.Engine(synth_small, 0)
    yield e
    e.close()


@pytest.fixture(scope="module")
def medium_engine(libs, synth_medium):
    pysilo, _ = libs
    e = pysilo.Engine(synth_medium, 0)
    yield e
    e.close()


@pytest.mark.parametrize("case", KATS, ids=lambda c: c["name"])
def test_operator_kat(libs, case):
    pysilo, _ = libs
    if case.get("throws"):
        with pytest.raises(pysilo.SiloError):
            pysilo.eval_optree(case["tree"])
        return
    assert pysilo.eval_optree(case["tree"]).tolist() == case["expected"]


def test_operator_kat_large_ids(libs):
    """containers of all three kinds, several 65,536-row keys, ids beyond row_count (quirk Q6)"""
    pysilo, oracle = libs
    rng = np.random.default_rng(5)
    dense = sorted(set(rng.integers(0, 200000, 90000).tolist()))
    sparse = sorted(set(rng.integers(0, 300000, 500).tolist()))
    runs = list(range(1000, 70000)) + list(range(140000, 150000))
    for tree in [
        {"op": "intersection", "row_count": 250000, "children": [{"op": "index_scan", "ids": dense}, {"op": "index_scan", "ids": runs, "run_optimize": True}], "negated": [{"op": "index_scan", "ids": sparse}]},
        {"op": "union", "row_count": 250000, "children": [{"op": "index_scan", "ids": sparse}, {"op": "complement", "child": {"op": "index_scan", "ids": dense}}]},
        {"op": "threshold", "row_count": 250000, "n": 2, "exact": True, "children": [{"op": "index_scan", "ids": dense}, {"op": "index_scan", "ids": runs, "run_optimize": True}], "negated": [{"op": "index_scan", "ids": sparse}, {"op": "full"}]},
        {"op": "threshold", "row_count": 250000, "n": 3, "exact": False, "children": [{"op": "index_scan", "ids": dense}, {"op": "index_scan", "ids": runs}, {"op": "range_selection", "ranges": [[5, 100000], [100000, 120000]]}, {"op": "index_scan", "ids": sparse}], "negated": [{"op": "empty"}]},
    ]:
        assert np.array_equal(pysilo.eval_optree(tree), oracle.eval_optree(tree))


@pytest.mark.parametrize("path", golden_queries(), ids=lambda p: p.stem)
def test_e2e_fixture(example_engine, path):
    case = load_case(path)
    assert example_engine.execute(case["query"]) == case["expectedQueryResult"]


@pytest.mark.parametrize("path", sorted((GOLDEN / "invalidQueries").glob("*.json")), ids=lambda p: p.stem)
def test_invalid_query_fixture(libs, example_engine, path):
    pysilo, _ = libs
    case = load_case(path)
    action = case["query"].get("action", {})
    with pytest.raises(pysilo.SiloError) as err:
        example_engine.execute(case["query"])
    if action.get("type") in ("Details", "Fasta", "FastaAligned", "Insertions", "AminoAcidInsertions"):
        pytest.skip("action outside the hot path")
    assert err.value.kind == "parse" and err.value.message == case["expectedError"]["message"]


FILTERS_EXAMPLE = [
    {"type": "True"},
    {"type": "NucleotideEquals", "position": 23403, "symbol": "G"},
    {"type": "NucleotideEquals", "position": 23403, "symbol": "A"},
    {"type": "NucleotideEquals", "position": 100, "symbol": "N"},
    {"type": "Not", "child": {"type": "NucleotideEquals", "position": 28280, "symbol": "-"}},
    {"type": "Maybe", "child": {"type": "NucleotideEquals", "position": 241, "symbol": "T"}},
    {"type": "HasNucleotideMutation", "position": 3037},
    {"type": "HasAminoAcidMutation", "sequenceName": "S", "position": 501},
    {"type": "AminoAcidEquals", "sequenceName": "S", "position": 501, "symbol": "X"},
    {"type": "N-Of", "numberOfMatchers": 2, "matchExactly": True, "children": [
        {"type": "NucleotideEquals", "position": 241, "symbol": "T"}, {"type": "NucleotideEquals", "position": 3037, "symbol": "T"},
        {"type": "NucleotideEquals", "position": 14408, "symbol": "T"}, {"type": "Not", "child": {"type": "NucleotideEquals", "position": 23403, "symbol": "G"}}]},
    {"type": "And", "children": [{"type": "DateBetween", "column": "date", "from": "2021-01-01", "to": "2021-03-31"},
                                 {"type": "IntBetween", "column": "age", "from": 20, "to": 60}, {"type": "FloatBetween", "column": "qc_value", "from": 0.9, "to": None}]},
    {"type": "Or", "children": [{"type": "PangoLineage", "column": "pango_lineage", "value": "B.1.1.7", "includeSublineages": True},
                                {"type": "StringEquals", "column": "country", "value": "Switzerland"}, {"type": "StringEquals", "column": "division", "value": "Aargau"}]},
]


@pytest.mark.parametrize("index", range(len(FILTERS_EXAMPLE)))
def test_filter_row_sets_example(libs, example_db, example_engine, index):
    _, oracle = libs
    f = FILTERS_EXAMPLE[index]
    for p in range(example_db.n_partitions):
        assert np.array_equal(example_engine.filter_ids(f, p), oracle.filter_ids(example_db, f, p)), (f, p)


def tables_equal(libs, db, engine, query, names_and_lengths, sigma, faithful=True):
    _, oracle = libs
    for name, length in names_and_lengths:
        got, got_any = engine.mutation_counts(query, name, sigma * length)
        want, want_any = oracle.mutation_counts(db, query, name, sigma * length, faithful=faithful)
        assert got_any == want_any, name
        mismatch = np.nonzero(got != want)[0]
        assert mismatch.size == 0, (name, mismatch[:5], got[mismatch[:5]], want[mismatch[:5]])
    assert engine.execute(query) == oracle.execute(db, query, faithful=faithful)


@pytest.mark.parametrize("index", range(len(FILTERS_EXAMPLE)))
def test_count_tables_example(libs, example_db, example_engine, index):
    f = FILTERS_EXAMPLE[index]
    tables_equal(libs, example_db, example_engine, {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": f}, [("main", 29903)], 16)
    if index % 3 == 0:
        tables_equal(libs, example_db, example_engine, {"action": {"type": "AminoAcidMutations", "minProportion": 0.05}, "filterExpression": f}, sorted(GENE_LENGTH.items()), 25)


def test_config_c1_thousand_sequences(libs, thousand_db, thousand_engine):
    """BASELINE config C1: nucleotide Mutations minProportion 0.05 under a date + lineage filter"""
    f = {"type": "And", "children": [{"type": "DateBetween", "column": "date", "from": "2020-01-01", "to": "2021-12-31"},
                                     {"type": "PangoLineage", "column": "pangoLineage", "value": "B.1", "includeSublineages": True}]}
    query = {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": f}
    tables_equal(libs, thousand_db, thousand_engine, query, [("main", 29903)], 16)
    rows = thousand_engine.execute(query)
    assert len(rows) > 0 and all(r["proportion"] >= 0.05 for r in rows)
    tables_equal(libs, thousand_db, thousand_engine, {"action": {"type": "AminoAcidMutations", "minProportion": 0.05}, "filterExpression": f}, sorted(GENE_LENGTH.items()), 25)
    tables_equal(libs, thousand_db, thousand_engine, {"action": {"type": "Mutations", "minProportion": 0.0}, "filterExpression": {"type": "True"}}, [("main", 29903)], 16)
    _, oracle = libs
    agg = {"action": {"type": "Aggregated"}, "filterExpression": f}
    assert thousand_engine.execute(agg) == oracle.execute(thousand_db, agg)


def synthetic_filters(db):
    clade = db.clade_mutations()[:10]
    nof = {"type": "N-Of", "numberOfMatchers": 3, "matchExactly": False,
           "children": [{"type": "NucleotideEquals", "position": p + 1, "symbol": s} for p, s in clade]}
    return [
        {"type": "True"},
        {"type": "False"},
        nof,
        {"type": "PangoLineage", "column": "pangoLineage", "value": "B.1.3", "includeSublineages": True},
        {"type": "And", "children": [{"type": "Or", "children": [{"type": "NucleotideEquals", "position": clade[0][0] + 1, "symbol": clade[0][1]},
                                                                {"type": "NucleotideEquals", "position": clade[1][0] + 1, "symbol": clade[1][1]}]},
                                     {"type": "Not", "child": {"type": "NucleotideEquals", "position": 30, "symbol": "-"}},
                                     {"type": "DateBetween", "column": "date", "from": "2021-01-01", "to": "2021-12-31"},
                                     {"type": "StringEquals", "column": "region", "value": "Europe"},
                                     {"type": "IntBetween", "column": "age", "from": 20, "to": 60}]},
        {"type": "Not", "child": {"type": "NucleotideEquals", "position": 12, "symbol": "N"}},
        {"type": "FloatBetween", "column": "qc", "from": 0.25, "to": 0.75},
    ]


def test_synthetic_small_all_stores(libs, synth_small, small_engine):
    """faithful oracle (the reference's loops) vs GPU on every store of a 3,000-row synthetic database"""
    _, oracle = libs
    for f in synthetic_filters(synth_small):
        for p in range(synth_small.n_partitions):
            assert np.array_equal(small_engine.filter_ids(f, p), oracle.filter_ids(synth_small, f, p))
        tables_equal(libs, synth_small, small_engine, {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": f}, [("main", 3000)], 16)
        tables_equal(libs, synth_small, small_engine, {"action": {"type": "AminoAcidMutations", "minProportion": 0.01}, "filterExpression": f}, sorted(GENE_LENGTH.items()), 25)
        agg = {"action": {"type": "Aggregated"}, "filterExpression": f}
        assert small_engine.execute(agg) == oracle.execute(synth_small, agg)


def test_synthetic_medium_full_genome(libs, synth_medium, medium_engine):
    """120,000 rows x 29,903 nt (BASELINE configs C2 / C3 shapes); oracle in fast mode (proven equal to faithful in the CPU suite)"""
    _, oracle = libs
    for f in synthetic_filters(synth_medium):
        for p in range(0, synth_medium.n_partitions, 5):
            assert np.array_equal(medium_engine.filter_ids(f, p), oracle.filter_ids(synth_medium, f, p))
        tables_equal(libs, synth_medium, medium_engine, {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": f}, [("main", 29903)], 16, faithful=False)
        agg = {"action": {"type": "Aggregated"}, "filterExpression": f}
        assert medium_engine.execute(agg) == oracle.execute(synth_medium, agg)


def test_size_independent_properties(libs, synth_medium, medium_engine):
    """coverage identity: for every position, sum over ALL symbols of counts = |F| - #missing; and the
    table under True equals table(F) + table(not F) (linearity over row sets)."""
    f = synthetic_filters(synth_medium)[3]
    q = lambda flt: {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": flt}
    t_all, _ = medium_engine.mutation_counts(q({"type": "True"}), "main", 16 * 29903)
    t_in, _ = medium_engine.mutation_counts(q(f), "main", 16 * 29903)
    t_out, _ = medium_engine.mutation_counts(q({"type": "Not", "child": f}), "main", 16 * 29903)
    assert np.array_equal(t_all, t_in + t_out)
    n_total = medium_engine.execute({"action": {"type": "Aggregated"}, "filterExpression": {"type": "True"}})[0]["count"]
    per_position = t_all.reshape(16, 29903).sum(axis=0)
    assert per_position.max() <= n_total and per_position[15000] > 0.9 * n_total


def test_concurrent_queries_are_isolated(libs, synth_medium, medium_engine):
    """Eight host threads issue different queries at once (one library lane each): every result equals its sequential result."""
    import threading
    filters = synthetic_filters(synth_medium)
    queries = [{"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": filters[k % len(filters)]} for k in range(8)]
    expected = [(medium_engine.mutation_counts(q, "main", 16 * 29903)[0], medium_engine.execute(q)) for q in queries]
    failures = []

    def work(k):
        try:
            for _ in range(5):
                table, _ = medium_engine.mutation_counts(queries[k], "main", 16 * 29903)
                if not np.array_equal(table, expected[k][0]) or medium_engine.execute(queries[k]) != expected[k][1]:
                    failures.append(k)
        except Exception as e:   # noqa: BLE001
            failures.append((k, repr(e)))

    threads = [threading.Thread(target=work, args=(k,)) for k in range(8)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert failures == []


@pytest.mark.parametrize("env", [{"SB_K1_CTAS": "1"}, {"SB_K1_CTAS": "3"}, {"SB_K1_CTAS": "7"}, {"SB_K1_CTAS": "4000"}, {"SB_ITEM_CAP_KB": "4"},
                                 {"SB_RUN_CAP_KB": "1"}, {"SB_RUN_CAP_KB": "4096"}, {"SB_TILE_POSITIONS": "512"}, {"SB_FLAT_MAX": "16"}, {"SB_FLAT_MAX": "64"}, {"SB_FLAT_MAX": "1024"},
                                 {"SB_K1_VARIANT": "1"}, {"SB_K1_VARIANT": "2", "SB_K1_CTAS": "5"}, {"SB_K1_VARIANT": "2", "SB_TILE_POSITIONS": "512", "SB_ITEM_CAP_KB": "4"}],
                         ids=lambda e: "-".join(f"{k}={v}" for k, v in e.items()))
def test_scan_kernel_variants(libs, synth_small, synth_medium, env, monkeypatch):
    """the container scan with odd numbers of persistent CTAs (one CTA takes every item, more CTAs than items), with tiny work
    items (many items per segment), one item or a whole slab per run, 512-position tiles, and in its other compiled shapes
    (without predicated increments, two posting loads in flight at 4 CTAs per SM), against the oracle"""
    pysilo, oracle = libs
    for k, v in env.items():
        monkeypatch.setenv(k, v)      # read when the device database is created
    for db, names, sigma_tables, faithful in ((synth_small, [("main", 3000)], True, True), (synth_medium, [("main", 29903)], False, False)):
        engine = pysilo.Engine(db, 0)
        try:
            for f in synthetic_filters(db)[2:5]:
                tables_equal(libs, db, engine, {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": f}, names, 16, faithful=faithful)
                if sigma_tables:
                    tables_equal(libs, db, engine, {"action": {"type": "AminoAcidMutations", "minProportion": 0.01}, "filterExpression": f}, sorted(GENE_LENGTH.items()), 25)
        finally:
            engine.close()


def test_prepared_queries_and_plan_cache(libs, synth_medium, medium_engine, synth_small, small_engine):
    """a prepared query (with and without a CUDA graph) and the engine's plan cache replay the same tables and texts as the
    one-shot path, run after run; Aggregated and AminoAcidMutations over all genes included"""
    pysilo, oracle = libs
    for db, engine, length in ((synth_medium, medium_engine, 29903), (synth_small, small_engine, 3000)):
        for f in synthetic_filters(db):
            q = {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": f}
            want_table, produced = oracle.mutation_counts(db, q, "main", 16 * length, faithful=False)
            want_rows = oracle.execute(db, q, faithful=False)
            for use_graph in (False, True):
                prepared = pysilo.Prepared(engine, q, use_graph=use_graph)
                try:
                    for _ in range(3):
                        assert prepared.run() == want_rows
                        if produced:
                            assert np.array_equal(prepared.table(0, 16 * length), want_table)
                finally:
                    prepared.close()
            engine.set_plan_cache(True)
            for _ in range(4):      # miss, prepare, hit, hit
                assert engine.execute(q) == want_rows
            agg = {"action": {"type": "Aggregated"}, "filterExpression": f}
            prepared = pysilo.Prepared(engine, agg)
            assert prepared.run() == oracle.execute(db, agg) and prepared.run_device() == oracle.execute(db, agg)[0]["count"]
            prepared.close()
    aa = {"action": {"type": "AminoAcidMutations", "minProportion": 0.01}, "filterExpression": synthetic_filters(synth_small)[2]}
    prepared = pysilo.Prepared(small_engine, aa)
    for _ in range(2):
        assert prepared.run() == oracle.execute(synth_small, aa)
    prepared.close()


def test_in_library_reduce_single_rank(libs, synth_small):
    """sb_comm_init with a world of one rank: the NCCL reduce inside the query's submission must leave the tables as they are
    (the multi-rank sum: tests/test_gpu_multirank.py on two GPUs and, on the CPU, tests/test_distributed_cpu.py)"""
    import torch      # before the library loads NCCL: the process keeps one libnccl.so.2, and torch needs its own
    pysilo, oracle = libs
    engine = pysilo.Engine(synth_small, 0)
    try:
        pysilo.comm_init(engine, pysilo.comm_unique_id(), 0, 1)
        for f in synthetic_filters(synth_small)[1:4]:
            q = {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": f}
            assert json.loads(engine.execute_sharded(json.dumps(q), True))["queryResult"] == oracle.execute(synth_small, q)
            counts = torch.zeros(16 * 3000, dtype=torch.int32, device="cuda:0")
            dq = pysilo.DistributedQuery(engine, q)
            produced = dq.local_table_device(0, counts.data_ptr(), 16 * 3000)
            want, want_any = oracle.mutation_counts(synth_small, q, "main", 16 * 3000)
            assert produced == want_any
            if produced:
                assert np.array_equal(counts.cpu().numpy().view(np.uint32), want)
            dq.close()
        pysilo.comm_free(engine)
    finally:
        engine.close()


def test_batch_of_prepared_queries_from_one_thread(libs, synth_medium, medium_engine):
    """sb_query_run_many: the queries of a batch are enqueued together; every response equals the query run on its own"""
    pysilo, oracle = libs
    queries = [{"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": f} for f in synthetic_filters(synth_medium)[:5]]
    queries.append({"action": {"type": "Aggregated"}, "filterExpression": synthetic_filters(synth_medium)[2]})
    prepared = [pysilo.Prepared(medium_engine, q, use_graph=(k % 2 == 0)) for k, q in enumerate(queries)]
    try:
        for _ in range(2):
            assert pysilo.run_many(medium_engine, prepared) == [medium_engine.execute(q) for q in queries]
        assert pysilo.run_many(medium_engine, prepared[:1])[0] == oracle.execute(synth_medium, queries[0])
    finally:
        for p in prepared:
            p.close()
