"""GPU parity at the BASELINE sizes: the databases and queries bench.py and tools/configs_report.py run (SURVEY 8d,
C3 / C4 / C5), compared bit for bit with the oracle in fast mode -- proven equal to the reference-shaped faithful mode on
partitions with bitset containers in tests/test_host_cpu.py. At these sizes every partition spans several container keys
and slabs and most set bits live in bitset containers: the KIND_BITSET branch of the scan, key_local > 0, slab > 0, the
bitset copy of the leaf kernels and the 1,024-value array pieces are all on the path of these tests."""
import json

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
L = 29903
REFS = json.loads((__import__("conftest").GOLDEN / "exampleDataset1000Sequences" / "reference_genomes.json").read_text())
GENE_LENGTH = {g["name"]: len(g["sequence"]) for g in REFS["genes"]}


def mutations(flt, action="Mutations", **extra):
    return {"action": {"type": action, "minProportion": 0.05, **extra}, "filterExpression": flt}


def count(flt):
    return {"action": {"type": "Aggregated"}, "filterExpression": flt}


def assert_tables_equal(got, want, length, what):
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, (what, "symbols", bad[:6] // length, "positions", bad[:6] % length, got[bad[:6]], want[bad[:6]])


@pytest.fixture(scope="module")
def c3(libs):
    """the bench database: 10 M sequences, seed 0x51100003 (bench.py), and its N-Of(3 of 10) query"""
    pysilo, _ = libs
    import bench
    db = pysilo.Database.synthetic(10_000_000, 0x51100003)
    engine = pysilo.Engine(db, 0)
    yield db, engine, bench.c3_query(db)
    engine.close()
    db.close()


def test_c3_bench_query_table_and_row_sets_10m(libs, c3):
    """the whole uint32[16][29903] table, the per-partition filter row sets, the Aggregated count and the response of the
    bench query against the oracle (mutations.cpp:66-136)"""
    _, oracle = libs
    db, engine, query = c3
    info = db.info()["sequences"]["main"]
    assert info["numberOfBitsetContainers"] > 20000 and max(db.partition_rows(p) for p in range(db.n_partitions)) > 262144
    flt = query["filterExpression"]
    got, got_any = engine.mutation_counts(query, "main", 16 * L)
    want, want_any = oracle.mutation_counts(db, query, "main", 16 * L, faithful=False)
    assert got_any and want_any
    assert_tables_equal(got, want, L, "C3 N-Of")
    total = 0
    for p in range(db.n_partitions):
        ids = engine.filter_ids(flt, p)
        assert np.array_equal(ids, oracle.filter_ids(db, flt, p)), p
        total += len(ids)
    assert engine.execute(count(flt)) == oracle.execute(db, count(flt)) == [{"count": total}]
    assert total == 255389        # known answer of this seed (generator and query are deterministic)
    assert engine.execute(query) == oracle.execute(db, query, faithful=False)


def test_c3_other_filters_10m(libs, c3):
    """complement (every partition partial or full), True (precomputed tables), a lineage, metadata predicates"""
    _, oracle = libs
    db, engine, query = c3
    nof = query["filterExpression"]
    lineage = sorted(n for n in db.lineage_names() if n.count(".") == 2)[0]
    filters = [{"type": "Not", "child": nof}, {"type": "True"},
               {"type": "PangoLineage", "column": "pangoLineage", "value": lineage, "includeSublineages": True},
               {"type": "And", "children": [{"type": "IntBetween", "column": "age", "from": 20, "to": 60}, {"type": "DateBetween", "column": "date", "from": "2021-01-01", "to": "2021-12-31"},
                                            {"type": "Not", "child": {"type": "NucleotideEquals", "position": 23403, "symbol": "N"}}]}]
    for flt in filters:
        got, _ = engine.mutation_counts(mutations(flt), "main", 16 * L)
        want, _ = oracle.mutation_counts(db, mutations(flt), "main", 16 * L, faithful=False)
        assert_tables_equal(got, want, L, flt["type"])
        assert engine.execute(count(flt)) == oracle.execute(db, count(flt))


def test_c3_size_independent_properties_10m(libs, c3):
    """linearity over row sets, count conservation, idempotence, device row selection == the host's M5 loop on the same table"""
    pysilo, _ = libs
    db, engine, query = c3
    f = query["filterExpression"]
    n = 10_000_000
    t_in, any_in = engine.mutation_counts(mutations(f), "main", 16 * L)
    t_again, _ = engine.mutation_counts(mutations(f), "main", 16 * L)
    assert any_in and np.array_equal(t_in, t_again)
    t_out, _ = engine.mutation_counts(mutations({"type": "Not", "child": f}), "main", 16 * L)
    t_all, _ = engine.mutation_counts(mutations({"type": "True"}), "main", 16 * L)
    assert np.array_equal(t_all, t_in + t_out)
    n_in = engine.execute(count(f))[0]["count"]
    assert n_in + engine.execute(count({"type": "Not", "child": f}))[0]["count"] == n == engine.execute(count({"type": "True"}))[0]["count"]
    per_position = t_in.reshape(16, L).sum(axis=0, dtype=np.uint64)
    assert per_position.max() <= n_in and per_position[15000] > 0.9 * n_in
    rows = engine.execute(mutations(f))
    assert len(rows) > 0 and rows == pysilo.rows_from_table(db, mutations(f), "main", t_in)


def test_c4_amino_acid_mutations_10m(libs):
    """BASELINE config C4: AminoAcidMutations under HasAminoAcidMutation S:501 on 10 M sequences with all twelve genes;
    tables of S, ORF1a (the longest gene) and N, and the whole response over all genes"""
    pysilo, oracle = libs
    db = pysilo.Database.synthetic(10_000_000, 0x51100004, with_amino_acids=True)
    engine = pysilo.Engine(db, 0)
    try:
        for flt in ({"type": "HasAminoAcidMutation", "sequenceName": "S", "position": 501}, {"type": "HasNucleotideMutation", "position": 23063}):
            for gene in ("S", "ORF1a", "N"):
                query = mutations(flt, "AminoAcidMutations", sequenceName=gene)
                got, got_any = engine.mutation_counts(query, gene, 25 * GENE_LENGTH[gene])
                want, want_any = oracle.mutation_counts(db, query, gene, 25 * GENE_LENGTH[gene], faithful=False)
                assert got_any == want_any
                assert_tables_equal(got, want, GENE_LENGTH[gene], (flt["type"], gene))
            query = mutations(flt, "AminoAcidMutations")
            assert engine.execute(query) == oracle.execute(db, query, faithful=False)
            assert engine.execute(count(flt)) == oracle.execute(db, count(flt))
    finally:
        engine.close()
        db.close()


def test_c5_lineage_partitions_40m(libs):
    """BASELINE config C5: 40 M sequences in lineage DatabasePartitions; Mutations under a top-level clade with sublineages
    (full in some partitions, partial in one, empty in the rest) and under the N-Of query -- whole tables against the oracle,
    and a full / a partial / an empty partition one by one through a device database that owns only that partition"""
    pysilo, oracle = libs
    import bench
    db = pysilo.Database.synthetic(40_000_000, 0x51100005)
    engine = pysilo.Engine(db, 0)
    try:
        rows = [db.partition_rows(p) for p in range(db.n_partitions)]
        # a lineage (with sublineages) that fills some partitions, cuts through at least one and misses the rest: lineages are
        # the partition key, so a parent lineage covers whole partitions and shares the partition of merged small lineages
        names = sorted({".".join(n.split(".")[:k]) for n in db.lineage_names() for k in range(1, n.count(".") + 2)}, key=lambda n: (n.count("."), n))
        lineage = cards = fallback = None
        for name in names:
            flt = {"type": "PangoLineage", "column": "pangoLineage", "value": name, "includeSublineages": True}
            c = [oracle.filter_cardinality(db, flt, p) for p in range(db.n_partitions)]
            kinds = {"full" if c[p] == rows[p] else "partial" if c[p] else "empty" for p in range(db.n_partitions) if rows[p]}
            if kinds == {"full", "partial", "empty"}:
                lineage, cards = flt, c
                break
            if kinds == {"partial", "empty"} and fallback is None:
                fallback = (flt, c)
        if lineage is None:
            lineage, cards = fallback
        whole = {"type": "PangoLineage", "column": "pangoLineage", "value": names[0], "includeSublineages": True}      # the root lineage: every partition full
        assert all(oracle.filter_cardinality(db, whole, p) == rows[p] for p in range(db.n_partitions))
        partial = [p for p in range(db.n_partitions) if 0 < cards[p] < rows[p]]
        empty = [p for p in range(db.n_partitions) if cards[p] == 0]
        full = [p for p in range(db.n_partitions) if cards[p] == rows[p] and rows[p]]
        assert partial and empty
        for flt in (lineage, whole, bench.c3_query(db)["filterExpression"]):
            got, _ = engine.mutation_counts(mutations(flt), "main", 16 * L)
            want, _ = oracle.mutation_counts(db, mutations(flt), "main", 16 * L, faithful=False)
            assert_tables_equal(got, want, L, flt["type"])
            assert engine.execute(count(flt)) == oracle.execute(db, count(flt))
            assert engine.execute(mutations(flt)) == oracle.execute(db, mutations(flt), faithful=False)
        engine.close()
        engine = None
        for p, flt in ((full[0] if full else 0, lineage if full else whole), (partial[0], lineage), (empty[0], lineage)):
            one = pysilo.Engine(db, 0, [p])
            try:
                got, got_any = one.mutation_counts(mutations(flt), "main", 16 * L)
                want, want_any = oracle.mutation_counts_subset(db, mutations(flt), "main", 16 * L, [p], faithful=False)
                assert got_any == want_any == (oracle.filter_cardinality(db, flt, p) > 0)
                assert_tables_equal(got, want, L, ("partition", p))
            finally:
                one.close()
    finally:
        if engine:
            engine.close()
        db.close()
