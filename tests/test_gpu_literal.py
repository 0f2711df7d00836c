"""GPU parity on hand-made stores: the branches of the kernels that the synthetic generator and the two real datasets
never reach -- a flipped symbol (with and without a deleted one at the same position, partial and full filters), run
containers of more than 16 values in container keys >= 1, bitset containers in the second slab (keys 4-7), arrays of
exactly 1,024 / 1,025 / 4,096 values, partitions of exactly 262,144 and 262,145 rows. Everything goes through the C ABI
(sb_store_add with flipped / deleted flags) and is compared bit for bit with the oracle; where the store is consistent
with a symbol matrix also with the plain histogram of that matrix."""
import ctypes

import numpy as np
import pytest

from literal_util import CODE, NUC, fill_partition, histogram_table, random_matrix

pytestmark = pytest.mark.gpu


def q(flt, min_proportion=0.05):
    return {"action": {"type": "Mutations", "minProportion": min_proportion}, "filterExpression": flt}


def check_tables(libs, db, engine, flt, length, matrices=None, masks=None, histogram=True):
    """GPU table == oracle table (fast and, when cheap, faithful) [== histogram]; filter row sets; result rows"""
    pysilo, oracle = libs
    got, got_any = engine.mutation_counts(q(flt), "main", 16 * length)
    want, want_any = oracle.mutation_counts(db, q(flt), "main", 16 * length, faithful=False)
    assert got_any == want_any
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, (flt, "sym", bad[:6] // length, "pos", bad[:6] % length, got[bad[:6]], want[bad[:6]])
    if matrices is not None and histogram:
        assert np.array_equal(got, histogram_table(matrices, masks)), flt
    for p in range(db.n_partitions):
        ids = engine.filter_ids(flt, p)
        assert np.array_equal(ids, oracle.filter_ids(db, flt, p)), (flt, p)
        if masks is not None:
            assert np.array_equal(ids, np.nonzero(masks[p])[0]), (flt, p)
    assert engine.execute(q(flt)) == oracle.execute(db, q(flt), faithful=False)
    agg = {"action": {"type": "Aggregated"}, "filterExpression": flt}
    assert engine.execute(agg) == oracle.execute(db, agg)


def sequence_filters(matrices, pos_a, sym_a, pos_b, sym_b):
    """(filter JSON, expected row mask per partition)"""
    a = [m[:, pos_a] == CODE[sym_a] for m in matrices]
    b = [m[:, pos_b] == CODE[sym_b] for m in matrices]
    ones = [np.ones(m.shape[0], bool) for m in matrices]
    fa = {"type": "NucleotideEquals", "position": pos_a + 1, "symbol": sym_a}
    fb = {"type": "NucleotideEquals", "position": pos_b + 1, "symbol": sym_b}
    return [({"type": "True"}, ones),
            (fa, a),
            ({"type": "Not", "child": fa}, [~x for x in a]),
            ({"type": "And", "children": [fa, {"type": "Not", "child": fb}]}, [x & ~y for x, y in zip(a, b)]),
            ({"type": "Or", "children": [fa, fb]}, [x | y for x, y in zip(a, b)])]


def most_numerous(matrix, pos):
    counts = np.bincount(matrix[:, pos], minlength=16)[:16].copy()
    counts[CODE["N"]] = 0
    return NUC[int(np.argmax(counts))]


@pytest.mark.parametrize("modes", ["flip", "flip+delete", "raw", "mixed"])
def test_flipped_and_deleted_symbols(libs, modes):
    """every position state over two partitions and two position tiles; filters over a flipped, a deleted and a plain symbol"""
    pysilo, _ = libs
    rng = np.random.default_rng(21)
    length = 600
    reference = "".join(rng.choice(list("ACGT"), length))
    db = pysilo.LiteralDatabase({"main": reference})
    matrices = []
    for n in (900, 1400):
        m = random_matrix(rng, n, length, reference)
        p = db.add_partition(n)
        per_pos = [modes] * length if modes != "mixed" else [("raw", "delete", "flip", "flip+delete")[(pos // 3) % 4] for pos in range(length)]
        fill_partition(db, p, "main", m, per_pos)
        matrices.append(m)
    engine = pysilo.Engine(db, 0)
    try:
        pos_a, pos_b = 257, 12
        top = most_numerous(matrices[0], pos_a)          # the flipped / deleted symbol of that position in partition 0
        other = "G" if top != "G" else "C"
        for sym_a in (top, other):
            for flt, masks in sequence_filters(matrices, pos_a, sym_a, pos_b, reference[pos_b]):
                # flipped and deleted at one position under a FULL filter: the reference subtracts the raw cardinality of the flipped
                # symbol's stored bitmap (mutations.cpp:131-132), so the table is not the histogram there; the oracle is the authority
                full_quirk = modes in ("flip+delete", "mixed") and flt["type"] == "True"
                check_tables(libs, db, engine, flt, length, matrices, masks, histogram=not full_quirk)
    finally:
        engine.close()
        db.close()


def big_partition_matrix(rng, n, length, reference):
    """a matrix whose stored bitmaps hold: bitsets in every key (positions 3, 200: ~35 % alternative symbol), run containers
    with long runs in keys >= 1 (positions 5, 280: contiguous row blocks), arrays of exactly 1,024 / 1,025 / 4,096 values in one
    key (positions 7, 8, 9) and the usual sparse mutations elsewhere"""
    m = random_matrix(rng, n, length, reference, p_mut=0.004, p_missing_edge=0.15, p_gap=0.001)
    ref = np.array([CODE[c] for c in reference], dtype=np.uint8)
    alt = lambda pos: CODE["T"] if reference[pos] != "T" else CODE["C"]
    for pos in (3, 200):
        m[:, pos] = np.where(rng.random(n) < 0.35, alt(pos), ref[pos])
    for pos, blocks in ((5, [(70000, 90000), (131072, 131072 + 40000), (n - 3000, n)]), (280, [(65536, 65536 + 500), (200000, 230000)])):
        m[:, pos] = ref[pos]
        for a, b in blocks:
            m[a:min(b, n), pos] = alt(pos)
    for pos, card in ((7, 1024), (8, 1025), (9, 4096)):
        m[:, pos] = ref[pos]
        rows = 65536 + np.sort(rng.choice(65536, card, replace=False))      # all in container key 1
        m[rows[rows < n], pos] = alt(pos)
    return m


@pytest.mark.parametrize("n_rows", [262144, 262145, 300000])
def test_slab_boundaries_runs_and_exact_cardinalities(libs, n_rows):
    """one partition of exactly one slab, one slab + one row, and 4.6 keys (bitsets in the second slab)"""
    pysilo, oracle = libs
    rng = np.random.default_rng(n_rows)
    length = 300
    reference = "".join(rng.choice(list("ACGT"), length))
    db = pysilo.LiteralDatabase({"main": reference})
    m = big_partition_matrix(rng, n_rows, length, reference)
    p = db.add_partition(n_rows)
    fill_partition(db, p, "main", m, ["delete" if pos % 5 else "flip" for pos in range(length)])
    small = random_matrix(rng, 5000, length, reference)
    fill_partition(db, db.add_partition(5000), "main", small, ["delete"] * length)
    matrices = [m, small]
    info = db.info()["sequences"]["main"]
    assert info["numberOfBitsetContainers"] >= 8 and info["numberOfRunContainers"] >= 4, info
    engine = pysilo.Engine(db, 0)
    try:
        alt3 = "T" if reference[3] != "T" else "C"
        alt5 = "T" if reference[5] != "T" else "C"
        for flt, masks in sequence_filters(matrices, 3, alt3, 5, alt5) + sequence_filters(matrices, 9, "T" if reference[9] != "T" else "C", 200, reference[200])[1:]:
            check_tables(libs, db, engine, flt, length, matrices, masks)
        # N-Of over dense, run-shaped and exact-cardinality leaves
        picks = [(3, alt3), (5, alt5), (200, "T" if reference[200] != "T" else "C"), (280, "T" if reference[280] != "T" else "C"), (7, "T" if reference[7] != "T" else "C")]
        nof = {"type": "N-Of", "numberOfMatchers": 2, "matchExactly": False, "children": [{"type": "NucleotideEquals", "position": pos + 1, "symbol": s} for pos, s in picks]}
        hits = [sum((mm[:, pos] == CODE[s]).astype(np.int32) for pos, s in picks) >= 2 for mm in matrices]
        check_tables(libs, db, engine, nof, length, matrices, hits)
    finally:
        engine.close()
        db.close()


def test_row_ids_beyond_sequence_count_are_rejected(libs):
    """a position bitmap that holds a row id in a container key beyond the partition (silo_b200.cu sb_store_add) must fail loudly"""
    pysilo, _ = libs
    db = pysilo.LiteralDatabase({"main": "ACGT"})
    p = db.add_partition(1000)
    db.set_bitmap(p, "main", 1, "T", [5, 70000])
    with pytest.raises(pysilo.SiloError) as err:
        pysilo.Engine(db, 0)
    assert "beyond sequence_count" in str(err.value)
    db.close()
