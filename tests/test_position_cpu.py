"""CPU tests of the Position writers and of the oracle on hand-made stores.

The eight known-answer tests of src/silo/storage/position.test.cpp:31-162 (flip / delete of the most numerous symbol
from every initial state) run against host/position.hpp, and the oracle's Mutations tables on stores in every state
(raw, deleted, flipped, flipped + deleted) are checked against a plain histogram of the symbol matrix the store was
made from -- an expected value that shares no code with the oracle or the kernels."""
import numpy as np
import pytest

from literal_util import NUC, fill_partition, histogram_table, random_matrix

ROWS_C, ROWS_A = [1, 2, 3], [0, 4]


def fresh(libs, **init):
    pysilo, _ = libs
    db = pysilo.LiteralDatabase({"main": "A"})
    db.add_partition(5)
    if "flipped" in init:
        db.set_flags(0, "main", 0, flipped=init["flipped"])
    if "deleted" in init:
        db.set_flags(0, "main", 0, deleted=init["deleted"])
    db.add_values(0, "main", 0, "C", ROWS_C, 0, 5)
    db.add_values(0, "main", 0, "A", ROWS_A, 0, 5)
    return db


def bitmaps(db):
    return db.get_bitmap(0, "main", 0, "C")[0], db.get_bitmap(0, "main", 0, "A")[0]


def test_flips_most_numerous_from_initially_unoptimized(libs):      # position.test.cpp:31-46
    db = fresh(libs)
    assert db.flip_most_numerous(0, "main", 0, 5) == "C"
    assert bitmaps(db) == ([0, 4], [0, 4])
    assert db.flip_most_numerous(0, "main", 0, 5) is None
    assert bitmaps(db) == ([0, 4], [0, 4])


def test_flips_most_numerous_from_initially_different_symbol_flipped(libs):   # :48-64
    db = fresh(libs, flipped="A")
    assert db.flip_most_numerous(0, "main", 0, 5) == "C"
    assert bitmaps(db) == ([0, 4], [0, 4])
    assert db.flip_most_numerous(0, "main", 0, 5) is None
    assert bitmaps(db) == ([0, 4], [0, 4])


def test_flips_most_numerous_from_initially_same_symbol_flipped(libs):        # :66-82
    db = fresh(libs, flipped="C")
    assert db.flip_most_numerous(0, "main", 0, 5) is None
    assert bitmaps(db) == ([0, 4], [0, 4])
    assert db.flip_most_numerous(0, "main", 0, 5) is None
    assert bitmaps(db) == ([0, 4], [0, 4])


def test_deletes_most_numerous_from_initially_unoptimized(libs):              # :84-99
    pysilo, _ = libs
    db = fresh(libs)
    assert db.delete_most_numerous(0, "main", 0, 5) == "C"
    assert bitmaps(db) == ([], [0, 4])
    with pytest.raises(pysilo.SiloError):
        db.delete_most_numerous(0, "main", 0, 5)
    assert bitmaps(db) == ([], [0, 4])


def test_deletes_most_numerous_from_initially_different_symbol_deleted(libs):  # :101-113
    pysilo, _ = libs
    db = fresh(libs, deleted="A")
    assert bitmaps(db) == ([1, 2, 3], [])
    with pytest.raises(pysilo.SiloError):
        db.delete_most_numerous(0, "main", 0, 5)


def test_deletes_most_numerous_from_initially_same_symbol_deleted(libs):       # :115-127
    pysilo, _ = libs
    db = fresh(libs, deleted="C")
    with pytest.raises(pysilo.SiloError):
        db.delete_most_numerous(0, "main", 0, 5)
    assert bitmaps(db) == ([], [0, 4])


def test_deletes_from_initially_different_symbol_flipped(libs):               # :129-143
    db = fresh(libs, flipped="A")
    assert bitmaps(db) == ([1, 2, 3], [1, 2, 3])
    assert db.delete_most_numerous(0, "main", 0, 5) == "C"
    assert bitmaps(db) == ([], [0, 4])
    assert db.get_bitmap(0, "main", 0, "A")[1:] == (None, "C")


def test_flip_then_delete_from_initially_different_symbol_flipped(libs):      # :145-162
    db = fresh(libs, flipped="A")
    assert bitmaps(db) == ([1, 2, 3], [1, 2, 3])
    assert db.flip_most_numerous(0, "main", 0, 5) == "C"
    assert bitmaps(db) == ([0, 4], [0, 4])
    assert db.delete_most_numerous(0, "main", 0, 5) == "C"
    assert bitmaps(db) == ([], [0, 4])


@pytest.mark.parametrize("state", ["flip", "delete", "none"])
def test_oracle_counts_on_position_kat_states(libs, state):
    """the five rows C,C,C / A,A of the KATs in every final state: the Mutations table must be the histogram {A: 2, C: 3}"""
    _, oracle = libs
    db = fresh(libs)
    if state == "flip":
        db.flip_most_numerous(0, "main", 0, 5)
    elif state == "delete":
        db.delete_most_numerous(0, "main", 0, 5)
    for flt, want_a, want_c in (({"type": "True"}, 2, 3), ({"type": "NucleotideEquals", "position": 1, "symbol": "C"}, 0, 3),
                                ({"type": "Not", "child": {"type": "NucleotideEquals", "position": 1, "symbol": "C"}}, 2, 0)):
        table, produced = oracle.mutation_counts(db, {"action": {"type": "Mutations", "minProportion": 0}, "filterExpression": flt}, "main", 16)
        assert produced and table[NUC.index("A")] == want_a and table[NUC.index("C")] == want_c and table.sum() == want_a + want_c, (state, flt)


def literal_small(libs, seed=11, rows=(700, 450), length=300):
    """two partitions over a 300 nt sequence, every position state in turn"""
    pysilo, _ = libs
    rng = np.random.default_rng(seed)
    reference = "".join(rng.choice(list("ACGT"), length))
    db = pysilo.LiteralDatabase({"main": reference})
    matrices = []
    for n in rows:
        m = random_matrix(rng, n, length, reference)
        p = db.add_partition(n)
        fill_partition(db, p, "main", m, [("raw", "delete", "flip", "delete")[pos % 4] for pos in range(length)])
        matrices.append(m)
    return db, matrices, reference


def test_oracle_equals_histogram_on_literal_stores(libs):
    """faithful and fast oracle vs the histogram of the matrix, under True (full partitions), a sequence filter and its complement"""
    _, oracle = libs
    db, matrices, reference = literal_small(libs)
    length = len(reference)
    pos, sym = 40, "G"
    filters = [({"type": "True"}, [np.ones(m.shape[0], bool) for m in matrices]),
               ({"type": "NucleotideEquals", "position": pos + 1, "symbol": sym}, [m[:, pos] == NUC.index(sym) for m in matrices]),
               ({"type": "Not", "child": {"type": "NucleotideEquals", "position": pos + 1, "symbol": sym}}, [m[:, pos] != NUC.index(sym) for m in matrices])]
    for flt, masks in filters:
        query = {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": flt}
        want = histogram_table(matrices, masks)
        for faithful in (True, False):
            got, produced = oracle.mutation_counts(db, query, "main", 16 * length, faithful=faithful)
            assert produced
            bad = np.nonzero(got != want)[0]
            assert bad.size == 0, (flt, faithful, bad[:5] // length, bad[:5] % length, got[bad[:5]], want[bad[:5]])
        for p, mask in enumerate(masks):
            assert np.array_equal(oracle.filter_ids(db, flt, p), np.nonzero(mask)[0])
