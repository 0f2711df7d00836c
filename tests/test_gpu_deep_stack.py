"""The filter interpreter's global-memory evaluation stack (eval_plans<true>: more than 24 planes) on partitions that span several
container keys and slabs: a 4 M-row synthetic database over a 1,000 nt genome (largest partition 272,796 rows: two slabs, five container keys), N-Of over 60 leaves with 25 .. 58
matchers and right-deep And / Or / N-Of alternations, row sets per partition against the oracle. The small-database version against
the row model is tests/test_filter_model.py::test_wide_and_deep_trees_against_the_model."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def four_million(libs):
    pysilo, _ = libs
    db = pysilo.Database.synthetic(4_000_000, 0x5110000D, genome_length=1000)
    engine = pysilo.Engine(db, 0)
    yield db, engine
    engine.close()
    db.close()


def leaves(db, n):
    """alternately 'carries the clade mutation' and 'carries something else there' for the first n clade mutations"""
    out = []
    for k, (pos, sym) in enumerate(db.clade_mutations()[:n]):
        leaf = {"type": "NucleotideEquals", "position": pos + 1, "symbol": sym}
        out.append(leaf if k % 2 == 0 else {"type": "Not", "child": leaf})
    return out


def assert_row_sets_equal(db, engine, oracle, flt):
    matched = 0
    for p in range(db.n_partitions):
        got = np.asarray(engine.filter_ids(flt, p))
        assert np.array_equal(got, np.asarray(oracle.filter_ids(db, flt, p))), p
        matched += len(got)
    return matched


def test_wide_thresholds_on_multi_slab_partitions(libs, four_million):
    _, oracle = libs
    db, engine = four_million
    assert max(db.partition_rows(p) for p in range(db.n_partitions)) > 262144
    children = leaves(db, 60)
    assert len(children) == 60
    nonempty = 0
    for k, exactly in ((25, False), (29, True), (31, True), (47, False)):
        nonempty += assert_row_sets_equal(db, engine, oracle, {"type": "N-Of", "numberOfMatchers": k, "matchExactly": exactly, "children": children}) > 0
    assert nonempty >= 3


def test_right_deep_trees_on_multi_slab_partitions(libs, four_million):
    _, oracle = libs
    db, engine = four_million
    pool = leaves(db, 90)
    e = pool[0]
    for level in range(1, 45):
        other = [pool[(2 * level) % len(pool)], pool[(2 * level + 1) % len(pool)]]
        if level % 3 == 0:
            e = {"type": "And", "children": [other[0], e]}
        elif level % 3 == 1:
            e = {"type": "Or", "children": other + [e]}
        else:
            e = {"type": "N-Of", "numberOfMatchers": 2, "matchExactly": level % 2 == 0, "children": other + [e]}
    matched = assert_row_sets_equal(db, engine, oracle, e)
    assert 0 < matched < 4_000_000
