"""Aggregated with groupByFields on the device (sb_aggregated_grouped) against the oracle's restatement of
aggregated.cpp:91-149 / tuple.cpp: the reference's eight fixtures run in test_gpu_parity.py::test_e2e_fixture; here the
three counting paths -- shared-memory bins, global bins, hash table -- on synthetic databases, every column type, several
columns at once, ordering and limits."""
import pytest

pytestmark = pytest.mark.gpu

CLADE_FILTER = {"type": "N-Of", "numberOfMatchers": 1, "matchExactly": False, "children": []}


def canonical(rows):
    return sorted(rows, key=lambda r: sorted((k, (v is None, v)) for k, v in r.items() if k != "count") + [("count", (False, r["count"]))])


def grouped(fields, flt, **extra):
    action = {"type": "Aggregated", "groupByFields": fields}
    action.update(extra)
    return {"action": action, "filterExpression": flt}


@pytest.fixture(scope="module")
def medium_engine(libs, synth_medium):
    pysilo, _ = libs
    engine = pysilo.Engine(synth_medium, 0, synth_medium.owned)
    yield engine
    engine.close()


def filters(db):
    clade = db.clade_mutations()
    some = {"type": "N-Of", "numberOfMatchers": 2, "matchExactly": False,
            "children": [{"type": "NucleotideEquals", "position": p + 1, "symbol": s} for p, s in clade[:6]]}
    return [{"type": "True"}, some, {"type": "Not", "child": some}, {"type": "IntBetween", "column": "age", "from": 30, "to": 31},
            {"type": "And", "children": [some, {"type": "Not", "child": {"type": "True"}}]}]


@pytest.mark.parametrize("fields", [["region"], ["pangoLineage"], ["age"], ["date"], ["qc"], ["region", "pangoLineage"], ["age", "region", "pangoLineage"],
                                    ["date", "age"], ["date", "qc", "age", "pangoLineage"]], ids=lambda f: "+".join(f))
def test_grouped_counts_equal_the_oracle(libs, synth_medium, medium_engine, fields):
    _, oracle = libs
    for flt in filters(synth_medium):
        query = grouped(fields, flt)
        got, want = medium_engine.execute(query), oracle.execute(synth_medium, query)
        assert canonical(got) == canonical(want)
        assert sum(r["count"] for r in got) == medium_engine.execute({"action": {"type": "Aggregated"}, "filterExpression": flt})[0]["count"]


def test_grouped_order_limit_offset(libs, synth_medium, medium_engine):
    _, oracle = libs
    flt = filters(synth_medium)[1]
    for extra in [{"orderByFields": ["pangoLineage", "region"]}, {"orderByFields": [{"field": "count", "order": "descending"}, "pangoLineage", "region"], "limit": 7},
                  {"orderByFields": ["region", {"field": "pangoLineage", "order": "descending"}], "limit": 5, "offset": 3}]:
        query = grouped(["region", "pangoLineage"], flt, **extra)
        assert medium_engine.execute(query) == oracle.execute(synth_medium, query)      # a total order: the lists are equal as they come


def test_grouped_errors_are_the_reference_s(libs, synth_medium, medium_engine):
    pysilo, _ = libs
    with pytest.raises(pysilo.SiloError) as err:
        medium_engine.execute(grouped(["nope"], {"type": "True"}))
    assert err.value.kind == "parse" and err.value.message == "Metadata field 'nope' to group by not found"
    with pytest.raises(pysilo.SiloError) as err:
        medium_engine.execute(grouped(["region"], {"type": "True"}, orderByFields=["age"]))
    assert err.value.kind == "parse" and err.value.message == "The orderByField 'age' cannot be ordered by, as it does not appear in the groupByFields."
