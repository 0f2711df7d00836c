"""Pins the oracle (CPU restatement, oracle/) on the reference's own fixtures:
 * the 71 hot-path end-to-end query fixtures of endToEndTests/test/queries (exampleDataset, 100 seq; 8 of them Aggregated
   with groupByFields),
 * the invalid-query fixtures (HTTP 400 messages),
 * the operator known-answer tests of src/silo/query_engine/operators/*.test.cpp,
 * the size / container goldens of src/silo/database.test.cpp:74-111 and endToEndTests/test/info.test.js:27-37."""
import json

import pytest

from conftest import GOLDEN, golden_queries, load_case

KATS = json.loads((GOLDEN / "operator_kats.json").read_text())


@pytest.mark.parametrize("path", golden_queries(), ids=lambda p: p.stem)
def test_e2e_fixture(libs, example_db, path):
    _, oracle = libs
    case = load_case(path)
    assert oracle.execute(example_db, case["query"]) == case["expectedQueryResult"]


@pytest.mark.parametrize("path", sorted((GOLDEN / "invalidQueries").glob("*.json")), ids=lambda p: p.stem)
def test_invalid_query_fixture(libs, example_db, path):
    _, oracle = libs
    case = load_case(path)
    action = case["query"].get("action", {}).get("type")
    with pytest.raises(oracle.OracleError) as err:
        oracle.execute(example_db, case["query"])
    if action in ("Details", "Fasta", "FastaAligned", "Insertions", "AminoAcidInsertions"):
        pytest.skip("action outside the hot path (tests/golden/queries_out_of_scope.txt)")
    assert err.value.kind == "parse"                      # QueryParseException -> HTTP 400
    assert err.value.message == case["expectedError"]["message"]


@pytest.mark.parametrize("case", KATS, ids=lambda c: c["name"])
def test_operator_kat(libs, case):
    _, oracle = libs
    if case.get("throws"):
        with pytest.raises(oracle.OracleError):
            oracle.eval_optree(case["tree"])
        return
    assert oracle.eval_optree(case["tree"]).tolist() == case["expected"]


def test_size_goldens(example_db):
    info = example_db.info()
    assert info["sequenceCount"] == 100
    assert info["numberOfPartitions"] == 11
    assert sorted(info["partitionRows"]) == sorted([6, 4, 5, 50, 7, 10, 5, 4, 4, 4, 1])
    # database.test.cpp:109-111 (both nucleotide segments of the dataset are counted)
    main = info["sequences"]["main"]
    assert main["numberOfArrayContainers"] == 3065 and main["numberOfValuesStoredInArrayContainers"] == 4377
    assert main["numberOfRunContainers"] == 3 and main["numberOfValuesStoredInRunContainers"] == 9
    assert main["numberOfBitsetContainers"] == 0
    assert main["totalBitmapSizeArrayContainers"] == 8754
    assert main["totalBitmapSizeComputed"] == 42136719
