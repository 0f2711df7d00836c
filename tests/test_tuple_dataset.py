"""The three-row column group of the reference's tuple unit test (src/silo/query_engine/actions/tuple.test.cpp:10-95: a string, an indexed
string, an int, a float and a date column; rows 0 and 2 equal) as a database directory, queried through Aggregated with groupByFields.
What the unit test asserts on Tuple -- the field values (:99-176), equality of rows 0 and 2 and inequality with row 1 (:188-201), the
comparator per order-by field and direction (:203-299) -- is asserted here on the response rows: equal tuples fall into one group,
orderByFields orders the groups. The CPU test pins the oracle, the GPU test the device path, both on the same expectations."""
import json

import pytest

LONG = ("some very long string some very long string some very long string some very long string "
        "some very long string some very long string some very long string some very long string "
        "some very long string ")
COLUMNS = ["dummy_string_column", "dummy_indexed_string_column", "dummy_int_column", "dummy_float_column", "dummy_date_column"]
ROWS = [("key0", "ABCD", "ABCD", "42", "42.1", "2023-01-01"),
        ("key1", LONG, LONG, "-12389172", "-12389172.24222", "2023-01-01"),
        ("key2", "ABCD", "ABCD", "42", "42.1", "2023-01-01")]
GROUP_A = {"dummy_string_column": "ABCD", "dummy_indexed_string_column": "ABCD", "dummy_int_column": 42, "dummy_float_column": 42.1,
           "dummy_date_column": "2023-01-01", "count": 2}
GROUP_B = {"dummy_string_column": LONG, "dummy_indexed_string_column": LONG, "dummy_int_column": -12389172, "dummy_float_column": -12389172.24222,
           "dummy_date_column": "2023-01-01", "count": 1}


@pytest.fixture(scope="module")
def tuple_db(libs, tmp_path_factory):
    pysilo, _ = libs
    d = tmp_path_factory.mktemp("tuple_dataset")
    (d / "database_config.yaml").write_text(
        "schema:\n  instanceName: tuple_test\n  metadata:\n    - name: key\n      type: string\n"
        "    - name: dummy_string_column\n      type: string\n"
        "    - name: dummy_indexed_string_column\n      type: string\n      generateIndex: true\n"
        "    - name: dummy_int_column\n      type: int\n"
        "    - name: dummy_float_column\n      type: float\n"
        "    - name: dummy_date_column\n      type: date\n"
        "  primaryKey: key\n")
    (d / "preprocessing_config.yaml").write_text('metadataFilename: "metadata.tsv"\nreferenceGenomeFilename: "reference_genomes.json"\n')
    (d / "reference_genomes.json").write_text(json.dumps({"nucleotideSequences": [{"name": "main", "sequence": "ACGTACGT"}], "genes": []}))
    (d / "metadata.tsv").write_text("\t".join(["key"] + COLUMNS) + "\n" + "".join("\t".join(r) + "\n" for r in ROWS))
    (d / "nuc_main.fasta").write_text("".join(f">{r[0]}\nACGTACG{'T' if i != 1 else 'A'}\n" for i, r in enumerate(ROWS)))
    return pysilo.Database.from_directory(d)


def grouped(fields, **extra):
    action = {"type": "Aggregated", "groupByFields": fields}
    action.update(extra)
    return {"action": action, "filterExpression": {"type": "True"}}


# (group-by fields, orderByFields, expected rows restricted to those fields) -- the comparator cases of tuple.test.cpp:203-299
CASES = [
    (COLUMNS, ["dummy_indexed_string_column"], [GROUP_A, GROUP_B]),                                             # :213-232 ascending: "ABCD" before "some ..."
    (COLUMNS, [{"field": "dummy_indexed_string_column", "order": "descending"}], [GROUP_B, GROUP_A]),
    (COLUMNS, ["dummy_date_column", {"field": "dummy_string_column", "order": "descending"}], [GROUP_B, GROUP_A]),   # :278-298 equal dates, then the string descending
    (COLUMNS, ["dummy_int_column"], [GROUP_B, GROUP_A]),
    (COLUMNS, [{"field": "dummy_float_column", "order": "descending"}], [GROUP_A, GROUP_B]),
    (["dummy_date_column"], ["dummy_date_column"], [{"dummy_date_column": "2023-01-01", "count": 3}]),          # :255-276 the date does not tell the rows apart
    (["dummy_int_column", "dummy_float_column"], ["dummy_float_column"], [GROUP_B, GROUP_A]),
]


def restrict(row, fields):
    return {k: v for k, v in row.items() if k in fields or k == "count"}


def check(execute):
    for fields, order, want in CASES:
        assert execute(grouped(fields, orderByFields=order)) == [restrict(r, fields) for r in want]
    unordered = execute(grouped(COLUMNS))                                                                        # :235-253 no order-by fields: any order
    assert sorted(unordered, key=lambda r: r["count"]) == [GROUP_B, GROUP_A]
    assert execute(grouped(COLUMNS, orderByFields=["dummy_int_column"], limit=1)) == [GROUP_B]
    assert execute(grouped(COLUMNS, orderByFields=["dummy_int_column"], offset=1)) == [GROUP_A]
    # the third row joins the first only while the filter lets it: rows whose last nucleotide is T are rows 0 and 2
    only_t = {"action": {"type": "Aggregated", "groupByFields": COLUMNS}, "filterExpression": {"type": "NucleotideEquals", "position": 8, "symbol": "T"}}
    assert execute(only_t) == [GROUP_A]
    only_a = {"action": {"type": "Aggregated", "groupByFields": COLUMNS}, "filterExpression": {"type": "NucleotideEquals", "position": 8, "symbol": "A"}}
    assert execute(only_a) == [GROUP_B]


def test_oracle_on_the_tuple_test_dataset(libs, tuple_db):
    _, oracle = libs
    check(lambda q: oracle.execute(tuple_db, q))


@pytest.mark.gpu
def test_device_on_the_tuple_test_dataset(libs, tuple_db):
    pysilo, _ = libs
    engine = pysilo.Engine(tuple_db, 0)
    try:
        check(engine.execute)
    finally:
        engine.close()
