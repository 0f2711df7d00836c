"""Known answers of the reference's column unit tests, asked through the query path: each test builds the column of the unit test
as a one-partition database directory (rows keep their insertion order: keys ascend, no dateToSortBy, no partitionBy) and asks
for what the unit test reads off the column with a filter (-> row ids) or with Aggregated + groupByFields (-> printed values).
  pango_lineage_column.test.cpp:7-104   filter / filterIncludingSublineages / lookupAliasedValue
  indexed_string_column.test.cpp:9-48   filter, values
  string.test.cpp:62-96                 equality of strings that agree on a prefix of up to 32 characters
  date.test.cpp:7-31                    stringToDate / dateToString
  insertion_column.test.cpp:31-62       search(sequence, position, pattern)
  string_column.test.cpp:8-25           values come back as inserted
and, worked out by hand from int_between.cpp / int_equals.cpp / float_between.cpp / float_equals.cpp, the null and
bound rules of the numeric selections.
The CPU tests pin the oracle on these answers, the GPU tests the device path on the same ones."""
import json
from pathlib import Path

import pytest

GOLDEN = Path(__file__).parent / "golden"


def make_db(pysilo, directory, column_name, column_yaml, values, alias_file=None, schema_tail=""):
    (directory / "database_config.yaml").write_text(
        "schema:\n  instanceName: column_kat\n  metadata:\n    - name: key\n      type: string\n"
        f"    - name: {column_name}\n{column_yaml}  primaryKey: key\n{schema_tail}")
    pre = 'metadataFilename: "metadata.tsv"\nreferenceGenomeFilename: "reference_genomes.json"\n'
    if alias_file is not None:
        (directory / "pangolineage_alias.json").write_text(alias_file.read_text())
        pre += 'pangoLineageDefinitionFilename: "pangolineage_alias.json"\n'
    (directory / "preprocessing_config.yaml").write_text(pre)
    (directory / "reference_genomes.json").write_text(json.dumps({"nucleotideSequences": [{"name": "main", "sequence": "ACGT"}], "genes": []}))
    keys = [f"k{i:04d}" for i in range(len(values))]
    (directory / "metadata.tsv").write_text(f"key\t{column_name}\n" + "".join(f"{k}\t{v}\n" for k, v in zip(keys, values)))
    (directory / "nuc_main.fasta").write_text("".join(f">{k}\nACGT\n" for k in keys))
    return pysilo.Database.from_directory(directory)


class OnOracle:
    def __init__(self, oracle, db):
        self.oracle, self.db = oracle, db

    def ids(self, flt):
        return list(self.oracle.filter_ids(self.db, flt, 0))

    def execute(self, query):
        return self.oracle.execute(self.db, query)

    def close(self):
        pass


class OnDevice:
    def __init__(self, pysilo, db):
        self.engine = pysilo.Engine(db, 0)

    def ids(self, flt):
        return list(self.engine.filter_ids(flt, 0))

    def execute(self, query):
        return self.engine.execute(query)

    def close(self):
        self.engine.close()


BACKENDS = [pytest.param("oracle", id="oracle"), pytest.param("device", id="device", marks=pytest.mark.gpu)]


@pytest.fixture(params=BACKENDS)
def backend(request, libs):
    pysilo, oracle = libs
    made = []

    def open_on(db):
        b = OnOracle(oracle, db) if request.param == "oracle" else OnDevice(pysilo, db)
        made.append(b)
        return b
    yield open_on
    for b in made:
        b.close()


def lineage(value, sub):
    return {"type": "PangoLineage", "column": "pango_lineage", "value": value, "includeSublineages": sub}


def grouped_values(b, field):
    rows = b.execute({"action": {"type": "Aggregated", "groupByFields": [field], "orderByFields": [field]}, "filterExpression": {"type": "True"}})
    return [(r[field], r["count"]) for r in rows]


PANGO_YAML = "      type: pango_lineage\n"
ALIAS = GOLDEN / "exampleDataset" / "pangolineage_alias.json"


def test_lineage_then_sublineage(libs, backend, tmp_path):
    """pango_lineage_column.test.cpp:7-31"""
    b = backend(make_db(libs[0], tmp_path, "pango_lineage", PANGO_YAML, ["A.1.2", "A.1.2", "A.1.2.3", "A.1.2.3.4", "A.1.2"], ALIAS))
    assert b.ids(lineage("A.1.2", False)) == [0, 1, 4]
    assert b.ids(lineage("A.1.2", True)) == [0, 1, 2, 3, 4]
    assert b.ids(lineage("A.1.2.3", False)) == [2]
    assert b.ids(lineage("A.1.2.3", True)) == [2, 3]


def test_sublineage_then_lineage(libs, backend, tmp_path):
    """pango_lineage_column.test.cpp:33-58"""
    b = backend(make_db(libs[0], tmp_path, "pango_lineage", PANGO_YAML, ["A.1.2.3", "A.1.2.3", "A.1", "A.1.2", "A.1.2.3"], ALIAS))
    assert b.ids(lineage("A.1.2", False)) == [3]
    assert b.ids(lineage("A.1.2", True)) == [0, 1, 3, 4]
    assert b.ids(lineage("A.1.2.3", False)) == [0, 1, 4]
    assert b.ids(lineage("A.1.2.3", True)) == [0, 1, 4]


def test_parent_lineage_that_was_never_inserted(libs, backend, tmp_path):
    """pango_lineage_column.test.cpp:60-77: filter() has no bitmap for "A.1" (-> no rows), the sublineage filter has"""
    b = backend(make_db(libs[0], tmp_path, "pango_lineage", PANGO_YAML, ["A.1.2.3", "A.1.2.3", "A.2", "A.1.2"], ALIAS))
    assert b.ids(lineage("A.1", False)) == []
    assert b.ids(lineage("A.1", True)) == [0, 1, 3]
    assert b.ids(lineage("Z.9", True)) == [] and b.ids(lineage("Z.9", False)) == []


def test_aliased_lookup_values_are_what_group_by_prints(libs, backend, tmp_path):
    """pango_lineage_column.test.cpp:79-103: rows B.1.1, C (= B.1.1.1), B.1.1.1.1 hold ids 2, 3, 4, printed as B.1.1, B.1.1.1, C.1;
    and a query for the alias finds what was inserted under the full name"""
    b = backend(make_db(libs[0], tmp_path, "pango_lineage", PANGO_YAML, ["B.1.1", "C", "B.1.1.1.1"], ALIAS))
    assert grouped_values(b, "pango_lineage") == [("B.1.1", 1), ("B.1.1.1", 1), ("C.1", 1)]
    assert b.ids(lineage("C", False)) == [1] and b.ids(lineage("B.1.1.1", False)) == [1]
    assert b.ids(lineage("C", True)) == [1, 2] and b.ids(lineage("C.1", False)) == [2]
    assert b.ids(lineage("B", True)) == [0, 1, 2] and b.ids(lineage("B", False)) == []


def test_indexed_string_column(libs, backend, tmp_path):
    """indexed_string_column.test.cpp:9-48"""
    b = backend(make_db(libs[0], tmp_path, "division", "      type: string\n      generateIndex: true\n", ["value 1", "value 2", "value 2", "value 3", "value 1"]))
    equals = lambda v: {"type": "StringEquals", "column": "division", "value": v}
    assert b.ids(equals("value 1")) == [0, 4]
    assert b.ids(equals("value 2")) == [1, 2]
    assert b.ids(equals("value that does not exist")) == []
    assert grouped_values(b, "division") == [("value 1", 2), ("value 2", 2), ("value 3", 1)]


def test_strings_that_share_a_prefix_of_up_to_32_characters(libs, backend, tmp_path):
    """string.test.cpp:62-96: prefix + "x" differs from prefix + "y" for every prefix length 0..31 (the inline part of String<N> and
    its dictionary suffix), and equals itself; also the very long values of :29-60"""
    value = "1234567890abcdefghijklmnopqrstuv"
    very_long = ("some very long value 1 2 3 4 5 6 7 8 9 0 1 2 3 4 5 6 7 8 9 0 1 2 3 4 5 6 7 8 9 0 1 2 3 4 5 "
                 "6 7 8 9 0 1 2 3 4 5 6 7 8 9 0 1 2 3 4 5 6 7 8 9 0 1 2 3 4 5 6 7 8 9 0 1 2 3 4 5 6 7 8 9 0")
    rows = []
    for i in range(len(value)):
        rows += [value[:i] + "x", value[:i] + "y"]
    rows += [very_long, very_long + " different", very_long, "value 1", "some longer value 1"]
    b = backend(make_db(libs[0], tmp_path, "name", "      type: string\n", rows))
    equals = lambda v: {"type": "StringEquals", "column": "name", "value": v}
    for i in range(len(value)):
        assert b.ids(equals(value[:i] + "x")) == [2 * i]
        assert b.ids(equals(value[:i] + "y")) == [2 * i + 1]
        assert b.ids(equals(value[:i] + "z")) == []
    n = 2 * len(value)
    assert b.ids(equals(very_long)) == [n, n + 2]
    assert b.ids(equals(very_long + " different")) == [n + 1]
    assert b.ids(equals(very_long[:-1])) == []
    assert b.ids(equals("value 1")) == [n + 3] and b.ids(equals("some longer value 1")) == [n + 4]
    assert b.ids({"type": "Not", "child": equals(very_long)}) == [r for r in range(len(rows)) if r not in (n, n + 2)]
    printed = dict(grouped_values(b, "name"))
    assert printed[very_long] == 2 and printed[very_long + " different"] == 1 and len(printed) == len(rows) - 1


def test_dates_parse_and_print_as_in_the_reference(libs, backend, tmp_path):
    """date.test.cpp:7-31: y-m-d with any digit counts parses, everything else is the null date; printing pads to yyyy-mm-dd"""
    rows = ["2020-01-01", "2023-1-01", "2010-12-03", "12-12-12", "", "?", "----", "31-31-31", "-1-", "31-31123123-31123123123123412", "31-0-1", "2023-01-1", "2010-12-3"]
    b = backend(make_db(libs[0], tmp_path, "date", "      type: date\n", rows))
    assert grouped_values(b, "date") == [(None, 7), ("0012-12-12", 1), ("2010-12-03", 2), ("2020-01-01", 1), ("2023-01-01", 2)]
    between = lambda lo, hi: {"type": "DateBetween", "column": "date", "from": lo, "to": hi}
    # a date column that is not the dateToSortBy column compiles to Selection{>= from, < to} (date_between.cpp:53-71): the upper bound
    # is exclusive there, unlike the RangeSelection of the sorted column (:88-93, upper_bound); null dates (0) lie below from = 1
    assert b.ids(between("2010-12-03", "2010-12-03")) == []
    assert b.ids(between("2010-12-03", "2010-12-04")) == [2, 12]
    assert b.ids(between("2020-01-01", "2023-01-01")) == [0]
    assert b.ids(between(None, "2010-12-03")) == [3]
    assert b.ids(between("2023-01-01", None)) == [1, 11]
    assert b.ids(between("2023-01-02", None)) == []
    assert b.ids(between(None, None)) == [0, 1, 2, 3, 11, 12]


def test_sorted_date_column_takes_the_range_selection(libs, backend, tmp_path):
    """the dateToSortBy column: rows are ordered by (date, key) with null dates first, DateBetween compiles to a RangeSelection from
    lower_bound(from or 1) to upper_bound(to) (date_between.cpp:73-96) -- the upper bound is inclusive here"""
    rows = ["2020-01-01", "2023-1-01", "2010-12-03", "12-12-12", "", "2010-12-3", "?"]
    b = backend(make_db(libs[0], tmp_path, "date", "      type: date\n", rows, schema_tail="  dateToSortBy: date\n"))
    # row ids after ordering: 0, 1 = null dates (k0004, k0006); 2 = 0012-12-12; 3, 4 = 2010-12-03 (k0002, k0005); 5 = 2020-01-01; 6 = 2023-01-01
    keys = lambda flt: [r["key"] for r in b.execute({"action": {"type": "Aggregated", "groupByFields": ["key"], "orderByFields": ["key"]}, "filterExpression": flt})]
    between = lambda lo, hi: {"type": "DateBetween", "column": "date", "from": lo, "to": hi}
    assert b.ids(between("2010-12-03", "2010-12-03")) == [3, 4] and keys(between("2010-12-03", "2010-12-03")) == ["k0002", "k0005"]
    assert b.ids(between("2020-01-01", "2023-01-01")) == [5, 6]
    assert b.ids(between(None, "2010-12-03")) == [2, 3, 4]
    assert b.ids(between("2023-01-02", None)) == []
    assert b.ids(between(None, None)) == [2, 3, 4, 5, 6] and keys({"type": "Not", "child": between(None, None)}) == ["k0004", "k0006"]
    assert grouped_values(b, "date") == [(None, 2), ("0012-12-12", 1), ("2010-12-03", 2), ("2020-01-01", 1), ("2023-01-01", 1)]


def test_insertion_column_search(libs, backend, tmp_path):
    """insertion_column.test.cpp:31-62"""
    rows = ["25701:ACCA", "25701:CCG", "25701:CCG", "25701:TTACAT,25701:ACCA", "25701:ACCA", "25701:TTACAT,25701:ACCA,25701:AGCTGTTCAG"]
    b = backend(make_db(libs[0], tmp_path, "nucleotideInsertions", "      type: insertion\n", rows))
    search = lambda position, pattern: {"type": "InsertionContains", "column": "nucleotideInsertions", "sequenceName": "main", "position": position, "value": pattern}
    assert b.ids(search(25701, ".*CC.*")) == [0, 1, 2, 3, 4, 5]
    assert b.ids(search(25701, ".*TTA.*CAT.*")) == [3, 5]
    assert b.ids(search(25701, ".*AGC.*TGT.*TCA.*G.*")) == [5]
    assert b.ids(search(25701, ".*AGC.*TG.*T.*T.*C.*AG.*")) == [5]
    assert b.ids(search(25701, ".*TTT.*AAA.*")) == []
    assert b.ids(search(100, ".*TTT.*AAA.*")) == []
    assert b.ids(search(100, ".*CC.*")) == []
    without_name = {"type": "InsertionContains", "column": "nucleotideInsertions", "position": 25701, "value": ".*TTA.*CAT.*"}      # the default sequence
    assert b.ids(without_name) == [3, 5]


def test_string_column_values_come_back(libs, backend, tmp_path):
    """string_column.test.cpp:8-25"""
    rows = ["value 1", "value 2", "value 2", "value 3", "some string that is a little longer 1", "value 1"]
    b = backend(make_db(libs[0], tmp_path, "name", "      type: string\n", rows))
    for value in set(rows):
        assert b.ids({"type": "StringEquals", "column": "name", "value": value}) == [r for r, v in enumerate(rows) if v == value]
    assert grouped_values(b, "name") == [("some string that is a little longer 1", 1), ("value 1", 2), ("value 2", 2), ("value 3", 1)]


def test_int_selections(libs, backend, tmp_path):
    """int_equals.cpp:65 (null = INT32_MIN), int_between.cpp:41-48 (lower bound INT32_MIN + 1 when open -> nulls never match; upper bound inclusive)"""
    rows = ["4", "50", "", "51", "-3", "50", "0"]
    b = backend(make_db(libs[0], tmp_path, "age", "      type: int\n", rows))
    equals = lambda v: {"type": "IntEquals", "column": "age", "value": v}
    between = lambda lo, hi: {"type": "IntBetween", "column": "age", "from": lo, "to": hi}
    assert b.ids(equals(50)) == [1, 5] and b.ids(equals(0)) == [6] and b.ids(equals(7)) == []
    assert b.ids(equals(None)) == [2]
    assert b.ids(between(4, 50)) == [0, 1, 5]
    assert b.ids(between(None, 4)) == [0, 4, 6]
    assert b.ids(between(50, None)) == [1, 3, 5]
    assert b.ids(between(None, None)) == [0, 1, 3, 4, 5, 6]
    assert b.ids(between(51, 50)) == []
    assert b.ids({"type": "Not", "child": between(None, None)}) == [2]
    assert grouped_values(b, "age") == [(None, 1), (-3, 1), (0, 1), (4, 1), (50, 2), (51, 1)]


def test_float_selections(libs, backend, tmp_path):
    """float_equals.cpp:67 (null = NaN, equal to nothing), float_between.cpp:46-62 (>= from, < to: the upper bound is exclusive; no bounds at
    all is "!= NaN", true for every row, nulls included)"""
    rows = ["0.5", "0.9", "", "0.98", "0.5", "0.0", "-1.5"]
    b = backend(make_db(libs[0], tmp_path, "qc", "      type: float\n", rows))
    equals = lambda v: {"type": "FloatEquals", "column": "qc", "value": v}
    between = lambda lo, hi: {"type": "FloatBetween", "column": "qc", "from": lo, "to": hi}
    assert b.ids(equals(0.5)) == [0, 4] and b.ids(equals(0.0)) == [5] and b.ids(equals(0.51)) == []
    assert b.ids(equals(None)) == []
    assert b.ids(between(0.5, 0.98)) == [0, 1, 4]
    assert b.ids(between(0.5, 0.5)) == []
    assert b.ids(between(None, 0.5)) == [5, 6]
    assert b.ids(between(0.9, None)) == [1, 3]
    assert b.ids(between(None, None)) == [0, 1, 2, 3, 4, 5, 6]
    assert b.ids({"type": "Not", "child": between(-100.0, 100.0)}) == [2]
    # Not of a Selection with ONE predicate inverts the comparator instead of complementing (selection.cpp:127-132, :201-224): the null
    # row, NaN, compares false both ways and is lost; with two predicates it is a Complement and the null row comes back (line above)
    assert b.ids({"type": "Not", "child": between(None, 0.9)}) == [1, 3]
    assert b.ids({"type": "Not", "child": between(0.5, None)}) == [5, 6]
    assert b.ids({"type": "Not", "child": {"type": "Not", "child": between(0.5, None)}}) == [0, 1, 3, 4]
    assert b.ids({"type": "Not", "child": {"type": "And", "children": [between(0.5, None), {"type": "True"}]}}) == [5, 6]
    assert b.ids({"type": "Not", "child": between(None, None)}) == []
    assert b.ids({"type": "Not", "child": equals(0.5)}) == [1, 2, 3, 5, 6]      # != is true for NaN
    assert b.ids({"type": "Not", "child": equals(None)}) == [0, 1, 2, 3, 4, 5, 6]
    assert grouped_values(b, "qc") == [(None, 1), (-1.5, 1), (0.0, 1), (0.5, 2), (0.9, 1), (0.98, 1)]


def test_a_database_without_rows(libs, backend, tmp_path):
    """no input rows at all: one partition of zero sequences; every action answers, nothing is counted, nothing divides by zero
    (aggregated.cpp:58-66 sums no cardinalities, mutations.cpp:36-63 keeps no partition, tuple maps stay empty)"""
    db = make_db(libs[0], tmp_path, "age", "      type: int\n", [])
    assert db.n_partitions == 1 and db.info()["sequenceCount"] == 0
    b = backend(db)
    everything, nothing = {"type": "True"}, {"type": "Not", "child": {"type": "True"}}
    leaf = {"type": "NucleotideEquals", "position": 1, "symbol": "A"}
    for flt in (everything, nothing, leaf, {"type": "Not", "child": leaf}, {"type": "IntBetween", "column": "age", "from": None, "to": 5},
                {"type": "N-Of", "numberOfMatchers": 1, "matchExactly": True, "children": [leaf, everything]}):
        assert b.ids(flt) == []
        assert b.execute({"action": {"type": "Aggregated"}, "filterExpression": flt}) == [{"count": 0}]
        assert b.execute({"action": {"type": "Aggregated", "groupByFields": ["age"], "orderByFields": ["age"]}, "filterExpression": flt}) == []
        assert b.execute({"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": flt}) == []
