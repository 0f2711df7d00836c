"""Random filter expressions against a row-by-row model written from the reference's documented semantics.

The FilterExpression -> Operator compile rules (and.cpp, or.cpp, negation.cpp, nof.cpp, the leaf expressions; flipped and deleted
symbols, Selection / Intersection merging, De Morgan rewrites, Full / Empty shortcuts) live in host code that the oracle and the device
path share. Here the expected row set of an expression does not come from that code: `matches()` below decides for ONE ROW whether
the expression holds, from the row's strings and values, carrying the ambiguity mode down the tree the way Expression::compile does
(expression.cpp:38-46, negation.cpp:32, maybe.cpp, exact.cpp). 48 random rows, several hundred random trees of depth <= 4.
The Mutations / AminoAcidMutations actions get the same treatment: `mutation_rows()` counts symbols over the rows the model selects and
applies the output rule of mutations.cpp:184-229 (coverage over the valid mutation symbols only, count > ceil(total * p) - 1, the
reference symbol never printed, positions ascending and symbols in VALID_MUTATION_SYMBOLS order).
CPU: oracle. GPU: device path."""
import json
import random

import pytest

BACKENDS = [pytest.param("oracle", id="oracle"), pytest.param("device", id="device", marks=pytest.mark.gpu)]
NUC_REF, GENE_REF = "ACGTAC", "MKLV"
N_ROWS = 48
# nucleotide_symbol_equals.cpp:28-73: the symbols that may stand for a base
COULD_BE = {"A": "ARMWDHVN", "C": "CYMSBHVN", "G": "GRKSBDVN", "T": "TYKWBDHN"}
AA_SYMBOLS = "-ACDEFGHIKLMNPQRSTVWYBZX*"


def make_rows(rng):
    rows = []
    for r in range(N_ROWS):
        nuc = "".join(rng.choices([ref, "ACGT"[rng.randrange(4)], "N", "-", "R", "Y", "K"], weights=[60, 18, 8, 5, 4, 3, 2])[0] for ref in NUC_REF)
        gene = "".join(rng.choices([ref, "AKLMV"[rng.randrange(5)], "X", "B", "-"], weights=[60, 25, 8, 4, 3])[0] for ref in GENE_REF)
        if r % 11 == 7:
            nuc, gene = "N" * len(NUC_REF), "X" * len(GENE_REF)
        rows.append({
            "key": f"k{r:03d}", "nuc": nuc, "gene": gene,
            "age": rng.choice([None, None, 3, 17, 17, 42, 50, 51, -4, 0]),
            "qc": rng.choice([None, 0.0, 0.25, 0.5, 0.5, 0.9, 0.98, -1.5]),
            "region": rng.choice(["", "Europe", "Europe", "Asia", "Africa"]),
            "name": rng.choice(["", "alpha", "alpha", "a rather long name that does not fit the inline part", "a rather long name that does not fit the inline parT", "beta"]),
            "lineage": rng.choice(["", "A", "B", "B.1", "B.1", "B.1.1", "B.1.1.7", "B.2", "C.3"]),
            "date": rng.choice(["", "2020-01-01", "2020-06-15", "2021-03-03", "2021-03-03", "2022-12-31"]),
        })
    return rows


def build_db(libs, d, rows, schema_tail=""):
    (d / "database_config.yaml").write_text(
        "schema:\n  instanceName: model\n  metadata:\n    - name: key\n      type: string\n    - name: age\n      type: int\n    - name: qc\n      type: float\n"
        "    - name: region\n      type: string\n      generateIndex: true\n    - name: name\n      type: string\n"
        "    - name: lineage\n      type: pango_lineage\n    - name: date\n      type: date\n  primaryKey: key\n" + schema_tail)
    (d / "preprocessing_config.yaml").write_text('metadataFilename: "metadata.tsv"\nreferenceGenomeFilename: "reference_genomes.json"\n')
    (d / "reference_genomes.json").write_text(json.dumps({"nucleotideSequences": [{"name": "main", "sequence": NUC_REF}], "genes": [{"name": "G", "sequence": GENE_REF}]}))
    text = lambda v: "" if v is None else str(v)
    (d / "metadata.tsv").write_text("key\tage\tqc\tregion\tname\tlineage\tdate\n" + "".join(
        "\t".join([r["key"], text(r["age"]), text(r["qc"]), r["region"], r["name"], r["lineage"], r["date"]]) + "\n" for r in rows))
    (d / "nuc_main.fasta").write_text("".join(f">{r['key']}\n{r['nuc']}\n" for r in rows))
    (d / "gene_G.fasta").write_text("".join(f">{r['key']}\n{r['gene']}\n" for r in rows))
    return libs[0].Database.from_directory(d)


@pytest.fixture(scope="module")
def model_db(libs, tmp_path_factory):
    rows = make_rows(random.Random(20240511))
    return build_db(libs, tmp_path_factory.mktemp("filter_model"), rows), rows


def swapped(mode):
    return {"NONE": "NONE", "UPPER": "LOWER", "LOWER": "UPPER"}[mode]


def matches(e, row, mode="NONE"):
    t = e["type"]
    if t == "True":
        return True
    if t == "False":
        return False
    if t == "And":
        return all(matches(c, row, mode) for c in e["children"])
    if t == "Or":
        return any(matches(c, row, mode) for c in e["children"])
    if t == "Not":
        return not matches(e["child"], row, swapped(mode))
    if t == "Maybe":
        return matches(e["child"], row, "UPPER")
    if t == "Exact":
        return matches(e["child"], row, "LOWER")
    if t == "N-Of":
        count = sum(matches(c, row, mode) for c in e["children"])
        return count == e["numberOfMatchers"] if e["matchExactly"] else count >= e["numberOfMatchers"]
    if t == "NucleotideEquals":
        have, want = row["nuc"][e["position"] - 1], "-" if e["symbol"] == "." else e["symbol"]
        if mode == "UPPER" and want in COULD_BE:
            return have in COULD_BE[want]
        return have == want
    if t == "AminoAcidEquals":
        have = row["gene"][e["position"] - 1]
        return have == (GENE_REF[e["position"] - 1] if e["symbol"] == "." else e["symbol"])
    if t == "HasNucleotideMutation":
        have, ref = row["nuc"][e["position"] - 1], NUC_REF[e["position"] - 1]
        if mode == "UPPER":
            return have != ref
        return have in ("ACGT" if ref == "T" else "ACGT".replace(ref, ""))      # std::remove without erase: quirk Q2
    if t == "HasAminoAcidMutation":
        have, ref = row["gene"][e["position"] - 1], GENE_REF[e["position"] - 1]
        if mode == "UPPER":
            return have != ref
        return have != "X" and (have != ref or ref == "*")                      # quirk Q3
    if t == "IntEquals":
        return row["age"] == e["value"]
    if t == "IntBetween":
        return row["age"] is not None and (e["from"] is None or row["age"] >= e["from"]) and (e["to"] is None or row["age"] <= e["to"])
    if t == "FloatEquals":
        return row["qc"] is not None and e["value"] is not None and row["qc"] == e["value"]
    if t == "FloatBetween":
        if e["from"] is None and e["to"] is None:
            return True
        return row["qc"] is not None and (e["from"] is None or row["qc"] >= e["from"]) and (e["to"] is None or row["qc"] < e["to"])
    if t == "StringEquals":
        return row[e["column"]] == e["value"]
    if t == "PangoLineage":
        want, have = e["value"].upper(), row["lineage"]
        return have == want or (e["includeSublineages"] and have.startswith(want + ".") and want != "")
    if t == "DateBetween":      # quirk Q14: on a column rows are not ordered by >= from, < to; on the dateToSortBy column `to` is inclusive
        below_to = (lambda d: d <= e["to"]) if row.get("date_is_sorted") else (lambda d: d < e["to"])
        return row["date"] != "" and (e["from"] is None or row["date"] >= e["from"]) and (e["to"] is None or below_to(row["date"]))
    raise AssertionError(t)


def random_leaf(rng):
    kind = rng.randrange(14)
    if kind <= 2:
        return {"type": "NucleotideEquals", "position": rng.randint(1, len(NUC_REF)), "symbol": rng.choice("ACGTACGTN-RYK.")}
    if kind <= 4:
        return {"type": "AminoAcidEquals", "sequenceName": "G", "position": rng.randint(1, len(GENE_REF)), "symbol": rng.choice("MKLVAXB-.")}
    if kind == 5:
        return {"type": "HasNucleotideMutation", "position": rng.randint(1, len(NUC_REF))}
    if kind == 6:
        return {"type": "HasAminoAcidMutation", "sequenceName": "G", "position": rng.randint(1, len(GENE_REF))}
    if kind == 7:
        lo, hi = rng.choice([None, -4, 3, 17, 42]), rng.choice([None, 17, 50, 51, 2])
        return rng.choice([{"type": "IntBetween", "column": "age", "from": lo, "to": hi}, {"type": "IntEquals", "column": "age", "value": rng.choice([None, 17, 42, 5])}])
    if kind == 8:
        # both bounds or none: with ONE bound the Selection holds a single predicate, and Not of such a Selection inverts the comparator
        # (selection.cpp:127-132, :201-224) instead of complementing the rows -- null (NaN) rows then match neither the filter nor its
        # negation; that rule is spelled out in tests/test_column_kats.py::test_float_selections, the model here stays row-wise
        lo, hi = rng.choice([(None, None), (-1.5, 0.5), (0.25, 0.9), (0.5, 0.98), (0.5, 1.0), (0.9, 0.5)])
        return rng.choice([{"type": "FloatBetween", "column": "qc", "from": lo, "to": hi}, {"type": "FloatEquals", "column": "qc", "value": rng.choice([None, 0.5, 0.9, 0.3])}])
    if kind == 9:
        return {"type": "StringEquals", "column": "region", "value": rng.choice(["Europe", "Asia", "Africa", "", "Oceania"])}
    if kind == 10:
        return {"type": "StringEquals", "column": "name", "value": rng.choice(["alpha", "beta", "", "gamma", "a rather long name that does not fit the inline part", "a rather long name that does not fit the inline parT"])}
    if kind == 11:
        return {"type": "PangoLineage", "column": "lineage", "value": rng.choice(["B", "b.1", "B.1.1", "A", "C", "C.3", "B.1.1.7", "B.3", ""]), "includeSublineages": rng.random() < 0.6}
    if kind == 12:
        lo, hi = rng.choice([None, "2020-01-01", "2020-06-15", "2021-03-03"]), rng.choice([None, "2020-06-15", "2021-03-03", "2021-03-04", "2023-01-01"])
        return {"type": "DateBetween", "column": "date", "from": lo, "to": hi}
    return {"type": rng.choice(["True", "False"])}


def random_tree(rng, depth):
    if depth == 0 or rng.random() < 0.25:
        return random_leaf(rng)
    kind = rng.randrange(10)
    if kind <= 2:
        return {"type": "And", "children": [random_tree(rng, depth - 1) for _ in range(rng.randint(1, 4))]}
    if kind <= 4:
        return {"type": "Or", "children": [random_tree(rng, depth - 1) for _ in range(rng.randint(1, 4))]}
    if kind == 5:
        return {"type": "Not", "child": random_tree(rng, depth - 1)}
    if kind == 6:
        return {"type": rng.choice(["Maybe", "Exact"]), "child": random_tree(rng, depth - 1)}
    children = [random_tree(rng, depth - 1) for _ in range(rng.randint(1, 5))]
    return {"type": "N-Of", "numberOfMatchers": rng.randint(0, len(children) + 1), "matchExactly": rng.random() < 0.5, "children": children}


@pytest.fixture(params=BACKENDS)
def ids(request, libs, model_db):
    pysilo, oracle = libs
    db, _ = model_db
    if request.param == "oracle":
        yield lambda flt: list(oracle.filter_ids(db, flt, 0))
    else:
        engine = pysilo.Engine(db, 0)
        yield lambda flt: list(engine.filter_ids(flt, 0))
        engine.close()


def test_every_leaf_kind_against_the_model(ids, model_db):
    _, rows = model_db
    rng = random.Random(7)
    seen = set()
    for _ in range(600):
        leaf = random_leaf(rng)
        text = json.dumps(leaf, sort_keys=True)
        if text in seen:
            continue
        seen.add(text)
        for wrap in (lambda e: e, lambda e: {"type": "Maybe", "child": e}, lambda e: {"type": "Not", "child": e}):
            e = wrap(leaf)
            assert ids(e) == [r for r in range(N_ROWS) if matches(e, rows[r])], json.dumps(e)
    assert len(seen) > 150


@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_random_trees_against_the_model(ids, model_db, seed):
    _, rows = model_db
    rng = random.Random(1000 + seed)
    nonempty = 0
    for _ in range(150):
        e = random_tree(rng, rng.randint(1, 4))
        want = [r for r in range(N_ROWS) if matches(e, rows[r])]
        assert ids(e) == want, json.dumps(e)
        nonempty += 0 < len(want) < N_ROWS
    assert nonempty > 40


VALID = {"nuc": "-ACGT", "gene": "-ACDEFGHIKLMNPQRSTVWY*"}      # nucleotide_symbols.h / aa_symbols.h:56-79, in output order


def mutation_rows(rows, flt, kind, sequence_name, reference, min_proportion):
    import math
    selected = [r for r in rows if matches(flt, r)]
    out = []
    for pos, ref in enumerate(reference):
        counts = {sym: sum(r[kind][pos] == sym for r in selected) for sym in VALID[kind]}
        total = sum(counts.values())
        if total == 0:
            continue
        threshold = 0 if min_proportion == 0 else math.ceil(total * min_proportion) - 1
        for sym in VALID[kind]:
            if sym != ref and counts[sym] > threshold:
                out.append({"count": counts[sym], "mutation": f"{ref}{pos + 1}{sym}", "proportion": counts[sym] / total, "sequenceName": sequence_name})
    return out


@pytest.fixture(params=BACKENDS)
def execute(request, libs, model_db):
    pysilo, oracle = libs
    db, _ = model_db
    if request.param == "oracle":
        yield lambda query: oracle.execute(db, query)
    else:
        engine = pysilo.Engine(db, 0)
        yield engine.execute
        engine.close()


def test_mutation_actions_against_the_model(execute, model_db):
    _, rows = model_db
    rng = random.Random(99)
    filters = [{"type": "True"}, {"type": "False"}] + [random_tree(rng, rng.randint(0, 3)) for _ in range(60)]
    emitted = 0
    for k, flt in enumerate(filters):
        for min_proportion in ((0.05, 1.0) if k % 2 else (0.3, 0.0)):
            for action, kind, name, reference in (("Mutations", "nuc", "main", NUC_REF), ("AminoAcidMutations", "gene", "G", GENE_REF)):
                query = {"action": {"type": action, "sequenceName": name, "minProportion": min_proportion}, "filterExpression": flt}
                want = mutation_rows(rows, flt, kind, name, reference, min_proportion)
                assert execute(query) == want, json.dumps(query)
                emitted += len(want)
        assert execute({"action": {"type": "Aggregated"}, "filterExpression": flt}) == [{"count": sum(matches(flt, r) for r in rows)}]
    assert emitted > 500


def test_grouped_counts_against_the_model(execute, model_db):
    """Aggregated with groupByFields: one output row per distinct tuple of the selected rows (aggregated.cpp:100-149), null for the empty
    string / missing number / missing date (tuple.cpp), ordered by the group-by fields with nulls first"""
    from collections import Counter
    _, rows = model_db
    rng = random.Random(5)
    column_of = {"age": "age", "qc": "qc", "region": "region", "name": "name", "lineage": "lineage", "date": "date"}
    shown = lambda v: None if v in ("", None) else v
    for _ in range(40):
        fields = rng.sample(sorted(column_of), rng.randint(1, 4))
        flt = random_tree(rng, rng.randint(0, 2))
        groups = Counter(tuple(shown(r[column_of[f]]) for f in fields) for r in rows if matches(flt, r))
        ordered = sorted(groups.items(), key=lambda kv: [(v is not None, v) for v in kv[0]])
        want = [dict(zip(fields, key), count=count) for key, count in ordered]
        query = {"action": {"type": "Aggregated", "groupByFields": fields, "orderByFields": fields}, "filterExpression": flt}
        assert execute(query) == want, json.dumps(query)
        limited = dict(query, action=dict(query["action"], limit=3, offset=1))
        assert execute(limited) == want[1:4]


@pytest.fixture(scope="module")
def partitioned_db(libs, tmp_path_factory):
    """the same kind of rows, 300 of them, partitioned by lineage (preprocessor.cpp:159-199: here 8 partitions) and ordered by date inside
    each partition: flipped and deleted symbols differ from partition to partition, DateBetween takes the RangeSelection"""
    global N_ROWS
    keep, N_ROWS = N_ROWS, 300
    try:
        rows = make_rows(random.Random(77))
    finally:
        N_ROWS = keep
    for r in rows:
        r["date_is_sorted"] = True
    db = build_db(libs, tmp_path_factory.mktemp("filter_model_partitioned"), rows, "  dateToSortBy: date\n  partitionBy: lineage\n")
    assert db.n_partitions >= 5
    return db, rows


@pytest.fixture(params=BACKENDS)
def execute_partitioned(request, libs, partitioned_db):
    pysilo, oracle = libs
    db, _ = partitioned_db
    if request.param == "oracle":
        yield lambda query: oracle.execute(db, query)
    else:
        engine = pysilo.Engine(db, 0)
        yield engine.execute
        engine.close()


def test_actions_over_several_partitions_against_the_model(execute_partitioned, partitioned_db):
    from collections import Counter
    _, rows = partitioned_db
    rng = random.Random(4242)
    date_filters = [{"type": "DateBetween", "column": "date", "from": lo, "to": hi} for lo, hi in
                    ((None, None), ("2020-06-15", "2021-03-03"), (None, "2020-06-15"), ("2021-03-03", None), ("2021-03-04", "2022-12-30"), ("2020-06-15", "2020-06-15"))]
    filters = [{"type": "True"}] + date_filters + [{"type": "Not", "child": f} for f in date_filters[:3]] + [random_tree(rng, rng.randint(0, 3)) for _ in range(50)]
    shown = lambda v: None if v in ("", None) else v
    for k, flt in enumerate(filters):
        selected = [r for r in rows if matches(flt, r)]
        assert execute_partitioned({"action": {"type": "Aggregated"}, "filterExpression": flt}) == [{"count": len(selected)}], json.dumps(flt)
        min_proportion = (0.02, 0.5, 0.0)[k % 3]
        for action, kind, name, reference in (("Mutations", "nuc", "main", NUC_REF), ("AminoAcidMutations", "gene", "G", GENE_REF)):
            query = {"action": {"type": action, "sequenceName": name, "minProportion": min_proportion}, "filterExpression": flt}
            assert execute_partitioned(query) == mutation_rows(rows, flt, kind, name, reference, min_proportion), json.dumps(query)
        fields = rng.sample(["age", "qc", "region", "name", "lineage", "date"], rng.randint(1, 3))
        groups = Counter(tuple(shown(r[f]) for f in fields) for r in selected)
        want = [dict(zip(fields, key), count=count) for key, count in sorted(groups.items(), key=lambda kv: [(v is not None, v) for v in kv[0]])]
        assert execute_partitioned({"action": {"type": "Aggregated", "groupByFields": fields, "orderByFields": fields}, "filterExpression": flt}) == want


def test_wide_and_deep_trees_against_the_model(ids, model_db):
    """N-Of over 60 children with up to 61 matchers, and right-deep alternations of And / Or / N-Of 40 levels down: the evaluation stack
    of the device interpreter outgrows shared memory (more than 24 planes) and moves to global memory (eval_plans<true>)"""
    _, rows = model_db
    rng = random.Random(31337)
    check = lambda e: ids(e) == [r for r in range(N_ROWS) if matches(e, rows[r])]
    for trial in range(3):
        children = [random_leaf(rng) for _ in range(60)]
        per_row = sorted(sum(matches(c, r) for c in children) for r in rows)
        for k in sorted({0, 1, per_row[0], per_row[len(per_row) // 2], per_row[-1], per_row[-1] + 1, 24, 25, 26, 30, 47, 48, 59, 60, 61}):
            for exactly in (False, True):
                e = {"type": "N-Of", "numberOfMatchers": k, "matchExactly": exactly, "children": children}
                assert check(e), (trial, k, exactly)
                assert check({"type": "Not", "child": e}) and check({"type": "Maybe", "child": e})
    for trial in range(6):
        e = random_leaf(rng)
        for level in range(40):
            other = [random_tree(rng, 1) for _ in range(rng.randint(1, 2))]
            kind = (level + trial) % 3
            if kind == 0:
                e = {"type": "And", "children": other + [e]}
            elif kind == 1:
                e = {"type": "Or", "children": other + [e]}
            else:
                e = {"type": "N-Of", "numberOfMatchers": 1 + level % 2, "matchExactly": level % 4 == 0, "children": other + [e]}
            if level % 7 == 6:
                e = {"type": "Not", "child": e}
        assert check(e), trial
