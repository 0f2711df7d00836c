"""Fuzzed queries on the reference's 100-sequence dataset: random filter trees over every expression type and the actions of the path,
then random damage -- keys dropped, values replaced by junk of every JSON type, texts cut short. The host must answer or refuse
every one of them (QueryParseException -> "parse", anything else -> "other": the reference's 400 / 500), never crash; and the device
path must give exactly the oracle's answer or exactly its refusal, text included.
CPU: the oracle survives 1,500 queries. GPU: 1,500 queries, device == oracle."""
import copy
import json
import random

import pytest

JUNK = [None, 0, -1, 1.5, "", "x", [], {}, True, 2 ** 40, "A" * 300, [1, 2], {"type": "True"}, {"type": 7}, "é", -0.0, 1e308]
KEYS = ["type", "children", "child", "position", "symbol", "column", "value", "from", "to", "numberOfMatchers", "matchExactly", "sequenceName", "includeSublineages"]
ACTIONS = [{"type": "Aggregated"}, {"type": "Mutations", "minProportion": 0.05}, {"type": "AminoAcidMutations", "minProportion": 0.05, "sequenceName": "S"},
           {"type": "Mutations", "minProportion": 0.3, "sequenceName": "testSecondSequence"},
           {"type": "Aggregated", "groupByFields": ["region", "age"], "orderByFields": [{"field": "count", "order": "descending"}, "region", "age"], "limit": 3, "offset": 1}]


def leaf(rng):
    return rng.choice([
        {"type": "NucleotideEquals", "position": rng.randint(1, 29903), "symbol": rng.choice("ACGTN-.")},
        {"type": "NucleotideEquals", "position": rng.choice([241, 3037, 14408, 23403]), "symbol": rng.choice("ACGT")},
        {"type": "AminoAcidEquals", "sequenceName": "S", "position": rng.randint(1, 1273), "symbol": rng.choice("ACDEX.")},
        {"type": "HasNucleotideMutation", "position": rng.randint(1, 29903)},
        {"type": "HasAminoAcidMutation", "sequenceName": "S", "position": rng.choice([501, 614, 681])},
        {"type": "PangoLineage", "column": "pango_lineage", "value": rng.choice(["B.1.1.7", "B.1", "b.1.617.2", "XA"]), "includeSublineages": rng.random() < 0.5},
        {"type": "DateBetween", "column": "date", "from": rng.choice([None, "2021-01-01", "2021-03-18"]), "to": rng.choice([None, "2021-03-18", "2021-06-30"])},
        {"type": "DateBetween", "column": "unsorted_date", "from": "2021-01-01", "to": rng.choice([None, "2021-04-30"])},
        {"type": "IntBetween", "column": "age", "from": rng.choice([None, 10, 50]), "to": rng.choice([None, 60])},
        {"type": "IntEquals", "column": "age", "value": rng.choice([None, 4, 50])},
        {"type": "FloatBetween", "column": "qc_value", "from": rng.choice([None, 0.9]), "to": rng.choice([None, 0.97])},
        {"type": "FloatEquals", "column": "qc_value", "value": rng.choice([None, 0.9, 0.98])},
        {"type": "StringEquals", "column": rng.choice(["country", "region", "division", "gisaid_epi_isl"]), "value": rng.choice(["Switzerland", "Europe", "Basel-Land", "EPI_ISL_1408408", ""])},
        {"type": "InsertionContains", "column": "nucleotideInsertions", "position": rng.choice([25701, 22339]), "value": rng.choice(["CC.*", ".*C.*G.*", "CCC"])},
        {"type": "InsertionContains", "column": "aminoAcidInsertions", "sequenceName": "S", "position": 214, "value": rng.choice(["E.*", ".*P.*"])},
        {"type": rng.choice(["True", "False"])},
    ])


def tree(rng, depth):
    if depth == 0 or rng.random() < 0.3:
        return leaf(rng)
    kind = rng.randrange(5)
    if kind == 0:
        return {"type": "And", "children": [tree(rng, depth - 1) for _ in range(rng.randint(1, 3))]}
    if kind == 1:
        return {"type": "Or", "children": [tree(rng, depth - 1) for _ in range(rng.randint(1, 3))]}
    if kind == 2:
        return {"type": "Not", "child": tree(rng, depth - 1)}
    if kind == 3:
        return {"type": rng.choice(["Maybe", "Exact"]), "child": tree(rng, depth - 1)}
    children = [tree(rng, depth - 1) for _ in range(rng.randint(1, 4))]
    return {"type": "N-Of", "numberOfMatchers": rng.randint(0, len(children)), "matchExactly": rng.random() < 0.5, "children": children}


def damage(rng, node):
    if isinstance(node, dict):
        node = dict(node)
        r = rng.random()
        if node and r < 0.15:
            node.pop(rng.choice(list(node)))
        elif r < 0.3:
            node[rng.choice(list(node) or ["type"])] = rng.choice(JUNK)
        elif r < 0.35:
            node[rng.choice(KEYS)] = rng.choice(JUNK)
        for k in list(node):
            if rng.random() < 0.3:
                node[k] = damage(rng, node[k])
        return node
    if isinstance(node, list):
        node = [damage(rng, x) if rng.random() < 0.3 else x for x in node]
        if node and rng.random() < 0.1:
            node.pop()
        if rng.random() < 0.05:
            node.append(rng.choice(JUNK))
        return node
    return rng.choice(JUNK) if rng.random() < 0.2 else node


def fuzzed_queries(seed, n):
    rng = random.Random(seed)
    for _ in range(n):
        query = {"action": copy.deepcopy(rng.choice(ACTIONS)), "filterExpression": tree(rng, 3)}
        r = rng.random()
        if r < 0.45:
            query = damage(rng, query)          # heavy: anything anywhere
        elif r < 0.7:
            query["filterExpression"] = damage(rng, query["filterExpression"])
        text = json.dumps(query)
        if rng.random() < 0.05:
            text = text[:rng.randrange(len(text))]
        yield text


def answer(call, text):
    try:
        return ("ok", call(text))
    except Exception as error:      # OracleError / SiloError: both carry kind and message
        return (error.kind, error.message)


def test_the_host_answers_or_refuses_every_fuzzed_query(libs, example_db):
    _, oracle = libs
    kinds = {}
    for text in fuzzed_queries(11, 1500):
        kind, _ = answer(lambda t: oracle.execute(example_db, t), text)
        kinds[kind] = kinds.get(kind, 0) + 1
    assert kinds.get("ok", 0) > 400 and kinds.get("parse", 0) > 300, kinds


@pytest.mark.gpu
def test_the_device_answers_and_refuses_as_the_oracle_does(libs, example_db):
    pysilo, oracle = libs
    engine = pysilo.Engine(example_db, 0)
    try:
        answered = nonempty = 0
        for text in fuzzed_queries(12, 1500):
            want = answer(lambda t: oracle.execute(example_db, t), text)
            got = answer(lambda t: json.loads(engine.execute_text(t.encode()))["queryResult"], text)
            assert got == want, text
            answered += want[0] == "ok"
            nonempty += want[0] == "ok" and len(want[1]) > 0 and want[1] != [{"count": 0}]
        assert answered > 400 and nonempty > 150
    finally:
        engine.close()


def thousand_leaf(rng):
    return rng.choice([
        {"type": "NucleotideEquals", "position": rng.randint(1, 29903), "symbol": rng.choice("ACGTN-")},
        {"type": "NucleotideEquals", "position": rng.choice([241, 3037, 14408, 23403, 28881]), "symbol": rng.choice("ACGT")},
        {"type": "AminoAcidEquals", "sequenceName": rng.choice(["S", "N", "ORF1a"]), "position": rng.randint(1, 400), "symbol": rng.choice("ADGLX.")},
        {"type": "HasNucleotideMutation", "position": rng.choice([241, 3037, 14408, 23403, rng.randint(1, 29903)])},
        {"type": "HasAminoAcidMutation", "sequenceName": "S", "position": rng.choice([222, 501, 614, 681])},
        {"type": "PangoLineage", "column": "pangoLineage", "value": rng.choice(["D.2", "A.2.2", "B.1", "B", "B.1.1.7"]), "includeSublineages": rng.random() < 0.5},
        {"type": "DateBetween", "column": "date", "from": rng.choice([None, "2020-06-01", "2021-01-01"]), "to": rng.choice([None, "2020-12-31", "2021-06-30"])},
        {"type": "IntBetween", "column": "age", "from": rng.choice([None, 20]), "to": rng.choice([None, 60])},
        {"type": "FloatBetween", "column": "nextcladeQcOverallScore", "from": rng.choice([None, 1.0]), "to": rng.choice([None, 50.0])},
        {"type": "StringEquals", "column": rng.choice(["country", "region", "division"]), "value": rng.choice(["Australia", "Oceania", "Victoria", "Europe", ""])},
    ])


@pytest.mark.gpu
def test_random_queries_on_the_thousand_sequence_dataset(libs, thousand_db):
    """BASELINE config C1's dataset (real sequences: run containers, 11 lineage partitions, rows ordered by date): 250 random trees
    under every action of the path, device == oracle in its reference-shaped (faithful) mode"""
    pysilo, oracle = libs
    rng = random.Random(2024)
    actions = [{"type": "Aggregated"}, {"type": "Mutations", "minProportion": 0.05}, {"type": "Mutations", "minProportion": 0.0},
               {"type": "AminoAcidMutations", "minProportion": 0.05, "sequenceName": "S"}, {"type": "AminoAcidMutations", "minProportion": 0.2},
               {"type": "Aggregated", "groupByFields": ["pangoLineage", "country"], "orderByFields": ["pangoLineage", "country"]},
               {"type": "Aggregated", "groupByFields": ["date"], "orderByFields": [{"field": "count", "order": "descending"}, "date"], "limit": 10}]

    def random_tree(depth):
        if depth == 0 or rng.random() < 0.3:
            return thousand_leaf(rng)
        kind = rng.randrange(5)
        if kind <= 1:
            return {"type": ("And", "Or")[kind], "children": [random_tree(depth - 1) for _ in range(rng.randint(1, 3))]}
        if kind == 2:
            return {"type": "Not", "child": random_tree(depth - 1)}
        if kind == 3:
            return {"type": rng.choice(["Maybe", "Exact"]), "child": random_tree(depth - 1)}
        children = [random_tree(depth - 1) for _ in range(rng.randint(1, 4))]
        return {"type": "N-Of", "numberOfMatchers": rng.randint(0, len(children)), "matchExactly": rng.random() < 0.5, "children": children}
    engine = pysilo.Engine(thousand_db, 0)
    try:
        nonempty = 0
        for k in range(250):
            query = {"action": actions[k % len(actions)], "filterExpression": random_tree(3)}
            want = answer(lambda q: oracle.execute(thousand_db, q), query)
            got = answer(engine.execute, query)
            assert got == want, json.dumps(query)
            nonempty += want[0] == "ok" and len(want[1]) > 0 and want[1] != [{"count": 0}]
        assert nonempty > 100
    finally:
        engine.close()
