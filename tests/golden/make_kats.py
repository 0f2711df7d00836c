"""Extracts the known-answer cases of the reference's operator unit tests into operator_kats.json.

Run in the build container only (reads /root/reference/src/silo/query_engine/operators/*.test.cpp).
Only DATA is taken: literal bitmaps, row counts, parameters and the expected result sets. Each case is a
literal operator tree in the JSON form lapis-silo_b200/host/optree_json.hpp understands, so the same
case runs through the oracle (CPU) and through the GPU path.
"""
import json
import re
from pathlib import Path

REF = Path("/root/reference/src/silo/query_engine/operators")
OUT = Path(__file__).resolve().parent / "operator_kats.json"

ROARING = re.compile(r"roaring::Roaring\((?:\{([^}]*)\})?\)")
CMP = {"EQUALS": "==", "NOT_EQUALS": "!=", "LESS": "<", "LESS_OR_EQUALS": "<=", "HIGHER": ">", "HIGHER_OR_EQUALS": ">="}


def ids(text):
    return [int(x) for x in text.replace(" ", "").split(",") if x] if text else []


def roarings(text):
    return [ids(m.group(1)) for m in ROARING.finditer(text)]


def tests(path):
    src = (REF / path).read_text()
    for m in re.finditer(r"TEST\((\w+), (\w+)\) \{(.*?)\n\}\n", src, re.S):
        yield f"{m.group(1)}.{m.group(2)}", m.group(3)


def named_vectors(body):
    out = {}
    for m in re.finditer(r"const std::vector<roaring::Roaring> (\w+)\((.*?)\);\n", body, re.S):
        out[m.group(1)] = roarings(m.group(2))
    return out


def row_count(body):
    m = re.search(r"const uint32_t row_count = (\d+);", body)
    return int(m.group(1)) if m else None


def scans(bitmaps):
    return [{"op": "index_scan", "ids": b} for b in bitmaps]


def expected(body, var, deref=r"\*"):
    m = re.search(deref + var + r"(?:\.|->)evaluate\(\), (roaring::Roaring\((?:\{[^}]*\})?\))", body)
    return roarings(m.group(1))[0] if m else None


def main():
    cases = []

    for name, body in tests("threshold.test.cpp"):
        vecs, rc = named_vectors(body), row_count(body)
        if "ASSERT_THROW" in body and rc is not None or "ASSERT_THROW" in body:
            m = re.search(r"const Threshold under_test\(std::move\(non_negated\), std::move\(negated\), (\d+), (true|false), (\d+)\)", body)
            if m:
                cases.append({"name": name, "tree": {"op": "threshold", "row_count": int(m.group(3)), "children": [], "negated": [], "n": int(m.group(1)), "exact": m.group(2) == "true"}, "throws": True})
            continue
        pat = re.compile(r"const Threshold (\w+)\(\s*(?:generateTestInput\((\w+), row_count\)|OperatorVector\(\)),\s*(?:generateTestInput\((\w+), row_count\)|OperatorVector\(\)),\s*(\d+),\s*(true|false),\s*row_count\s*\);")
        for m in pat.finditer(body):
            var, pos, neg, n, exact = m.groups()
            exp = expected(body, var)
            if exp is None:
                continue
            cases.append({"name": f"{name}.{var}", "tree": {"op": "threshold", "row_count": rc, "children": scans(vecs.get(pos, [])), "negated": scans(vecs.get(neg, [])),
                                                            "n": int(n), "exact": exact == "true"}, "expected": exp})

    for name, body in tests("intersection.test.cpp"):
        vecs, rc = named_vectors(body), row_count(body)
        if "under_test" not in body or "type()" in body:
            continue
        tree = {"op": "intersection", "row_count": rc, "children": scans(vecs.get("test_bitmaps", [])), "negated": scans(vecs.get("test_negated_bitmaps", []))}
        if "ASSERT_THROW" in body:
            cases.append({"name": name, "tree": tree, "throws": True})
        else:
            cases.append({"name": name, "tree": tree, "expected": expected(body, "under_test")})

    for name, body in tests("union.test.cpp"):
        vecs, rc = named_vectors(body), row_count(body)
        if "type()" in body:
            continue
        cases.append({"name": name, "tree": {"op": "union", "row_count": rc, "children": scans(vecs.get("test_bitmaps", []))}, "expected": expected(body, "under_test")})

    for name, body in tests("complement.test.cpp"):
        m = re.search(r"const roaring::Roaring test_bitmap\((.*?)\);", body)
        if not m or "type()" in body:
            continue
        rc = row_count(body)
        cases.append({"name": name, "tree": {"op": "complement", "row_count": rc, "child": {"op": "index_scan", "ids": roarings(m.group(1))[0]}}, "expected": expected(body, "under_test")})

    for name, body in tests("bitmap_selection.test.cpp"):
        vecs = named_vectors(body)
        m = re.search(r"BitmapSelection::(CONTAINS|NOT_CONTAINS), (\d+)", body)
        if not m or "type()" in body:
            continue
        tree = {"op": "bitmap_selection", "bitmaps": vecs["test_bitmaps"], "contains": m.group(1) == "CONTAINS", "value": int(m.group(2))}
        cases.append({"name": name, "tree": tree, "expected": expected(body, "under_test")})
        neg = expected(body, "negated")
        if neg is not None:
            cases.append({"name": name + ".negated", "tree": dict(tree, negate=True), "expected": neg})

    for name, body in tests("range_selection.test.cpp"):
        rc = row_count(body)
        if "type()" in body or rc is None:
            continue
        ranges = [[int(a), int(b)] for a, b in re.findall(r"RangeSelection::Range\{(\d+), (\d+)\}", body)]
        tree = {"op": "range_selection", "row_count": rc, "ranges": ranges}
        cases.append({"name": name, "tree": tree, "expected": expected(body, "under_test")})
        neg = expected(body, "negated")
        if neg is not None:
            cases.append({"name": name + ".negated", "tree": dict(tree, negate=True), "expected": neg})

    for name, body in tests("selection.test.cpp"):
        m = re.search(r"const std::vector<int32_t> test_column\(\{\{([^}]*)\}\}\);", body)
        c = re.search(r"Comparator::(\w+),\s*(-?\d+)", body)
        exp = expected(body, "under_test")
        if not m or not c or exp is None:
            continue
        column = ids(m.group(1))
        tree = {"op": "selection", "row_count": len(column), "predicates": [{"kind": "int", "column": column, "cmp": CMP[c.group(1)], "value": int(c.group(2))}]}
        cases.append({"name": name, "tree": tree, "expected": exp})
        neg = expected(body, "negated")
        if neg is not None:
            cases.append({"name": name + ".negated", "tree": dict(tree, negate=True), "expected": neg})

    for name, body in tests("index_scan.test.cpp"):
        m = re.search(r"const roaring::Roaring test_bitmap\((.*?)\);", body)
        exp = expected(body, "under_test")
        if m and exp is not None:
            cases.append({"name": name, "tree": {"op": "index_scan", "row_count": row_count(body) or 5, "ids": roarings(m.group(1))[0]}, "expected": exp})
    for name, body in tests("bitmap_producer.test.cpp"):
        m = re.search(r"const roaring::Roaring test_bitmap\((.*?)\);", body)
        exp = expected(body, "under_test")
        if m and exp is not None:
            tree = {"op": "bitmap_producer", "row_count": row_count(body) or 5, "ids": roarings("roaring::Roaring(" + m.group(1) + ")")[0]}
            cases.append({"name": name, "tree": dict(tree, negate=True) if ".negate()" in body else tree, "expected": exp})
    for name, body in tests("empty.test.cpp"):
        exp = expected(body, "under_test")
        if exp is not None:
            cases.append({"name": name, "tree": {"op": "empty", "row_count": row_count(body) or 3}, "expected": exp})
    for name, body in tests("full.test.cpp"):
        exp = expected(body, "under_test")
        m = re.search(r"Full under_test\((\d+)\)", body)
        if exp is not None and m:
            cases.append({"name": name, "tree": {"op": "full", "row_count": int(m.group(1))}, "expected": exp})

    OUT.write_text(json.dumps(cases, indent=1) + "\n")
    by_op = {}
    for c in cases:
        by_op[c["tree"]["op"]] = by_op.get(c["tree"]["op"], 0) + 1
    print(len(cases), "cases:", by_op)


if __name__ == "__main__":
    main()
