"""Extracts the texts of the reference's query validation errors (the string literals of CHECK_SILO_QUERY in the filter expressions,
the actions on the path and query.cpp) into query_error_messages.json. Run in the build container only (reads /root/reference);
tests/test_host_cpu.py::test_query_error_messages_are_the_reference_s checks the host against the list. Texts only, no code."""
import glob
import json
import re
from pathlib import Path

REF = Path("/root/reference/src/silo/query_engine")
FILES = sorted(glob.glob(str(REF / "filter_expressions" / "*.cpp"))) + [str(REF / "actions" / f) for f in ("action.cpp", "mutations.cpp", "aggregated.cpp")] + [str(REF / "query.cpp")]


def main():
    out = []
    for f in FILES:
        if f.endswith(".test.cpp"):
            continue
        src = Path(f).read_text()
        for m in re.finditer(r"CHECK_SILO_QUERY\(", src):
            i, depth = m.end(), 1
            while depth and i < len(src):
                depth += src[i] == "("
                depth -= src[i] == ")"
                i += 1
            body = re.sub(r'"\s*\n?\s*"', "", src[m.end():i - 1])      # adjacent literals are one string
            for text in re.findall(r'"((?:[^"\\]|\\.)*)"', body):
                text = text.split(": \\")[0]      # a literal that ends in an escaped quote before `+ value +`: keep the fixed part
                if len(text) > 12:
                    out.append({"file": Path(f).name, "text": text.replace('\\"', '"')})
    (Path(__file__).parent / "query_error_messages.json").write_text(json.dumps(out, indent=1) + "\n")
    print(len(out), "messages")


if __name__ == "__main__":
    main()
