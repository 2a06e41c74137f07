#!/usr/bin/env python
"""Extracts the literal known-answer tables of the reference's own plugin tests into tests/golden/ref_test_tables.json
(container only: reads /root/reference; the JSON travels, the reference does not).

For each listed test function, every outermost list / tuple literal with at least four numbers is taken in source order
(`ast.literal_eval`), e.g. the Mitsuba-0.6 tables of src/render/tests/test_microfacet.py:33-285.  tests/test_oracle_ref_tests.py
restates the inputs of each test next to the table it compares with.

    python oracle/gen_golden_tables.py
"""
import ast
import json
import os

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SOURCES = {
    "src/render/tests/test_microfacet.py": ["test02_eval_pdf_beckmann", "test03_smith_g1_beckmann", "test04_sample_beckmann",
                                            "test03_smith_g1_ggx", "test05_sample_ggx"],
}


def count_numbers(node):
    return sum(isinstance(n, ast.Constant) and isinstance(n.value, (int, float)) for n in ast.walk(node))


def tables_of(fn):
    found = []

    def visit(node):
        if isinstance(node, (ast.List, ast.Tuple)) and count_numbers(node) >= 4:
            try:
                found.append((node.lineno, ast.literal_eval(node)))
                return
            except ValueError:
                pass   # mixed literal (e.g. a tuple of np.array(...) calls): descend
        for c in ast.iter_child_nodes(node):
            visit(c)
    visit(fn)
    return found


def main():
    out = {}
    for rel, names in SOURCES.items():
        tree = ast.parse(open(os.path.join(REF, rel)).read())
        for fn in tree.body:
            if isinstance(fn, ast.FunctionDef) and fn.name in names:
                out[f"{rel}::{fn.name}"] = [{"line": ln, "values": v} for ln, v in tables_of(fn)]
    path = os.path.join(ROOT, "tests", "golden", "ref_test_tables.json")
    json.dump(out, open(path, "w"), indent=0)
    for k, v in out.items():
        print(k, [(t["line"], len(json.dumps(t["values"]).split(","))) for t in v])


if __name__ == "__main__":
    main()
