#!/usr/bin/env python
"""Extracts the known-answer vectors of the reference's own principled test (src/bsdfs/tests/test_principled.py:124-200,
test05_eval_pdf: 20 outgoing directions, pdf_true / evaluate_true lists and the plugin parameters) into
tests/golden/principled_eval_pdf.json.  Container only (needs /root/reference):  python oracle/gen_golden_principled.py"""
import ast, json, os, re, sys
HERE = os.path.dirname(os.path.abspath(__file__))
ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
src = open(os.path.join(ref, "src/bsdfs/tests/test_principled.py")).read()
body = src[src.index("def test05_eval_pdf"):]
lists = {}
for name in ("pdf_true", "evaluate_true"):
    m = re.search(name + r"\s*=\s*(\[[^\]]*\])", body)
    lists[name] = ast.literal_eval(m.group(1))
params = {k: float(v) for k, v in re.findall(r'<float name="(\w+)" value="([-\d.eE]+)"/>', body)}
out = {"_generator": "oracle/gen_golden_principled.py from src/bsdfs/tests/test_principled.py:124-200 (scalar_rgb variant, "
                     "default base_color 0.5, wi = normalize(1,0,1) *unnormalised in the test*: si.wi = [1,0,1])",
       "params": params, "wi": [1.0, 0.0, 1.0], "n": 20, **lists}
path = os.path.join(HERE, "..", "tests", "golden", "principled_eval_pdf.json")
json.dump(out, open(path, "w"), indent=1)
print(params, len(lists["pdf_true"]), len(lists["evaluate_true"]))
