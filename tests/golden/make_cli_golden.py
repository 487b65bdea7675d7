"""Extract the reference's command-line surface (flag names, defaults, types, choices) with ``ast``.

The reference's train.py / sample.py parse their flags at import time and import TensorFlow, so
they cannot be imported here; their ``add_argument`` calls are read from the source instead.
Run in the build container:  python tests/golden/make_cli_golden.py   -> cli_golden.json
"""
import ast
import json
import os

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def flags_of(path):
    tree = ast.parse(open(path).read())
    out = {}
    for node in ast.walk(tree):
        if isinstance(node, ast.Call) and getattr(node.func, "attr", "") == "add_argument":
            names = [ast.literal_eval(a) for a in node.args]
            kw = {}
            for k in node.keywords:
                if k.arg in ("default", "choices", "action", "required"):
                    kw[k.arg] = ast.literal_eval(k.value)
                elif k.arg == "type":
                    kw["type"] = getattr(k.value, "id", getattr(k.value, "attr", None))
            out[names[-1]] = {"names": names, **kw}
    return out


if __name__ == "__main__":
    gold = {f: flags_of(os.path.join(REF, f)) for f in ("train.py", "sample.py", "data_splitter.py")}
    json.dump(gold, open(os.path.join(HERE, "cli_golden.json"), "w"), indent=1, sort_keys=True)
    print({k: len(v) for k, v in gold.items()})
