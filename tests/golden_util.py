"""Loader for tests/golden/qnn_reference_goldens.json (made by tests/golden/gen_from_reference_tests.py)."""
import base64
import json
import os

import numpy as np

from tachikoma_b200 import relay as R

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "qnn_reference_goldens.json")


def dec(d):
    shape, dtype = d["shape"], d["dtype"]
    if "fill" in d:
        return np.full(shape, d["fill"], dtype=dtype)
    if "b64" in d:
        return np.frombuffer(base64.b64decode(d["b64"]), dtype=dtype).reshape(shape).copy()
    return np.array(d["data"], dtype=dtype).reshape(shape)


def load_cases():
    with open(_PATH) as f:
        doc = json.load(f)
    out = []
    for i, c in enumerate(doc["cases"]):
        g = c["graph"]
        nodes = []
        for n in g["nodes"]:
            n = dict(n)
            if n["kind"] == "const":
                n["data"] = dec(n).reshape(-1).tolist()
            nodes.append(n)
        func = R.from_dict(dict(nodes=nodes, params=g["params"], body=g["body"]))
        out.append(dict(id=f"{i:03d}-{c['source'].split('/')[-1]}", func=func,
                        inputs={k: dec(v) for k, v in c["inputs"].items()}, golden=dec(c["golden"])))
    return out
