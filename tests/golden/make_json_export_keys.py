"""Generator of tests/golden/json_export_keys.json: the key tree (names, nesting, order, value kinds) of one packet entry of
the reference's own example export, /root/reference/json/json_examples/worker_pool_0000_0000000000, which its MATLAB /
Python viewers read.  Run in the build container (the reference tree does not exist on the GPU box)."""
import json
import os

SRC = "/root/reference/json/json_examples/worker_pool_0000_0000000000"


def tree(x):
    if isinstance(x, dict):
        return {k: tree(v) for k, v in x.items()}
    if isinstance(x, list):
        return ["%s[%d]" % (type(x[0]).__name__ if x else "none", len(x))]
    return type(x).__name__


if __name__ == "__main__":
    d = json.load(open(SRC))
    first = d[next(iter(d))]
    out = {"source": "json/json_examples/worker_pool_0000_0000000000 of the reference", "entry_prefix": "packet_",
           "entries_per_file": len(d), "tree": tree(first)}
    json.dump(out, open(os.path.join(os.path.dirname(__file__), "json_export_keys.json"), "w"), indent=1)
