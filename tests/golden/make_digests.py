#!/usr/bin/env python
"""Writes tests/golden/oracle_digests.json: digests of the oracle's trajectories on the
golden cases (tests/golden/cases.py).  Run from the repository root; needs only gcc."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

from oracle import pyoracle  # noqa: E402
import cases  # noqa: E402


def main():
    out = {}
    for name, spec in cases.CASES.items():
        o = pyoracle.Oracle(spec["seed"], threads=os.cpu_count() or 1)
        out[name] = cases.run_case(o, spec, lambda n, l, e: o.create_population(n, l, e))
        o.close()
        print(name, out[name]["best"])
    with open(os.path.join(HERE, "oracle_digests.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
