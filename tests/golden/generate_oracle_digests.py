"""
Writes tests/golden/oracle_digests.json: the hits of the parity scenarios of tests/digests.py, traced by the CPU oracle
of this repo (test infrastructure) and reduced to a SHA-256 each. Run from the repository root after a DELIBERATE change
of the results (new draw order, new arithmetic contract) and commit the file with that change:

    python tests/golden/generate_oracle_digests.py
"""

import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import digests as D  # noqa: E402
import harness as H  # noqa: E402


def main() -> None:
    out = {"_comment": "outputs of this repo's oracle (not of the reference): regression anchor, see tests/digests.py",
           "n_paths": D.N_PATHS, "cases": {}}
    for name in D.CASES:
        cfg = D.build(name, "cpu")
        H.runOracle(cfg)
        out["cases"][name] = D.digestOf(H.sortedHits(cfg))
        print(name, out["cases"][name])
    with open(D.PATH, "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")


if __name__ == "__main__":
    main()
