"""Generates tests/golden/ms_golden.json by RUNNING THE REFERENCE's generate_miller_schupp_presentations
(ac_solver/search/miller_schupp/miller_schupp.py:20-83, imported from /root/reference; build container only):
the full dictionary for (n, max_w_len) in {(1, 3), (2, 5), (3, 7)} and a sha256 of its JSON for (7, 7), (1, 8).

    python tests/golden/make_ms_golden.py
"""
import hashlib
import importlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref_loader  # noqa: E402


def canon(d):
    return {str(k): [[int(v) for v in p] for p in ps] for k, ps in sorted(d.items())}


def main():
    ref_loader._install_gymnasium_stand_in()
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "ac_solver" or k.startswith("ac_solver.")}
    sys.path.insert(0, ref_loader.REFERENCE_ROOT)
    try:
        ms = importlib.import_module("ac_solver.search.miller_schupp.miller_schupp")
        out = {"full": {}, "sha256": {}}
        for n, w in ((1, 3), (2, 5), (3, 7)):
            out["full"]["%d_%d" % (n, w)] = canon(ms.generate_miller_schupp_presentations(n, w))
        for n, w in ((7, 7), (1, 8)):
            d = canon(ms.generate_miller_schupp_presentations(n, w))
            out["sha256"]["%d_%d" % (n, w)] = hashlib.sha256(json.dumps(d, separators=(",", ":")).encode()).hexdigest()
            out["sha256"]["%d_%d_counts" % (n, w)] = {k: len(v) for k, v in d.items()}
    finally:
        sys.path.remove(ref_loader.REFERENCE_ROOT)
        for k in [k for k in sys.modules if k == "ac_solver" or k.startswith("ac_solver.")]:
            sys.modules.pop(k)
        sys.modules.update(saved)
    with open(os.path.join(HERE, "ms_golden.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote ms_golden.json", out["sha256"])


if __name__ == "__main__":
    main()
