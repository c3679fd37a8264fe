"""generate_miller_schupp_presentations of the drop-in package, with its enumeration kernel functor
(csrc/acx_ms.cuh) compiled for the CPU, against fixtures produced by the reference's own generator
(tests/golden/ms_golden.json) and against the 1190-row dataset the reference ships."""
import hashlib
import json
import os

import numpy as np
import pytest

from host_harness import harness


@pytest.fixture(scope="module")
def golden(golden_dir):
    with open(os.path.join(golden_dir, "ms_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def ms():
    from ac_solver.search.miller_schupp import miller_schupp

    return miller_schupp


def canon(d):
    return {str(k): [[int(v) for v in p] for p in ps] for k, ps in sorted(d.items())}


def check_against_golden(ms, golden, backend):
    for tag, want in golden["full"].items():
        n, w = (int(t) for t in tag.split("_"))
        got = ms.generate_miller_schupp_presentations(n, w, _backend=backend)
        assert canon(got) == want, tag
        assert all(type(v) is int for ps in got.values() for p in ps for v in p)
    for tag in ("7_7", "1_8"):
        n, w = (int(t) for t in tag.split("_"))
        d = canon(ms.generate_miller_schupp_presentations(n, w, _backend=backend))
        assert {k: len(v) for k, v in d.items()} == golden["sha256"][tag + "_counts"]
        assert hashlib.sha256(json.dumps(d, separators=(",", ":")).encode()).hexdigest() == golden["sha256"][tag]


def test_generator_matches_the_reference(ms, golden):
    check_against_golden(ms, golden, harness.HostMsBackend())


def test_generator_reproduces_the_shipped_dataset(ms, golden_dir):
    # all_presentations.txt = MS(n, w), n = 1..7, len(w) = 1..7, each at its own width (1190 rows)
    from helpers import dataset_rows

    rows = dataset_rows(golden_dir)
    mine = []
    for n in range(1, 8):
        d = ms.generate_miller_schupp_presentations(n, 7, _backend=harness.HostMsBackend())
        for lenw in range(1, 8):
            mine += d[lenw]
    assert len(mine) == len(rows) == 1190

    def relators(p):
        p = np.asarray(p)
        h = len(p) // 2
        return p[:h][p[:h] != 0].tolist(), p[h:][p[h:] != 0].tolist()

    assert sorted(map(relators, mine)) == sorted(map(relators, rows))


def test_argument_checks(ms):
    with pytest.raises(AssertionError):
        ms.generate_miller_schupp_presentations(0, 3, _backend=harness.HostMsBackend())
    with pytest.raises(ValueError):
        ms.generate_miller_schupp_presentations(1, 16, _backend=harness.HostMsBackend())
