"""Generates tests/golden/barcode_golden.json by RUNNING THE REFERENCE's C++ programs (oracle/_ref/, built by
`make -C oracle ref` from the sources under /root/reference/barcode_analysis; build container only):

  * sha256 + node / edge counts of the four files `ac_bfs n` writes, classic and prime moves, n = 2..11
    (barcode_analysis/simplex_data_generation/*/ac_bfs.cpp);
  * the full node-size and edge lists for n = 5 (small enough to commit);
  * radius-5 ball sizes of neibourhoods.cpp for the README's test_input.txt and for every 40th row of the two
    Miller-Schupp files next to it (5_steps_neibourhoods/*.txt), classic and prime moves.

    python tests/golden/make_barcode_golden.py
"""
import hashlib
import json
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import barcode_oracle as bo  # noqa: E402

REF = "/root/reference/barcode_analysis/5_steps_neibourhoods"


def read_rows(path):
    return [[int(t) for t in re.findall(r"-?\d+", line)] for line in open(path) if line.strip()]


def main():
    assert bo.have_ref(), "run `make -C oracle ref` first"
    out = {"ac_bfs": {}, "neighbourhoods": {}}
    for moves in ("classic", "prime"):
        for n in range(2, 12):
            files = bo.ref_ac_bfs_files(n, moves == "classic")
            entry = {k: hashlib.sha256(v).hexdigest() for k, v in files.items()}
            entry["n_nodes"] = files["zero_filtrations"].count(b",")
            entry["n_edges"] = files["one_filtrations"].count(b",")
            if n == 5:
                entry["sizes"] = json.loads(files["zero_filtrations"])["0-filt"][:-1]
                entry["edges"] = json.loads(files["one_simplices"])["1-simplices"][:-1]
                entry["efilt"] = json.loads(files["one_filtrations"])["1-filt"][:-1]
            out["ac_bfs"]["%s_%d" % (moves, n)] = entry
    rows = read_rows(os.path.join(REF, "test_input.txt"))
    rows += read_rows(os.path.join(REF, "solved_miller_schupp_presentations.txt"))[::40]
    rows += read_rows(os.path.join(REF, "unsolved_miller_schupp_presentations.txt"))[::40]
    out["neighbourhoods"]["rows"] = rows
    for moves in ("classic", "prime"):
        out["neighbourhoods"][moves] = bo.ref_neighbourhoods(rows, 5, moves == "classic")
    with open(os.path.join(HERE, "barcode_golden.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote barcode_golden.json:", len(rows), "neighbourhood rows")


if __name__ == "__main__":
    main()
