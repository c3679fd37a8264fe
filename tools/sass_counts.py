#!/usr/bin/env python
"""SASS mnemonic counts per kernel of libacx.so (cuobjdump -sass): the evidence that the hot kernels move tiles with
bulk async copies (UBLKCP), complete them on mbarriers (SYNCS), use 256-bit global accesses and device atomics.

    python tools/sass_counts.py > profiles/r02_sass_counts.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "ac-solver-caltech_b200", "acx", "libacx.so")
PATTERNS = [("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS"), ("UTMACMDFLUSH", r"\bUTMACMDFLUSH"), ("LDG.256", r"\bLDG\.E\.[A-Z0-9.]*256"),
            ("STG.256", r"\bSTG\.E\.[A-Z0-9.]*256"), ("LDG.128", r"\bLDG\.E\.[A-Z0-9.]*128"), ("STG.128", r"\bSTG\.E\.[A-Z0-9.]*128"),
            ("ATOMG", r"\bATOMG"), ("REDG", r"\bREDG"), ("ATOMS", r"\bATOMS"), ("VOTE", r"\bVOTE"), ("PRMT", r"\bPRMT"),
            ("LDS", r"\bLDS"), ("STS", r"\bSTS"), ("BAR", r"\bBAR\."), ("MEMBAR", r"\bMEMBAR"), ("ACQBULK/DEPBAR", r"\bDEPBAR|\bACQBULK")]
WANT = sys.argv[1:] or ["k_env_tiles", "k_rows_tma", "k_rows_simple", "k_expand_warp", "k_pbfs_expand2", "k_pbfs_probe3", "k_pbfs_commit3",
                        "k_pbfs_decide", "k_pbfs_expand<", "k_search", "k_path_replay", "k_graph_blocks"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", out)), capture_output=True, text=True).stdout.split("\n")
    blocks = re.split(r"\n\s*Function : \S+\n", out)[1:]
    print("SASS mnemonic counts per kernel, %s (sm_100a), cuobjdump -sass" % os.path.relpath(SO, ROOT))
    print("%-86s %6s " % ("kernel", "instr") + " ".join("%7s" % p[0][:7] for p in PATTERNS))
    for name, body in zip(names, blocks):
        if not any(w in name for w in WANT):
            continue
        lines = [l for l in body.split("\n") if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l)]
        text = "\n".join(lines)
        short = re.sub(r"\(.*", "", name.replace("void acx::", "").replace("(int)", ""))
        print("%-86s %6d " % (short[:86], len(lines)) + " ".join("%7d" % len(re.findall(rx, text)) for _, rx in PATTERNS))


if __name__ == "__main__":
    main()
