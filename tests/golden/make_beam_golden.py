"""Generates tests/golden/beam_golden.json by RUNNING THE REFERENCE's model-scored searches
(value_search/value_guided_search.py:253-561: beam_search and value_guided_greedy_search with a value network),
imported from /root/reference; build container only.

    python tests/golden/make_beam_golden.py

The networks are the reference's own classes (value_search/value_network.py FeatureMLP / SequenceValueNet) with
FIXED-SEED weights that are small integers (final layer: integers / 1024) and feature statistics that are integers /
powers of two, so that every score is computed exactly in float32 whatever the order of the additions: the reference
scores on the CPU here, the GPU tests score with the same weights on the device, and the ranking cannot depend on the
rounding of either.  (The two ratio features of the MLP get zero weight; the sequence network is run on
max_relator_length 32 rows, where the global average pool divides by 64.)  The state_dicts travel in the fixture.
"""
import importlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_loader  # noqa: E402
from helpers import dataset_rows  # noqa: E402


def int_weights(module, rng, lo=-2, hi=2, final_scale=1.0 / 1024):
    """replace every parameter of `module` by small integers (the last Linear by integers * final_scale)"""
    import torch

    linears = [m for m in module.modules() if isinstance(m, torch.nn.Linear)]
    with torch.no_grad():
        for p in module.parameters():
            p.copy_(torch.from_numpy(rng.integers(lo, hi + 1, size=tuple(p.shape)).astype(np.float32)))
        last = linears[-1]
        last.weight.mul_(final_scale)
        last.bias.mul_(final_scale)


def state_dict_to_json(sd):
    return {k: {"shape": list(v.shape), "data": [float(x) for x in v.flatten().tolist()]} for k, v in sd.items()}


def main():
    import torch

    ref_loader._install_gymnasium_stand_in()
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "ac_solver" or k.startswith("ac_solver.")}
    sys.path.insert(0, ref_loader.REFERENCE_ROOT)
    try:
        vgs = importlib.import_module("value_search.value_guided_search")
        vnet = importlib.import_module("value_search.value_network")
        rng = np.random.default_rng(2026)
        mlp = vnet.FeatureMLP(input_dim=14, hidden_dims=[16, 8], dropout=0.0).eval()
        int_weights(mlp, rng)
        with torch.no_grad():
            first = [m for m in mlp.modules() if isinstance(m, torch.nn.Linear)][0]
            first.weight[:, 12:] = 0.0  # the two ratio features are not exactly representable: no weight
        seq = vnet.SequenceValueNet(embed_dim=4, num_conv_layers=1, conv_channels=4, kernel_size=3, mlp_hidden=4, max_seq_len=64,
                                    dropout=0.0).eval()
        int_weights(seq, rng, -1, 1)
        with torch.no_grad():
            seq.embedding.weight[2] = 0.0  # padding_idx row
        # a third network whose score is a monotone function of the total length (feature 0): a beam that solves things
        lenm = vnet.FeatureMLP(input_dim=14, hidden_dims=[4], dropout=0.0).eval()
        with torch.no_grad():
            for p_ in lenm.parameters():
                p_.zero_()
            lins = [m for m in lenm.modules() if isinstance(m, torch.nn.Linear)]
            lins[0].weight[0, 0] = 1.0
            lins[1].weight[0, 0] = 1.0 / 64
        feat_mean = np.array([20, 10, 10, 3, 3, 3, 3, 3, 3, 3, 3, 0, 0, 0], dtype=np.float32)
        feat_std = np.array([8, 4, 4, 2, 2, 2, 2, 2, 2, 2, 2, 4, 1, 1], dtype=np.float32)
        rows = dataset_rows(HERE)
        ak2 = np.array([1, 1, -2, -2, -2, 0, 0, 1, 2, 1, -2, -1, -2, 0], dtype=np.int8)
        rsel = np.random.default_rng(5)
        pool = [ak2] + [rows[int(i)] for i in rsel.choice(len(rows), size=22, replace=False)]
        pool32 = [r for r in rows if len(r) == 64]
        pool32 = [pool32[int(i)] for i in rsel.choice(len(pool32), size=10, replace=False)]
        out = {"mlp_state_dict": state_dict_to_json(mlp.state_dict()), "seq_state_dict": state_dict_to_json(seq.state_dict()),
               "len_state_dict": state_dict_to_json(lenm.state_dict()), "len_hidden_dims": [4],
               "feat_mean": feat_mean.tolist(), "feat_std": feat_std.tolist(),
               "mlp_hidden_dims": [16, 8], "seq_kwargs": dict(embed_dim=4, num_conv_layers=1, conv_channels=4, kernel_size=3,
                                                            mlp_hidden=4, max_seq_len=64),
               "beam": [], "vgreedy": []}

        def stats_json(st):
            return {k: (bool(v) if isinstance(v, (bool, np.bool_)) else (int(v) if float(v).is_integer() else float(v)))
                    for k, v in st.items() if k != "time"}

        zero, one = np.zeros(14, dtype=np.float32), np.ones(14, dtype=np.float32)
        for name, arch, model, cases in (("mlp", "mlp", mlp, pool), ("seq", "seq", seq, pool32), ("len", "mlp", lenm, pool[:14])):
            fm, fs = (zero, one) if name == "len" else (feat_mean, feat_std)
            for idx, pres in enumerate(cases):
                for width in (1, 10, 50):
                    for budget in (100, 3000):
                        cyc = bool((idx + width + budget // 100) % 2)
                        ok, path, st = vgs.beam_search(pres, model, architecture=arch, feat_mean=fm, feat_std=fs,
                                                       beam_width=width, max_nodes_to_explore=budget, device="cpu",
                                                       cyclically_reduce_after_moves=cyc)
                        out["beam"].append({"model": name, "arch": arch, "presentation": np.asarray(pres).tolist(), "beam_width": width, "budget": budget,
                                            "cyclical": cyc, "solved": bool(ok), "path": [[int(a), int(t)] for a, t in path],
                                            "stats": stats_json(st)})
                for budget in (100, 1500):
                    cyc = bool((idx + budget // 100) % 2)
                    ok, path, st = vgs.value_guided_greedy_search(pres, model, architecture=arch, feat_mean=fm, feat_std=fs,
                                                                  max_nodes_to_explore=budget, device="cpu",
                                                                  cyclically_reduce_after_moves=cyc)
                    out["vgreedy"].append({"model": name, "arch": arch, "presentation": np.asarray(pres).tolist(), "budget": budget, "cyclical": cyc,
                                           "solved": bool(ok), "path": [[int(a), int(t)] for a, t in path], "stats": stats_json(st)})
    finally:
        sys.path.remove(ref_loader.REFERENCE_ROOT)
        for k in [k for k in sys.modules if k == "ac_solver" or k.startswith("ac_solver.") or k.startswith("value_search")]:
            sys.modules.pop(k)
        sys.modules.update(saved)
    with open(os.path.join(HERE, "beam_golden.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print({"beam": len(out["beam"]), "beam_solved": sum(c["solved"] for c in out["beam"]), "vgreedy": len(out["vgreedy"]),
           "vgreedy_solved": sum(c["solved"] for c in out["vgreedy"])})


if __name__ == "__main__":
    main()
