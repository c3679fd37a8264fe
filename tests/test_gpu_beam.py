"""GPU parity of acx.beam: against the reference itself (beam_golden.json, bottom of this file) and -- as a
debugging aid with arbitrary float models -- against a literal restatement of the reference's beam_search
(value_search/value_guided_search.py:417-561) that uses the oracle's ACMove, numpy compute_features
(value_search/feature_extraction.py:11-91) and the same torch model on the same device for scoring."""
import numpy as np
import pytest

from helpers import AK2, AK3, dataset_rows, pad_row, random_rows, repad

pytestmark = pytest.mark.gpu


def _features_np(p, m):
    r1, r2 = p[:m], p[m:]
    l1, l2 = int(np.count_nonzero(r1)), int(np.count_nonzero(r2))
    tot = l1 + l2
    a, b = r1[:l1], r2[:l2]
    c = [int(np.sum(a == v)) for v in (1, -1, 2, -2)] + [int(np.sum(b == v)) for v in (1, -1, 2, -2)]
    ex = (c[0] - c[1]) + (c[4] - c[5])
    ratio = l1 / tot if tot > 0 else 0.5
    mn, mx = min(l1, l2), max(l1, l2)
    mm = mx / mn if mn > 0 else mx
    return np.array([tot, l1, l2] + c + [ex, ratio, mm], dtype=np.float32)


def _score(states, model, arch, mean, std, dev, dim):
    import torch

    if arch == "mlp":
        f = np.stack([_features_np(s, len(s) // 2) for s in states])
        f = (f - mean) / std
        t = torch.tensor(f, dtype=torch.float32, device=dev)
    else:
        tok = np.full((len(states), dim), 2, dtype=np.int64)
        for k, s in enumerate(states):
            tok[k, : len(s)] = s.astype(np.int64) + 2
        t = torch.tensor(tok, dtype=torch.long, device=dev)
    with torch.no_grad():
        return torch.expm1(model(t).squeeze(-1)).cpu().numpy().tolist()


def _beam_oracle(oracle, presentation, model, arch, mean, std, width, budget, dev, cyc, cache=None):
    from acx.solution_cache import _check_cache_with_rotations  # pinned by reference vectors (test_solution_cache.py)

    presentation = np.array(presentation, dtype=np.int8)
    m = len(presentation) // 2
    if cache is not None:
        hit = _check_cache_with_rotations(tuple(presentation.tolist()), m, cache)
        if hit is not None:
            return True, hit, {"nodes_explored": 1, "steps": 0, "cache_hit": True}
    l1, l2 = int(np.count_nonzero(presentation[:m])), int(np.count_nonzero(presentation[m:]))
    parent_map = {0: (None, l1 + l2, None)}
    next_id = 1
    beam = [(presentation.copy(), [l1, l2], 0)]
    visited = {tuple(presentation)}
    total_nodes, step = 1, 0

    def recon(nid):
        path = []
        while nid is not None:
            a, ln, par = parent_map[nid]
            if par is None:
                break
            path.append((a, ln))
            nid = par
        path.reverse()
        return path

    while beam and total_nodes < budget:
        step += 1
        children = []
        for state, wl, pid in beam:
            for action in range(12):
                ns, nl = oracle.ACMove(action, state, m, wl, cyc)
                new_length = int(sum(nl))
                tup = tuple(ns)
                if new_length == 2:
                    return True, recon(pid) + [(action, new_length)], {"nodes_explored": total_nodes, "steps": step}
                if cache is not None:
                    hit = _check_cache_with_rotations(tuple(int(v) for v in ns), m, cache)
                    if hit is not None:
                        return True, recon(pid) + [(action, new_length)] + list(hit), {"nodes_explored": total_nodes, "steps": step,
                                                                                      "cache_hit": True}
                if tup not in visited:
                    visited.add(tup)
                    total_nodes += 1
                    parent_map[next_id] = (action, new_length, pid)
                    children.append((ns, nl, next_id))
                    next_id += 1
                    if total_nodes >= budget:
                        break
            if total_nodes >= budget:
                break
        if not children:
            break
        scores = _score([c[0] for c in children], model, arch, mean, std, dev, len(presentation))
        scored = sorted(zip(scores, range(len(children))), key=lambda x: x[0])
        beam = [children[k] for _, k in scored[:width]]
    return False, [], {"nodes_explored": total_nodes, "steps": step}


def _models(dev):
    import torch

    torch.manual_seed(7)
    mlp = torch.nn.Sequential(torch.nn.Linear(14, 32), torch.nn.ReLU(), torch.nn.Linear(32, 1)).to(dev).eval()

    class Seq(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.emb = torch.nn.Embedding(5, 8)
            self.conv = torch.nn.Conv1d(8, 8, 3, padding=1)
            self.out = torch.nn.Linear(8, 1)

        def forward(self, t):
            x = self.emb(t).transpose(1, 2)
            return self.out(torch.relu(self.conv(x)).mean(dim=2))

    return mlp, Seq().to(dev).eval()


def test_features_vs_numpy(cuda):
    import torch

    from acx.beam import features

    rng = np.random.default_rng(3)
    for m in (7, 18, 36, 128):
        rows = random_rows(rng, 600, m, kinds=(0, 1, 2, 3, 4))
        rows[0] = 0
        rows[1, m:] = 0
        got = features(torch.from_numpy(rows).to(cuda)).cpu().numpy()
        want = np.stack([_features_np(r, m) for r in rows])
        assert got.dtype == np.float32 and (got.view(np.uint32) == want.view(np.uint32)).all()


@pytest.mark.parametrize("arch", ["mlp", "seq", "len"])
def test_beam_search_vs_reference_restatement(cuda, oracle, golden_dir, arch):
    import torch

    from acx.beam import beam_search

    mlp, seq = _models(cuda)
    model = mlp if arch == "mlp" else seq
    if arch == "len":  # score = a monotone function of the total length: a beam that actually solves things
        model = torch.nn.Linear(14, 1).to(cuda).eval()
        with torch.no_grad():
            model.weight.zero_()
            model.bias.zero_()
            model.weight[0, 0] = 1.0
    mean = np.linspace(1.0, 14.0, 14).astype(np.float32)
    std = np.linspace(2.0, 5.0, 14).astype(np.float32)
    rows = dataset_rows(golden_dir)
    rng = np.random.default_rng(11)
    cases = [(repad(AK2, 8), 4, 300, False), (pad_row(AK3[0], AK3[1], 36), 50, 3000, True)]
    for k in rng.choice(len(rows), size=10, replace=False):
        cases.append((rows[int(k)], int(rng.choice([1, 5, 50])), int(rng.choice([2, 40, 700, 2500])), bool(rng.integers(0, 2))))
    if arch == "len":
        cases += [(rows[k], 10, 2000, bool(k % 2)) for k in range(0, 40, 4)]
    a = "seq" if arch == "seq" else "mlp"
    n_solved = 0
    for row, width, budget, cyc in cases:
        want = _beam_oracle(oracle, row, model, a, mean, std, width, budget, cuda, cyc)
        got = beam_search(row, model, architecture=a, feat_mean=mean, feat_std=std, beam_width=width,
                          max_nodes_to_explore=budget, device=str(cuda), cyclically_reduce_after_moves=cyc)
        assert got == want, (len(row), width, budget, cyc)
        n_solved += int(want[0])
    assert n_solved >= 1 or arch != "len"


def _vgreedy_oracle(oracle, presentation, model, arch, mean, std, budget, dev, use_len, cyc):
    """literal restatement of value_guided_greedy_search (value_guided_search.py:253-414) without the cache"""
    import heapq

    presentation = np.array(presentation, dtype=np.int8)
    m = len(presentation) // 2
    wl = [int(np.count_nonzero(presentation[:m])), int(np.count_nonzero(presentation[m:]))]
    total0 = sum(wl)
    init = float(total0) if use_len else _score([presentation], model, arch, mean, std, dev, len(presentation))[0]
    parent_map = {0: (None, total0, None)}
    next_id = 1
    state_to_id = {tuple(presentation): 0}
    heap = [(init, 0, 0, tuple(presentation), tuple(wl))]
    min_length = total0

    def recon(nid):
        path = []
        while nid is not None:
            a, ln, par = parent_map[nid]
            if par is None:
                break
            path.append((a, ln))
            nid = par
        path.reverse()
        return path

    while heap:
        _, plen, cur, state_tuple, wl_tuple = heapq.heappop(heap)
        state = np.array(state_tuple, dtype=np.int8)
        children = []
        for action in range(12):
            ns, nl = oracle.ACMove(action, state, m, list(wl_tuple), cyc)
            new_length = int(sum(nl))
            tup = tuple(ns)
            min_length = min(min_length, new_length)
            if new_length == 2:
                return True, recon(cur) + [(action, new_length)], {"nodes_explored": len(state_to_id), "min_length": min_length}
            if tup not in state_to_id:
                children.append((ns, nl, action, new_length, tup))
        if children and not use_len:
            scores = _score([c[0] for c in children], model, arch, mean, std, dev, len(presentation))
        else:
            scores = [c[3] for c in children]
        for (ns, nl, action, new_length, tup), score in zip(children, scores):
            nid = next_id
            next_id += 1
            parent_map[nid] = (action, new_length, cur)
            state_to_id[tup] = nid
            heapq.heappush(heap, (score, plen + 1, nid, tup, tuple(nl)))
        if len(state_to_id) >= budget:
            break
    return False, [], {"nodes_explored": len(state_to_id), "min_length": min_length}


@pytest.mark.parametrize("mode", ["mlp", "seq", "length"])
def test_value_guided_greedy_vs_reference_restatement(cuda, oracle, golden_dir, mode):
    from acx.beam import value_guided_greedy_search

    mlp, seq = _models(cuda)
    model = {"mlp": mlp, "seq": seq, "length": None}[mode]
    arch = "seq" if mode == "seq" else "mlp"
    mean = np.linspace(1.0, 14.0, 14).astype(np.float32)
    std = np.linspace(2.0, 5.0, 14).astype(np.float32)
    rows = dataset_rows(golden_dir)
    rng = np.random.default_rng(23)
    cases = [(repad(AK2, 8), 400, False), (pad_row(AK3[0], AK3[1], 36), 1500, True)]
    for k in rng.choice(len(rows), size=8, replace=False):
        cases.append((rows[int(k)], int(rng.choice([1, 30, 500, 1500])), bool(rng.integers(0, 2))))
    if mode == "length":
        cases += [(rows[k], 3000, bool(k % 2)) for k in range(0, 40, 5)]
    n_solved = 0
    for row, budget, cyc in cases:
        want = _vgreedy_oracle(oracle, row, model, arch, mean, std, budget, cuda, mode == "length", cyc)
        got = value_guided_greedy_search(row, model, architecture=arch, feat_mean=mean, feat_std=std, max_nodes_to_explore=budget,
                                         device=str(cuda), use_length_priority=(mode == "length"),
                                         cyclically_reduce_after_moves=cyc)
        assert got == want, (len(row), budget, cyc)
        n_solved += int(want[0])
    assert n_solved >= 1 or mode != "length"


# ---- parity with the reference ITSELF: tests/golden/beam_golden.json (made by tests/golden/make_beam_golden.py) -------
def _golden_models(g, dev):
    """the value networks of the fixture, re-built from their state_dicts (the layer stacks of value_search/value_network.py
    FeatureMLP / SequenceValueNet with dropout 0: Linear-ReLU-...-Linear, Embedding-Conv1d-ReLU-mean-Linear-ReLU-Linear)"""
    import torch

    def tensor(e):
        return torch.tensor(e["data"], dtype=torch.float32).reshape(e["shape"])

    def mlp(sd, hidden):
        layers, prev = [], 14
        for h in hidden:
            layers += [torch.nn.Linear(prev, h), torch.nn.ReLU(), torch.nn.Dropout(0.0)]
            prev = h
        layers.append(torch.nn.Linear(prev, 1))
        net = torch.nn.Sequential(*layers)
        net.load_state_dict({k[len("net."):]: tensor(v) for k, v in sd.items()})
        return net.to(dev).eval()

    class Seq(torch.nn.Module):
        def __init__(self, embed_dim, num_conv_layers, conv_channels, kernel_size, mlp_hidden, max_seq_len):
            super().__init__()
            self.embedding = torch.nn.Embedding(5, embed_dim, padding_idx=2)
            conv, cin = [], embed_dim
            for _ in range(num_conv_layers):
                conv += [torch.nn.Conv1d(cin, conv_channels, kernel_size, padding=kernel_size // 2), torch.nn.ReLU(), torch.nn.Dropout(0.0)]
                cin = conv_channels
            self.conv = torch.nn.Sequential(*conv)
            self.head = torch.nn.Sequential(torch.nn.Linear(conv_channels, mlp_hidden), torch.nn.ReLU(), torch.nn.Dropout(0.0),
                                            torch.nn.Linear(mlp_hidden, 1))

        def forward(self, t):
            x = self.conv(self.embedding(t).transpose(1, 2))
            return self.head(x.mean(dim=2))

    seq = Seq(**g["seq_kwargs"])
    seq.load_state_dict({k: tensor(v) for k, v in g["seq_state_dict"].items()})
    return {"mlp": mlp(g["mlp_state_dict"], g["mlp_hidden_dims"]), "len": mlp(g["len_state_dict"], g["len_hidden_dims"]),
            "seq": seq.to(dev).eval()}


def _golden_stats(name, g):
    if name == "len":
        return np.zeros(14, dtype=np.float32), np.ones(14, dtype=np.float32)
    return np.array(g["feat_mean"], dtype=np.float32), np.array(g["feat_std"], dtype=np.float32)


def test_beam_search_matches_the_reference(cuda, golden_dir):
    """value_search.beam_search of the reference, run with fixed-weight value networks (exact float32 scores), against
    acx.beam.beam_search with the same weights on the GPU: (solved, path, stats) for k in {1, 10, 50}, budgets {100, 3000}"""
    import json
    import os

    from acx.beam import beam_search

    with open(os.path.join(golden_dir, "beam_golden.json")) as f:
        g = json.load(f)
    models = _golden_models(g, cuda)
    n_solved = 0
    for c in g["beam"]:
        mean, std = _golden_stats(c["model"], g)
        got = beam_search(np.array(c["presentation"], dtype=np.int8), models[c["model"]], architecture=c["arch"], feat_mean=mean,
                          feat_std=std, beam_width=c["beam_width"], max_nodes_to_explore=c["budget"], device=str(cuda),
                          cyclically_reduce_after_moves=c["cyclical"])
        want = (c["solved"], [tuple(p) for p in c["path"]], c["stats"])
        assert (got[0], [tuple(p) for p in got[1]], {k: v for k, v in got[2].items() if k != "time"}) == want, \
            (c["model"], c["beam_width"], c["budget"], c["cyclical"], c["presentation"])
        n_solved += int(c["solved"])
    assert n_solved >= 10 and len(g["beam"]) >= 200


def test_value_guided_greedy_matches_the_reference(cuda, golden_dir):
    """value_guided_greedy_search of the reference with a model (value_guided_search.py:253-414) against acx.beam"""
    import json
    import os

    from acx.beam import value_guided_greedy_search

    with open(os.path.join(golden_dir, "beam_golden.json")) as f:
        g = json.load(f)
    models = _golden_models(g, cuda)
    n_solved = 0
    for c in g["vgreedy"]:
        mean, std = _golden_stats(c["model"], g)
        got = value_guided_greedy_search(np.array(c["presentation"], dtype=np.int8), models[c["model"]], architecture=c["arch"],
                                         feat_mean=mean, feat_std=std, max_nodes_to_explore=c["budget"], device=str(cuda),
                                         cyclically_reduce_after_moves=c["cyclical"])
        want = (c["solved"], [tuple(p) for p in c["path"]], c["stats"])
        assert (got[0], [tuple(p) for p in got[1]], {k: v for k, v in got[2].items() if k != "time"}) == want, \
            (c["model"], c["budget"], c["cyclical"], c["presentation"])
        n_solved += int(c["solved"])
    assert n_solved >= 3
