"""The reference's native C++ programs on the GPU graph engine (csrc/acx_graph.cuh).

* ``ac_graph_bfs(n, moves)`` / ``write_simplex_files`` -- barcode_analysis/simplex_data_generation/
  {classic,prime}_moves/ac_bfs.cpp:13-97: BFS from (a, b) over unordered relator pairs of total size <= n;
  node names, sizes (0-filtration), edges and their filtration values, in the reference's order, and the
  same four output files byte for byte.
* ``neighbourhood_sizes`` / ``read_do_and_write`` -- barcode_analysis/5_steps_neibourhoods/
  neibourhoods.cpp:18-99: size of the radius-r ball around each presentation, all instances of a batch
  expanded together (the instance number is part of the visited-set key).

Level-synchronous: the frontier of a level is a contiguous range of node ids, every step expands a chunk of
it with one thread per (parent, move), resolves first discoveries by candidate order and appends them, so ids
equal the reference's FIFO names.  There is no CPU implementation: without CUDA the calls raise.
"""
import ctypes
import io
import os
import re

import numpy as np
import torch

from . import _lib
from .batch import _prep

N_MOVES = {True: 14, False: 12}


class _Step(ctypes.Structure):  # acx_graph_step_t (include/acx.h)
    _fields_ = [("width_words", ctypes.c_int32), ("classic", ctypes.c_int32), ("max_total", ctypes.c_int32),
                ("want_edges", ctypes.c_int32), ("allow_empty", ctypes.c_int32), ("nodes", ctypes.c_void_p), ("n_nodes", ctypes.c_int64),
                ("parent_first", ctypes.c_int64), ("n_parents", ctypes.c_int64), ("table", ctypes.c_void_p),
                ("tcap", ctypes.c_int64), ("cand", ctypes.c_void_p), ("ref", ctypes.c_void_p),
                ("csize", ctypes.c_void_p), ("flag", ctypes.c_void_p), ("excl", ctypes.c_void_p),
                ("scan_tmp", ctypes.c_void_p), ("scan_tmp_bytes", ctypes.c_int64), ("node_size", ctypes.c_void_p),
                ("edges", ctypes.c_void_p), ("edge_filt", ctypes.c_void_p), ("counts", ctypes.c_void_p)]


class _CudaBackend:
    """libacx.so on a CUDA device (the only backend the product has)."""

    def __init__(self, device):
        self.dev = torch.device(device)
        self.stream = _prep(self.dev)  # raises without CUDA
        self.L = _lib.lib()
        self.prefix = "acx_graph_"

    def call(self, name, *args):
        _lib.check(getattr(self.L, self.prefix + name)(*args, self.stream), self.prefix + name)

    def scan_bytes(self, n):
        r = self.L.acx_graph_scan_bytes(n)
        if r < 0:
            _lib.check(int(r), "acx_graph_scan_bytes")
        return int(r)


def _next_pow2(x):
    p = 1024
    while p < x:
        p <<= 1
    return p


class GraphEngine:
    """Node store + visited table + per-step scratch on one device."""

    def __init__(self, width_words, classic, max_total=0, want_edges=False, allow_empty=False, chunk=1 << 18,
                 device="cuda:0", _backend=None):
        if width_words not in (2, 4, 8):
            raise ValueError("width_words must be 2, 4 or 8")
        self.be = _backend if _backend is not None else _CudaBackend(device)
        self.W, self.classic, self.max_total, self.want_edges = width_words, bool(classic), int(max_total), bool(want_edges)
        self.allow_empty = bool(allow_empty)
        self.cap_letters = 32 * width_words - 8
        if self.max_total > self.cap_letters + 1:
            raise ValueError(f"max size {max_total} needs wider planes (width_words={width_words} holds relators of "
                             f"{self.cap_letters} letters)")
        self.nm = N_MOVES[self.classic]
        self.chunk = int(chunk)
        d = self.be.dev
        C = self.chunk * self.nm
        self.cand = torch.empty((C, 4 * self.W), dtype=torch.int32, device=d)
        self.ref = torch.empty((C,), dtype=torch.int64, device=d)
        self.csize = torch.empty((C,), dtype=torch.int16, device=d)
        self.flag = torch.empty((C,), dtype=torch.uint8, device=d)
        self.excl = torch.empty((C,), dtype=torch.int32, device=d)
        self.scan_tmp = torch.empty((self.be.scan_bytes(C),), dtype=torch.uint8, device=d)
        self.edges = torch.empty((C if self.want_edges else 1, 2), dtype=torch.int64, device=d)
        self.edge_filt = torch.empty((C if self.want_edges else 1,), dtype=torch.int16, device=d)
        self.counts = torch.zeros((4,), dtype=torch.int64, device=d)
        self.capacity = 0
        self.nodes = torch.empty((0, 4 * self.W), dtype=torch.int32, device=d)
        self.node_size = torch.empty((0,), dtype=torch.int16, device=d)
        self.table = None
        self.tcap = 0
        self.n_nodes = 0
        self.edge_parts, self.filt_parts = [], []
        self.n_edges = 0

    # -- storage -------------------------------------------------------------------------------------
    def _reserve(self, extra):
        need = self.n_nodes + extra
        if need > self.capacity:
            cap = max(need, 2 * self.capacity, 1 << 16)
            nodes = torch.empty((cap, 4 * self.W), dtype=torch.int32, device=self.be.dev)
            size = torch.empty((cap,), dtype=torch.int16, device=self.be.dev)
            nodes[: self.n_nodes].copy_(self.nodes[: self.n_nodes])
            size[: self.n_nodes].copy_(self.node_size[: self.n_nodes])
            self.nodes, self.node_size, self.capacity = nodes, size, cap
        if self.tcap < 2 * need:
            self.tcap = _next_pow2(4 * need)
            self.table = torch.zeros((self.tcap,), dtype=torch.int64, device=self.be.dev)
            self.be.call("rehash", self.nodes.data_ptr(), self.n_nodes, self.W, self.table.data_ptr(), self.tcap)

    def seed(self, rows, inst=None):
        """rows int8 [n, 2m] ([r1 | r2], zero padded) become nodes 0..n-1 (sorted pairs)."""
        rows = np.ascontiguousarray(rows, dtype=np.int8)
        n, m = rows.shape[0], rows.shape[1] // 2
        d = self.be.dev
        if self.n_nodes:
            raise RuntimeError("seed() must come first")
        self._reserve(n)
        rt = torch.from_numpy(rows).to(d)
        it = None if inst is None else torch.as_tensor(np.asarray(inst, dtype=np.int32)).to(d)
        ok = torch.empty((n,), dtype=torch.uint8, device=d)
        self.be.call("pack", rt.data_ptr(), n, m, 0 if it is None else it.data_ptr(), self.W, self.nodes.data_ptr(),
                     ok.data_ptr())
        bad = (ok == 0).nonzero().flatten().cpu().numpy()
        if len(bad):
            raise ValueError(f"presentation {int(bad[0])} is outside the engine's domain (letters +-1/+-2, freely reduced, "
                             f"non-empty relators of at most {self.cap_letters} letters)")
        self.node_size[:n] = torch.from_numpy(np.count_nonzero(rows, axis=1).astype(np.int16)).to(d)
        self.n_nodes = n
        self.tcap = 0  # force a table that holds the roots
        self._reserve(0)

    # -- one chunk of parents ------------------------------------------------------------------------
    def step(self, first, n_parents):
        C = n_parents * self.nm
        self._reserve(C)
        s = _Step(self.W, int(self.classic), self.max_total, int(self.want_edges), int(self.allow_empty),
                  self.nodes.data_ptr(), self.n_nodes,
                  first, n_parents, self.table.data_ptr(), self.tcap, self.cand.data_ptr(), self.ref.data_ptr(),
                  self.csize.data_ptr(), self.flag.data_ptr(), self.excl.data_ptr(),
                  self.scan_tmp.data_ptr(), self.scan_tmp.numel(), self.node_size.data_ptr(), self.edges.data_ptr(),
                  self.edge_filt.data_ptr(), self.counts.data_ptr())
        self.be.call("step", ctypes.byref(s))
        new, ne, err = (int(v) for v in self.counts[:3].cpu().numpy())
        if err:
            raise OverflowError("a relator outgrew the bit planes or cancelled completely "
                                f"(width_words={self.W}: at most {self.cap_letters} letters per relator)")
        self.n_nodes += new
        if self.want_edges and ne:
            self.edge_parts.append(self.edges[:ne].clone())
            self.filt_parts.append(self.edge_filt[:ne].clone())
            self.n_edges += ne
        return new

    def expand_level(self, lo, hi):
        for off in range(lo, hi, self.chunk):
            self.step(off, min(self.chunk, hi - off))

    def rows(self, m, first=0, count=None):
        """Nodes as zero-padded int8 rows [count, 2m] (relators in the stored, sorted order) and their instance."""
        count = self.n_nodes - first if count is None else count
        d = self.be.dev
        out = torch.empty((count, 2 * m), dtype=torch.int8, device=d)
        inst = torch.empty((count,), dtype=torch.int32, device=d)
        self.be.call("unpack", self.nodes[first:].data_ptr(), count, self.W, m, out.data_ptr(), inst.data_ptr())
        return out.cpu().numpy(), inst.cpu().numpy()


def ac_graph_bfs(n_max, moves="classic", chunk=1 << 18, width_words=2, device="cuda:0", with_rows=False, _backend=None):
    """ac_bfs.cpp main(): returns dict(sizes int16[N] (zero_filtrations), edges int32[E,2] (one_simplices; int64 beyond
    2^31 nodes), efilt uint8[E] (one_filtrations), levels=[frontier sizes], rows=int8[N, 2(n_max-1)] if with_rows)."""
    if moves not in ("classic", "prime"):
        raise ValueError("moves must be 'classic' or 'prime'")
    n_max = int(n_max)
    eng = GraphEngine(width_words, moves == "classic", max_total=max(n_max, 1), want_edges=True, chunk=chunk, device=device,
                      _backend=_backend)
    eng.seed(np.array([[1, 2]], dtype=np.int8))  # Presentation({1}, {2}), ac_bfs.cpp:40
    lo, hi, levels = 0, 1, []
    if n_max < 2:  # the root is named unconditionally (:44-49); every child is over the limit
        hi = 0
    while lo < hi:
        levels.append(hi - lo)
        eng.expand_level(lo, hi)
        lo, hi = hi, eng.n_nodes
    edges, efilt = _gather_edges(eng)
    out = dict(sizes=eng.node_size[: eng.n_nodes].cpu().numpy(), edges=edges, efilt=efilt, levels=levels)
    if with_rows:
        out["rows"] = eng.rows(max(n_max - 1, 1))[0]
    return out


def _gather_edges(eng):
    """Edge list and filtration of all steps on the host: int32 pairs while the names fit (int64 beyond 2^31 nodes) and
    uint8 sizes, copied straight from the per-step device buffers into pinned memory."""
    E = eng.n_edges
    cuda = eng.be.dev.type == "cuda"
    idt = torch.int32 if eng.n_nodes < (1 << 31) else torch.int64
    edges = torch.empty((E, 2), dtype=idt, pin_memory=cuda)
    efilt = torch.empty((E,), dtype=torch.uint8, pin_memory=cuda)
    pos = 0
    for e, f in zip(eng.edge_parts, eng.filt_parts):
        k = e.shape[0]
        edges[pos:pos + k].copy_(e.to(idt), non_blocking=True)
        efilt[pos:pos + k].copy_(f.to(torch.uint8), non_blocking=True)
        pos += k
    if cuda:
        torch.cuda.current_stream(eng.be.dev).synchronize()
    return edges.numpy(), efilt.numpy()


def _join(fmt, arr):
    if len(arr) == 0:
        return b""
    buf = io.BytesIO()
    np.savetxt(buf, arr, fmt=fmt, newline=",")
    return buf.getvalue()


def simplex_file_contents(result):
    """The four files ac_bfs.cpp writes (:22-37, :47-48, :65-66, :76-77, :84-87), as bytes."""
    n = len(result["sizes"])
    return {
        "zero_simplices": b'{"0-simplices":[' + _join("[%d]", np.arange(n, dtype=np.int64)) + b"[]]}",
        "zero_filtrations": b'{"0-filt":[' + _join("%d", np.asarray(result["sizes"], dtype=np.int64)) + b"-5]}",
        "one_simplices": b'{"1-simplices":[' + _join("[%d,%d]", np.asarray(result["edges"], dtype=np.int64)) + b"[]]}",
        "one_filtrations": b'{"1-filt":[' + _join("%d", np.asarray(result["efilt"], dtype=np.int64)) + b"-5]}",
    }


def write_simplex_files(result, n_max, outdir="."):
    """Write zero_simplices_n, zero_filtrations_n, one_simplices_n, one_filtrations_n like `a.out n`."""
    for k, v in simplex_file_contents(result).items():
        with open(os.path.join(outdir, "%s_%d" % (k, n_max)), "wb") as f:
            f.write(v)


# ---- neibourhoods.cpp ----------------------------------------------------------------------------
def _split(data):
    """neibourhoods.cpp:74-86: first half -> rel1, second half -> rel2, zeros dropped."""
    data = [int(v) for v in data]
    h = len(data) // 2
    return [v for v in data[:h] if v != 0], [v for v in data[h:] if v != 0]


def _length_bound(l1, l2, radius):
    lo, hi = sorted((l1, l2))
    for _ in range(radius):
        lo, hi = hi, hi + max(lo, 2)
    return hi


def neighbourhood_sizes(presentations, radius=5, classic=True, width_words=None, instances_per_batch=128,
                        chunk=1 << 18, device="cuda:0", _backend=None):
    """neibourhood() (neibourhoods.cpp:18-54) for every presentation (sequences of ints, [r1 | r2] halves,
    zeros ignored).  Returns int64 [n] ball sizes."""
    pairs = [_split(p) for p in presentations]
    if not pairs:
        return np.zeros((0,), dtype=np.int64)
    for k, (a, b) in enumerate(pairs):
        if not a or not b:
            raise ValueError(f"presentation {k} has an empty relator")
    if width_words is None:
        bound = max(_length_bound(len(a), len(b), radius) for a, b in pairs)
        width_words = next((w for w in (2, 4, 8) if 32 * w - 8 >= bound), None)
        if width_words is None:
            raise ValueError(f"relators may reach {bound} letters within radius {radius}: beyond the 248-letter planes")
    m = max(max(len(a), len(b)) for a, b in pairs)
    out = np.zeros((len(pairs),), dtype=np.int64)
    per = min(int(instances_per_batch), 65536)
    for b0 in range(0, len(pairs), per):
        batch = pairs[b0:b0 + per]
        rows = np.zeros((len(batch), 2 * m), dtype=np.int8)
        for k, (a, b) in enumerate(batch):
            rows[k, :len(a)] = a
            rows[k, m:m + len(b)] = b
        eng = GraphEngine(width_words, classic, max_total=0, want_edges=False, allow_empty=True, chunk=chunk, device=device,
                          _backend=_backend)
        eng.seed(rows, inst=np.arange(len(batch), dtype=np.int32))
        lo, hi = 0, eng.n_nodes
        for _ in range(int(radius)):  # only nodes at distance < radius are expanded (:41)
            eng.expand_level(lo, hi)
            lo, hi = hi, eng.n_nodes
        counts = torch.zeros((len(batch),), dtype=torch.int64, device=eng.be.dev)
        eng.be.call("count_instances", eng.nodes.data_ptr(), eng.n_nodes, eng.W, counts.data_ptr(), len(batch))
        out[b0:b0 + len(batch)] = counts.cpu().numpy()
    return out


def neighbourhood_sizes_sharded(presentations, radius=5, classic=True, device=None, _backend=None, **kw):
    """neighbourhood_sizes over the ranks of an initialised torch.distributed group (one process per GPU): balls are
    independent, so rank r simply takes presentations r, r + world, ... -- no data-path collective -- and the sizes meet
    in one all-gather.  Every rank returns the full int64 [n] array."""
    import torch.distributed as dist

    rank, world = dist.get_rank(), dist.get_world_size()
    n = len(presentations)
    if device is None:
        device = "cuda:%d" % torch.cuda.current_device() if torch.cuda.is_available() else "cpu"
    mine = list(range(rank, n, world))
    local = neighbourhood_sizes([presentations[i] for i in mine], radius, classic, device=device, _backend=_backend, **kw)
    per = (n + world - 1) // world
    comm_dev = torch.device(device) if dist.get_backend() == "nccl" else torch.device("cpu")
    buf = torch.full((per,), -1, dtype=torch.int64, device=comm_dev)
    buf[: len(mine)] = torch.from_numpy(local).to(comm_dev)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    out = np.zeros((n,), dtype=np.int64)
    for r, part in enumerate(parts):
        idx = list(range(r, n, world))
        out[idx] = part[: len(idx)].cpu().numpy()
    return out


_LETTER = {1: "a", -1: "A", 2: "b", -2: "B"}


def read_do_and_write(sourcefile, outfile, r, classic, verbose=True, **kw):
    """neibourhoods.cpp:57-99: one presentation per line of `sourcefile` ("[1, 0, ..., 2, 0, ...]"), one ball size
    per line of `outfile`; prints each sorted presentation like the reference (:88)."""
    pres = []
    with open(sourcefile) as f:
        for line in f:
            pres.append([int(t) for t in re.findall(r"-?\d+", line)])
    sizes = neighbourhood_sizes(pres, radius=r, classic=classic, **kw)
    with open(outfile, "w") as f:
        for p, s in zip(pres, sizes):
            if verbose:
                a, b = _split(p)
                if not (len(a), a) < (len(b), b):  # sort_ (AC_UTILS_no_hash.cpp:129-136): only r1 < r2 keeps the order
                    a, b = b, a
                print("(%s, %s)" % ("".join(_LETTER[v] for v in a), "".join(_LETTER[v] for v in b)))
            f.write("%d\n" % int(s))
    return 0


def _main(argv=None):
    """Command lines of the two reference programs (`a.out n`; `a.out` in a directory with the two .txt files)."""
    import argparse
    import time

    ap = argparse.ArgumentParser(prog="python -m acx.graph")
    sub = ap.add_subparsers(dest="cmd", required=True)
    a = sub.add_parser("ac_bfs", help="simplex_data_generation/*/ac_bfs.cpp")
    a.add_argument("n", type=int)
    a.add_argument("--moves", choices=("classic", "prime"), default="classic")
    a.add_argument("--outdir", default=".")
    a.add_argument("--device", default="cuda:0")
    b = sub.add_parser("neighbourhoods", help="5_steps_neibourhoods/neibourhoods.cpp")
    b.add_argument("sourcefile", nargs="?")
    b.add_argument("outfile", nargs="?")
    b.add_argument("--radius", type=int, default=5)
    b.add_argument("--moves", choices=("classic", "prime"), default="prime")
    b.add_argument("--device", default="cuda:0")
    args = ap.parse_args(argv)
    if args.cmd == "ac_bfs":
        print(args.n)  # ac_bfs.cpp:16
        t0 = time.perf_counter()
        write_simplex_files(ac_graph_bfs(args.n, args.moves, device=args.device), args.n, args.outdir)
        print("time %d" % int((time.perf_counter() - t0) * 1e6))  # :95 (microseconds)
        return 0
    if args.sourcefile:
        if not args.outfile:
            ap.error("neighbourhoods: give both sourcefile and outfile, or neither")
        return read_do_and_write(args.sourcefile, args.outfile, args.radius, args.moves == "classic", device=args.device)
    # neibourhoods.cpp main() (:101-124): both files, classic first, then prime
    solved, unsolved = "solved_miller_schupp_presentations.txt", "unsolved_miller_schupp_presentations.txt"
    read_do_and_write(solved, "solved_clasic_moves", args.radius, True, device=args.device)
    read_do_and_write(unsolved, "unsolved_clasic_moves", args.radius, True, device=args.device)
    read_do_and_write(solved, "solved_prime_moves", args.radius, False, device=args.device)
    read_do_and_write(unsolved, "unsolved_prime_moves", args.radius, False, device=args.device)
    return 0


if __name__ == "__main__":
    raise SystemExit(_main())
