"""Runs the UNMODIFIED reference (prairiedog + LemonGraph) in this container.

TEST INFRASTRUCTURE -- only tests/golden/make_golden.py and the ``not gpu`` oracle pinning
tests import this.  It needs /root/reference and therefore never runs on the GPU box.

The reference Python is imported from where it lies (``/root/reference`` and
``/root/reference/lemongraph`` on ``sys.path``); the compiled cffi module comes from
``oracle/_ref`` (oracle/build_ref.py).  Four pure-Python dependencies of the reference are
not installed here and there is no network, so minimal stand-ins are injected into
``sys.modules`` (they carry no hot-path logic):

* ``lazy``          -- lemongraph/LemonGraph/__init__.py:19 (a cached-property descriptor)
* ``coloredlogs``   -- prairiedog/logger.py:2   (install() + ColoredFormatter)
* ``pyinstrument``  -- prairiedog/profiler.py:25 (Profiler)
* ``pydgraph``      -- prairiedog/dgraph.py:9   (only if something imports the Dgraph backend)

``prairiedog/config.py:19`` lists ``samples/`` relative to the CWD at import time and
``prairiedog/logger.py:13`` opens ``prairiedog.log`` in the CWD, so the import happens inside a
scratch directory.
"""
import glob
import importlib.util
import logging
import os
import sys
import tempfile
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("PDG_REFERENCE", "/root/reference")
_loaded = {}


def available() -> bool:
    return os.path.isdir(os.path.join(REF, "prairiedog")) and bool(
        glob.glob(os.path.join(HERE, "_ref", "_lemongraph_cffi*.so")))


def _inject_stubs():
    if "lazy" not in sys.modules:
        m = types.ModuleType("lazy")

        class lazy(object):  # noqa: N801 - mirrors the package's name
            def __init__(self, func):
                self.func = func
                self.__name__ = func.__name__

            def __get__(self, inst, owner):
                if inst is None:
                    return self
                v = self.func(inst)
                inst.__dict__[self.__name__] = v
                return v
        m.lazy = lazy
        sys.modules["lazy"] = m
    if "coloredlogs" not in sys.modules:
        m = types.ModuleType("coloredlogs")
        m.install = lambda *a, **k: None
        m.ColoredFormatter = logging.Formatter
        sys.modules["coloredlogs"] = m
    if "pyinstrument" not in sys.modules:
        m = types.ModuleType("pyinstrument")

        class Profiler(object):
            def start(self):
                pass

            def stop(self):
                pass

            def output_text(self, *a, **k):
                return ""
        m.Profiler = Profiler
        sys.modules["pyinstrument"] = m


def load():
    """Import the reference; returns a namespace with Kmers, SubgraphRef, GraphRef, LGGraph, ..."""
    if _loaded:
        return _loaded["ns"]
    if not available():
        raise RuntimeError("live reference unavailable (needs /root/reference and oracle/_ref)")
    _inject_stubs()
    # the cffi module is looked up as LemonGraph._lemongraph_cffi (LemonGraph/__init__.py:3)
    so = glob.glob(os.path.join(HERE, "_ref", "_lemongraph_cffi*.so"))[0]
    spec = importlib.util.spec_from_file_location("LemonGraph._lemongraph_cffi", so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    sys.modules["LemonGraph._lemongraph_cffi"] = mod
    for p in (os.path.join(REF, "lemongraph"), REF):
        if p not in sys.path:
            sys.path.insert(0, p)
    scratch = tempfile.mkdtemp(prefix="pdg_liveref_")
    os.makedirs(os.path.join(scratch, "samples"))
    cwd = os.getcwd()
    os.chdir(scratch)
    try:
        sys.dont_write_bytecode = True
        import prairiedog  # noqa: F401
        logging.getLogger("prairiedog").setLevel(logging.ERROR)
        logging.getLogger().setLevel(logging.ERROR)
        from prairiedog.kmers import Kmers
        from prairiedog.subgraph_ref import SubgraphRef
        from prairiedog.graph_ref import GraphRef
        from prairiedog.lemon_graph import LGGraph
        from prairiedog.node import Node, concat_values
        from prairiedog.edge import Edge
        from prairiedog.graph import Graph
    finally:
        os.chdir(cwd)
    ns = types.SimpleNamespace(Kmers=Kmers, SubgraphRef=SubgraphRef, GraphRef=GraphRef,
                               LGGraph=LGGraph, Node=Node, Edge=Edge, Graph=Graph,
                               concat_values=concat_values, scratch=scratch)
    _loaded["ns"] = ns
    return ns


def new_lggraph(path=None):
    ns = load()
    if path is None:
        fd, path = tempfile.mkstemp(prefix="pdg_lg_", dir=ns.scratch)
        os.close(fd)
    return ns.LGGraph(path, delete_on_exit=True)


def build_graph(fasta_paths, k=11):
    """Reference pipeline (Snakefile:51-54,62-111 rule bodies): Kmers -> update_graph -> save."""
    ns = load()
    g = new_lggraph()
    sg = ns.SubgraphRef(g)
    gr = ns.GraphRef()
    stats = []
    cwd = os.getcwd()
    os.chdir(ns.scratch)
    try:
        for p in fasta_paths:
            km = ns.Kmers(os.path.abspath(os.path.join(cwd, p)), k)
            gr.index_kmers(km)
            n = len(km)
            uniq = km.unique_kmers
            c = sg.update_graph(km, gr)
            sg.save(None)
            stats.append({"file": os.path.basename(p), "len": n, "unique": uniq, "returned": c})
    finally:
        os.chdir(cwd)
    return g, stats


def canonical(g):
    """Canonical node / edge lists by VALUE (LemonGraph ids are unusable, R1)."""
    nodes = {}
    for n in g.nodes:
        nodes[n.db_id] = n.value
    edges = set()
    for e in g.edges:
        edges.add((nodes[e.src], nodes[e.tgt], e.edge_type, int(e.edge_value)))
    g.save()
    return sorted(set(nodes.values())), sorted(edges)


def query(g, a, b):
    """Returns (connected, n_start_edges, sorted [(string, edge_type)]) through the reference."""
    ns = load()
    connected, starts = g.connected(a, b)
    paths, metas = g.path(a, b)
    hits = sorted((ns.concat_values(p), m["edge_type"]) for p, m in zip(paths, metas))
    return bool(connected), len(starts), hits
