"""prairiedog_b200: the pangenome-graph construction hot path of superphy/prairiedog on B200.

Python host code mirrors the reference's plugin API (Kmers, Graph, Node, Edge, SubgraphRef,
Prairiedog); all graph work runs in hand-written sm_100a CUDA kernels behind the C ABI of
``libpdgpu.so`` (include/pdgpu.h), reached through ctypes.  There is no CPU fallback.
"""
from .node import Node, concat_values, DEFAULT_NODE_TYPE  # noqa: F401
from .edge import Edge  # noqa: F401
from .graph import Graph  # noqa: F401

__version__ = "0.1.0"


def __getattr__(name):
    # the CUDA-backed classes are imported lazily so that `import prairiedog_b200` works before
    # the library is built (build() needs to import the package)
    if name == "Kmers":
        from .kmers import Kmers
        return Kmers
    if name == "B200Graph":
        from .b200_graph import B200Graph
        return B200Graph
    if name == "SubgraphRef":
        from .subgraph_ref import SubgraphRef
        return SubgraphRef
    if name == "GraphRef":
        from .graph_ref import GraphRef
        return GraphRef
    if name == "Prairiedog":
        from .prairiedog import Prairiedog
        return Prairiedog
    raise AttributeError(name)
