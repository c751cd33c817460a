"""The backend plugin interface -- API mirror of prairiedog/graph.py:11-113 (the drop-in boundary)."""
import abc
import logging
import typing

from .edge import Edge
from .node import Node

log = logging.getLogger("prairiedog")


class Graph(metaclass=abc.ABCMeta):
    """A directed graph backend (prairiedog/graph.py:11-80)."""

    @abc.abstractmethod
    def upsert_node(self, node: Node, echo: bool = True) -> typing.Optional[Node]:
        pass

    @abc.abstractmethod
    def add_edge(self, edge: Edge, echo: bool = True) -> typing.Optional[Edge]:
        pass

    @abc.abstractmethod
    def clear(self):
        pass

    @property
    @abc.abstractmethod
    def nodes(self) -> typing.Set[Node]:
        pass

    @property
    @abc.abstractmethod
    def edges(self) -> typing.Set[Edge]:
        pass

    @abc.abstractmethod
    def get_labels(self, node: str) -> dict:
        pass

    @abc.abstractmethod
    def save(self, f: str = None):
        pass

    @property
    @abc.abstractmethod
    def edgelist(self) -> typing.Generator:
        pass

    @abc.abstractmethod
    def set_graph_labels(self, labels: dict):
        pass

    @abc.abstractmethod
    def filter(self):
        pass

    @abc.abstractmethod
    def __len__(self):
        pass

    @abc.abstractmethod
    def connected(self, node_a: str, node_b: str) -> typing.Tuple[bool, typing.Tuple]:
        pass

    @abc.abstractmethod
    def path(self, node_a: str, node_b: str) -> typing.Tuple[tuple, tuple]:
        pass

    @staticmethod
    def matching_edges(src_edges: typing.Tuple[Edge],
                       tgt_edges: typing.Tuple[Edge]) -> typing.Tuple[bool, typing.Tuple[Edge]]:
        """Start edges from which a target edge can be reached: same edge_type and
        src.edge_value <= tgt.edge_value (prairiedog/graph.py:82-113).  Host-side helper kept for
        API parity; B200Graph.connected evaluates the same predicate on the device."""
        best = {}
        for t in tgt_edges:
            if t.edge_type not in best or t.edge_value > best[t.edge_type]:
                best[t.edge_type] = t.edge_value
        st = set(s for s in src_edges if s.edge_type in best and s.edge_value <= best[s.edge_type])
        if not st:
            log.warning("No connected edges found")
        return len(st) > 0, tuple(st)
