"""Query front end -- API mirror of prairiedog/prairiedog.py:13-33 and pretty_hits.py:4-51."""
import logging

from .graph import Graph
from .node import concat_values
from .subgraph_ref import uncouple_edge_type

log = logging.getLogger("prairiedog")


class Hit:
    def __init__(self, string: str, edge_type: str):
        self.sequence = string
        self.source = edge_type
        self.sample, self.contig = uncouple_edge_type(edge_type)


class PrettyHits:
    """sample -> contig -> set(sequence)  (pretty_hits.py:13-51)."""

    def __init__(self, list_hits: list):
        self.sample_map = {}
        for hit in (Hit(string=h['string'], edge_type=h['edge_type']) for h in list_hits):
            self.sample_map.setdefault(hit.sample, {}).setdefault(hit.contig, set()).add(hit.sequence)

    def __str__(self):
        s = "\n"
        for sample, contig_map in self.sample_map.items():
            s += "Sample {} has:\n".format(sample)
            for contig, variant_set in contig_map.items():
                s += "Contig {} with variant(s):\n".format(contig)
                for variant in variant_set:
                    s += "{}\n".format(variant)
        return s


class Prairiedog:
    def __init__(self, g: Graph):
        self.g = g

    def query(self, src: str, dst: str):
        """All strings between src and dst (prairiedog.py:17-33); returns the hit list the
        reference only logs."""
        paths, paths_meta = self.g.path(src, dst)
        list_hits = [{'string': concat_values(p), **m} for p, m in zip(paths, paths_meta)]
        log.info("Found: {}".format(PrettyHits(list_hits)))
        return list_hits
