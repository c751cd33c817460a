"""SubgraphRef: the graph-building loop -- API mirror of prairiedog/subgraph_ref.py:17-125."""
import logging
import os
import typing

from .edge import Edge
from .graph import Graph
from .node import Node

log = logging.getLogger("prairiedog")

ET_DELIMITER = ' in '


def uncouple_edge_type(edge_type: str) -> typing.Tuple[str, str]:
    """'{header} in {file}' -> (file, header)  (subgraph_ref.py:20-23)."""
    split = edge_type.split(ET_DELIMITER)
    assert len(split) == 2
    return split[1], split[0]


class SubgraphRef:
    def __init__(self, graph: Graph):
        self.graph = graph
        self.out_file = None

    def __str__(self):
        return "SubgraphRef"

    def update_graph(self, km, gr, encode: bool = False, buffer: int = 333) -> int:
        """Adds every consecutive k-mer pair of ``km`` as an edge labelled
        '{header} in {file}' / index-in-contig (subgraph_ref.py:39-115); returns the number of
        k-mers covered.  With a B200Graph and a well-formed Kmers the whole genome goes to the GPU
        in one call; otherwise the reference's per-edge loop runs against the Graph API."""
        self.out_file = os.path.join('outputs/', km.filepath + '.rdf')
        from .b200_graph import B200Graph
        if isinstance(self.graph, B200Graph) and not encode and getattr(km, "well_formed", False):
            c = self.graph.add_genome(km)
            # the reference leaves the iterator exhausted (subgraph_ref.py:51); the contig lengths are in the
            # layout already -- km.sequences would read the whole genome back from the device and decode it
            lens = km.layout.seq_len
            km.li, km.pi = int(lens.size) - 1, int(lens[-1]) - km.k + 1
            return c
        implicit_nodes = isinstance(self.graph, B200Graph)   # like LGGraph: add_edge upserts endpoints
        c = 0
        while km.has_next:
            header1, kmer1 = km.next()
            node1_label = gr.node_label(kmer1) if encode else kmer1
            if not implicit_nodes:
                self.graph.upsert_node(Node(value=node1_label))
            c += 1
            edge_c = 0
            while km.contig_has_next:
                header2, kmer2 = km.next()
                node2_label = gr.node_label(kmer2) if encode else kmer2
                if not implicit_nodes:
                    self.graph.upsert_node(Node(node2_label))
                self.graph.add_edge(
                    Edge(src=node1_label, tgt=node2_label,
                         edge_type='{}{}{}'.format(header2, ET_DELIMITER, str(km)), edge_value=edge_c),
                    echo=False)
                node1_label = node2_label
                c += 1
                edge_c += 1
                if c % buffer == 0 and not implicit_nodes:
                    self.graph.save(self.out_file)
        return c

    def save(self, f: str = None):
        self.graph.save(self.out_file)
