"""Exception types -- API mirror of prairiedog/errors.py:8-27."""
import logging

log = logging.getLogger("prairiedog")


class Error(Exception):
    pass


class GraphException(Error):
    """Generic graph problem; dumps the graph like the reference (prairiedog/errors.py:14-27)."""

    def __init__(self, g):
        nodes, edges = g.nodes, g.edges
        super(GraphException, self).__init__(
            "Problem with Graph with nodes {} and edges {}".format(nodes, edges))
        for node in nodes:
            log.error(node)
        for edge in edges:
            log.error(edge)
