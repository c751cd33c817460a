"""Node value type and sequence reconstruction -- API mirror of prairiedog/node.py:6-37."""
import typing

DEFAULT_NODE_TYPE = 'km'


class Node:
    """A graph node: ``value`` is the k-mer string (prairiedog/node.py:6-16)."""

    def __init__(self, value: str, node_type: str = DEFAULT_NODE_TYPE,
                 db_id: int = None, labels: dict = None):
        self.value = value
        self.node_type = node_type
        self.db_id = db_id
        self.labels = labels

    def __str__(self):
        return "prairiedog.node.Node with vars {}".format(vars(self))


def concat_values(nodes: typing.Sequence[Node], additional: int = 1) -> str:
    """Rebuilds the sequence spelled by a node path: the first value plus the last ``additional``
    characters of every following value (prairiedog/node.py:22-37)."""
    if len(nodes) == 0:
        return ""
    parts = [nodes[0].value]
    for node in nodes[1:]:
        v = node.value
        parts.append(v[len(v) - additional:])
    return "".join(parts)
