"""Node value type of the Graph plugin API and the path -> sequence rule.

Same constructor, attributes and text form as the reference's Node / concat_values
(prairiedog/node.py:3-37).
"""
import typing

DEFAULT_NODE_TYPE = 'km'          # prairiedog/node.py:3
_FIELDS = ("value", "node_type", "db_id", "labels")


class Node(object):
    """``value`` is the k-mer string (or any other node value of the per-edge API), ``db_id`` the
    backend's numeric id, ``labels`` extra properties."""
    __slots__ = _FIELDS

    def __init__(self, value: str, node_type: str = DEFAULT_NODE_TYPE, db_id: int = None, labels: dict = None):
        self.value, self.node_type, self.db_id, self.labels = value, node_type, db_id, labels

    def as_dict(self) -> dict:
        return {f: getattr(self, f) for f in _FIELDS}

    def __str__(self):
        return "prairiedog.node.Node with vars {}".format(self.as_dict())

    __repr__ = __str__


def concat_values(nodes: typing.Sequence[Node], additional: int = 1) -> str:
    """The sequence a node path spells: consecutive k-mers overlap in all but ``additional`` characters,
    so it is the first value followed by the tail of every later one (prairiedog/node.py:22-37;
    "ABC", "BCD" -> "ABCD")."""
    if not nodes:
        return ""
    head, rest = nodes[0].value, nodes[1:]
    return head + "".join(n.value[len(n.value) - additional:] for n in rest)
