"""Edge value type of the Graph plugin API.

Same constructor, attributes and text form as the reference's Edge (prairiedog/edge.py:1-28), written
around one rule: the value of an edge is always an int, and an edge built from LemonGraph properties
carries it in the 'incr' label instead of the edge_value argument.
"""

_FIELDS = ("src", "tgt", "edge_type", "edge_value", "labels", "db_id")
_UNSET = -1


def _resolve_value(edge_value, labels):
    """int value of an edge.  An unset value (-1) is taken from labels['incr'], which is consumed
    (prairiedog/edge.py:13-15); anything that is not an int is converted (edge.py:18-19)."""
    if edge_value == _UNSET and labels is not None and 'incr' in labels:
        edge_value = labels.pop('incr')
    return edge_value if isinstance(edge_value, int) else int(edge_value)


class Edge(object):
    __slots__ = _FIELDS

    def __init__(self, src: str, tgt: str, edge_type: str = 'e', edge_value: int = _UNSET,
                 labels: dict = None, db_id=None):
        self.src, self.tgt, self.edge_type = src, tgt, edge_type
        self.edge_value = _resolve_value(edge_value, labels)
        self.labels, self.db_id = labels, db_id

    @property
    def origin(self) -> str:
        """the label that says where the edge comes from: '{contig header} in {file}'"""
        return self.edge_type

    def as_dict(self) -> dict:
        return {f: getattr(self, f) for f in _FIELDS}

    def __str__(self):
        return "prairiedog.edge.Edge with vars {}".format(self.as_dict())

    __repr__ = __str__
