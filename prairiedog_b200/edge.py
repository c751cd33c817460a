"""Edge value type -- API mirror of prairiedog/edge.py:1-28."""


class Edge:
    """A directed labelled edge.  ``edge_value`` is coerced to int; when it is left at -1 and the
    labels carry 'incr', that label becomes the value and is removed (prairiedog/edge.py:6-21)."""

    def __init__(self, src: str, tgt: str, edge_type: str = 'e',
                 edge_value: int = -1, labels: dict = None, db_id=None):
        self.src = src
        self.tgt = tgt
        self.edge_type = edge_type
        if edge_value == -1 and labels is not None and 'incr' in labels:
            edge_value = labels.pop('incr')
        self.edge_value = edge_value if isinstance(edge_value, int) else int(edge_value)
        self.labels = labels
        self.db_id = db_id

    @property
    def origin(self) -> str:
        return self.edge_type

    def __str__(self):
        return "prairiedog.edge.Edge with vars {}".format(vars(self))
