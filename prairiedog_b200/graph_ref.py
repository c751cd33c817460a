"""GraphRef: per-run bookkeeping -- API mirror of prairiedog/graph_ref.py:17-103, gref.py:9-19
(without the pandas MIC-label table, which is outside the hot path)."""
import os


class GRef:
    @staticmethod
    def _upsert_map(d: dict, value, prev_value: int = 0, zero=False) -> int:
        """First-seen order ids (gref.py:9-19)."""
        if value not in d:
            d[value] = len(d) + prev_value + 1 - (1 if zero else 0)
        return d[value]


class GraphRef(GRef):
    def __init__(self, mic_csv=None):
        self.max_num_nodes = 0
        self.MIC_DF = None
        self.mic_map = {}
        self.kmer_map = {}
        self.file_map = {}

    @staticmethod
    def get_short_name(km):
        return os.path.basename(km.filepath).split('.')[0]

    def index_kmers(self, km):
        """graph_ref.py:78-88."""
        if km.unique_kmers > self.max_num_nodes:
            self.max_num_nodes = km.unique_kmers

    def node_label(self, kmer: str) -> int:
        return self._upsert_map(self.kmer_map, kmer, zero=True)

    def edge_label(self, km) -> int:
        return self._upsert_map(self.file_map, GraphRef.get_short_name(km), zero=True)
