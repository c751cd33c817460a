"""GraphRef: per-run bookkeeping -- API mirror of prairiedog/graph_ref.py:17-103, gref.py:9-19:
first-seen-order ids of k-mers / files (encode=True) and the MIC graph-label table."""
import os

import numpy as np

MIC_CSV = 'samples/public_mic_class_dataframe.csv'      # prairiedog/config.py:9


class GRef:
    @staticmethod
    def _upsert_map(d: dict, value, prev_value: int = 0, zero=False) -> int:
        """First-seen order ids (gref.py:9-19)."""
        if value not in d:
            d[value] = len(d) + prev_value + 1 - (1 if zero else 0)
        return d[value]


class GraphRef(GRef):
    def __init__(self, mic_csv=None):
        """graph_ref.py:20-41: the MIC table is read when the csv exists (index = genome short name,
        one column per drug), else MIC_DF is None and graph labels are unavailable."""
        self.max_num_nodes = 0
        self.MIC_DF = None
        path = mic_csv if mic_csv is not None and os.path.isfile(mic_csv) else (MIC_CSV if os.path.isfile(MIC_CSV) else None)
        if path is not None:
            import pandas as pd
            self.MIC_DF = pd.read_csv(path, index_col=0)
            self.MIC_COLUMNS = self.MIC_DF.columns
        self.mic_map = {}      # MIC value -> int, first-seen order starting at 1
        self.kmer_map = {}
        self.file_map = {}

    @staticmethod
    def _one_hot(node_label: int, num_unique_node_labels: int) -> np.ndarray:
        """graph_ref.py:43-47"""
        v = np.zeros(num_unique_node_labels, dtype=np.int64)
        v[node_label] = 1
        return v

    def _record_graph_label(self, km):
        """graph_ref.py:66-76: every MIC value of the genome's row gets an id."""
        series = self.MIC_DF.loc[GraphRef.get_short_name(km), :]
        for label in self.MIC_COLUMNS:
            self._upsert_map(self.mic_map, series[label])

    def get_graph_label(self, km, column: str) -> int:
        """graph_ref.py:90-95"""
        series = self.MIC_DF.loc[GraphRef.get_short_name(km), :]
        return self.mic_map[series[column]]

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
