"""CPU restatement of the reference's pangenome-graph hot path (pure Python, small cases).

TEST INFRASTRUCTURE.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` leg may import this module; the product package ``prairiedog_b200`` never does.

Parity status: PINNED.  tests/test_oracle_pinning.py checks this file against (a) the golden
vectors the reference's own tests hold (tests/test_kmers.py:65-69, tests/test_prebuilt.py:18-75,
tests/test_connected.py:26-380, the golden DB tests/pangenome.lemongraph dumped to
tests/golden/prebuilt_pangenome.json) and (b) outputs of the unmodified reference run in the
build container (tests/golden/*.json, made by tests/golden/make_golden.py).

Each function names the reference lines it restates (paths relative to /root/reference).
"""
import os

ET_DELIMITER = " in "          # prairiedog/subgraph_ref.py:17
DEFAULT_LIMIT = 1000           # sys.getrecursionlimit() in prairiedog/lemon_graph.py:229-236


def load_fasta(path):
    """prairiedog/kmers.py:51-77 (_load): rstrip every line, '>' lines are headers kept verbatim,
    other lines are concatenated; a sequence is flushed when the next header arrives (only if
    non-empty) and once more, unconditionally, at EOF."""
    headers, sequences, seq = [], [], ""
    with open(path) as f:
        for line in f:
            ln = line.rstrip()
            if ln.startswith(">"):
                if len(seq) != 0:
                    sequences.append(seq)
                    seq = ""
                headers.append(ln)
            else:
                seq += ln
    sequences.append(seq)
    return headers, sequences


def emit_kmers(headers, sequences, k):
    """prairiedog/kmers.py:79-124 (has_next / contig_has_next / next) as a generator of
    (contig_index, header, kmer).  Reproduces the skip-one-short-contig rule and raises
    IndexError exactly where the reference does."""
    li, pi = 0, 0
    while not (li == len(sequences) - 1 and pi + k > len(sequences[li])):
        if pi + k > len(sequences[li]):          # kmers.py:106-108
            li += 1
            pi = 0
        if k > len(sequences[li]) - 1:           # kmers.py:111-117 (IndexError if li runs off)
            li += 1
            pi = 0
        yield li, headers[li], sequences[li][pi:pi + k]
        pi += 1


def kmer_counts(headers, sequences, k):
    """prairiedog/kmers.py:134-150 (_count_unique): (len(km), km.unique_kmers)."""
    st, c = set(), 0
    for _, _, kmer in emit_kmers(headers, sequences, k):
        st.add(kmer)
        c += 1
    return c, len(st)


class _Cursor(object):
    """State machine of prairiedog/kmers.py:79-124 over already loaded contigs."""

    def __init__(self, headers, sequences, k):
        self.h, self.s, self.k, self.li, self.pi = headers, sequences, k, 0, 0

    def _end(self):                                # kmers.py:79-80
        return self.pi + self.k > len(self.s[self.li])

    def has_next(self):                            # kmers.py:82-88
        return not (self.li == len(self.s) - 1 and self._end())

    def contig_has_next(self):                     # kmers.py:90-94
        return not self._end()

    def next(self):                                # kmers.py:96-124
        if not self.has_next():
            return "", ""
        if self._end():
            self.li, self.pi = self.li + 1, 0
        if self.k > len(self.s[self.li]) - 1:
            self.li, self.pi = self.li + 1, 0
        out = self.h[self.li], self.s[self.li][self.pi:self.pi + self.k]
        self.pi += 1
        return out


def genome_edges(path, k, lemongraph_nodes=True):
    """prairiedog/subgraph_ref.py:51-102 (update_graph): returns (nodes, edges, c).

    edges: list of (src, tgt, edge_type, edge_value) with edge_type = header + ' in ' + basename
    (header = the one returned with the TARGET k-mer, subgraph_ref.py:63,75-76) and edge_value =
    0-based index of the edge inside its run.  nodes: with the LemonGraph backend only edge
    endpoints become nodes (subgraph_ref.py:55-57,66-68; lemon_graph.py:115-121); other backends
    also upsert the first k-mer of every run.  c = number of emitted k-mers."""
    headers, sequences = load_fasta(path)
    base = os.path.basename(path)
    cur = _Cursor(headers, sequences, k)
    nodes, edges, c = set(), [], 0
    while cur.has_next():                          # subgraph_ref.py:51
        _, kmer1 = cur.next()
        if not lemongraph_nodes:
            nodes.add(kmer1)
        c += 1
        edge_c = 0
        while cur.contig_has_next():               # subgraph_ref.py:62
            header2, kmer2 = cur.next()
            edges.append((kmer1, kmer2, header2 + ET_DELIMITER + base, edge_c))
            nodes.add(kmer1)
            nodes.add(kmer2)
            kmer1 = kmer2
            c += 1
            edge_c += 1
    return nodes, edges, c


class SpecGraph(object):
    """Value-level model of the graph content and queries.

    nodes: set of value strings (node_type is always 'km' on this path).
    edges: set of (src, tgt, edge_type, edge_value) -- LemonGraph upserts edges uniquely on
    (type, value, src, tgt) (lemongraph/lib/lemongraph.c:434-443,674-685)."""

    def __init__(self):
        self.nodes = set()
        self.edges = set()

    def upsert_node(self, value):                 # prairiedog/lemon_graph.py:99-113
        self.nodes.add(value)

    def add_edge(self, src, tgt, edge_type="e", edge_value=-1):   # lemon_graph.py:115-128
        self.nodes.add(src)
        self.nodes.add(tgt)
        self.edges.add((src, tgt, edge_type, int(edge_value)))

    def add_genome(self, path, k=11):
        nodes, edges, c = genome_edges(path, k)
        self.nodes |= nodes
        self.edges |= set(edges)
        return c

    def canonical(self):
        return sorted(self.nodes), sorted(self.edges)

    def filter(self, bottom_percentile=0.9):
        """Graph.filter() as the one backend that implements it defines it (prairiedog/networkx_graph.py:53-75;
        LGGraph.filter is `pass`, lemon_graph.py:172-173): drop the nodes of low degree, then the isolates.  The
        labelled multigraph is read as the nx.Graph that backend keeps (networkx_graph.py:11-26: undirected, one
        edge per node pair whatever the labels), nodes go together with every labelled edge that touches them.
        Returns the node values that were removed."""
        keep = filter_keep(self.nodes, ((e[0], e[1]) for e in self.edges), bottom_percentile)
        gone = self.nodes - keep
        self.nodes = set(keep)
        self.edges = set(e for e in self.edges if e[0] in keep and e[1] in keep)
        return gone

    def out_edges(self, a):
        return [e for e in self.edges if e[0] == a]

    def in_edges(self, b):
        return [e for e in self.edges if e[1] == b]

    def connected(self, a, b):
        """prairiedog/lemon_graph.py:178-219 + graph.py:82-113 (matching_edges): start edges are
        the out-edges of a that share an edge_type with some in-edge of b whose value is >=."""
        src_edges, tgt_edges = self.out_edges(a), self.in_edges(b)
        if not src_edges or not tgt_edges:
            return False, ()
        st = set()
        for s in src_edges:
            for t in tgt_edges:
                if s[2] == t[2] and s[3] <= t[3]:
                    st.add(s)
        return (len(st) > 0), tuple(sorted(st))

    def _chain(self, edge_type, i, j):
        """prairiedog/lemon_graph.py:244-264 (_find_path): every chain of edges
        (edge_type, i), (edge_type, i+1) ... (edge_type, j) whose endpoints connect."""
        by_val = {}
        for e in self.edges:
            if e[2] == edge_type and i <= e[3] <= j:
                by_val.setdefault(e[3], []).append(e)
        chains = [[e[0], e[1]] for e in by_val.get(i, [])]
        for v in range(i + 1, j + 1):
            nxt = []
            for ch in chains:
                for e in by_val.get(v, []):
                    if e[0] == ch[-1]:
                        nxt.append(ch + [e[1]])
            chains = nxt
        return chains

    def path(self, a, b, limit=DEFAULT_LIMIT):
        """prairiedog/lemon_graph.py:275-334: returns sorted [(edge_type, i, j, (node values))].
        A pair is skipped when j - i > limit (lemon_graph.py:229-236)."""
        ok, src_edges = self.connected(a, b)
        if not ok:
            return []
        hits = []
        for s in src_edges:
            tgt_edges = [t for t in self.in_edges(b) if t[2] == s[2]]
            for t in tgt_edges:
                if s[3] > t[3]:
                    continue
                if t[3] - s[3] > limit:
                    continue
                chains = self._chain(s[2], s[3], t[3])
                assert len(chains) == 1, "lemon_graph.py:254-258 asserts exactly one chain"
                hits.append((s[2], s[3], t[3], tuple(chains[0])))
        return sorted(hits)

    def query(self, a, b, limit=DEFAULT_LIMIT):
        """prairiedog/prairiedog.py:17-33 + node.py:22-37: sorted [(string, edge_type)]."""
        out = []
        for edge_type, _, _, nodes in self.path(a, b, limit):
            out.append((concat_values(nodes), edge_type))
        return sorted(out)


def filter_keep(nodes, pairs, bottom_percentile=0.9):
    """Node values that survive prairiedog/networkx_graph.py:53-75 on the undirected simple graph with the given
    nodes and (a, b) pairs.
      :53-67  _filter_degree: max_degrees = max degree; bottom_degrees = int(bottom_percentile * max_degrees);
              remove every node with degree <= bottom_degrees
      :69-71  _filter_isolates: remove every node left without a neighbour
    Degree is networkx's: number of distinct neighbours, a self-loop counting twice (nx.Graph.degree)."""
    nodes = set(nodes)
    adj = dict((v, set()) for v in nodes)
    for a, b in pairs:
        adj.setdefault(a, set()).add(b)
        adj.setdefault(b, set()).add(a)
    nodes = set(adj)
    if not nodes:
        return set()

    def degree(v, alive):
        return sum(2 if u == v else 1 for u in adj[v] if u in alive)

    max_degrees = max(degree(v, nodes) for v in nodes)
    bottom_degrees = int(bottom_percentile * max_degrees)
    keep = set(v for v in nodes if degree(v, nodes) > bottom_degrees)
    return set(v for v in keep if degree(v, keep) > 0)


def concat_values(values, additional=1):
    """prairiedog/node.py:22-37 on plain value strings."""
    if len(values) == 0:
        return ""
    s = values[0]
    for v in values[1:]:
        s += v[len(v) - additional:]
    return s
