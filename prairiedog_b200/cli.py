"""Console entry point: the reference's `prairiedog` CLI (prairiedog/cli.py:68-90) and the graphing
rules of its Snakefile (Snakefile:22-25,46-111) for the b200 backend.

    python -m prairiedog_b200.cli pangenome samples/ --k 11 --out outputs/pangenome.pdg
    python -m prairiedog_b200.cli query ATACGACGCCA CGTCCGGACGT --graph outputs/pangenome.pdg
    python -m prairiedog_b200.cli query-batch pairs.txt --graph outputs/pangenome.pdg
    python -m prairiedog_b200.cli export --graph outputs/pangenome.pdg --format rdf --out outputs/pangenome.rdf

`pangenome` is Snakemake's `kmers` + `pangenome` + `done` rules in one process: every FASTA file of
the samples directory (the reference's suffix list) goes to the GPU in one batch; a file the bulk path
cannot represent exactly runs through Kmers + SubgraphRef.update_graph, the reference's own loop.
"""
import json
import logging
import os
import sys

import click
import numpy as np

log = logging.getLogger("prairiedog")

SUFFIXES = ('.fna', '.fasta', '.fa')        # Snakefile:22-24
DEFAULT_GRAPH = os.path.join('outputs', 'pangenome.pdg')


def sample_files(samples_dir):
    """The Snakefile's INPUTS, in sorted order so that type ids are reproducible."""
    return sorted(os.path.join(samples_dir, f) for f in os.listdir(samples_dir) if f.endswith(SUFFIXES))


def build_pangenome(paths, k=11, device=0, graph=None):
    """One pangenome graph from FASTA files; returns (graph, {path: k-mers covered})."""
    from .b200_graph import B200Graph
    from .graph_ref import GraphRef
    from .kmers import Kmers
    from .subgraph_ref import SubgraphRef
    g = graph if graph is not None else B200Graph(k=k, device=device)
    files = [(os.path.basename(p), np.fromfile(p, dtype=np.uint8), None) for p in paths]
    try:
        counts = g.add_fasta_batch(files)
        failed = []
    except ValueError as e:
        if not hasattr(e, "failed"):
            raise
        counts, failed = e.counts, e.failed
    if failed:
        gr, sg = GraphRef(), SubgraphRef(g)
        for i in failed:
            log.info("%s takes the per-edge path", paths[i])
            km = Kmers(paths[i], k)
            gr.index_kmers(km)
            counts[i] = sg.update_graph(km, gr)
    g.save()
    return g, dict(zip(paths, counts))


def connect(graph_file, samples_dir, k, device=0):
    """cli.py:24-31 connect_lemongraph: the saved pangenome if there is one, else build it now."""
    from .b200_graph import B200Graph
    if graph_file and os.path.exists(graph_file):
        return B200Graph.load(graph_file, device=device)
    if samples_dir and os.path.isdir(samples_dir):
        g, _ = build_pangenome(sample_files(samples_dir), k=k, device=device)
        return g
    raise click.ClickException("no graph at {} and no samples directory {}".format(graph_file, samples_dir))


@click.group()
@click.option('--debug/--no-debug', default=False)
def cli(debug):
    click.echo('Debug mode is %s' % ('on' if debug else 'off'))
    logging.basicConfig(level=logging.DEBUG if debug else logging.INFO, format="%(message)s")


@cli.command()
@click.argument('samples_dir', nargs=1)
@click.option('--k', default=11, show_default=True, help='k-mer length (config.yaml: k)')
@click.option('--out', default=DEFAULT_GRAPH, show_default=True, help='snapshot file to write')
@click.option('--device', default=0, show_default=True)
def pangenome(samples_dir, k, out, device):
    """Create a pan-genome from every FASTA file of SAMPLES_DIR."""
    paths = sample_files(samples_dir)
    if not paths:
        raise click.ClickException("no {} files in {}".format('/'.join(SUFFIXES), samples_dir))
    g, counts = build_pangenome(paths, k=k, device=device)
    for p, c in counts.items():
        click.echo("{}\t{} k-mers".format(os.path.basename(p), c))
    n_nodes, n_edges, _ = g.engine.counts()
    g.snapshot(out)
    click.echo("pangenome of {} genomes: {} nodes, {} edges -> {}".format(len(paths), n_nodes, n_edges, out))


@cli.command()
@click.argument('src', nargs=1)
@click.argument('dst', nargs=1)
@click.option('--backend', default='b200', help='Backend graph database')
@click.option('--graph', 'graph_file', default=DEFAULT_GRAPH, show_default=True)
@click.option('--samples', default='samples/', show_default=True, help='used when the graph file is missing')
@click.option('--k', default=11, show_default=True)
def query(src, dst, backend, graph_file, samples, k):
    """Query the pan-genome for a path between two k-mers."""
    if backend != 'b200':
        raise click.ClickException("this build only has the b200 backend (got {})".format(backend))
    from .prairiedog import Prairiedog, PrettyHits
    g = connect(graph_file, samples, k)
    hits = Prairiedog(g=g).query(src, dst)
    click.echo("Found: {}".format(PrettyHits(hits)))


@cli.command('query-batch')
@click.argument('pairs_file', nargs=1)
@click.option('--graph', 'graph_file', default=DEFAULT_GRAPH, show_default=True)
@click.option('--samples', default='samples/', show_default=True)
@click.option('--k', default=11, show_default=True)
@click.option('--strings/--no-strings', default=True, help='spell every hit (else only type, i, j)')
def query_batch(pairs_file, graph_file, samples, k, strings):
    """All paths for every 'SRC DST' line of PAIRS_FILE in one GPU batch; one JSON line per pair."""
    pairs = []
    with open(pairs_file) as f:
        for ln in f:
            ln = ln.split()
            if len(ln) >= 2 and not ln[0].startswith('#'):
                pairs.append((ln[0], ln[1]))
    g = connect(graph_file, samples, k)
    for (a, b), hits in zip(pairs, g.query_many(pairs, strings=strings)):
        sys.stdout.write(json.dumps({"src": a, "dst": b, "hits": hits}) + "\n")


@cli.command()
@click.option('--graph', 'graph_file', default=DEFAULT_GRAPH, show_default=True)
@click.option('--format', 'fmt', type=click.Choice(['rdf', 'edgelist']), default='rdf', show_default=True)
@click.option('--out', required=True)
def export(graph_file, fmt, out):
    """Write the graph as Dgraph RDF N-Quads (dgraph.py:576-613) or a networkx edgelist."""
    from . import snapshot
    from .b200_graph import B200Graph
    g = B200Graph.load(graph_file)
    n = snapshot.write_rdf(g, out) if fmt == 'rdf' else snapshot.write_edgelist(g, out)
    click.echo("{} edges -> {}".format(n, out))


def main():
    cli()     # pylint: disable=no-value-for-parameter


if __name__ == "__main__":
    sys.exit(main())
