"""Command-line face of the drop-in for the three commands on the hot path (reference: cphasing/cli.py):

    python -m cphasing_b200 hypergraph     <table[.gz]> <contigsizes> <out.hg> [--pairs] ...     (cli.py:3709-3941)
    python -m cphasing_b200 hyperpartition <in.hg|table> <contigsizes> <out.clusters.txt> ...    (cli.py:3945-5360)
    python -m cphasing_b200 kprune         <AlleleTable> <contacts.tsv> -o <out.prune.table> [-n -m -fc -t]   (cli.py:3525-3669)
    python -m cphasing_b200 utils hg2contacts <in.hg> <contigsizes> -o <out.contacts>            (cli.py:9290-9401)

Same positional arguments, option names and defaults (cphasing/_config.py:30-55) for the options that belong to the
accelerated path (HCR BED, contig splitting, prune / allele tables included); the reference's remaining `hyperpartition`
options parse as well -- those that steer stages out of scope (fasta alleles, ultra-long reads, collapsed rescue, the
experimental post-passes) raise NotImplementedError when used, never silently.  This is a thin argparse layer over
HyperExtractor / Extractor / HyperPartition / KPrunerRust.
"""
from __future__ import annotations

import argparse
import logging
import re
import sys

import msgspec
import pandas as pd

logger = logging.getLogger("cphasing_b200")

MIN_CONTACTS, MIN_CIS_WEIGHT, MIN_WEIGHT = 25, 20, 2.0            # _config.py:43-45
MIN_SCAFFOLD_LENGTH, MIN_LENGTH, MAX_ORDER = 2e6, 10_000, 50       # _config.py:38, 54-55


def _size(s):
    """'5m' / '500k' / '2e6' -> int (cli.py humanized sizes)."""
    s = str(s).strip().lower()
    mult = {"k": 1e3, "m": 1e6, "g": 1e9}.get(s[-1])
    return int(float(s[:-1]) * mult) if mult else int(float(s))


def read_contigsizes(path):
    """two-column `contig<TAB>length` file -> DataFrame indexed by contig (utilities.read_chrom_sizes)."""
    df = pd.read_csv(path, sep="\t", header=None, index_col=0, names=["contig", "length"], usecols=[0, 1], comment="#")
    df.index = df.index.astype(str)
    return df


def load_hg(path):
    """cli.py:4811-4834; the binary sidecar next to the file is preferred when it is current (cphasing_b200/hgio.py)"""
    from . import hgio
    return hgio.load(path)


def auto_split_length(sizes, max_split_len=20_000_000, min_split_len=1_000_000, outlier_factor=5.0):
    """`--split-length auto` (utilities.py:1709-1744): split only when one contig is an outlier -- longer than 5 x the 95th
    length percentile and than 20 Mb; the bin length is then that percentile, kept within [1 Mb, 20 Mb] and rounded up to
    the next whole Mb.  None: no splitting."""
    ln = sizes["length"]
    if len(ln) < 10:
        return None
    longest, p95 = ln.max(), ln.quantile(0.95)
    if not (longest > p95 * outlier_factor and longest > max_split_len):
        return None
    sl = int(max(int(min(p95, max_split_len)), min_split_len))
    return (sl // 1_000_000 + 1) * 1_000_000


def split_contigs(sizes, split_length, trim_length=0):
    """cli.py:4776-4790 / utilities.binnify (:1430-1480): contigs, in NAME order, cut into `split_length` bins named
    `ctg|start_end`; with `trim_length` (hyperpartition: --alleles-trim-length) the bin edges of a contig longer than three
    times that are pulled `trim_length` in from both ends.  -> (split sizes DataFrame, dict original contig -> index of its
    first bin)"""
    names, lens, first = [], [], {}
    for c, L in sorted(zip(sizes.index, sizes["length"].astype(int)), key=lambda x: x[0]):
        first[c] = len(names)
        nb = max(int(-(-L // split_length)), 1)
        edges = [min(k * split_length, L) for k in range(nb)] + [L]
        if trim_length and L > 3 * trim_length:
            edges = [min(max(e, trim_length), L - trim_length) for e in edges]
        for a, b in zip(edges[:-1], edges[1:]):
            names.append(f"{c}|{a}_{b}")
            lens.append(b - a)
    return pd.DataFrame({"length": lens}, index=pd.Index(names, name="contig")), first


def _index_contigs(args, sizes):
    """which contigs (or split bins) get a row of H, and in which order -> (sizes of the indexed rows, name -> row,
    split length or None, dict original contig -> row of its first bin or None).
    As in the reference's `hypergraph` command (cli.py:3885-3908) the whitelist only acts when the contigs are NOT split:
    with `--split-length` the reference replaces the contig list by ALL split bins after the whitelist was applied, and
    the row of an alignment is `first bin of its contig + mid // split_length` in that unfiltered numbering, so filtering
    the bins afterwards would put alignments into the wrong rows."""
    sl = getattr(args, "split_length", "none")
    if sl is None or str(sl).lower() in ("none", "no", "null", "0"):
        sl = None
    elif str(sl).lower() == "auto":
        sl = auto_split_length(sizes)                                           # cli.py:4757-4758
        logger.info(f"--split-length auto: {'no splitting' if not sl else sl}")
    else:
        sl = _size(sl)
    boundarys = None
    wl = getattr(args, "extract_whitelist", None)
    if sl:
        sizes, boundarys = split_contigs(sizes, sl, trim_length=getattr(args, "alleles_trim_length", 0) or 0)
        if wl:
            logger.warning("--whitelist has no effect together with --split-length (every split bin is indexed, as in the "
                           "reference)")
    elif wl:                                                                    # cli.py:3890-3892: only whitelisted contigs are indexed
        keep = {l.strip() for l in open(wl) if l.strip()}
        sizes = sizes[[c in keep for c in sizes.index]]
    idx = {c: i for i, c in enumerate(sizes.index)}
    return sizes, idx, sl, boundarys


def _extract(args, sizes):
    """-> (extractor, contig sizes actually indexed: the split bins when --split-length is on)"""
    from . import hypergraph as hg
    sizes, idx, sl, boundarys = _index_contigs(args, sizes)
    sz = sizes["length"].to_dict()
    el = _size(args.edge_length) if str(args.edge_length).lower() not in ("none", "", "0") else 0
    tables = [args.table]
    if getattr(args, "fofn", False):                                            # a file of file names
        tables = [l.strip() for l in open(args.table) if l.strip()]
    pre = dict(hcr_bed=getattr(args, "hcr_bed", None), hcr_invert=getattr(args, "hcr_invert", False), split_length=sl,
               split_contig_boundarys=boundarys)
    if args.pairs:
        return hg.Extractor(tables, idx, sz, min_quality=args.min_quality, edge_length=el, threads=args.threads,
                            **pre), sizes
    return hg.HyperExtractor(tables, idx, sz, min_order=args.min_order, max_order=args.max_order,
                             min_alignments=getattr(args, "min_alignments", 30), min_quality=args.min_quality, edge_length=el,
                             threads=args.threads, **pre), sizes


def cmd_hypergraph(args):
    sizes = read_contigsizes(args.contigsizes)
    ex, _ = _extract(args, sizes)
    ex.save(args.output)
    return 0


def _allele_table(path):
    from .kprune import allele_frame

    class AT:
        data = allele_frame(path)
        filename = path
    return AT()


def cmd_hyperpartition(args):
    from .hyperpartition import HyperPartition
    sizes = read_contigsizes(args.contigsizes)
    split = False
    if args.porec or args.pairs:
        args.table = args.hypergraph
        ex, sizes2 = _extract(args, sizes)
        split = sizes2 is not sizes
        sizes, edges = sizes2, ex.edges
    else:
        edges = load_hg(args.hypergraph)
    n = [int(x) if x else 0 for x in re.split(r":|x|\|", args.n)] if args.n else None      # cli.py:4637
    read_list = lambda p: [l.strip().split()[0] for l in open(p) if l.strip()] if p else None
    at = _allele_table(args.alleletable) if args.alleletable else None
    mode = args.mode
    if mode is None:
        mode = "phasing" if (n and len(n) > 1 and n[1] and n[1] > 1) else ("basal_withprune" if (at or args.prunetable)
                                                                          else "basal")
    if not n and args.resolution1 < 0:
        args.resolution1 = args.init_resolution1          # no group number to search for: one clustering at the start value
    # options that steer stages outside the accelerated path: accepted where they change nothing here, refused loudly where
    # they would (cli.py:3945-4597 defines them all; a pipeline script must not die on an unknown flag)
    if args.ultra_long:
        raise NotImplementedError("-ul / --ultra-long: ultra-long hypergraphs are outside the accelerated path")
    if args.collapsed_contigs:
        raise NotImplementedError("--collapsed-contigs (collapsed rescue) is outside the accelerated path")
    if args.contacts:
        logger.info("--contacts is not needed: the in-process kprune reads its contacts off the hypergraph "
                    "(hyperpartition.py:399-421)")
    if args.kprune_norm_method not in ("none", "cis"):
        logger.warning(f"--kprune-norm-method {args.kprune_norm_method}: defined by the cphasing-rs binary only; the in-process "
                       "kprune scores raw contacts")
    groups_1based = lambda v: [int(x) for x in str(v).split(",") if x.strip()] if v else None
    if args.output_hg and (args.porec or args.pairs):
        ex.save(re.sub(r"\.(porec|pairs)(\.gz)?$|\.pqs$", "", args.hypergraph) + ".hg")                 # cli.py:4796-4834
    hp = HyperPartition(edges, sizes, k=n, alleletable=at, prunetable=args.prunetable, normalize=args.normalize, split=split,
                        fasta=args.fasta, kprune_norm_method=args.kprune_norm_method,
                        merge_use_allele=not args.disable_merge_use_allele,
                        allelic_positive_factor=float(args.allelic_positive_factor),
                        exclude_group_to_second=groups_1based(args.exclude_group_to_second),
                        exclude_group_from_first=groups_1based(args.exclude_group_from_first),
                        is_recluster_contigs=args.recluster_contigs, refine=args.refine,
                        disable_recluster_by_linkage=args.disable_recluster_by_linkage,
                        allelic_factor=args.allelic_factor, cross_allelic_factor=args.cross_allelic_factor,
                        allelic_similarity=args.allelic_similarity, min_allelic_overlap=args.min_allelic_overlap,
                        disable_merge_in_first=args.disable_merge_in_first, whitelist=read_list(args.whitelist),
                        blacklist=read_list(args.blacklist), min_contacts=args.min_contacts,
                        min_length=_size(args.min_length), resolution1=args.resolution1, resolution2=args.resolution2,
                        init_resolution1=args.init_resolution1, init_resolution2=args.init_resolution2,
                        min_weight=args.min_weight, min_cis_weight=args.min_cis_weight,
                        cluster_method=args.cluster_method, min_quality1=args.min_quality1,
                        min_quality2=args.min_quality2, min_scaffold_length=_size(args.min_scaffold_length),
                        threshold=args.threshold, max_round=args.max_round, threads=args.threads)
    if args.merge_cluster:                                                                  # cli.py:4885-4892, 5005-5018
        if mode == "phasing":
            logger.error("Imcomplete function of merge cluster in phasing mode.")            # the reference's own message
            sys.exit(-1)
        groups = [l.rstrip("\n").split("\t")[2].split(" ") for l in open(args.merge_cluster) if l.strip()]
        hp.merge_cluster(groups)
        hp.to_cluster(args.output)
        return 0
    if mode == "phasing" or args.incremental:                                               # cli.py:5334-5346
        first = None
        if args.first_cluster:
            first = [l.rstrip("\n").split("\t")[2].split(" ") for l in open(args.first_cluster) if l.strip()]
        hp.incremental_partition(n, first_cluster=first)
    else:                                                                                    # cli.py:5347-5352
        hp.single_partition(k=n[0] if n else None, is_kprune=(mode == "basal_withprune" and at is not None))
    hp.to_cluster(args.output)
    return 0


def cmd_kprune(args):
    """cli.py:3642-3669: AlleleTable + contacts -> prune table, through the KPrunerRust-shaped class (in-process, no binary)"""
    from .kprune import KPrunerRust
    kp = KPrunerRust(args.alleletable, args.contacts, args.output, method=args.method, first_cluster=args.first_cluster,
                     count_re=args.count_re, norm_method=args.norm_method, threads=args.threads).run()
    logger.info(f"Wrote {len(kp.allelic_pairs)} allelic and {len(kp.weak_pairs)} cross-allelic pairs to `{args.output}`.")
    return 0


def cmd_hg2contacts(args):
    """cli.py:9350-9401: .hg -> H -> unweighted clique expansion -> 3-column contacts (what `kprune` reads)"""
    from .algorithms.hypergraph import HyperGraph
    sizes = read_contigsizes(args.contigsizes)
    edges = load_hg(args.hypergraph)
    HG = HyperGraph(edges, min_quality=args.min_quality, mapq_filter=False)
    H = HG.incidence_matrix(args.min_contacts)
    A = HyperGraph.clique_expansion_init(H, min_weight=args.min_weight, use_high_order=False)    # hyperpartition.py:1566
    HyperGraph.to_contacts(A, HG.nodes, min_weight=args.min_weight, output=args.output)
    return 0


def build_parser():
    ap = argparse.ArgumentParser(prog="cphasing_b200", description=__doc__.split("\n\n")[0])
    sub = ap.add_subparsers(dest="command", required=True)

    def extract_opts(p):
        p.add_argument("-prs", "--pairs", action="store_true", help="input is a 4DN pairs file")
        p.add_argument("-pct", "--porec", action="store_true", help="input is a Pore-C table (default for `hypergraph`)")
        p.add_argument("-e", "--edge-length", default="5m")
        p.add_argument("--min-order", type=int, default=2)
        p.add_argument("--max-order", type=int, default=MAX_ORDER)
        p.add_argument("-q", "--min-quality", type=int, default=1)
        p.add_argument("-t", "--threads", type=int, default=4)
        p.add_argument("-hcr", "--hcr-bed", dest="hcr_bed", default=None, help="BED of high-confidence regions")
        p.add_argument("-hcr-invert", "--hcr-invert", dest="hcr_invert", action="store_true")
        p.add_argument("-sl", "--split-length", dest="split_length", default=None,
                       help="split contigs into bins of this length; `auto` (the default of hyperpartition, as in the "
                            "reference): only when one contig is an outlier in length; `none`: off")

    g = sub.add_parser("hypergraph", aliases=["hg"])
    g.add_argument("table"); g.add_argument("contigsizes"); g.add_argument("output")
    extract_opts(g)
    g.add_argument("-ma", "--min-alignments", dest="min_alignments", type=int, default=30)
    g.add_argument("--fofn", action="store_true", help="the first argument is a file of contact-table file names")
    g.add_argument("-wl", "--whitelist", dest="extract_whitelist", default=None)
    g.set_defaults(fn=cmd_hypergraph)

    h = sub.add_parser("hyperpartition", aliases=["hp"])
    h.add_argument("hypergraph"); h.add_argument("contigsizes"); h.add_argument("output")
    extract_opts(h)
    h.add_argument("--mode", choices=["basal", "haploid", "phasing", "basal_withprune"], default=None)
    h.add_argument("-n", default=None, help="group numbers, e.g. 12 or 12:4")
    h.add_argument("-at", "--alleletable", default=None)
    h.add_argument("-pt", "--prunetable", default=None)
    h.add_argument("-norm", "--normalize", action="store_true")
    h.add_argument("-af", "--allelic-factor", type=float, default=0)
    h.add_argument("-caf", "--cross-allelic-factor", type=float, default=0)
    h.add_argument("-as", "--allelic-similarity", type=float, default=0.85)
    h.add_argument("-mao", "--min-allelic-overlap", type=float, default=0.3)
    h.add_argument("-fc", "--first-cluster", default=None)
    h.add_argument("--disable-merge-in-first", action="store_true")
    h.add_argument("-wl", "--whitelist", default=None)
    h.add_argument("-bl", "--blacklist", default=None)
    h.add_argument("-mc", "--min-contacts", type=int, default=MIN_CONTACTS)
    h.add_argument("-ml", "--min-length", default=str(MIN_LENGTH))
    h.add_argument("-r1", "--resolution1", type=float, default=-1.0)
    h.add_argument("-r2", "--resolution2", type=float, default=-1.0)
    h.add_argument("-ir1", "--init-resolution1", type=float, default=1.0)
    h.add_argument("-ir2", "--init-resolution2", type=float, default=1.0)
    h.add_argument("-mw", "--min-weight", type=float, default=MIN_WEIGHT)
    h.add_argument("-mcw", "--min-cis-weight", type=float, default=MIN_CIS_WEIGHT)
    h.add_argument("-q1", "--min-quality1", type=int, default=1)
    h.add_argument("-q2", "--min-quality2", type=int, default=1)
    h.add_argument("-ms", "--min-scaffold-length", default=str(MIN_SCAFFOLD_LENGTH))
    h.add_argument("--cluster-method", choices=["louvain", "leiden"], default="louvain")
    h.add_argument("--threshold", type=float, default=0.01)
    h.add_argument("--max-round", type=int, default=1)
    # the rest of cli.py:3945-4597, so that a command line written for the reference parses; see cmd_hyperpartition for what
    # each of them does here
    h.add_argument("-c", "--contacts", default=None)
    h.add_argument("-ul", "--ultra-long", dest="ultra_long", default=None)
    h.add_argument("-ul-w", "--ul-weight", dest="ul_weight", type=int, default=1)
    h.add_argument("-f", "--fasta", default=None)
    h.add_argument("-alleles-k", "--alleles-k", "--alleles-kmer-size", dest="alleles_kmer_size", type=int, default=19)
    h.add_argument("-alleles-w", "--alleles-w", "--alleles-window-size", dest="alleles_window_size", type=int, default=19)
    h.add_argument("-alleles-m", "--alleles-m", "--alleles-minimum-similarity", dest="alleles_minimum_similarity", type=float,
                   default=0.2)
    h.add_argument("-alleles-d", "--alleles-d", "--alleles-diff-thres", dest="alleles_diff_thres", type=float, default=0.1)
    h.add_argument("-alleles-tl", "--alleles-trim-length", dest="alleles_trim_length", type=int, default=25000)
    h.add_argument("-inc", "--incremental", action="store_true")
    h.add_argument("-knm", "--kprune-norm-method", dest="kprune_norm_method", choices=["none", "re", "cis", "cis_density"],
                   default="cis")
    h.add_argument("--merge-cluster", dest="merge_cluster", default=None)
    h.add_argument("--disable-merge-use-allele", "--disable-merge-first-use-allele", dest="disable_merge_use_allele",
                   action="store_true")
    h.add_argument("-apf", "--allelic-positive-factor", dest="allelic_positive_factor", default="3.0")
    h.add_argument("--exclude-group-to-second", "--exclude-to-second", dest="exclude_group_to_second", default=None)
    h.add_argument("--exclude-from-first", "--exclude-group-from-first", dest="exclude_group_from_first", default=None)
    h.add_argument("--enable-misassembly-remove", dest="enable_misassembly_remove", action="store_true", default=True)
    h.add_argument("--is-recluster-contigs", "--recluster-contigs", dest="recluster_contigs", action="store_true")
    h.add_argument("--disable-recluster-by-linkage", dest="disable_recluster_by_linkage", action="store_true")
    h.add_argument("--refine", action="store_true")
    h.add_argument("--collapsed-contigs", dest="collapsed_contigs", default=None)
    h.add_argument("--output-hg", dest="output_hg", action="store_true")
    h.set_defaults(fn=cmd_hyperpartition, split_length="auto")

    k = sub.add_parser("kprune")                                                   # cli.py:3525-3645
    k.add_argument("alleletable", metavar="AlleleTable")
    k.add_argument("contacts", metavar="INPUT_CONTACTS_PATH")
    k.add_argument("-o", "--output", default="prune.contig.table")
    k.add_argument("-m", "--method", choices=["fast", "precise"], default="precise",
                   help="`greedy` exists only inside the cphasing-rs binary (parity unpinned) and is not offered")
    k.add_argument("-fc", "--first-cluster", dest="first_cluster", default=None)
    k.add_argument("-cr", "--countre", "--count-re", dest="count_re", default=None)
    k.add_argument("-n", "--norm-method", dest="norm_method", choices=["none", "cis", "cis_unique", "auto"], default="none")
    k.add_argument("-t", "--threads", type=int, default=4)
    k.set_defaults(fn=cmd_kprune)

    u = sub.add_parser("utils")
    usub = u.add_subparsers(dest="util", required=True)
    c = usub.add_parser("hg2contacts")                                             # cli.py:9290-9401
    c.add_argument("hypergraph"); c.add_argument("contigsizes")
    c.add_argument("-o", "--output", default="hypergraph.expansion.contacts")
    c.add_argument("-mc", "--min-contacts", type=int, default=MIN_CONTACTS)
    c.add_argument("-mw", "--min-weight", type=float, default=1.0)
    c.add_argument("-q", "--min-quality", type=int, default=1)
    c.set_defaults(fn=cmd_hg2contacts)
    return ap


def main(argv=None):
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(name)s %(levelname)s %(message)s", stream=sys.stderr)
    args = build_parser().parse_args(argv)
    return args.fn(args)


if __name__ == "__main__":
    sys.exit(main())
