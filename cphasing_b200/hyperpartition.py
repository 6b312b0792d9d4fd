"""Drop-in for the device-relevant core of ``cphasing.hyperpartition.HyperPartition``
(reference: cphasing/hyperpartition.py).

Covered: construction (filter -> H -> normalisation weights ->
prune pairs), ``single_partition`` with the resolution auto-search, ``filter_cluster``, ``_merge``,
``recluster_by_linkage``, in-process kprune on A, ``to_cluster``.  All matrix arithmetic runs in libcph; this module
is the reference's control flow over n- and g-sized host objects.  Not covered here: the second round
(``incremental_partition``), contig splitting, ``alleles()`` (sequence similarity), ultra-complex re-partitioning,
``raw_sort``: they either sit outside the path or are the next rows of SURVEY.md 8f.
"""
from __future__ import annotations

import logging
from collections import defaultdict
from pathlib import Path

import numpy as np
import pandas as pd
import torch

from . import device as D
from . import graphops
from ._lib import call_ws, ptr
from .algorithms.hypergraph import HyperGraph, IRMM, extract_incidence_matrix2
from .kprune import KPruneHyperGraph

logger = logging.getLogger(__name__)


def group_sums(A, K, want_S=False, want_V=False):
    """device S = A M (n x g) and / or V = M^T A M (g x g) for the groups K (lists of vertex ids)."""
    a = graphops.upload_adjacency(A)
    n, g = a.n, len(K)
    lab = np.full(n, -1, dtype=np.int32)
    for gid, grp in enumerate(K):
        lab[np.asarray(list(grp), dtype=np.int64)] = gid
    dev = a.indptr.device
    labd = D.to_dev(lab, torch.int32, dev)
    S = torch.empty((n, g), dtype=torch.float64, device=dev) if want_S else None
    V = torch.empty((g, g), dtype=torch.float64, device=dev) if want_V else None
    d64 = a.data64
    d32 = a.data32 if d64 is None else None
    call_ws("cph_group_sums", dev, ptr(a.indptr), ptr(a.indices), ptr(d32), ptr(d64), n, a.nnz, ptr(labd), g, ptr(S), ptr(V))
    return (None if S is None else S.cpu().numpy()), (None if V is None else V.cpu().numpy())


class HyperPartition:
    """Same constructor arguments as the reference (hyperpartition.py:114-166); options that belong to stages
    outside the accelerated path are accepted and stored."""

    def __init__(self, edges, contigsizes, ultra_long=None, ul_weight=1.0, k=None, fasta=None, alleletable=None,
                 prunetable=None, normalize=False, split=False, contacts=None, kprune_norm_method="auto", count_re=None,
                 allelic_factor=-1, cross_allelic_factor=0.3, allelic_similarity=0.8, min_allelic_overlap=0.1,
                 ultra_complex=5.0, disable_merge_in_first=False, merge_use_allele=False, allelic_positive_factor=3.0,
                 exclude_group_to_second=None, exclude_group_from_first=None, whitelist=None, blacklist=None,
                 min_contacts=3, min_length=10000, resolution1=1.0, resolution2=1.0, init_resolution1=0.8,
                 init_resolution2=0.8, min_weight=1.0, min_cis_weight=1.0, cluster_method='louvain', min_quality1=1,
                 min_quality2=2, mapq_filter_1_to_2=False, min_scaffold_length=10000, is_remove_misassembly=False,
                 is_recluster_contigs=False, disable_recluster_by_linkage=False, refine=False,
                 collapsed_contigs_df=None, threshold=0.01, max_round=1, threads=4, chunksize=None, **kw):
        if split or ultra_long is not None or chunksize:
            raise NotImplementedError("split contigs / ultra-long edges / chunked H are outside the accelerated path")
        self.edges = edges
        if isinstance(contigsizes, dict):
            contigsizes = pd.DataFrame({"length": pd.Series(contigsizes)})
        self.contigsizes = contigsizes
        self.k = k
        self.fasta = fasta
        self.prunetable = prunetable
        self.alleletable = alleletable
        self.normalize = normalize
        self.split = split
        self.allelic_factor = allelic_factor
        self.cross_allelic_factor = cross_allelic_factor
        self.allelic_similarity = allelic_similarity
        self.min_allelic_overlap = min_allelic_overlap
        self.disable_merge_in_first = disable_merge_in_first
        self.merge_use_allele = merge_use_allele
        self.whitelist = whitelist
        self.blacklist = blacklist
        self.min_contacts = min_contacts
        self.min_length = min_length
        self.resolution1 = resolution1
        self.resolution2 = resolution2
        self.init_resolution1 = init_resolution1
        self.init_resolution2 = init_resolution2
        self.min_weight = min_weight
        self.min_cis_weight = min_cis_weight
        self.cluster_method = cluster_method
        self.min_quality1 = min_quality1
        self.min_quality2 = min_quality2
        self.min_scaffold_length = int(min_scaffold_length)
        self.is_recluster_by_linkage = not disable_recluster_by_linkage
        self.threshold = threshold
        self.max_round = max_round
        self.threads = threads
        self.contig_sizes = self.contigsizes.to_dict()['length']
        self.contigs = self.contigsizes.index.values.tolist()
        self.K = []
        self.cluster_assignments = None
        if not mapq_filter_1_to_2:                                                  # :246-255
            self.HG = HyperGraph(self.edges, min_quality=self.min_quality1, mapq_filter=False)
        else:
            self.HG = HyperGraph(self.edges, min_quality=self.min_quality2,
                                 mapq_filter=self.min_quality2 > self.min_quality1)
        self.filter_hypergraph()
        self.H, self.vertices = self.get_hypergraph()
        self.NW = self.get_normalize_weight() if self.normalize else None
        if self.prunetable is not None:
            self.P_allelic_idx, self.P_weak_idx, self.prune_pair_df = self.get_prune_pairs()
        else:
            self.P_allelic_idx, self.P_weak_idx, self.prune_pair_df = None, None, None
        if self.P_allelic_idx:
            df = self.prune_pair_df
            self.allelic_idx_set = set(map(tuple, df[(df['type'] == 0) & (df['similarity'] >= self.allelic_similarity)]
                                           [['contig1', 'contig2']].values))

    # -- hyperpartition.py:284-302
    @property
    def vertices_idx(self):
        return dict(zip(self.vertices, range(len(self.vertices))))

    @property
    def idx_to_vertices(self):
        return dict(zip(range(len(self.vertices)), self.vertices))

    @property
    def vertices_idx_sizes(self):
        return {j: self.contig_sizes[i] for i, j in self.vertices_idx.items()}

    # -- hyperpartition.py:305-358
    def get_hypergraph(self):
        H = self.HG.incidence_matrix(self.min_contacts)
        self.cis_counts = None
        vertices = self.HG.nodes
        logger.info(f"Generated filtered hypergraph that containing {H.shape[0]:,} vertices"
                    f" and {H.shape[1]:,} hyperedges.")
        return H, vertices

    # -- hyperpartition.py:360-397
    def filter_hypergraph(self, remove_contigs=None):
        remove_contigs = set() if remove_contigs is None else set(remove_contigs)
        short = self.contigsizes[self.contigsizes < self.min_length].dropna().index.tolist()
        remove_contigs.update(short)
        if self.whitelist:
            remove_contigs.update(set(self.contigs) - set(self.whitelist))
        if self.blacklist:
            remove_contigs.update(set(self.blacklist))
        if len(remove_contigs) == 0:
            return
        self.HG.remove_rows(remove_contigs)
        logger.info(f"Total {len(remove_contigs):,} contigs were removed,")
        logger.info(f"\tbecause it's length too short (<{self.min_length}) or your specified in blacklist or not in whitelist.")

    # -- hyperpartition.py:496-570 (hard-wired method "svc"); returned as the VECTOR w (NW = outer(w, w) is never formed)
    def get_normalize_weight(self):
        from .algorithms.hypergraph import upload_incidence
        a = self.contigsizes.loc[self.vertices]['length'].astype(np.float64).values.copy()
        a[a <= 0] = 1.0
        dH = upload_incidence(self.H)
        rs = torch.empty(max(dH.n_rows, 1), dtype=torch.int64, device=dH.device)
        from ._lib import check, lib, stream
        check(lib().cph_row_sums(ptr(dH.pem_row), ptr(dH.pem_pk), None, dH.nnz, max(dH.n_rows, 1), 0, None, 1, ptr(rs),
                                 None, stream()))
        deg = rs[:dH.n_rows].cpu().numpy().astype(np.float64)                         # H.sum(axis=1)
        deg[deg <= 0] = 1.0
        global_density = np.sum(deg) / np.sum(a)
        return 1.0 / np.sqrt(deg + 0.5 * a * global_density)

    # -- hyperpartition.py:423-461; `prunetable` = path of a 7-column prune table (core.py:711-723) or a DataFrame
    def get_prune_pairs(self):
        cols = ['contig1', 'contig2', 'mz1', 'mz2', 'mzShared', 'similarity', 'type']
        if isinstance(self.prunetable, pd.DataFrame):
            pair_df = self.prunetable.copy()
        else:
            pair_df = pd.read_csv(self.prunetable, sep='\t', header=None, names=cols, comment="#")
        vs = set(self.vertices)
        pair_df = pair_df[pair_df['contig1'].isin(vs) & pair_df['contig2'].isin(vs)]
        if pair_df.empty:
            logger.warning("No valid prune pairs found in the provided prunetable.")
            return None, None, None
        vi = self.vertices_idx
        pair_df = pair_df.assign(contig1=pair_df['contig1'].map(vi.get), contig2=pair_df['contig2'].map(vi.get))
        pair_df = pair_df.dropna(axis=0).astype({'contig1': 'int', 'contig2': 'int', 'type': 'int'})
        sym = pair_df.rename(columns={'contig1': 'contig2', 'contig2': 'contig1', 'mz1': 'mz2', 'mz2': 'mz1'})
        pair_df = pd.concat([pair_df, sym[pair_df.columns]], axis=0)                    # PruneTable.symmetric_table
        pair_df = pair_df.drop_duplicates(subset=['contig1', 'contig2']).reset_index(drop=True)
        Pa = [pair_df.loc[pair_df['type'] == 0, 'contig1'], pair_df.loc[pair_df['type'] == 0, 'contig2']]
        Pw = [pair_df.loc[pair_df['type'] == 1, 'contig1'], pair_df.loc[pair_df['type'] == 1, 'contig2']]
        return Pa, Pw, pair_df

    # -- hyperpartition.py:399-421 (the in-tree Python kprune; the production `cphasing-rs kprune` binary is not in the
    #    reference tree).  `alleletable`: object with `.data` (core.AlleleTable) or a sequence of (name1, name2).
    def get_prune_pairs_kprune(self):
        A = HyperGraph.clique_expansion_init(self.H, NW=self.get_normalize_weight() if self.normalize else None)
        kph = KPruneHyperGraph(self.alleletable, A, self.vertices_idx)
        kph.run()
        kph.prunetable = kph.get_prune_table()
        al = pd.DataFrame(kph.allelic_prune_list, columns=['contig1', 'contig2'])
        wk = pd.DataFrame(kph.weak_prune_list, columns=['contig1', 'contig2'])
        return [al['contig1'], al['contig2']], [wk['contig1'], wk['contig2']], kph.prunetable

    def kprune(self, alleletable=None, **kw):
        """hyperpartition.py:1563-1595 shells out to `cphasing-rs kprune`; here the table is produced in-process."""
        if alleletable is not None:
            self.alleletable = alleletable
        _, _, table = self.get_prune_pairs_kprune()
        self.prunetable = table
        self.P_allelic_idx, self.P_weak_idx, self.prune_pair_df = self.get_prune_pairs()
        return table

    # -- hyperpartition.py:4012-4044
    def filter_cluster(self, verbose=1):
        size_dict = self.vertices_idx_sizes
        _K = [g for g in self.K if sum(size_dict.get(x, 0) for x in g) >= self.min_scaffold_length]
        if verbose == 1:
            logger.info(f"Removed groups less than {self.min_scaffold_length} in length. (--min-scaffold-length)")
        logger.info(f"After filtering, {len(_K)} groups remain.")
        return _K

    def _irmm(self, A, resolution):
        return IRMM(self.H, A, self.NW, self.P_allelic_idx, self.P_weak_idx, self.allelic_factor, self.cross_allelic_factor,
                    resolution, self.min_weight, self.threshold, self.max_round, threads=self.threads,
                    method=self.cluster_method)

    # -- hyperpartition.py:911-1237 (basal / basal_withprune mode)
    def single_partition(self, k=None, sort_by_length=True, sort_group=False, is_kprune=False):
        logger.info("Start hyperpartition ...")
        if sort_group:
            raise NotImplementedError("raw_sort (scaffolding order) is outside the accelerated path")
        if is_kprune and self.alleletable is not None and self.prunetable is None:
            self.kprune(self.alleletable)
        A = self.HG.clique_expansion_init(self.H, allelic_factor=self.allelic_factor, min_weight=self.min_weight)
        # NB the reference passes the idx-keyed size dict with name-valued `nodes` (:989): no name maps, every size is
        # NaN and the filter keeps everything.  Reproduced as is, so groups stay identical to the reference's.
        retain_idx1 = np.asarray(HyperGraph.filter_adjacency_matrix(A, self.vertices, self.vertices_idx_sizes,
                                                                    invert=True)).astype(np.int64)
        dia = A.diagonal()
        raw_contig_counts = A.shape[0]
        retain_idx = np.intersect1d(np.where(dia >= self.min_cis_weight)[0], retain_idx1)
        if len(retain_idx) < raw_contig_counts:                                           # :1005-1042
            idx_to_vertices = self.idx_to_vertices
            self.H, _, _ = extract_incidence_matrix2(self.H, retain_idx)
            self.vertices = self.vertices[retain_idx]
            if self.NW is not None:
                self.NW = self.NW[retain_idx]
            if self.prune_pair_df is not None:
                df = self.prune_pair_df
                keep = set(retain_idx.tolist())
                df = df[df['contig1'].isin(keep) & df['contig2'].isin(keep)].copy()
                vi = self.vertices_idx
                df['contig1'] = df['contig1'].map(idx_to_vertices.get).map(vi.get)
                df['contig2'] = df['contig2'].map(idx_to_vertices.get).map(vi.get)
                df = df.dropna(axis=0).reset_index(drop=True).astype({'contig1': 'int', 'contig2': 'int'})
                self.prune_pair_df = df
                self.P_allelic_idx = [df.loc[df['type'] == 0, 'contig1'], df.loc[df['type'] == 0, 'contig2']]
                self.P_weak_idx = [df.loc[df['type'] == 1, 'contig1'], df.loc[df['type'] == 1, 'contig2']]
            logger.info(f"Removed {raw_contig_counts - len(retain_idx):,} contigs that self edge weight < "
                        f"{self.min_cis_weight} (--min-cis-weight).")
            A = self.HG.clique_expansion_init(self.H, allelic_factor=self.allelic_factor, min_weight=self.min_weight)
        A = HyperGraph.normalize_adjacency_matrix(A, self.NW) if self.NW is not None else A
        self.H = None                                                                      # :1047
        self.resolutions_tried = []
        if self.resolution1 == -1:                                                         # :1049-1104
            result_K_length = len(self.K)
            tmp_resolution = self.init_resolution1
            max_round, round_count = 100, 0
            if not k:
                logger.warning("To automatic search best resolution, the `-n` must be specified.")
            if result_K_length < k:
                while result_K_length < k:
                    if round_count >= max_round:
                        logger.warning("Maximum rounds reached during automatic resolution search.")
                        break
                    raw_A, _A, self.cluster_assignments, self.K = self._irmm(A, tmp_resolution)
                    self.K = list(map(list, self.K))
                    self.K = self.filter_cluster(verbose=0)
                    result_K_length = len(self.K)
                    self.resolutions_tried.append((tmp_resolution, result_K_length))
                    logger.info(f"Generated `{result_K_length}` groups at resolution `{tmp_resolution:.2f}`.")
                    tmp_resolution += 0.2
                    round_count += 1
            if result_K_length > k and not self.merge_use_allele and not self.disable_merge_in_first:
                while result_K_length > k:
                    if tmp_resolution <= 0 or round_count >= max_round:
                        logger.warning("Minimum resolution reached during automatic resolution search.")
                        break
                    raw_A, _A, self.cluster_assignments, self.K = self._irmm(A, tmp_resolution)
                    self.K = list(map(list, self.K))
                    self.K = self.filter_cluster(verbose=0)
                    result_K_length = len(self.K)
                    self.resolutions_tried.append((tmp_resolution, result_K_length))
                    logger.info(f"Generated `{result_K_length}` groups at resolution `{tmp_resolution:.2f}`.")
                    tmp_resolution -= 0.1
                    round_count += 1
        else:
            raw_A, _A, self.cluster_assignments, self.K = self._irmm(A, self.resolution1)
        A = _A
        self.K = list(map(list, self.K))
        self.K = self.filter_cluster()
        vertices_idx_sizes = pd.DataFrame(self.vertices_idx_sizes, index=['length']).T
        if k and len(self.K) > k:                                                          # :1148-1178
            logger.info(f"Merging {len(self.K)} groups into {k} groups ...")
            self.K = HyperPartition._merge(A, self.K, vertices_idx_sizes, k, allelic_idx_set=set(),
                                           min_allelic_overlap=self.min_allelic_overlap, allele_use_method="conflict",
                                           method='mean')
        reclustered_count = 1 if self.is_recluster_by_linkage else 0
        iter_count = 0
        while reclustered_count > 0 and iter_count < 5:                                    # :1188-1192
            iter_count += 1
            reclustered_count = self.recluster_by_linkage(A=raw_A, min_ratio=1.5, iter_count=iter_count)
        if sort_by_length:
            sizes = self.vertices_idx_sizes
            self.K = sorted(self.K, key=lambda x: sum(sizes[i] for i in x), reverse=True)
        logger.info(f"Hyperpartition result {len(self.K)} groups:")
        return self.K

    # -- hyperpartition.py:2288-2471: greedy merge of groups down to k.  V = M^T A M comes from the device ONCE
    #    (cph_group_sums); merging two groups adds their rows and columns, which is what recomputing M^T A M gives.
    @staticmethod
    def _merge(A, K, vertices_idx_sizes, k=None, allelic_idx_set=set(), alleletable_df=None, min_allelic_overlap=0.1,
               allele_use_method="conflict", method='mean'):
        if not k:
            return K
        if alleletable_df is not None:
            raise NotImplementedError("_merge with minimizer-weighted allele overlaps (alleletable_df) is not implemented")
        if hasattr(vertices_idx_sizes, 'to_dict'):
            size_db = vertices_idx_sizes.to_dict()
            if 'length' in size_db:
                size_db = size_db['length']
        else:
            size_db = dict(vertices_idx_sizes)
        K = [list(g) for g in K]
        if len(K) <= k:
            return K
        _, V = group_sums(A, K, want_V=True)
        V = np.array(V, dtype=np.float64)
        iter_round = 0
        while k and len(K) > k:
            g = len(K)
            if iter_round > (g - k + 100):
                break
            group_lengths = np.array([sum(size_db.get(c, 0.0) for c in grp) for grp in K], dtype=np.float64)
            flag = np.ones((g, g))
            overlap = np.zeros((g, g))
            if allelic_idx_set:                                                           # :2349-2392
                c2g = {c: i for i, grp in enumerate(K) for c in grp}
                s1, s2 = defaultdict(set), defaultdict(set)
                for c1, c2 in allelic_idx_set:
                    g1, g2 = c2g.get(c1), c2g.get(c2)
                    if g1 is not None and g2 is not None and g1 != g2:
                        if g1 > g2:
                            g1, g2, c1, c2 = g2, g1, c2, c1
                        s1[(g1, g2)].add(c1)
                        s2[(g1, g2)].add(c2)
                for (i, j), a in s1.items():
                    l1 = float(sum(size_db.get(c, 0.0) for c in a))
                    l2 = float(sum(size_db.get(c, 0.0) for c in s2.get((i, j), set())))
                    o1 = l1 / group_lengths[i] if group_lengths[i] > 0 else 0
                    o2 = l2 / group_lengths[j] if group_lengths[j] > 0 else 0
                    if o1 > min_allelic_overlap or o2 > min_allelic_overlap:
                        flag[i, j] = 0
                        overlap[i, j] = max(o1, o2)
            value = V.copy()
            np.fill_diagonal(value, 0)
            if method == "mean":
                sizes = np.array([len(grp) for grp in K], dtype=np.float32)
                so = np.outer(sizes, sizes)
                so[so == 0] = 1.0
                value = value / so
            total = max(np.triu(value, 1).sum(), 1e-9)
            rs = value.sum(axis=1)
            Q = value - np.outer(rs, rs) / total
            if allele_use_method == "conflict":
                Q = np.where(flag > 0, Q, -2.0 ** 64)
            else:
                Q = np.where(flag > 0, Q, np.maximum(overlap * Q, 0))
            Q[np.tril_indices(g)] = -np.inf                                               # pairs i < j only
            i1, i2 = np.unravel_index(int(np.argmax(Q)), Q.shape)                        # first maximum, row-major
            K[i1].extend(K[i2])
            K.pop(i2)
            V[i1, :] += V[i2, :]
            V[:, i1] += V[:, i2]
            V = np.delete(np.delete(V, i2, axis=0), i2, axis=1)
            iter_round += 1
        return K

    # -- hyperpartition.py:2860-2968: S = A M on the device, decisions on the host as in the reference
    def recluster_by_linkage(self, A=None, min_ratio=1.2, iter_count=1):
        if A is None:
            A = HyperGraph.clique_expansion_init(self.H, min_weight=self.min_weight, use_high_order=False)
        n = A.shape[0]
        if len(self.K) == 0:
            return 0
        contig_to_group = np.full(n, -1, dtype=int)
        for gid, group in enumerate(self.K):
            contig_to_group[list(group)] = gid
        allelic_lookup = defaultdict(set)
        if getattr(self, 'allelic_idx_set', None):
            for c1, c2 in self.allelic_idx_set:
                allelic_lookup[c1].add(c2)
                allelic_lookup[c2].add(c1)
        itv = self.idx_to_vertices
        contig_lens = np.array([self.contig_sizes.get(itv[i], 1) for i in range(n)], dtype=np.float32)
        S, _ = group_sums(A, self.K, want_S=True)
        best_gids = np.argmax(S, axis=1)
        best_S = S[np.arange(n), best_gids]
        curr_S = np.zeros(n)
        valid = contig_to_group != -1
        curr_S[valid] = S[np.where(valid)[0], contig_to_group[valid]]
        total = np.sum(S, axis=1)
        spec = np.divide(best_S, total, out=np.zeros(n), where=total > 0)
        mask = (best_gids != contig_to_group) & (best_S > curr_S * min_ratio) & (spec > 0.9) & (total > self.min_weight)
        candidates = np.where(mask)[0]
        if len(candidates) == 0:
            return 0
        candidates = sorted(candidates, key=lambda x: contig_lens[x], reverse=True)
        count = 0
        new_K = [set(g) for g in self.K]
        for c in candidates:
            bg, cg = best_gids[c], contig_to_group[c]
            if S[c, bg] / (total[c] + 1e-9) < 0.25:
                continue
            if allelic_lookup.get(c, set()).isdisjoint(new_K[bg]):
                if cg != -1:
                    new_K[cg].discard(c)
                new_K[bg].add(c)
                contig_to_group[c] = bg
                count += 1
        if count > 0:
            logger.debug(f"Round [{iter_count}]: reclustered {count} contigs based on linkage strength.")
            self.K = [sorted(int(x) for x in g) for g in new_K]
        return count

    # -- hyperpartition.py:4238-4296 (first-round / basal output: `ChrNN\t<count>\t<contigs>`)
    def to_cluster(self, output, merge=True):
        itv = self.idx_to_vertices
        clusters = [[itv[x] for x in grp] for grp in self.K]
        clusters = [c for c in clusters if c]
        with open(output, 'w') as out:
            for i, group in enumerate(clusters, 1):
                print(f'Chr{i:0>2}\t{len(group)}\t{" ".join(group)}', file=out)
        logger.info(f"Successful output hyperpartition results in `{output}`.")
