"""TEST INFRASTRUCTURE ONLY -- a SECOND, independent restatement of the reference's table -> triplets step, in pandas.

Why it exists: the reference does this step with polars (hypergraph.py:1395-1481 for Pore-C tables, :960-1044 for pairs)
and polars cannot be installed here, so `hypergraph_oracle.porec_edges` / `pairs_edges` (numpy: argsort + reduceat) could
not be pinned against a run of the reference.  This module restates the same lines a different way -- expression by
expression on the pandas operations that mirror the polars ones (`join(how="inner")` -> `merge(how="inner")`,
`n_unique().over("read_idx")` -> `groupby("read_idx").transform("nunique")`, `group_by([...]).agg(count, max)` ->
`groupby([...]).agg(...)`, `with_row_count("col")` -> running number) -- on NAMED contigs, as the reference sees them,
not on the integer codes the numpy restatement takes.  tests/test_oracle_groupby_pandas.py asserts the two agree on
generated tables (seeded and hypothesis-made); it does NOT pin either against polars itself (parity for this step stays
"unpinned"), it removes the single-restatement risk the way networkx does for the clustering oracle.

Nothing under cphasing_b200/ imports this file.
"""
import numpy as np
import pandas as pd


def porec_triplets(df, contig_idx, contigsizes=None, min_quality=1, min_order=2, max_order=50, edge_length=0):
    """df: read_idx, chrom (names), start, end, mapping_quality.  contig_idx: dict name -> row index; contigsizes: dict
    name -> length.  -> DataFrame read_idx, chrom_idx, mapping_quality, chrom_idx_count (hypergraph.py:1479), unordered."""
    df = df[df["mapping_quality"] >= min_quality] if min_quality > 0 else df                 # :1774-1775
    mapping_df = pd.DataFrame({"chrom": list(contig_idx.keys()), "chrom_idx": list(contig_idx.values())})
    df = df.merge(mapping_df, on="chrom", how="inner")                                        # :1433-1436
    if edge_length:                                                                           # :1438-1457
        mapping_len = pd.DataFrame({"chrom": list(contigsizes.keys()), "length": list(contigsizes.values())})
        df = df.merge(mapping_len, on="chrom", how="inner")
        pos = df["start"].astype(np.int64) + (df["end"].astype(np.int64) - df["start"].astype(np.int64)) // 2
        df = df[(pos < edge_length) | ((df["length"].astype(np.int64) - pos) < edge_length)]
    df = df[["read_idx", "chrom_idx", "mapping_quality"]]                                     # :1459
    nunique = df.groupby("read_idx")["chrom_idx"].transform("nunique")                        # :1461-1463
    df = df[(nunique >= min_order) & (nunique <= max_order)]                                  # :1465-1468
    out = (df.groupby(["read_idx", "chrom_idx"], sort=False)                                  # :1470-1477
             .agg(chrom_idx_count=("chrom_idx", "count"), mapping_quality=("mapping_quality", "max"))
             .reset_index())
    return out[["read_idx", "chrom_idx", "mapping_quality", "chrom_idx_count"]]               # :1479


def porec_canonical(t):
    """the triplets as (row i32, col i32, count u32, mapq i8) sorted by (row = contig, col = read): the form
    hypergraph_oracle.porec_edges returns (hypergraph.py:1877-1889: sort by contig, casts)"""
    t = t.sort_values(["chrom_idx", "read_idx"], kind="stable")
    return (t["chrom_idx"].to_numpy().astype(np.int32), t["read_idx"].to_numpy().astype(np.int32),
            t["chrom_idx_count"].to_numpy().astype(np.uint32), t["mapping_quality"].to_numpy().astype(np.int8))


def pairs_triplets(p, contig_idx, contigsizes=None, min_mapq=1, edge_length=0):
    """p: chrom1, pos1, chrom2, pos2, mapq (names in chrom1/chrom2).  -> DataFrame row, col, mapq sorted by row
    (hypergraph.py:968-1044); count is one per entry (:1287-1292)."""
    if min_mapq > 0:
        p = p[p["mapq"] >= min_mapq]                                                          # :969-970
    if edge_length:                                                                           # :976-992
        ml = pd.DataFrame({"chrom": list(contigsizes.keys()), "length": list(contigsizes.values())})
        p = (p.merge(ml.rename(columns={"chrom": "chrom1", "length": "length1"}), on="chrom1", how="inner")
              .merge(ml.rename(columns={"chrom": "chrom2", "length": "length2"}), on="chrom2", how="inner"))
        p = p[((p["pos1"] < edge_length) | (p["pos1"] > (p["length1"] - edge_length)))
              & ((p["pos2"] < edge_length) | (p["pos2"] > (p["length2"] - edge_length)))]
        p = p[["chrom1", "pos1", "chrom2", "pos2", "mapq"]]
    idx = pd.DataFrame({"chrom": list(contig_idx.keys()), "idx": list(contig_idx.values())})   # _process_df_pl :773-784
    p = (p.merge(idx.rename(columns={"chrom": "chrom1"}), on="chrom1", how="inner").drop(columns="chrom1")
          .rename(columns={"idx": "chrom1"})
          .merge(idx.rename(columns={"chrom": "chrom2"}), on="chrom2", how="inner").drop(columns="chrom2")
          .rename(columns={"idx": "chrom2"}))
    p = p.reset_index(drop=True)
    p["col"] = np.arange(len(p))                                                              # with_row_count("col") :1031
    df1 = p[["chrom1", "col", "mapq"]].rename(columns={"chrom1": "row"})                      # :1033-1042
    df2 = p[["chrom2", "col", "mapq"]].rename(columns={"chrom2": "row"})
    return pd.concat([df1, df2]).sort_values("row", kind="stable")                            # :1044


def pairs_canonical(t):
    t = t.sort_values(["row", "col"], kind="stable")
    return (t["row"].to_numpy().astype(np.int32), t["col"].to_numpy().astype(np.int32),
            np.ones(len(t), np.uint32), t["mapq"].to_numpy().astype(np.int8))
