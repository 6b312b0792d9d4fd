"""Import the UNMODIFIED reference package for test / bench use.

The reference (wangyibin/CPhasing 0.3.2) is pure Python; it was installed here with
    python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
        --target baseline/_ref /tmp/<copy of /root/reference>
(`--no-deps`: biopython, polars, python-igraph, cooler, ... are not in the offline wheelhouse; the install needs a
writable source tree, hence the copy).  baseline/_ref is git-ignored and travels to the GPU box with the snapshot.

Its modules import polars / igraph / cooler / ... at import time; none of them can be installed, so they are replaced
by inert stubs (an attribute of a stub is a stub; calling one returns it).  Only reference code that never touches
those modules can be *executed*: HyperEdges, HyperGraph.get_data / remove_rows / incidence_matrix /
clique_expansion_init / normalize / filter, extract_incidence_matrix2, find_csr_data_offsets and the control flow of
HyperPartition (single_partition, incremental_partition, _merge, recluster_by_linkage, filter_cluster) -- SURVEY.md 8c.
This file is test/bench infrastructure, never imported by the product (cphasing_b200/).
"""
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIRS = (os.path.join(HERE, "_ref"), "/root/reference")

STUBBED = ["igraph", "polars", "cooler", "pytools", "natsort", "pyranges", "pandarallel", "Bio", "Bio.SeqIO", "Bio.Seq",
           "rich_click", "pyfaidx", "pysam", "dask", "dask.array", "hicmatrix", "portion", "line_profiler", "seaborn",
           "colormaps", "intervaltree", "ncls", "pypdf", "pyfastx"]


class _Stub(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Stub(self.__name__ + "." + name)

    def __call__(self, *a, **k):
        return self


def available():
    return any(os.path.isdir(os.path.join(d, "cphasing")) for d in REF_DIRS)


def where():
    for d in REF_DIRS:
        if os.path.isdir(os.path.join(d, "cphasing")):
            return d
    return None


def load():
    """-> the reference's ``cphasing.algorithms.hypergraph`` module (importing it makes ``cphasing.*`` importable)."""
    d = where()
    if d is None:
        raise ImportError("the reference is neither under baseline/_ref nor /root/reference")
    for m in STUBBED:
        if m in sys.modules:
            continue
        try:
            __import__(m)
        except Exception:
            sys.modules[m] = _Stub(m)
    if d not in sys.path:
        sys.path.insert(0, d)
    import cphasing.algorithms.hypergraph as ah
    return ah


def stubbed_modules():
    return sorted(m for m in STUBBED if isinstance(sys.modules.get(m), _Stub))
