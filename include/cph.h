/* libcph -- C-ABI of the B200-native hyperpartition hot path (sm_100a).
 *
 * The reference (wangyibin/CPhasing, pure Python) has no FFI: the seam is its Python API
 * (SURVEY.md 8b).  Every entry point below names the reference code it replaces
 * (paths under /root/reference/cphasing/).  Conventions:
 *   - plain pointers and sizes only; every non-`h_` pointer is a DEVICE pointer owned by the caller
 *     (a torch tensor on the Python side); `h_` pointers are host memory;
 *   - every function returns 0 (CPH_OK) or a positive CPH_ERR_* code, never throws;
 *     cph_last_error() gives the message of the calling thread's last failure;
 *   - no function allocates device memory: temporaries are carved from `ws` (ws_bytes big);
 *     calling with ws == NULL returns the required size in *h_ws_need and does nothing else;
 *   - functions that report data-dependent sizes write them to a host array and synchronise
 *     `stream` before returning; all other functions are asynchronous on `stream`.
 *
 * Device formats
 *   RAW EDGE-MAJOR (rem): the hyperedge triplets of cphasing.algorithms.hypergraph.HyperEdges
 *     (algorithms/hypergraph.py:100-123) ordered by (col, arrival), i.e. the CSC view of H:
 *       rem_row[t]   uint32  contig index | (1u<<31 if t is the first entry of its hyperedge)
 *       rem_col[t]   uint32  hyperedge (read / pair) id
 *       rem_count[t] uint32  alignment count      rem_mapq[t] int8  max MAPQ
 *   PACKED EDGE-MAJOR (pem): two uint32 per entry after the row/column filters:
 *       pem_row = new row index | head bit   pem_pk = (D_e << 24) | count,  count < 2^24,
 *       D_e = max(int8(k_e) - 1, 1), the reference's hyperedge degree term (algorithms/hypergraph.py:385,418)
 *   CSR of H (contig-major): indptr int64[n+1], indices int32 (renumbered hyperedge id), pk uint32 (same .y)
 *   ACC: packed upper triangle (i<=j) of A as unsigned 64-bit fixed point with `frac_bits` fraction bits:
 *       acc[i*(2n-i+1)/2 + (j-i)]
 */
#ifndef CPH_H
#define CPH_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CPH_OK 0
#define CPH_ERR_CUDA 1
#define CPH_ERR_ARG 2
#define CPH_ERR_WORKSPACE 3
#define CPH_ERR_TOO_LARGE 4
#define CPH_ERR_NOT_CONVERGED 5

#define CPH_ROW_MASK 0x7fffffffu
#define CPH_HEAD_BIT 0x80000000u
#define CPH_COUNT_MASK 0x00ffffffu

typedef void* cph_stream_t; /* cudaStream_t */

int cph_version(void);
const char* cph_last_error(void);
/* number of kernel launches issued by this library on the calling process so far */
uint64_t cph_launch_count(void);

/* ---- stage 1a: Hi-C pairs -> raw edge-major -------------------------------------------------
 * Replaces Extractor.generate_edges (hypergraph.py:895-1300): mapq filter (:970-971), both-ends
 * edge filter (:977-992), col = running number of the surviving pairs (:1031), two triplets per
 * pair with count 1 (:1033-1044, 1287-1292).  Outputs need capacity 2*n_pairs.
 * h_counts[0] = nnz (2 * surviving pairs), h_counts[1] = number of columns.  rem_count = (1 << 24) | 1 for every entry: count 1
 * with D_e = 1 of the two-entry hyperedge in the top byte (the packed form, see cph_porec_to_edges). */
int cph_pairs_to_edges(const uint32_t* chrom1, const uint32_t* chrom2, const int8_t* mapq,
                       const uint32_t* pos1, const uint32_t* pos2, const int64_t* contig_len,
                       uint64_t n_pairs, uint32_t n_contigs, int min_mapq, uint64_t edge_length,
                       uint32_t* rem_row, uint32_t* rem_col, uint32_t* rem_count, int8_t* rem_mapq,
                       uint64_t* h_counts, void* ws, size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);

/* ---- stage 1b: Pore-C alignment records -> raw edge-major -----------------------------------
 * Replaces process_pore_c_table (hypergraph.py:1395-1481) + the mapq filter of
 * _load_and_filter_single (:1774-1775): edge filter (:1438-1457), distinct-contig order filter
 * (:1461-1468), group_by(read, contig) -> count, max(mapq) (:1470-1477); read ids are shifted by
 * `read_offset` (generate_edges :1871-1875).  start/end/contig_len may be NULL when edge_length==0.
 * `sorted_by_read` != 0 promises read_idx is non-decreasing (skips the global sort).
 * Outputs need capacity n_records.  h_counts has SIX entries: [0] = nnz, [1] = max col + 1 (0 if empty), [4] = number of
 * hyperedges emitted (0 when the general sort path ran), [5] = 1 when rem_count carries D_e of the entry's hyperedge in
 * its top byte (D_e << 24 | count, what cph_expand_* read as `pem_pk`): then rem_row / rem_count are the packed incidence
 * of the table itself, and consumers that want plain counts pass `packed = 1` (cph_row_sums, cph_incidence_build) or mask
 * with 0x00ffffff.  [2], [3]: see cph_porec_to_edges_host. */
int cph_porec_to_edges(const uint32_t* read_idx, const uint32_t* contig_idx, const int8_t* mapq,
                       const uint32_t* start, const uint32_t* end, const int64_t* contig_len,
                       uint64_t n_records, uint32_t n_contigs, int min_mapq, uint32_t min_order,
                       uint32_t max_order, uint64_t edge_length, uint32_t read_offset, int sorted_by_read,
                       uint32_t* rem_row, uint32_t* rem_col, uint32_t* rem_count, int8_t* rem_mapq,
                       uint64_t* h_counts, void* ws, size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);

/* ---- extractor pre-filters (all device pointers; NULL / 0 = off) ------------------------------
 * HCR BED: keep a Pore-C record iff its alignment midpoint start + (end - start) / 2 lies inside the LAST BED interval
 * of the same contig whose start is <= the midpoint (polars join_asof(strategy="backward") + pos < end_bed,
 * hypergraph.py:1721-1752; hcr_invert keeps the others); a pair iff EITHER end does (:786-893).
 * split: the contig column(s) carry codes of the ORIGINAL contig names (0 <= code < n_names), the hyperedge row is
 * split_init[code] + position / split_length with position = the midpoint (Pore-C, :1406-1424) or pos1 / pos2 (pairs,
 * :994-1023); n_contigs is then the number of split bins and no edge filter applies (the reference's else-branch).
 * With either filter on, bed_ptr / split_init index the same code space of n_names names. */
typedef struct cph_prefilter {
    const uint64_t* bed_ptr;    /* [n_names + 1] offsets into bed_start / bed_end, intervals sorted by (name, start) */
    const uint32_t* bed_start;
    const uint32_t* bed_end;
    int hcr_invert;
    uint64_t split_length;      /* 0 = no splitting */
    const uint32_t* split_init; /* [n_names] first split bin of each original contig; 0xffffffff = not in the map */
    uint32_t n_names;
} cph_prefilter;

/* cph_pairs_to_edges / cph_porec_to_edges with the pre-filters above (`prefilter` is a HOST pointer, may be NULL).
 * pos1/pos2 resp. start/end must be given when a filter is on. */
int cph_pairs_to_edges_pf(const uint32_t* chrom1, const uint32_t* chrom2, const int8_t* mapq,
                          const uint32_t* pos1, const uint32_t* pos2, const int64_t* contig_len,
                          uint64_t n_pairs, uint32_t n_contigs, int min_mapq, uint64_t edge_length,
                          const cph_prefilter* prefilter,
                          uint32_t* rem_row, uint32_t* rem_col, uint32_t* rem_count, int8_t* rem_mapq,
                          uint64_t* h_counts, void* ws, size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);
int cph_porec_to_edges_pf(const uint32_t* read_idx, const uint32_t* contig_idx, const int8_t* mapq,
                          const uint32_t* start, const uint32_t* end, const int64_t* contig_len,
                          uint64_t n_records, uint32_t n_contigs, int min_mapq, uint32_t min_order,
                          uint32_t max_order, uint64_t edge_length, uint32_t read_offset, int sorted_by_read,
                          const cph_prefilter* prefilter,
                          uint32_t* rem_row, uint32_t* rem_col, uint32_t* rem_count, int8_t* rem_mapq,
                          uint64_t* h_counts, void* ws, size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);

/* Host-table form of cph_porec_to_edges for tables grouped by read (what _load_and_filter_single,
 * hypergraph.py:1700-1789, hands to process_pore_c_table): h_* are HOST columns, d_* caller-owned device
 * staging columns of n_records elements.  The columns are copied in chunks of `chunk_records` (0 = default)
 * on a private stream and the group-by kernel runs on the tiles of every chunk that has landed, i.e. under
 * the transfer of the next one.  pack_threads = 0: chunks are copied as they are (9 B per record; pinned
 * memory gives the full PCIe rate).  pack_threads > 0: that many host threads first re-encode every chunk
 * into a pinned staging ring owned by the library as (read-id delta u8 | mapq | contig u16 [u32 above 65536
 * contigs] | one u32 anchor per 1024 records) = 4 B per record, and a kernel restores the device columns
 * bit for bit; a chunk whose read ids step by more than 255 is copied as it is; the h_* arrays may be
 * pageable.  pack_threads < 0: two fewer than the host has cores, at most 16.  No edge filter
 * (edge_length == 0 only).  h_counts has SIX entries ([4], [5] as above): [0] = nnz, [1] = max col + 1, [3] = bytes copied to the
 * device, [2] = 1 if the table is
 * NOT grouped by read: the d_* columns are complete, the outputs are not -- call cph_porec_to_edges on them. */
int cph_porec_to_edges_host(const uint32_t* h_read_idx, const uint32_t* h_contig_idx, const int8_t* h_mapq,
                            uint32_t* d_read_idx, uint32_t* d_contig_idx, int8_t* d_mapq,
                            uint64_t n_records, uint32_t n_contigs, int min_mapq, uint32_t min_order,
                            uint32_t max_order, uint32_t read_offset, uint64_t chunk_records, int pack_threads,
                            uint32_t* rem_row, uint32_t* rem_col, uint32_t* rem_count, int8_t* rem_mapq,
                            uint64_t* h_counts, void* ws, size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);

/* frees the pinned staging ring that the packed form of cph_porec_to_edges_host keeps between calls */
int cph_release_host_staging(void);

/* ---- HyperEdges wire order <-> raw edge-major --------------------------------------------------
 * cph_edges_to_triplets: the `sort by chrom_idx` + casts of generate_edges (hypergraph.py:1877-1889,
 * 1044): (row, col, count, mapq) sorted by (row, col).
 * cph_triplets_to_edges: the inverse, for HyperEdges that arrive from the host (.hg files,
 * algorithms/hypergraph.py:66-79) or from scipy CSR matrices; input order is free. */
int cph_edges_to_triplets(const uint32_t* rem_row, const uint32_t* rem_col, const uint32_t* rem_count,
                          const int8_t* rem_mapq, uint64_t nnz, uint32_t n_contigs,
                          int32_t* row, int32_t* col, uint32_t* count, int8_t* mapq,
                          void* ws, size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);
int cph_triplets_to_edges(const int32_t* row, const int32_t* col, const uint32_t* count, const int8_t* mapq,
                          uint64_t nnz, uint32_t n_cols,
                          uint32_t* rem_row, uint32_t* rem_col, uint32_t* rem_count, int8_t* rem_mapq,
                          void* ws, size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);

/* ---- stage 1c: filters + incidence matrix ------------------------------------------------------
 * cph_row_sums: per-contig sum of counts over the entries that pass the HyperGraph.get_data mapq
 * filter (algorithms/hypergraph.py:175-184; disabled when rem_mapq == NULL) and whose contig is not
 * flagged in row_drop (remove_rows :207-223; may be NULL): the `matrix.sum(axis=1)` of :287.
 * rowsum (uint64[n_contigs]) is overwritten.  `packed` != 0: counts carry D_e in the top byte.
 * max_count (device uint32, may be NULL) receives the largest count seen; with the row sums it bounds every
 * A_ij (<= max_count * min(rowsum_i, rowsum_j)), which is how the caller picks frac_bits without a CSR pass. */
int cph_row_sums(const uint32_t* rem_row, const uint32_t* rem_count, const int8_t* rem_mapq, uint64_t nnz,
                 uint32_t n_contigs, int min_mapq, const uint8_t* row_drop, int packed,
                 uint64_t* rowsum, uint32_t* max_count, cph_stream_t stream);

/* cph_incidence_build: HyperGraph.incidence_matrix (algorithms/hypergraph.py:249-314) and
 * extract_incidence_matrix2 (:938-954) on the edge-major form.  row_map[c] = new row of contig c or
 * -1 to drop it; entries with mapq < min_mapq are dropped when rem_mapq != NULL; a hyperedge
 * survives when the summed count of its surviving entries >= col_min_sum (2 in the reference; 0
 * keeps every hyperedge AND the original column ids, the `min_contacts=0` branch :309-311).
 * With packed != 0 the input is itself a pem (rem_row = pem_row, rem_count = pem_pk, rem_col may be
 * NULL when col_min_sum > 0).  Outputs (capacity nnz): pem_row, pem_pk; CSR indptr[n_rows_out+1], indices, pk.
 * indptr == NULL skips the CSR view (the clique expansion only needs the pem; cph_incidence_csr builds it later).
 * h_counts[0] = nnz_out, h_counts[1] = number of columns of H. */
int cph_incidence_build(const uint32_t* rem_row, const uint32_t* rem_col, const uint32_t* rem_count,
                        const int8_t* rem_mapq, uint64_t nnz, uint32_t n_contigs, int min_mapq,
                        const int32_t* row_map, uint32_t n_rows_out, uint32_t col_min_sum, int packed,
                        uint32_t* pem_row, uint32_t* pem_pk, int64_t* indptr, int32_t* indices, uint32_t* pk,
                        uint64_t* h_counts, void* ws, size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);

/* CSR view (sorted indices) of a pem, built on demand: what scipy hands back from incidence_matrix (:282-306).
 * pem_col: explicit hyperedge id per entry, or NULL = running number of the hyperedge heads. */
int cph_incidence_csr(const uint32_t* pem_row, const uint32_t* pem_pk, const uint32_t* pem_col, uint64_t nnz,
                      uint32_t n_rows, int64_t* indptr, int32_t* indices, uint32_t* pk,
                      void* ws, size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);

/* cph_incidence_csr_fixed: cph_incidence_csr for callers that cannot supply a workspace.  It has no counterpart in the
 * reference (whose scipy calls allocate freely); it exists for two kinds of user of THIS library:
 *   - code that runs as a stage inside another program (an operator callback of a workflow engine, a plug-in) and is
 *     given input and output buffers only, on a GPU whose remaining memory the host program's own pool has claimed;
 *   - long-running services that phase several assemblies on one GPU and must keep cudaMalloc / cudaFree, which stall
 *     every stream of the device, out of the request path.
 * The temporaries sit in a 16 MiB array compiled into libcph.so (present on every GPU the library is loaded on, until
 * the process ends).  Same inputs, outputs and results as cph_incidence_csr; no ws arguments; nothing is allocated.
 *   Limit : nnz <= CPH_INCIDENCE_CSR_FIXED_MAX_NNZ, otherwise CPH_ERR_TOO_LARGE and a message naming both numbers.
 *   Set-up: the first call on each device looks up the array's device address (the CUDA device that is current at
 *           the call; `stream` must belong to it).  Make that first call before starting any stream capture.
 *   Sharing: all calls on a device use the one array and find the previous call's bytes in it.  Nothing in it needs
 *           to start at zero (every temporary is written before it is read), so nothing is cleared; but two calls
 *           whose kernels overlap on the GPU (different streams, or different host threads on different streams)
 *           overwrite each other's temporaries.  Order them on one stream or with events. */
#define CPH_INCIDENCE_CSR_FIXED_MAX_NNZ 524288ull
int cph_incidence_csr_fixed(const uint32_t* pem_row, const uint32_t* pem_pk, const uint32_t* pem_col, uint64_t nnz,
                            uint32_t n_rows, int64_t* indptr, int32_t* indices, uint32_t* pk, cph_stream_t stream);

/* float32 `data` array of the scipy csr_matrix the reference returns (:282): data[t] = count. */
int cph_csr_counts_f32(const uint32_t* pk, uint64_t nnz, float* data, cph_stream_t stream);

/* ---- stage 2: clique expansion  A = H diag(d) H^T  ----------------------------------------------
 * Replaces HyperGraph.clique_expansion_init (algorithms/hypergraph.py:358-485): d_e = 1/D_e
 * (use_high_order) or 1 (:385-395, 418), A_ij = sum_e d_e c_ie c_je (:432-446), entries below
 * min_weight dropped (:453-455).  Accumulation is 64-bit fixed point (deterministic, exact for the
 * integer contact counts); the caller picks frac_bits from cph_row_sq_sums so that nothing overflows:
 *     frac_bits <= min(52, 61 - ceil(log2(max_i rowsq[i] + 1))).
 * cph_row_sq_sums  : rowsq[i] = sum_e c_ie^2  (an upper bound of A_ii, hence of every A_ij)
 * cph_diag_terms   : diag_fx[i] = sum_e round(c_ie^2 2^frac_bits / D_e)   (no atomics: CSR row reduce)
 * cph_diag_hist    : the same from the pem (no CSR needed): shared-memory u64 histogram, one flush per CTA
 * cph_expand_accumulate: acc[i,j] += round(c_ie c_je 2^frac_bits / D_e) for every unordered pair of
 *     distinct entries of a hyperedge (x2 when both entries sit on the same contig: the duplicated
 *     (row,col) of intra-contig Hi-C pairs, :277-282).  acc is zeroed first when zero_first != 0.
 *     entry_w (float per pem entry, constant inside a hyperedge; NULL = 1) is the W of :427-430.
 * Between accumulate and count the caller may sum acc / diag_fx over ranks (NCCL all-reduce).
 * cph_expand_count : thresholds and counts; leaves per-(row, column tile) offsets in ws; *h_nnz = nnz(A)
 * cph_expand_emit  : writes CSR(A) (sorted indices); ws must be the buffer cph_expand_count filled.
 *     node_w (double[n], may be NULL): A_ij *= float(node_w[i]*node_w[j]) -- the NW of
 *     hyperpartition.py:554-568 applied as in normalize_adjacency_matrix (:590-602) AFTER thresholding.
 *     data32 / data64: either may be NULL. */
int cph_row_sq_sums(const int64_t* indptr, const uint32_t* pk, uint32_t n, uint64_t* rowsq, cph_stream_t stream);
int cph_diag_terms(const int64_t* indptr, const uint32_t* pk, const float* csr_entry_w, uint32_t n,
                   int use_high_order, int frac_bits, uint64_t* diag_fx, cph_stream_t stream);
int cph_diag_hist(const uint32_t* pem_row, const uint32_t* pem_pk, const float* entry_w, uint64_t nnz, uint32_t n,
                  int use_high_order, int frac_bits, uint64_t* diag_fx, cph_stream_t stream);
int cph_expand_accumulate(const uint32_t* pem_row, const uint32_t* pem_pk, const float* entry_w, uint64_t nnz, uint32_t n,
                          int use_high_order, int frac_bits, int zero_first, uint64_t* acc, cph_stream_t stream);
/* cph_diag_hist + cph_expand_accumulate in ONE pass over the pem (the hot-path form): diag_fx is zeroed and filled,
 * acc is zeroed first when zero_first != 0. */
int cph_expand_accumulate_diag(const uint32_t* pem_row, const uint32_t* pem_pk, const float* entry_w, uint64_t nnz,
                               uint32_t n, int use_high_order, int frac_bits, int zero_first, uint64_t* acc,
                               uint64_t* diag_fx, cph_stream_t stream);
/* Binned form of cph_expand_accumulate_diag for accumulators beyond L2 (n > ~5.6k, no entry weights): the pair updates
 * are first written out grouped by accumulator block (contig-block pairs of <= 48 MB), then applied block after block so
 * that the cells being updated stay in L2; integer sums are order-free, so acc / diag_fx come out bit-identical to the
 * direct kernel.  Two calls with the SAME workspace: phase 1 zeroes, counts the pairs per block, fills diag_fx and
 * returns the number of records in *h_pairs; the caller allocates pair_records (u64 each) and calls phase 2, which
 * writes and applies them.  n <= 115368 (46 block rows of <= 50 MB blocks).
 * phase 3 / 4 run the two halves of phase 2 apart (record fill only / apply only) for timing. */
/* Swept form (round 2) for the same regime, nothing materialised per pair.  cph_expand_sweep_prep sorts every hyperedge
 * run by contig (stable) and packs each entry to 4 bytes -- head(1) | D_e(7) | count(24 - rb) | row(rb), rb = bits(n - 1)
 * -- into pe4[nnz]; *flag (device u32) becomes non-zero when a count does not fit its field or a run is longer than 64
 * entries (the caller then falls back to the direct kernel).  cph_expand_sweep_band handles ONE row band
 * [row_lo, row_hi) of the packed triangle: it zeroes the band (zero_first), streams pe4 once, and adds every pair whose
 * smaller contig lies in the band with one RED into the (L2-resident) band, plus the diagonal terms of the band's rows.
 * Calling it for a partition of [0, n) into bands gives acc / diag_fx bit-identical to cph_expand_accumulate_diag.
 * Between two band calls the caller may start the exchange of the finished band (multi-GPU). n <= 2^20. */
int cph_expand_sweep_prep(const uint32_t* pem_row, const uint32_t* pem_pk, uint64_t nnz, uint32_t n, int use_high_order,
                          uint32_t* pe4, uint32_t* flag, cph_stream_t stream);
int cph_expand_sweep_band(const uint32_t* pe4, uint64_t nnz, uint32_t n, uint32_t row_lo, uint32_t row_hi, int frac_bits,
                          int zero_first, uint64_t* acc, uint64_t* diag_fx, cph_stream_t stream);
int cph_expand_accumulate_binned(int phase, const uint32_t* pem_row, const uint32_t* pem_pk, uint64_t nnz, uint32_t n,
                                 int use_high_order, int frac_bits, int zero_first, uint64_t* acc, uint64_t* diag_fx,
                                 uint64_t* pair_records, uint64_t* h_pairs, void* ws, size_t ws_bytes,
                                 size_t* h_ws_need, cph_stream_t stream);
int cph_expand_count(const uint64_t* acc, const uint64_t* diag_fx, uint32_t n, int frac_bits, float min_weight,
                     uint64_t* h_nnz, void* ws, size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);
int cph_expand_emit(const uint64_t* acc, const uint64_t* diag_fx, uint32_t n, int frac_bits, float min_weight,
                    const double* node_w, const void* ws, int64_t* indptr, int32_t* indices,
                    float* data32, double* data64, cph_stream_t stream);

/* ---- CSR(A) epilogues ------------------------------------------------------------------------------
 * A is CSR with SORTED column indices (what cph_expand_emit writes).  Pair lists are int64 device arrays.
 * cph_csr_find_pairs   : find_csr_data_offsets (algorithms/hypergraph.py:81-97): offsets[t] = position of
 *                        (pair_r[t], pair_c[t]) in indices/data, or -1 when it is not a stored entry.
 * cph_csr_reweight_pairs: stored entries addressed by the allelic list are multiplied by allelic_factor, those
 *                        addressed by the weak list by weak_factor (a factor of 0 stores an exact 0); a pair listed
 *                        twice is applied once and weak wins over allelic, as in the reference's fancy-index
 *                        assignments (algorithms/hypergraph.py:464-481, 632-709 numeric branch, 1158-1170).
 *                        mark: nnz bytes of scratch.
 * cph_csr_scale_nodes  : normalize_adjacency_matrix (algorithms/hypergraph.py:590-602) with NW = outer(w, w) of
 *                        hyperpartition.py:554-568 formed on the fly: data[t] *= float(w[i] * w[j]).
 * cph_csr_abs_degree   : deg[i] = sum_j |A_ij| (algorithms/hypergraph.py:545) (diagonal skipped if skip_diag),
 *                        diag[i] = A_ii (may be NULL). */
int cph_csr_find_pairs(const int64_t* indptr, const int32_t* indices, uint32_t n, const int64_t* pair_r,
                       const int64_t* pair_c, uint64_t n_pairs, int64_t* offsets, cph_stream_t stream);
int cph_csr_reweight_pairs(const int64_t* indptr, const int32_t* indices, uint32_t n, uint64_t nnz,
                           const int64_t* al_r, const int64_t* al_c, uint64_t n_al, float allelic_factor,
                           const int64_t* wk_r, const int64_t* wk_c, uint64_t n_wk, float weak_factor,
                           float* data32, double* data64, uint8_t* mark, cph_stream_t stream);
int cph_csr_scale_nodes(const int64_t* indptr, const int32_t* indices, uint32_t n, const double* node_w,
                        float* data32, double* data64, cph_stream_t stream);
int cph_csr_abs_degree(const int64_t* indptr, const int32_t* indices, uint32_t n, const float* data32,
                       const double* data64, int skip_diag, double* deg, double* diag, cph_stream_t stream);

/* ---- stage 3: kprune on A ---------------------------------------------------------------------------
 * Replaces KPruneHyperGraph.run (kprune.py:407-517) and the l1 x l2 decision of KPruner (kprune.py:269-335).
 * A: CSR with sorted indices (diagonal ignored, :419).  row_of[t] = row of stored entry t (cph_csr_row_ids).
 * ag_ptr/ag_idx: allele lists per contig (CSR-like; empty = the contig has no allele group); at most
 * max_alleles (1..5) of them are used (1 = "first listed allele", kprune.py:437-447).
 * al_r/al_c: allele-table pairs as contig indices, table order; only rows with al_r < al_c count (:456).
 * flag (nnz bytes, overwritten): 1 = allelic stored upper entry (:455-468), 2 = weak / cross-allelic upper
 * entry (:470-517), 0 otherwise.  best_with / best_all (double[nnz], may be NULL): weight of the best matching
 * that pairs c1 with c2, and of the best matching overall, for every scored candidate. */
int cph_csr_row_ids(const int64_t* indptr, uint32_t n, int32_t* row_of, cph_stream_t stream);
int cph_kprune(const int64_t* indptr, const int32_t* indices, const float* data32, const double* data64, uint32_t n,
               uint64_t nnz, const int32_t* row_of, const int64_t* ag_ptr, const int32_t* ag_idx, int max_alleles,
               const int64_t* al_r, const int64_t* al_c, uint64_t n_al, uint8_t* flag, double* best_with,
               double* best_all, cph_stream_t stream);

/* ---- stage 4: Louvain / Leiden modularity clustering ---------------------------------------------------
 * Replaces the igraph calls of IRMM (algorithms/hypergraph.py:1239-1252): community_multilevel(weights,
 * resolution) [method 0] and community_leiden(weights, resolution_parameter, objective_function="modularity")
 * [method 1, n_iterations passes, python-igraph default 2] on the undirected graph of A's off-diagonal entries
 * (:1157-1244; A must be symmetric CSR, both triangles stored; the diagonal and non-finite values are ignored).
 * Quality = sum_c [ in_c / 2m - resolution (tot_c / 2m)^2 ].  Weights may be negative (allelic_factor < 0).
 * Deterministic: weights are converted to 64-bit fixed point, ties go to the lowest community id.
 * n_phases (<= 0: 8) = interleaved vertex phases per local-moving sweep, max_sweeps (<= 0: 64).
 * init_membership (device int32[n], labels in [0, n), or NULL = singletons): igraph's initial_membership (:1366-1373).
 * membership (device int32[n], out): groups numbered 0..g-1 in order of first appearance.
 * h_quality (host, may be NULL), h_stats (host uint32[4], may be NULL: groups, levels, sweeps, refinement rounds).
 * Synchronises `stream`.
 * cph_modularity: the same quality function for a given membership (labels in [0, n)). */
int cph_community(const int64_t* indptr, const int32_t* indices, const float* data32, const double* data64, uint32_t n,
                  uint64_t nnz, double resolution, int method, int n_iterations, int n_phases, int max_sweeps,
                  const int32_t* init_membership, int32_t* membership, double* h_quality, uint32_t* h_stats, void* ws,
                  size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);
int cph_modularity(const int64_t* indptr, const int32_t* indices, const float* data32, const double* data64, uint32_t n,
                   uint64_t nnz, double resolution, const int32_t* membership, double* h_quality, void* ws, size_t ws_bytes,
                   size_t* h_ws_need, cph_stream_t stream);

/* ---- group-level affinity sums (partition driver post-processing) ---------------------------------
 * labels: device int32[n], group of each contig in [0, n_groups) or -1 (in no group).
 * S (device double[n * n_groups], row-major, may be NULL) = A M: linkage of every contig to every group
 *     (recluster_by_linkage, hyperpartition.py:2889-2892; the diagonal of A is included, as there).
 * V (device double[n_groups * n_groups], may be NULL) = M^T A M: value_matrix of _merge (hyperpartition.py:2402-2410).
 * Deterministic (64-bit fixed-point accumulation).  Synchronises `stream` once. */
int cph_group_sums(const int64_t* indptr, const int32_t* indices, const float* data32, const double* data64,
                   uint32_t n, uint64_t nnz, const int32_t* labels, uint32_t n_groups, double* S, double* V, void* ws,
                   size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);

/* ---- synthetic workload generators (bench infrastructure; device twin of cphasing_b200/synth.py) ---
 * Seeded, counter-based tables of the shapes BASELINE.json names, written straight to HBM.
 * cph_synth_porec_count: nrec_off (u64[m+1]) <- exclusive prefix of the records per concatemer, *h_total = N.
 * cph_synth_porec_fill : Pore-C alignment records of concatemers [first, first+m), sorted by read_idx
 *                        (start/end may be NULL).  law: 0 = 2+Poisson(2), 1 = 2+floor(Exp(8)), 2 = pairs.
 * cph_synth_pairs      : Hi-C pairs (pos1/pos2 may be NULL); `intra` = fraction of intra-contig pairs. */
int cph_synth_porec_count(uint64_t seed, uint64_t first, uint64_t m, int law, uint64_t* nrec_off, uint64_t* h_total,
                          void* ws, size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream);
int cph_synth_porec_fill(uint64_t seed, uint64_t first, uint64_t m, int law, uint32_t n, uint32_t G, uint32_t ploidy,
                         const double* cum, const double* gcum, const int32_t* gstart, const int32_t* order,
                         const int64_t* length, const uint64_t* nrec_off, uint32_t* read_idx, uint32_t* contig_idx,
                         uint32_t* start, uint32_t* end, int8_t* mapq, cph_stream_t stream);
int cph_synth_pairs(uint64_t seed, uint64_t first, uint64_t m, double intra, uint32_t n, uint32_t G, uint32_t ploidy,
                    const double* cum, const double* gcum, const int32_t* gstart, const int32_t* order,
                    const int64_t* length, uint32_t* chrom1, uint32_t* chrom2, uint32_t* pos1, uint32_t* pos2,
                    int8_t* mapq, cph_stream_t stream);

/* ---- IRMM rounds 2.. : iterative hyperedge re-weighting (algorithms/hypergraph.py:1258-1397) ------------------
 * cph_irmm_edge_weights: per hyperedge, from the cluster labels of its contigs (labels[n], n_clusters = c):
 *     K_el over the ORIGINAL entries (:1282), e_c over the entries not yet masked (:1283), w_new (:1285-1289),
 *     purity rules (:1291-1313; suppressed entries get mask = 1), W update (:1316-1317).  mask (u8) and entry_w
 *     (float, constant inside a hyperedge) are per-pem-entry state carried from round to round (start: 0 / 1.0f).
 *     Two masks, because of an aliasing effect in the reference: `H_T = H.T` (:1135) and `H.tocoo()` (:1298) share H's
 *     data array, so the zeros of the FIRST suppression round also land in H_T; after that H gets fresh arrays (:1308-1310)
 *     and H_T stays as it is.  mask = zeros of H (all rounds), mask_t = zeros of H_T; `alias` != 0 while no suppression has
 *     happened yet (the caller clears it after the first call that reports n_suppressed > 0).  K_el reads H_T (mask_t).
 * cph_expand_full: A_new = H' W D_e^-1 H_T (:1321) into a FULL n x n fixed-point matrix (zeroed first): the left factor
 *     of a term must be unmasked in `mask`, the right one in `mask_t`, (H' W) is rounded to float32 as in the reference.
 * cph_dense_from_csr: CSR(A) -> dense fp64 n x n.   cph_irmm_blend: A <- (1-damping) A + damping thr(A_new), optional
 *     w_i w_j normalisation, zero diagonal (:1322-1329).   cph_dense_scale_pairs: A[r,c] *= factor (0 -> exact 0) for a
 *     DEDUPLICATED pair list (:1330-1349).   cph_dense_to_csr_count/emit: dense -> CSR; mode 0 = A as is, mode 1 = the
 *     undirected graph igraph builds from triu(A) (:1351-1366), mirrored into both triangles. */
int cph_irmm_edge_weights(const uint32_t* pem_row, const uint32_t* pem_pk, uint64_t nnz, const int32_t* labels,
                          uint32_t n_clusters, uint8_t* mask, uint8_t* mask_t, int alias, float* entry_w,
                          uint64_t* n_suppressed, cph_stream_t stream);
int cph_expand_full(const uint32_t* pem_row, const uint32_t* pem_pk, const float* entry_w, const uint8_t* mask,
                    const uint8_t* mask_t, uint64_t nnz, uint32_t n, int frac_bits, uint64_t* acc_full,
                    cph_stream_t stream);
int cph_dense_from_csr(const int64_t* indptr, const int32_t* indices, const float* data32, const double* data64,
                       uint32_t n, double* dense, cph_stream_t stream);
int cph_irmm_blend(const uint64_t* acc_full, int frac_bits, double min_weight, const double* node_w, uint32_t n,
                   double damping, double* dense, cph_stream_t stream);
int cph_dense_scale_pairs(double* dense, uint32_t n, const int64_t* rows, const int64_t* cols, uint64_t n_pairs,
                          double factor, cph_stream_t stream);
int cph_dense_to_csr_count(const double* dense, uint32_t n, int mode, uint64_t* h_nnz, void* ws, size_t ws_bytes,
                           size_t* h_ws_need, cph_stream_t stream);
int cph_dense_to_csr_emit(const double* dense, uint32_t n, int mode, const void* ws, int64_t* indptr, int32_t* indices,
                          float* data32, double* data64, cph_stream_t stream);

/* ---- host-side table decode (replaces the polars CSV reader of hypergraph.py:1706-1717, 904-909) -------
 * Multi-threaded TSV field parser, host memory in and out (no device work).  A block of whole lines is decoded into
 * columnar arrays: n_fields selected columns in increasing column order, field_kind 0 = unsigned -> u32, 1 = contig name
 * -> index through the open-addressing table (name_slots[h & (n_slots-1)] = index + 1, 0 = empty, linear probing, keyed
 * by cph_fnv1a; unknown names -> n_names), 2 = integer -> i8 (saturating).  Lines starting with '#' and empty lines are
 * skipped; malformed lines are skipped and counted in *n_bad.  out[k] must hold `cap` rows (cph_tsv_count_rows). */
uint64_t cph_fnv1a(const char* p, uint64_t len);
int cph_tsv_count_rows(const char* buf, uint64_t len, int n_threads, uint64_t* n_rows);
int cph_tsv_parse(const char* buf, uint64_t len, int n_threads, int n_fields, const int* field_col, const int* field_kind,
                  void* const* out, uint64_t cap, const char* names_blob, const uint64_t* names_off, uint32_t n_names,
                  const uint32_t* name_slots, uint32_t n_slots, uint64_t* n_rows, uint64_t* n_bad);

#ifdef __cplusplus
}
#endif
#endif
