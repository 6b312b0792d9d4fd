// Stage 1c: row / column filters and the incidence matrix H, on the edge-major form.
//
// Replaces (paths under /root/reference/cphasing/):
//   algorithms/hypergraph.py:175-184   HyperGraph.get_data      entry filter  mapq >= min_quality
//   algorithms/hypergraph.py:207-223   HyperGraph.remove_rows   (row_drop / row_map < 0)
//   algorithms/hypergraph.py:277-306   incidence_matrix         CSR build, row filter sum>=min_contacts,
//                                                               column filter sum>=2 (after the row filter)
//   algorithms/hypergraph.py:938-954   extract_incidence_matrix2 (same machinery, row_map = position in idx)
// scipy builds CSR, copies it for the row slice and again for the column slice (3 x O(nnz), one
// thread).  Here: one flag pass over the hyperedge runs, one scan, one compaction that writes the
// packed edge-major entries (consumed by the clique expansion) and the sort keys, a 1-2 pass radix
// sort by row for the CSR view.
#include "common.cuh"

namespace {

constexpr int TPB = 256;
constexpr u32 RS_MAX_SMEM_BINS = 49152;      // 192 KB of u32 bins
constexpr u32 RS_CHUNK = 32768;              // entries per CTA; count < 2^16 each -> u32 bins cannot overflow

__device__ __forceinline__ size_t gtid() { return (size_t)blockIdx.x * blockDim.x + threadIdx.x; }

__device__ __forceinline__ bool entry_ok(const int8_t* mapq, int min_mapq, size_t t) {
    return mapq == nullptr || (int)mapq[t] >= min_mapq;
}

// per-contig sum of counts; shared-memory privatised histogram, one flush per CTA
__global__ void __launch_bounds__(TPB) row_sums_kernel(const u32* __restrict__ rem_row, const u32* __restrict__ rem_count,
                                                       const int8_t* __restrict__ rem_mapq, u64 nnz, u32 n, int min_mapq,
                                                       const u8* __restrict__ row_drop, u32 cmask, int use_smem,
                                                       unsigned long long* __restrict__ rowsum, u32* __restrict__ max_count) {
    extern __shared__ u32 bins[];
    if (use_smem) {
        for (u32 i = threadIdx.x; i < n; i += TPB) bins[i] = 0;
        __syncthreads();
    }
    const size_t lo = (size_t)blockIdx.x * RS_CHUNK;
    const size_t hi = min((size_t)nnz, lo + RS_CHUNK);
    u32 mx = 0;
    for (size_t t = lo + threadIdx.x; t < hi; t += TPB) {
        if (!entry_ok(rem_mapq, min_mapq, t)) continue;
        u32 r = rem_row[t] & CPH_ROWM;
        if (r >= n || (row_drop && row_drop[r])) continue;
        u32 c = rem_count[t] & cmask;
        mx = max(mx, c);
        if (use_smem && c < 65536u) atomicAdd(&bins[r], c);
        else atomicAdd(&rowsum[r], (unsigned long long)c);
    }
    if (max_count) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if ((threadIdx.x & 31) == 0 && mx) atomicMax(max_count, mx);
    }
    if (use_smem) {
        __syncthreads();
        for (u32 i = threadIdx.x; i < n; i += TPB) {
            u32 v = bins[i];
            if (v) atomicAdd(&rowsum[i], (unsigned long long)v);
        }
    }
}

// One thread per hyperedge head: decide the fate of the whole run.
//   flag[t] = 0 (dropped) or  1 | newhead<<1 | D<<8   ;   sc[t] = kept | newhead<<32  (scan input)
__global__ void __launch_bounds__(TPB) edge_flag_kernel(const u32* __restrict__ rem_row, const u32* __restrict__ rem_count,
                                                        const int8_t* __restrict__ rem_mapq, u64 nnz, u32 n, int min_mapq,
                                                        const i32* __restrict__ row_map, u32 col_min_sum, u32 cmask,
                                                        u32* __restrict__ flag, u64* __restrict__ sc, u32* __restrict__ bad) {
    size_t t = gtid();
    if (t > nnz) return;
    if (t == nnz) {
        sc[t] = 0;   // sentinel
        return;
    }
    u32 r0 = rem_row[t];
    if (!(r0 & CPH_HEAD) && t != 0) return;
    // pass 1: surviving entries of the run
    u64 colsum = 0;
    u32 k = 0;
    size_t e = t;
    do {
        u32 r = rem_row[e] & CPH_ROWM;
        bool ok = entry_ok(rem_mapq, min_mapq, e) && r < n && row_map[r] >= 0;
        if (ok) {
            u32 c = rem_count[e];
            if (cmask == 0xffffffffu && c > CPH_CNTM) *bad = 1u;   // not representable (the reference stores float32)
            colsum += c & cmask;
            ++k;
        }
        ++e;
    } while (e < nnz && !(rem_row[e] & CPH_HEAD));
    const bool edge_keep = k > 0 && colsum >= (u64)col_min_sum;     // algorithms/hypergraph.py:305-306
    // D_e = max(int8(k) - 1, 1) with numpy's int8 wrap-around (algorithms/hypergraph.py:385, 418)
    int d8 = (int)(int8_t)(u8)((k - 1u) & 0xffu);
    const u32 D = (u32)max(d8, 1);
    bool first = true;
    for (size_t u = t; u < e; ++u) {
        u32 r = rem_row[u] & CPH_ROWM;
        bool ok = edge_keep && entry_ok(rem_mapq, min_mapq, u) && r < n && row_map[r] >= 0;
        u32 f = 0;
        u64 s = 0;
        if (ok) {
            f = 1u | (first ? 2u : 0u) | (D << 8);
            s = 1ull | (first ? (1ull << 32) : 0ull);
            first = false;
        }
        flag[u] = f;
        sc[u] = s;
    }
}

__global__ void __launch_bounds__(TPB) compact_kernel(const u32* __restrict__ rem_row, const u32* __restrict__ rem_col,
                                                      const u32* __restrict__ rem_count, u64 nnz, const i32* __restrict__ row_map,
                                                      u32 cmask, int renumber, const u32* __restrict__ flag,
                                                      const u64* __restrict__ sc, u32* __restrict__ pem_row, u32* __restrict__ pem_pk,
                                                      u64* __restrict__ keys, u32* __restrict__ vals) {
    size_t t = gtid();
    if (t >= nnz) return;
    u32 f = flag[t];
    if (!f) return;
    u64 s = sc[t];
    u32 pos = (u32)s;
    u32 heads_before = (u32)(s >> 32);
    bool head = (f & 2u) != 0;
    u32 newcol = renumber ? (head ? heads_before : heads_before - 1u) : rem_col[t];
    u32 newrow = (u32)row_map[rem_row[t] & CPH_ROWM];
    u32 pk = ((f >> 8) << 24) | (rem_count[t] & cmask);
    pem_row[pos] = newrow | (head ? CPH_HEAD : 0u);
    pem_pk[pos] = pk;
    if (keys) {
        keys[pos] = ((u64)newrow << 32) | newcol;
        vals[pos] = pk;
    }
}

__global__ void __launch_bounds__(TPB) csr_finish_kernel(const u64* __restrict__ keys, const u32* __restrict__ vals, u64 nnz,
                                                         u32 n_rows, i64* __restrict__ indptr, i32* __restrict__ indices,
                                                         u32* __restrict__ pk) {
    size_t t = gtid();
    if (t >= nnz) return;
    u64 k = keys[t];
    indices[t] = (i32)(u32)k;
    pk[t] = vals[t];
    i64 r = (i64)(k >> 32);
    i64 rprev = t ? (i64)(keys[t - 1] >> 32) : -1;
    for (i64 q = rprev + 1; q <= r; ++q) indptr[q] = (i64)t;
    if (t == nnz - 1)
        for (i64 q = r + 1; q <= (i64)n_rows; ++q) indptr[q] = (i64)nnz;
}

__global__ void __launch_bounds__(TPB) counts_f32_kernel(const u32* __restrict__ pk, u64 nnz, float* __restrict__ data) {
    size_t t = gtid();
    if (t < nnz) data[t] = (float)(pk[t] & CPH_CNTM);
}

// CSR-view keys from a pem: (row << 32) | col, col = explicit id or running number of hyperedge heads
__global__ void __launch_bounds__(TPB) pem_keys_kernel(const u32* __restrict__ pem_row, const u32* __restrict__ pem_pk,
                                                       const u32* __restrict__ pem_col, const u32* __restrict__ heads_incl, u64 nnz,
                                                       u64* __restrict__ keys, u32* __restrict__ vals) {
    size_t t = gtid();
    if (t >= nnz) return;
    u32 r = pem_row[t];
    // heads_incl is the EXCLUSIVE scan of the head flags: heads before t; entry t belongs to hyperedge (that + head(t)) - 1
    u32 col = pem_col ? pem_col[t] : heads_incl[t] + ((r & CPH_HEAD) ? 1u : 0u) - 1u;
    keys[t] = ((u64)(r & CPH_ROWM) << 32) | col;
    vals[t] = pem_pk[t];
}

__global__ void __launch_bounds__(TPB) head_flags_kernel(const u32* __restrict__ pem_row, u64 nnz, u32* __restrict__ f) {
    size_t t = gtid();
    if (t < nnz) f[t] = (pem_row[t] & CPH_HEAD) ? 1u : 0u;
}

// diag_fx[i] += round(c^2 w 2^F / D) over the pem entries; shared-memory u64 bins, one flush per CTA
__global__ void __launch_bounds__(TPB) diag_hist_kernel(const u32* __restrict__ pem_row, const u32* __restrict__ pem_pk,
                                                        const float* __restrict__ ew, u64 nnz, u32 n, int use_high_order,
                                                        int frac_bits, int use_smem, unsigned long long* __restrict__ diag) {
    extern __shared__ unsigned long long dbins[];
    __shared__ double scale[128];
    for (int d = threadIdx.x; d < 128; d += TPB) scale[d] = d ? ldexp(1.0, frac_bits) / (double)d : 0.0;
    if (use_smem)
        for (u32 i = threadIdx.x; i < n; i += TPB) dbins[i] = 0ull;
    __syncthreads();
    const size_t lo = (size_t)blockIdx.x * RS_CHUNK;
    const size_t hi = min((size_t)nnz, lo + RS_CHUNK);
    for (size_t t = lo + threadIdx.x; t < hi; t += TPB) {
        u32 p = pem_pk[t];
        u32 r = pem_row[t] & CPH_ROWM;
        double c = (double)(p & CPH_CNTM);
        double sc = scale[use_high_order ? (p >> 24) : 1u];
        if (ew) sc *= (double)ew[t];
        unsigned long long v = __double2ull_rn((c * sc) * c);
        if (use_smem) atomicAdd(&dbins[r], v);
        else atomicAdd(&diag[r], v);
    }
    if (use_smem) {
        __syncthreads();
        for (u32 i = threadIdx.x; i < n; i += TPB) {
            unsigned long long v = dbins[i];
            if (v) atomicAdd(&diag[i], v);
        }
    }
}

}  // namespace

extern "C" int cph_row_sums(const u32* rem_row, const u32* rem_count, const int8_t* rem_mapq, u64 nnz, u32 n_contigs, int min_mapq,
                            const u8* row_drop, int packed, u64* rowsum, u32* max_count, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(rowsum && n_contigs > 0);
    ARG_CHECK(nnz == 0 || (rem_row && rem_count));
    CUDA_TRY(cudaMemsetAsync(rowsum, 0, (size_t)n_contigs * sizeof(u64), st));
    if (max_count) CUDA_TRY(cudaMemsetAsync(max_count, 0, sizeof(u32), st));
    if (nnz == 0) return CPH_OK;
    const int use_smem = n_contigs <= RS_MAX_SMEM_BINS;
    const size_t smem = use_smem ? (size_t)n_contigs * sizeof(u32) : 0;
    if (smem > 48 * 1024) CUDA_TRY(cudaFuncSetAttribute(row_sums_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    row_sums_kernel<<<grid_for(nnz, RS_CHUNK), TPB, smem, st>>>(rem_row, rem_count, rem_mapq, nnz, n_contigs, min_mapq, row_drop,
                                                               packed ? CPH_CNTM : 0xffffffffu, use_smem,
                                                               (unsigned long long*)rowsum, max_count);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_incidence_build(const u32* rem_row, const u32* rem_col, const u32* rem_count, const int8_t* rem_mapq, u64 nnz,
                                   u32 n_contigs, int min_mapq, const i32* row_map, u32 n_rows_out, u32 col_min_sum, int packed,
                                   u32* pem_row, u32* pem_pk, i64* indptr, i32* indices, u32* pk, u64* h_counts, void* ws,
                                   size_t ws_bytes, size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(nnz < (1ull << 32) - 8192);
    ARG_CHECK(n_contigs > 0 && n_contigs < CPH_HEAD);
    const int renumber = col_min_sum > 0;
    const u32 cmask = packed ? CPH_CNTM : 0xffffffffu;
    WS_BEGIN(ar);
    u32* flag = ar.take<u32>(nnz);
    u64* sc = ar.take<u64>(nnz + 1);
    u64* ka = ar.take<u64>(nnz);
    u64* kb = ar.take<u64>(nnz);
    u32* va = ar.take<u32>(nnz);
    u32* vb = ar.take<u32>(nnz);
    u64* total = ar.take<u64>(1);
    u32* bad = ar.take<u32>(1);
    if (!ar.dry) {
        ARENA_CHECK(ar);
        ARG_CHECK(row_map && pem_row && pem_pk && h_counts);
        ARG_CHECK(indptr == nullptr || (indices && pk));
        ARG_CHECK(nnz == 0 || (rem_row && rem_count));
        ARG_CHECK(renumber || rem_col || nnz == 0);
        if (indptr) CUDA_TRY(cudaMemsetAsync(indptr, 0, ((size_t)n_rows_out + 1) * sizeof(i64), st));
        CUDA_TRY(cudaMemsetAsync(bad, 0, sizeof(u32), st));
        if (nnz > 0) {
            edge_flag_kernel<<<grid_for(nnz + 1, TPB), TPB, 0, st>>>(rem_row, rem_count, rem_mapq, nnz, n_contigs, min_mapq, row_map,
                                                                     col_min_sum, cmask, flag, sc, bad);
            KERNEL_CHECK();
        }
    }
    TRY(scan_u64(sc, nnz > 0 ? nnz + 1 : 0, total, ar, st));
    u64 h_total = 0;
    u32 h_bad = 0;
    if (!ar.dry && nnz > 0) {
        compact_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(rem_row, rem_col, rem_count, nnz, row_map, cmask, renumber, flag, sc,
                                                           pem_row, pem_pk, indptr ? ka : nullptr, va);
        KERNEL_CHECK();
        CUDA_TRY(cudaMemcpyAsync(&h_total, total, sizeof(u64), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaMemcpyAsync(&h_bad, bad, sizeof(u32), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        if (h_bad) {
            cph_set_error(__FILE__, __LINE__, "an alignment count >= 2^24 cannot be represented exactly (the reference stores counts as float32)");
            return CPH_ERR_TOO_LARGE;
        }
    }
    const u64 nnz_out = h_total & 0xffffffffull;
    const u64 n_heads = h_total >> 32;
    int in_b = 0;
    // dry mode reserves for the worst case (nothing filtered, CSR wanted)
    if (ar.dry || indptr)
        TRY(radix_sort_u64(ka, kb, va, vb, 4, ar.dry ? nnz : nnz_out, 32, 32 + bits_for(n_rows_out > 1 ? n_rows_out - 1 : 1), &in_b, ar, st));
    WS_QUERY_RETURN(ar);
    if (nnz_out > 0 && indptr) {
        csr_finish_kernel<<<grid_for(nnz_out, TPB), TPB, 0, st>>>(in_b ? kb : ka, in_b ? vb : va, nnz_out, n_rows_out, indptr, indices,
                                                                  pk);
        KERNEL_CHECK();
    }
    h_counts[0] = nnz_out;
    h_counts[1] = n_heads;   // number of surviving hyperedges (= columns of H when renumbering)
    return CPH_OK;
}

extern "C" int cph_csr_counts_f32(const u32* pk, u64 nnz, float* data, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (nnz == 0) return CPH_OK;
    ARG_CHECK(pk && data);
    counts_f32_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(pk, nnz, data);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_incidence_csr(const u32* pem_row, const u32* pem_pk, const u32* pem_col, u64 nnz, u32 n_rows, i64* indptr,
                                 i32* indices, u32* pk, void* ws, size_t ws_bytes, size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(nnz < (1ull << 32) - 8192);
    WS_BEGIN(ar);
    u32* hf = ar.take<u32>(nnz + 1);
    u64* ka = ar.take<u64>(nnz);
    u64* kb = ar.take<u64>(nnz);
    u32* va = ar.take<u32>(nnz);
    u32* vb = ar.take<u32>(nnz);
    if (!ar.dry) {
        ARENA_CHECK(ar);
        ARG_CHECK(indptr && n_rows > 0);
        ARG_CHECK(nnz == 0 || (pem_row && pem_pk && indices && pk));
        CUDA_TRY(cudaMemsetAsync(indptr, 0, ((size_t)n_rows + 1) * sizeof(i64), st));
        if (nnz > 0 && !pem_col) {
            head_flags_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(pem_row, nnz, hf);
            KERNEL_CHECK();
        }
    }
    TRY(scan_u32(hf, (ar.dry || !pem_col) ? nnz : 0, nullptr, ar, st));
    if (!ar.dry && nnz > 0) {
        pem_keys_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(pem_row, pem_pk, pem_col, hf, nnz, ka, va);
        KERNEL_CHECK();
    }
    int in_b = 0;
    TRY(radix_sort_u64(ka, kb, va, vb, 4, nnz, 32, 32 + bits_for(n_rows > 1 ? n_rows - 1 : 1), &in_b, ar, st));
    WS_QUERY_RETURN(ar);
    if (nnz > 0) {
        csr_finish_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(in_b ? kb : ka, in_b ? vb : va, nnz, n_rows, indptr, indices, pk);
        KERNEL_CHECK();
    }
    return CPH_OK;
}

extern "C" int cph_diag_hist(const u32* pem_row, const u32* pem_pk, const float* entry_w, u64 nnz, u32 n, int use_high_order,
                             int frac_bits, u64* diag_fx, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(diag_fx && n > 0);
    ARG_CHECK(frac_bits >= 0 && frac_bits <= 52);
    CUDA_TRY(cudaMemsetAsync(diag_fx, 0, (size_t)n * sizeof(u64), st));
    if (nnz == 0) return CPH_OK;
    ARG_CHECK(pem_row && pem_pk);
    const int use_smem = n <= RS_MAX_SMEM_BINS / 2;
    const size_t smem = use_smem ? (size_t)n * sizeof(u64) : 0;
    if (smem > 48 * 1024) CUDA_TRY(cudaFuncSetAttribute(diag_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    diag_hist_kernel<<<grid_for(nnz, RS_CHUNK), TPB, smem, st>>>(pem_row, pem_pk, entry_w, nnz, n, use_high_order, frac_bits, use_smem,
                                                                (unsigned long long*)diag_fx);
    KERNEL_CHECK();
    return CPH_OK;
}
