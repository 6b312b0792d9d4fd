// Small CSR(A) epilogues that the reference does with scipy fancy indexing.
//
// Replaces (paths under /root/reference/cphasing/):
//   algorithms/hypergraph.py:81-97     find_csr_data_offsets  (argsort of nnz(A) 64-bit keys + searchsorted
//                                      per call -> one binary search per pair inside its sorted CSR row)
//   algorithms/hypergraph.py:464-481   allelic / weak overwrite in clique_expansion_init
//   algorithms/hypergraph.py:632-709   reweight_adjacency_matrix
//   algorithms/hypergraph.py:1158-1170 IRMM allelic_factor / weak zeroing
//   algorithms/hypergraph.py:590-602   normalize_adjacency_matrix (dense n x n NW gather -> w_i * w_j on the fly)
//   algorithms/hypergraph.py:545-551   filter_adjacency_matrix degree = sum_j |A_ij|
// All HBM bound: bytes = 12 nnz(A) for the sweeps, 16 B + log2(row length) probes per pair.
#include "common.cuh"

namespace {

constexpr int TPB = 256;

__device__ __forceinline__ size_t gtid() { return (size_t)blockIdx.x * blockDim.x + threadIdx.x; }

__device__ __forceinline__ i64 find_in_row(const i64* __restrict__ indptr, const i32* __restrict__ indices, i64 r, i32 c) {
    i64 lo = indptr[r], hi = indptr[r + 1];
    while (lo < hi) {
        i64 mid = (lo + hi) >> 1;
        i32 v = indices[mid];
        if (v < c) lo = mid + 1;
        else hi = mid;
    }
    return (lo < indptr[r + 1] && indices[lo] == c) ? lo : -1;
}

__global__ void __launch_bounds__(TPB) find_pairs_kernel(const i64* __restrict__ indptr, const i32* __restrict__ indices, u32 n,
                                                         const i64* __restrict__ pr, const i64* __restrict__ pc, u64 np,
                                                         i64* __restrict__ off) {
    size_t t = gtid();
    if (t >= np) return;
    i64 r = pr[t], c = pc[t];
    off[t] = (r >= 0 && r < (i64)n && c >= 0 && c < (i64)n) ? find_in_row(indptr, indices, r, (i32)c) : -1;
}

// mark[off] = 1 for every stored entry addressed by a pair (duplicated pairs mark once)
__global__ void __launch_bounds__(TPB) mark_pairs_kernel(const i64* __restrict__ indptr, const i32* __restrict__ indices, u32 n,
                                                         const i64* __restrict__ pr, const i64* __restrict__ pc, u64 np,
                                                         u8* __restrict__ mark, u8 tag) {
    size_t t = gtid();
    if (t >= np) return;
    i64 r = pr[t], c = pc[t];
    if (r < 0 || r >= (i64)n || c < 0 || c >= (i64)n) return;
    i64 o = find_in_row(indptr, indices, r, (i32)c);
    if (o >= 0) mark[o] = tag;
}

__global__ void __launch_bounds__(TPB) apply_marks_kernel(const u8* __restrict__ mark, u64 nnz, float f1, float f2,
                                                          float* __restrict__ d32, double* __restrict__ d64) {
    size_t t = gtid();
    if (t >= nnz) return;
    u8 m = mark[t];
    if (!m) return;
    // tag 1: multiply by f1 (allelic) ; tag 2: multiply by f2 (weak / cross-allelic); factor 0 -> exact 0
    float f = (m == 1) ? f1 : f2;
    if (d32) d32[t] = (f == 0.0f) ? 0.0f : d32[t] * f;
    if (d64) d64[t] = (f == 0.0f) ? 0.0 : d64[t] * (double)f;
}

// warp per row
__global__ void __launch_bounds__(TPB) scale_nodes_kernel(const i64* __restrict__ indptr, const i32* __restrict__ indices, u32 n,
                                                          const double* __restrict__ w, float* __restrict__ d32,
                                                          double* __restrict__ d64) {
    const u32 row = (u32)(gtid() >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    const double wi = w[row];
    for (i64 t = indptr[row] + lane, e = indptr[row + 1]; t < e; t += 32) {
        float nw = (float)(wi * w[indices[t]]);       // hyperpartition.py:568 NW stored as float32
        if (d32) d32[t] *= nw;
        if (d64) d64[t] *= (double)nw;
    }
}

__global__ void __launch_bounds__(TPB) abs_degree_kernel(const i64* __restrict__ indptr, const i32* __restrict__ indices, u32 n,
                                                         const float* __restrict__ d32, const double* __restrict__ d64,
                                                         int skip_diag, double* __restrict__ deg, double* __restrict__ diag) {
    const u32 row = (u32)(gtid() >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    double s = 0.0, dg = 0.0;
    for (i64 t = indptr[row] + lane, e = indptr[row + 1]; t < e; t += 32) {
        double v = d64 ? d64[t] : (double)d32[t];
        if ((u32)indices[t] == row) {
            dg += v;
            if (skip_diag) continue;
        }
        s += fabs(v);
    }
    s = warp_sum(s);
    dg = warp_sum(dg);
    if (lane == 0) {
        deg[row] = s;
        if (diag) diag[row] = dg;
    }
}

}  // namespace

extern "C" int cph_csr_find_pairs(const i64* indptr, const i32* indices, u32 n, const i64* pair_r, const i64* pair_c, u64 n_pairs,
                                  i64* offsets, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n_pairs == 0) return CPH_OK;
    ARG_CHECK(indptr && indices && pair_r && pair_c && offsets && n > 0);
    find_pairs_kernel<<<grid_for(n_pairs, TPB), TPB, 0, st>>>(indptr, indices, n, pair_r, pair_c, n_pairs, offsets);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_csr_reweight_pairs(const i64* indptr, const i32* indices, u32 n, u64 nnz, const i64* al_r, const i64* al_c,
                                      u64 n_al, float allelic_factor, const i64* wk_r, const i64* wk_c, u64 n_wk,
                                      float weak_factor, float* data32, double* data64, u8* mark, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (nnz == 0 || (n_al == 0 && n_wk == 0)) return CPH_OK;
    ARG_CHECK(indptr && indices && mark && n > 0 && (data32 || data64));
    CUDA_TRY(cudaMemsetAsync(mark, 0, nnz, st));
    if (n_al) {
        ARG_CHECK(al_r && al_c);
        mark_pairs_kernel<<<grid_for(n_al, TPB), TPB, 0, st>>>(indptr, indices, n, al_r, al_c, n_al, mark, 1);
        KERNEL_CHECK();
    }
    if (n_wk) {   // weak is applied after allelic in the reference, so it wins on overlap
        ARG_CHECK(wk_r && wk_c);
        mark_pairs_kernel<<<grid_for(n_wk, TPB), TPB, 0, st>>>(indptr, indices, n, wk_r, wk_c, n_wk, mark, 2);
        KERNEL_CHECK();
    }
    apply_marks_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(mark, nnz, allelic_factor, weak_factor, data32, data64);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_csr_scale_nodes(const i64* indptr, const i32* indices, u32 n, const double* node_w, float* data32,
                                   double* data64, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(indptr && indices && node_w && n > 0 && (data32 || data64));
    scale_nodes_kernel<<<grid_for((size_t)n * 32, TPB), TPB, 0, st>>>(indptr, indices, n, node_w, data32, data64);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_csr_abs_degree(const i64* indptr, const i32* indices, u32 n, const float* data32, const double* data64,
                                  int skip_diag, double* deg, double* diag, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(indptr && indices && deg && n > 0 && (data32 || data64));
    abs_degree_kernel<<<grid_for((size_t)n * 32, TPB), TPB, 0, st>>>(indptr, indices, n, data32, data64, skip_diag, deg, diag);
    KERNEL_CHECK();
    return CPH_OK;
}
