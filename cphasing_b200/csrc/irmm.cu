// IRMM rounds 2.. : iterative hyperedge re-weighting (max_round > 1).
//
// Replaces the loop body of IRMM, /root/reference/cphasing/algorithms/hypergraph.py:1258-1397:
//   :1274-1285  K = H^T C (hyperedge x cluster counts, a DENSE m x c float array on the host), e_c = column sums,
//               w_new = sum_c 1/(1+K_ec) * (e_c + c) / max(2 e_c, 1)
//   :1288-1311  purity rules: in a concatemer that is >= 75 % one cluster, the monomers of other clusters are zeroed in H
//   :1312-1317  low-purity penalty, W <- (w_new + w_prev)/2 for e_c >= 2
//   :1321       A_new = H' W D_e^-1 H^T   -- H' is the re-weighted matrix, H^T still the ORIGINAL transpose, so the
//               product is not symmetric; igraph later reads its upper triangle only (:1354)
//   :1322-1329  threshold, A <- A/2 + A_new/2, normalisation, zero diagonal
// Here K never exists: a thread walks its hyperedge in the packed edge-major form and only meets the clusters that
// occur in it (sum_c 1/(1+K_ec) = (c - #distinct) + sum over the distinct ones).  A_new is accumulated in 64-bit fixed
// point into a full n x n matrix (two REDs per unordered pair, each gated by the H' mask of its left factor), blended
// into a dense fp64 A and re-emitted as CSR for the clustering.  Per-entry state: mask (u8) and weight (f32, constant
// inside a hyperedge) arrays of nnz elements.  HBM bound: 13 nnz per weight pass, 13 nnz + REDs per expansion,
// 16 n^2 per blend.
#include "common.cuh"

namespace {

constexpr int TPB = 256;

__device__ __forceinline__ size_t gtid() { return (size_t)blockIdx.x * blockDim.x + threadIdx.x; }

// one thread per hyperedge (at its head entry)
__global__ void __launch_bounds__(TPB) irmm_edge_kernel(const u32* __restrict__ pem_row, const u32* __restrict__ pem_pk, u64 nnz,
                                                        const i32* __restrict__ label, u32 n_clusters, u8* __restrict__ mask,
                                                        u8* __restrict__ maskT, int alias, float* __restrict__ w,
                                                        unsigned long long* __restrict__ n_suppressed) {
    const size_t t = gtid();
    if (t >= nnz) return;
    if (!(pem_row[t] & CPH_HEAD)) return;
    size_t e = t + 1;
    while (e < nnz && !(pem_row[e] & CPH_HEAD)) ++e;
    const double c = (double)n_clusters;
    double ec = 0.0;      // column sum of the CURRENT (re-weighted) H  (:1283)
    for (size_t j = t; j < e; ++j)
        if (!mask[j]) ec += (double)(pem_pk[j] & CPH_CNTM);
    // K_el over the entries of H_T (:1282).  H_T shares H's data array until H is rebuilt (:1298-1310), so it carries
    // the zeros of the FIRST suppression round and none of the later ones: maskT.
    int nd = 0, maj = 0x7fffffff;
    double inv = 0.0, maxK = -1.0;
    for (size_t j = t; j < e; ++j) {
        const i32 l = label[pem_row[j] & CPH_ROWM];
        bool first = true;
        for (size_t q = t; q < j; ++q)
            if (label[pem_row[q] & CPH_ROWM] == l) {
                first = false;
                break;
            }
        if (!first) continue;
        double K = 0.0;
        for (size_t q = j; q < e; ++q)
            if (label[pem_row[q] & CPH_ROWM] == l && !maskT[q]) K += (double)(pem_pk[q] & CPH_CNTM);
        ++nd;
        inv += 1.0 / (1.0 + K);
        if (K > maxK || (K == maxK && l < maj)) {      // np.argmax: first (lowest) cluster among ties
            maxK = K;
            maj = l;
        }
    }
    const double inv_sum = (c - (double)nd) + inv;
    double w_new = inv_sum * (ec + c) / fmax(ec * 2.0, 1.0);
    const double purity = maxK / fmax(ec, 1.0);
    if (purity >= 0.75 && ec > maxK && ec >= 3.0) {     // :1290-1310
        u32 hit = 0;
        for (size_t j = t; j < e; ++j)
            if (!mask[j] && label[pem_row[j] & CPH_ROWM] != maj) {
                mask[j] = 1;
                if (alias) maskT[j] = 1;
                ++hit;
            }
        if (hit && n_suppressed) atomicAdd(n_suppressed, 1ull);
    }
    if (ec >= 4.0 && purity < 0.4) w_new *= 0.1;        // :1312-1313
    const float w_prev = w[t];
    const float w_out = ec >= 2.0 ? (float)(0.5 * ((double)(float)w_new + (double)w_prev)) : w_prev;   // :1316-1317
    for (size_t j = t; j < e; ++j) w[j] = w_out;
}

// thread per entry a: ordered pairs with the later entries b of its hyperedge, into the FULL matrix
__global__ void __launch_bounds__(TPB) expand_full_kernel(const u32* __restrict__ pem_row, const u32* __restrict__ pem_pk,
                                                          const float* __restrict__ w, const u8* __restrict__ maskL,
                                                          const u8* __restrict__ maskT, u64 nnz, u32 n, int frac_bits,
                                                          unsigned long long* __restrict__ acc) {
    __shared__ double scale[128];
    for (int d = threadIdx.x; d < 128; d += TPB) scale[d] = d ? ldexp(1.0, frac_bits) / (double)d : 0.0;
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * TPB;
    for (size_t a = gtid(); a < nnz; a += stride) {
        size_t b = a + 1;
        if (b >= nnz) break;
        u32 rb = pem_row[b];
        if (rb & CPH_HEAD) continue;
        const u32 pa = pem_pk[a];
        const u64 i = pem_row[a] & CPH_ROWM;
        const bool mla = maskL && maskL[a], mta = maskT && maskT[a];
        const double sc = scale[pa >> 24];
        const float wa = w ? w[a] : 1.0f;
        const float ca = (float)(pa & CPH_CNTM);
        // `H @ W @ D_e_inv @ H_T` runs left to right: (H @ W) is a float32 product, the rest float64 (:1321)
        const double fa = (double)(ca * wa) * sc;
        for (;;) {
            const u64 j = rb & CPH_ROWM;
            const float cb = (float)(pem_pk[b] & CPH_CNTM);
            if (i != j) {                                   // the diagonal is zeroed afterwards (:1329)
                // A_new[i,j] += H'[i,e] w_e d_e H_T[e,j] : left factor from the re-weighted H, right one from H_T
                if (!mla && !(maskT && maskT[b])) atomicAdd(&acc[i * n + j], __double2ull_rn(fa * (double)cb));
                if (!(maskL && maskL[b]) && !mta) atomicAdd(&acc[j * n + i], __double2ull_rn((double)(cb * wa) * sc * (double)ca));
            }
            if (++b >= nnz) break;
            rb = pem_row[b];
            if (rb & CPH_HEAD) break;
        }
    }
}

// ---- dense fp64 A (n x n, row-major) ----------------------------------------------------------
__global__ void __launch_bounds__(TPB) dense_from_csr_kernel(const i64* __restrict__ indptr, const i32* __restrict__ indices,
                                                             const float* __restrict__ d32, const double* __restrict__ d64, u32 n,
                                                             double* __restrict__ A) {
    const u32 row = (u32)(gtid() >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    for (i64 t = indptr[row] + lane, e = indptr[row + 1]; t < e; t += 32)
        A[(size_t)row * n + (u32)indices[t]] = d64 ? d64[t] : (double)d32[t];
}

// A <- (1 - damping) A + damping thr(A_new) ; optional w_i w_j ; zero diagonal   (:1322-1329)
__global__ void __launch_bounds__(TPB) irmm_blend_kernel(const unsigned long long* __restrict__ acc, double inv_scale,
                                                         double min_weight, const double* __restrict__ node_w, u32 n, double damping,
                                                         double* __restrict__ A) {
    const size_t t = gtid();
    if (t >= (size_t)n * n) return;
    const u32 i = (u32)(t / n), j = (u32)(t % n);
    double an = (double)acc[t] * inv_scale;
    if (min_weight > 0.0 && an < min_weight) an = 0.0;
    double v = A[t] * (1.0 - damping) + an * damping;
    if (node_w) v *= (double)(float)(node_w[i] * node_w[j]);      // NW is a float32 matrix (hyperpartition.py:568)
    A[t] = (i == j) ? 0.0 : v;
}

__global__ void __launch_bounds__(TPB) dense_scale_pairs_kernel(double* __restrict__ A, u32 n, const i64* __restrict__ pr,
                                                                const i64* __restrict__ pc, u64 np, double factor) {
    const size_t t = gtid();
    if (t >= np) return;
    const i64 r = pr[t], c = pc[t];
    if (r < 0 || r >= (i64)n || c < 0 || c >= (i64)n) return;
    double& x = A[(size_t)r * n + (size_t)c];
    x = (factor == 0.0) ? 0.0 : x * factor;
}

// mode 0: A as is.  mode 1: the undirected graph igraph builds from triu(A) (:1354-1366), mirrored.
__device__ __forceinline__ double dense_view(const double* __restrict__ A, u32 n, u32 i, u32 j, int mode) {
    if (mode == 0) return A[(size_t)i * n + j];
    if (i == j) return 0.0;
    return i < j ? A[(size_t)i * n + j] : A[(size_t)j * n + i];
}

__global__ void __launch_bounds__(TPB) dense_row_count_kernel(const double* __restrict__ A, u32 n, int mode, u64* __restrict__ cnt) {
    const u32 row = (u32)(gtid() >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    u32 c = 0;
    for (u32 j0 = 0; j0 < n; j0 += 32) {
        const u32 j = j0 + lane;
        const bool k = j < n && dense_view(A, n, row, j, mode) != 0.0;
        c += __popc(__ballot_sync(0xffffffffu, k));
    }
    if (lane == 0) cnt[row] = c;
}

__global__ void __launch_bounds__(TPB) dense_row_emit_kernel(const double* __restrict__ A, u32 n, int mode, const u64* __restrict__ off,
                                                             i64* __restrict__ indptr, i32* __restrict__ indices,
                                                             float* __restrict__ d32, double* __restrict__ d64) {
    const u32 row = (u32)(gtid() >> 5), lane = threadIdx.x & 31;
    if (row > n) return;
    if (row == n) {
        if (lane == 0) indptr[n] = (i64)off[n];
        return;
    }
    u64 pos = off[row];
    if (lane == 0) indptr[row] = (i64)pos;
    const u32 lt = lanemask_lt();
    for (u32 j0 = 0; j0 < n; j0 += 32) {
        const u32 j = j0 + lane;
        const double v = j < n ? dense_view(A, n, row, j, mode) : 0.0;
        const u32 m = __ballot_sync(0xffffffffu, v != 0.0);
        if (v != 0.0) {
            const u64 p = pos + __popc(m & lt);
            indices[p] = (i32)j;
            if (d32) d32[p] = (float)v;
            if (d64) d64[p] = v;
        }
        pos += __popc(m);
    }
}

}  // namespace

extern "C" int cph_irmm_edge_weights(const u32* pem_row, const u32* pem_pk, u64 nnz, const i32* labels, u32 n_clusters, u8* mask,
                                     u8* mask_t, int alias, float* entry_w, u64* n_suppressed, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (nnz == 0) return CPH_OK;
    ARG_CHECK(pem_row && pem_pk && labels && mask && mask_t && entry_w && n_clusters > 0);
    irmm_edge_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(pem_row, pem_pk, nnz, labels, n_clusters, mask, mask_t, alias, entry_w,
                                                         (unsigned long long*)n_suppressed);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_expand_full(const u32* pem_row, const u32* pem_pk, const float* entry_w, const u8* mask, const u8* mask_t,
                               u64 nnz, u32 n, int frac_bits, u64* acc_full, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(acc_full && n > 0);
    ARG_CHECK(frac_bits >= 0 && frac_bits <= 52);
    CUDA_TRY(cudaMemsetAsync(acc_full, 0, (size_t)n * n * sizeof(u64), st));
    if (nnz == 0) return CPH_OK;
    ARG_CHECK(pem_row && pem_pk);
    size_t want = ceil_div(nnz, (size_t)TPB);
    unsigned grid = (unsigned)(want < 148 * 8 ? want : 148 * 8);
    expand_full_kernel<<<grid, TPB, 0, st>>>(pem_row, pem_pk, entry_w, mask, mask_t, nnz, n, frac_bits,
                                             (unsigned long long*)acc_full);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_dense_from_csr(const i64* indptr, const i32* indices, const float* data32, const double* data64, u32 n,
                                  double* dense, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(indptr && dense && n > 0 && (data32 || data64));
    CUDA_TRY(cudaMemsetAsync(dense, 0, (size_t)n * n * sizeof(double), st));
    dense_from_csr_kernel<<<grid_for((size_t)n * 32, TPB), TPB, 0, st>>>(indptr, indices, data32, data64, n, dense);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_irmm_blend(const u64* acc_full, int frac_bits, double min_weight, const double* node_w, u32 n, double damping,
                              double* dense, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(acc_full && dense && n > 0);
    irmm_blend_kernel<<<grid_for((size_t)n * n, TPB), TPB, 0, st>>>((const unsigned long long*)acc_full, ldexp(1.0, -frac_bits),
                                                                    min_weight, node_w, n, damping, dense);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_dense_scale_pairs(double* dense, u32 n, const i64* rows, const i64* cols, u64 n_pairs, double factor,
                                     void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n_pairs == 0) return CPH_OK;
    ARG_CHECK(dense && rows && cols && n > 0);
    dense_scale_pairs_kernel<<<grid_for(n_pairs, TPB), TPB, 0, st>>>(dense, n, rows, cols, n_pairs, factor);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_dense_to_csr_count(const double* dense, u32 n, int mode, u64* h_nnz, void* ws, size_t ws_bytes,
                                      size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(n > 0);
    WS_BEGIN(ar);
    u64* cnt = ar.take<u64>((size_t)n + 1);
    u64* total = ar.take<u64>(1);
    if (!ar.dry) {
        ARENA_CHECK(ar);
        ARG_CHECK(dense && h_nnz);
        CUDA_TRY(cudaMemsetAsync(cnt, 0, ((size_t)n + 1) * sizeof(u64), st));
        dense_row_count_kernel<<<grid_for((size_t)n * 32, TPB), TPB, 0, st>>>(dense, n, mode, cnt);
        KERNEL_CHECK();
    }
    TRY(scan_u64(cnt, (size_t)n + 1, total, ar, st));
    WS_QUERY_RETURN(ar);
    CUDA_TRY(cudaMemcpyAsync(h_nnz, total, sizeof(u64), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return CPH_OK;
}

extern "C" int cph_dense_to_csr_emit(const double* dense, u32 n, int mode, const void* ws, i64* indptr, i32* indices,
                                     float* data32, double* data64, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(dense && ws && indptr && indices && n > 0);
    dense_row_emit_kernel<<<grid_for(((size_t)n + 1) * 32, TPB), TPB, 0, st>>>(dense, n, mode, (const u64*)ws, indptr, indices,
                                                                              data32, data64);
    KERNEL_CHECK();
    return CPH_OK;
}
