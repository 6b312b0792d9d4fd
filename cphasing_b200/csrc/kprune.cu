// Stage 3: kprune -- allelic / cross-allelic ("weak") contig-pair classification read straight off A.
//
// Replaces (paths under /root/reference/cphasing/):
//   kprune.py:407-517  KPruneHyperGraph: allelic = allele-table pairs (i<j) that are stored entries of
//                      triu(A) (:455-468, their UPPER entry is zeroed); weak = every other stored (c1<c2)
//                      whose contigs both have an allele group, scored on the block [c1,*al(c1)] x [c2,*al(c2)]
//                      of A and called weak unless the maximum-weight matching pairs c1 with c2 (:470-517;
//                      one igraph call of ~0.3 ms per pair there).
//   kprune.py:269-335  KPruner: the same decision on general l1 x l2 blocks.
// (The production `cphasing-rs kprune` binary is not in the reference tree -- SURVEY.md 2.2.)
// Here: one thread per stored upper entry, binary-search gathers of at most 6 x 6 scores (L1/L2
// resident: A is at most a few hundred MB), exact max-weight assignment by subset DP in registers.
// Tie rule: a tie between the best matching through (c1,c2) and the best overall counts as strong.
// Algorithmic bytes: 12 B per gathered A entry + 17 B per candidate pair.
#include "common.cuh"

namespace {

constexpr int TPB = 128;
constexpr int KP_MAXL = 6;     // contig + at most 5 alleles (hexaploid)

__device__ __forceinline__ size_t gtid() { return (size_t)blockIdx.x * blockDim.x + threadIdx.x; }

__device__ __forceinline__ i64 find_entry(const i64* __restrict__ indptr, const i32* __restrict__ indices, i64 r, i32 c) {
    i64 lo = indptr[r], hi = indptr[r + 1];
    const i64 end = hi;
    while (lo < hi) {
        i64 mid = (lo + hi) >> 1;
        if (indices[mid] < c) lo = mid + 1;
        else hi = mid;
    }
    return (lo < end && indices[lo] == c) ? lo : -1;
}

struct KpIn {
    const i64* indptr;
    const i32* indices;
    const float* d32;
    const double* d64;
    u32 n;
    const i64* ag_ptr;     // [n+1] allele lists per contig (CSR-like); empty list = contig has no allele group
    const i32* ag_idx;
    int max_alleles;       // use at most this many alleles per contig (1 = KPruneHyperGraph's "first listed allele")
};

// flag[t] = 1 for every stored upper entry that is an allele pair (kprune.py:455-468)
__global__ void __launch_bounds__(TPB) kp_mark_allelic_kernel(KpIn in, const i64* __restrict__ al_r, const i64* __restrict__ al_c,
                                                              u64 n_al, u8* __restrict__ flag) {
    size_t t = gtid();
    if (t >= n_al) return;
    i64 a = al_r[t], b = al_c[t];
    if (a < 0 || b < 0 || a >= (i64)in.n || b >= (i64)in.n || a >= b) return;     // only rows with contig1 < contig2
    i64 o = find_entry(in.indptr, in.indices, a, (i32)b);
    if (o >= 0) flag[o] = 1;
}

__device__ __forceinline__ double kp_score(const KpIn& in, const u8* __restrict__ flag, i32 i, i32 j) {
    if (i == j) return 0.0;                                   // setdiag(0), kprune.py:419
    i64 o = find_entry(in.indptr, in.indices, i, j);
    if (o < 0) return 0.0;
    if (i < j && flag[o] == 1) return 0.0;                    // zeroed allelic (upper entry only, :466)
    return in.d64 ? in.d64[o] : (double)in.d32[o];
}

// warp-free: one thread per stored entry of A
__global__ void __launch_bounds__(TPB) kp_weak_kernel(KpIn in, u64 nnz, const i32* __restrict__ row_of, u8* __restrict__ flag,
                                                      double* __restrict__ best_with, double* __restrict__ best_all) {
    size_t t = gtid();
    if (t >= nnz) return;
    const i32 c1 = row_of[t], c2 = in.indices[t];
    if (c1 >= c2 || flag[t] == 1) return;                     // upper triangle, not allelic
    const double v12 = in.d64 ? in.d64[t] : (double)in.d32[t];
    if (v12 == 0.0) return;                                   // explicit zeros are not contacts
    i64 p1 = in.ag_ptr[c1], p2 = in.ag_ptr[c2];
    int l1 = (int)min((i64)in.max_alleles, in.ag_ptr[c1 + 1] - p1), l2 = (int)min((i64)in.max_alleles, in.ag_ptr[c2 + 1] - p2);
    if (l1 <= 0 || l2 <= 0) return;                           // KeyError branch: a contig without allele group
    l1 = min(l1, KP_MAXL - 1);
    l2 = min(l2, KP_MAXL - 1);
    i32 g1[KP_MAXL], g2[KP_MAXL];
    g1[0] = c1;
    g2[0] = c2;
    for (int k = 0; k < l1; ++k) g1[k + 1] = in.ag_idx[p1 + k];
    for (int k = 0; k < l2; ++k) g2[k + 1] = in.ag_idx[p2 + k];
    ++l1;
    ++l2;
    // rows = the shorter side (kprune.py:277-283); index 0 of both sides is (c1, c2) either way
    const bool tr = l1 > l2;
    const int R = tr ? l2 : l1, Cn = tr ? l1 : l2;
    double S[KP_MAXL][KP_MAXL];
    for (int a = 0; a < R; ++a)
        for (int b = 0; b < Cn; ++b) {
            i32 x = tr ? g1[b] : g1[a], y = tr ? g2[a] : g2[b];
            S[a][b] = (x == c1 && y == c2) ? v12 : kp_score(in, flag, x, y);
        }
    // exact assignment by DP over column subsets: dp[mask] = best sum with the first popcount(mask) rows on `mask`
    double dp[1 << KP_MAXL], dq[1 << KP_MAXL];                // dq: row 0 forced onto column 0
    const int full = 1 << Cn;
    const double NEG = -1.0e300;
    for (int m = 0; m < full; ++m) {
        dp[m] = NEG;
        dq[m] = NEG;
    }
    dp[0] = 0.0;
    dq[0] = 0.0;
    double ball = NEG, bwith = NEG;
    for (int m = 0; m < full; ++m) {
        const int r = __popc(m);
        if (r >= R) {
            if (r == R) {
                ball = fmax(ball, dp[m]);
                bwith = fmax(bwith, dq[m]);
            }
            continue;
        }
        for (int b = 0; b < Cn; ++b) {
            if (m & (1 << b)) continue;
            const int m2 = m | (1 << b);
            if (dp[m] > NEG) dp[m2] = fmax(dp[m2], dp[m] + S[r][b]);
            if (dq[m] > NEG && (r != 0 || b == 0)) dq[m2] = fmax(dq[m2], dq[m] + S[r][b]);
        }
    }
    if (best_with) best_with[t] = bwith;
    if (best_all) best_all[t] = ball;
    if (ball > bwith + 1e-12) flag[t] = 2;                    // c1 is not matched to c2 -> weak
}

__global__ void __launch_bounds__(256) row_of_kernel(const i64* __restrict__ indptr, u32 n, i32* __restrict__ row_of) {
    const u32 row = (u32)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    for (i64 t = indptr[row] + lane, e = indptr[row + 1]; t < e; t += 32) row_of[t] = (i32)row;
}

}  // namespace

extern "C" int cph_csr_row_ids(const i64* indptr, u32 n, i32* row_of, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(indptr && row_of && n > 0);
    row_of_kernel<<<grid_for((size_t)n * 32, 256), 256, 0, st>>>(indptr, n, row_of);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_kprune(const i64* indptr, const i32* indices, const float* data32, const double* data64, u32 n, u64 nnz,
                          const i32* row_of, const i64* ag_ptr, const i32* ag_idx, int max_alleles, const i64* al_r, const i64* al_c,
                          u64 n_al, u8* flag, double* best_with, double* best_all, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(n > 0 && flag);
    ARG_CHECK(max_alleles >= 1 && max_alleles <= KP_MAXL - 1);
    if (nnz == 0) return CPH_OK;
    ARG_CHECK(indptr && indices && (data32 || data64) && row_of && ag_ptr);
    CUDA_TRY(cudaMemsetAsync(flag, 0, nnz, st));
    KpIn in{indptr, indices, data32, data64, n, ag_ptr, ag_idx, max_alleles};
    if (n_al) {
        ARG_CHECK(al_r && al_c);
        kp_mark_allelic_kernel<<<grid_for(n_al, TPB), TPB, 0, st>>>(in, al_r, al_c, n_al, flag);
        KERNEL_CHECK();
    }
    kp_weak_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(in, nnz, row_of, flag, best_with, best_all);
    KERNEL_CHECK();
    return CPH_OK;
}
