// Group-level affinity sums on CSR(A): the heavy products of the partition driver's post-processing.
//
// Replaces (paths under /root/reference/cphasing/):
//   hyperpartition.py:2402-2410   _merge:                value_matrix = M^T A M   (M = n x g group indicator)
//   hyperpartition.py:2889-2892   recluster_by_linkage:  S = A M                  (contig x group linkage)
// Both are segmented sums of A's stored entries keyed by the group of the column (and of the row).  To make them
// independent of thread order they are accumulated as 64-bit fixed point (scale 2^s from nnz * max|a|, like the
// clustering kernels) with integer REDs and converted to double at the end.  One warp per row: the row's entries
// are read once, coalesced; S's row is private to the warp, V's g x g block lives in L2.
#include "common.cuh"
#include "cph.h"
#include <math.h>
#include <string.h>

namespace {

typedef unsigned long long ull;
constexpr int GO_THREADS = 256;
constexpr int GO_WPB = GO_THREADS / 32;

__global__ void __launch_bounds__(GO_THREADS) go_absmax_kernel(const i64* __restrict__ indptr, const float* __restrict__ d32,
                                                               const double* __restrict__ d64, u32 n, ull* __restrict__ out_bits) {
    u32 v = blockIdx.x * GO_WPB + (threadIdx.x >> 5);
    if (v >= n) return;
    double mx = 0;
    for (i64 t = indptr[v] + lane_id(); t < indptr[v + 1]; t += 32) {
        double x = fabs(d64 ? d64[t] : (double)d32[t]);
        if (x == x && !isinf(x) && x > mx) mx = x;
    }
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane_id() == 0 && mx > 0) atomicMax(out_bits, (ull)__double_as_longlong(mx));
}

// S_fx[i, lab[j]] += fx(A_ij) for every stored entry with lab[j] >= 0
__global__ void __launch_bounds__(GO_THREADS) go_rows_kernel(const i64* __restrict__ indptr, const i32* __restrict__ indices,
                                                             const float* __restrict__ d32, const double* __restrict__ d64, u32 n,
                                                             const i32* __restrict__ lab, u32 g, double scale,
                                                             i64* __restrict__ S_fx) {
    u32 v = blockIdx.x * GO_WPB + (threadIdx.x >> 5);
    if (v >= n) return;
    i64* row = S_fx + (size_t)v * g;
    for (i64 t = indptr[v] + lane_id(); t < indptr[v + 1]; t += 32) {
        int c = lab[indices[t]];
        if (c < 0) continue;
        double x = d64 ? d64[t] : (double)d32[t];
        if (x != x || isinf(x)) continue;
        atomicAdd((ull*)&row[c], (ull)llrint(x * scale));
    }
}

// V_fx[lab[i], :] += S_fx[i, :]
__global__ void go_groups_kernel(const i64* __restrict__ S_fx, const i32* __restrict__ lab, u32 n, u32 g, i64* __restrict__ V_fx) {
    size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)n * g) return;
    u32 i = (u32)(idx / g), c = (u32)(idx % g);
    int gi = lab[i];
    i64 x = S_fx[idx];
    if (gi >= 0 && x != 0) atomicAdd((ull*)&V_fx[(size_t)gi * g + c], (ull)x);
}

__global__ void go_to_double_kernel(const i64* __restrict__ fx, size_t count, double inv_scale, double* __restrict__ out) {
    size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < count) out[idx] = (double)fx[idx] * inv_scale;
}

}  // namespace

extern "C" int cph_group_sums(const int64_t* indptr, const int32_t* indices, const float* data32, const double* data64,
                              uint32_t n, uint64_t nnz, const int32_t* labels, uint32_t n_groups, double* S, double* V, void* ws,
                              size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream) {
    ARG_CHECK(n > 0 && n_groups > 0);
    cudaStream_t st = (cudaStream_t)stream;
    WS_BEGIN(ar);
    ull* bits = ar.take<ull>(1);
    i64* S_fx = ar.take<i64>((size_t)n * n_groups);
    i64* V_fx = ar.take<i64>((size_t)n_groups * n_groups);
    WS_QUERY_RETURN(ar);
    ARG_CHECK(indptr && labels && (S || V) && (nnz == 0 || (indices && (data32 || data64))));
    CUDA_TRY(cudaMemsetAsync(bits, 0, sizeof(ull), st));
    CUDA_TRY(cudaMemsetAsync(S_fx, 0, sizeof(i64) * (size_t)n * n_groups, st));
    CUDA_TRY(cudaMemsetAsync(V_fx, 0, sizeof(i64) * (size_t)n_groups * n_groups, st));
    double scale = 1.0;
    if (nnz > 0) {
        go_absmax_kernel<<<grid_for(n, GO_WPB), GO_THREADS, 0, st>>>(indptr, data32, data64, n, bits);
        KERNEL_CHECK();
        ull hb = 0;
        CUDA_TRY(cudaMemcpyAsync(&hb, bits, sizeof(ull), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        double mx;
        memcpy(&mx, &hb, sizeof(double));
        if (mx > 0) {
            int e;
            frexp(mx * (double)nnz, &e);
            int s = 60 - e;
            if (s > 52) s = 52;
            scale = ldexp(1.0, s);
            go_rows_kernel<<<grid_for(n, GO_WPB), GO_THREADS, 0, st>>>(indptr, indices, data32, data64, n, labels, n_groups, scale,
                                                                      S_fx);
            KERNEL_CHECK();
        }
    }
    if (V) {
        go_groups_kernel<<<grid_for((size_t)n * n_groups, 256), 256, 0, st>>>(S_fx, labels, n, n_groups, V_fx);
        KERNEL_CHECK();
        go_to_double_kernel<<<grid_for((size_t)n_groups * n_groups, 256), 256, 0, st>>>(V_fx, (size_t)n_groups * n_groups,
                                                                                       1.0 / scale, V);
        KERNEL_CHECK();
    }
    if (S) {
        go_to_double_kernel<<<grid_for((size_t)n * n_groups, 256), 256, 0, st>>>(S_fx, (size_t)n * n_groups, 1.0 / scale, S);
        KERNEL_CHECK();
    }
    return CPH_OK;
}
