// Device-wide exclusive scan (reduce -> spine -> apply).  No inter-CTA spinning: three launches,
// each one bandwidth bound; used on tile-count tables, so its traffic is a few % of the data pass.
#include "common.cuh"

namespace {

constexpr int SC_THREADS = 256;
constexpr int SC_ITEMS = 16;
constexpr int SC_TILE = SC_THREADS * SC_ITEMS;

template <typename T>
__global__ void __launch_bounds__(SC_THREADS) scan_reduce_kernel(const T* __restrict__ data, size_t n, T* __restrict__ block_sums) {
    __shared__ T sm[SC_THREADS / 32];
    size_t base = (size_t)blockIdx.x * SC_TILE;
    T s = 0;
#pragma unroll
    for (int j = 0; j < SC_ITEMS; ++j) {
        size_t i = base + (size_t)j * SC_THREADS + threadIdx.x;
        if (i < n) s += data[i];
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        T v = (threadIdx.x < SC_THREADS / 32) ? sm[threadIdx.x] : (T)0;
        v = warp_sum(v);
        if (threadIdx.x == 0) block_sums[blockIdx.x] = v;
    }
}

// single block: exclusive scan of nb values in place, total to *total_out (optional)
template <typename T>
__global__ void __launch_bounds__(1024) scan_spine_kernel(T* __restrict__ sums, size_t nb, T* __restrict__ total_out) {
    __shared__ T sm[1024 / 32 + 1];
    T carry = 0;
    for (size_t start = 0; start < nb; start += 1024) {
        size_t i = start + threadIdx.x;
        T v = (i < nb) ? sums[i] : (T)0;
        T tot;
        T ex = block_excl_scan<T, 1024>(v, sm, &tot);
        if (i < nb) sums[i] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry;
}

template <typename T>
__global__ void __launch_bounds__(SC_THREADS) scan_apply_kernel(T* __restrict__ data, size_t n, const T* __restrict__ block_offsets) {
    __shared__ T sm[SC_THREADS / 32 + 1];
    size_t base = (size_t)blockIdx.x * SC_TILE;
    T carry = block_offsets[blockIdx.x];
#pragma unroll 1
    for (int j = 0; j < SC_ITEMS; ++j) {
        size_t i = base + (size_t)j * SC_THREADS + threadIdx.x;
        T v = (i < n) ? data[i] : (T)0;
        T tot;
        T ex = block_excl_scan<T, SC_THREADS>(v, sm, &tot);
        if (i < n) data[i] = carry + ex;
        carry += tot;
    }
}

template <typename T>
int scan_impl(T* data, size_t n, T* total_out, Arena& ar, cudaStream_t st) {
    size_t nb = ceil_div(n > 0 ? n : 1, SC_TILE);
    T* sums = ar.take<T>(nb);
    if (ar.dry) return CPH_OK;
    ARENA_CHECK(ar);
    if (n == 0) {
        if (total_out) CUDA_TRY(cudaMemsetAsync(total_out, 0, sizeof(T), st));
        return CPH_OK;
    }
    scan_reduce_kernel<T><<<(unsigned)nb, SC_THREADS, 0, st>>>(data, n, sums);
    KERNEL_CHECK();
    scan_spine_kernel<T><<<1, 1024, 0, st>>>(sums, nb, total_out);
    KERNEL_CHECK();
    scan_apply_kernel<T><<<(unsigned)nb, SC_THREADS, 0, st>>>(data, n, sums);
    KERNEL_CHECK();
    return CPH_OK;
}

}  // namespace

int scan_u32(u32* data, size_t n, u32* total_out, Arena& ar, cudaStream_t st) { return scan_impl<u32>(data, n, total_out, ar, st); }
int scan_u64(u64* data, size_t n, u64* total_out, Arena& ar, cudaStream_t st) { return scan_impl<u64>(data, n, total_out, ar, st); }
