// Stable LSD radix sort of 64-bit keys (optionally carrying 4- or 8-byte values), 8-bit digits.
//
// Per pass:  upsweep (per-tile digit counts)  ->  exclusive scan of the digit-major count table
//            ->  downsweep (stable in-tile ranking with warp match, shared-memory reorder so the
//                global stores leave the CTA as contiguous per-digit runs).
// No CTA waits on another CTA (no decoupled look-back), so a bug cannot hang the device.
// Traffic per pass: 2 key reads + 1 key write (+ value read/write) + ~3 % for the count table.
//
// Replaces the polars sort / group_by engine underneath cphasing/hypergraph.py:1470-1477, 1877
// and the argsort of cphasing/algorithms/hypergraph.py:269-275.
#include "common.cuh"

namespace {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;   // 4096 keys per CTA
constexpr int RADIX = 256;

// Tile layout: warp w owns keys [w*512, (w+1)*512) of the tile; item j of lane l is key
// w*512 + j*32 + l, so loads are coalesced and the (warp, item, lane) order is the key order.

__device__ __forceinline__ void warp_rank_items(const u64 (&key)[RS_ITEMS], const bool (&valid)[RS_ITEMS], int shift, u32 mask,
                                                u32* __restrict__ wh /* this warp's 256 counters */,
                                                u32 (&rank)[RS_ITEMS]) {
    const u32 lane = lane_id();
    const u32 lt = lanemask_lt();
#pragma unroll
    for (int j = 0; j < RS_ITEMS; ++j) {
        // invalid (past-the-end) items match among themselves under a key no digit can take
        u32 d = valid[j] ? ((u32)(key[j] >> shift) & mask) : 0xffffffffu;
        u32 peers = __match_any_sync(0xffffffffu, d);
        u32 leader = __ffs(peers) - 1;
        u32 prev = 0;
        if (lane == leader && valid[j]) {
            prev = wh[d];
            wh[d] = prev + __popc(peers);
        }
        prev = __shfl_sync(0xffffffffu, prev, leader);
        rank[j] = prev + __popc(peers & lt);
        __syncwarp();
    }
}

__global__ void __launch_bounds__(RS_THREADS) rs_upsweep_kernel(const u64* __restrict__ keys, u32 n, int shift, u32 mask,
                                                                  u32* __restrict__ table, u32 num_tiles) {
    __shared__ u32 wh[RS_WARPS][RADIX];
    const u32 warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < RS_WARPS * RADIX; i += RS_THREADS) (&wh[0][0])[i] = 0;
    __syncthreads();
    const u32 tile = blockIdx.x;
    const size_t base = (size_t)tile * RS_TILE + warp * (RS_ITEMS * 32) + lane;
    u64 key[RS_ITEMS];
    bool valid[RS_ITEMS];
#pragma unroll
    for (int j = 0; j < RS_ITEMS; ++j) {
        size_t i = base + (size_t)j * 32;
        valid[j] = i < n;
        key[j] = valid[j] ? keys[i] : 0;
    }
    u32 rank[RS_ITEMS];
    warp_rank_items(key, valid, shift, mask, wh[warp], rank);
    __syncthreads();
    {
        u32 d = threadIdx.x;   // RS_THREADS == RADIX
        u32 s = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) s += wh[w][d];
        table[(size_t)d * num_tiles + tile] = s;
    }
}

template <typename V>
struct ValIO {
    static __device__ __forceinline__ V load(const V* p, size_t i) { return p[i]; }
    static __device__ __forceinline__ void store(V* p, size_t i, V v) { p[i] = v; }
};
struct NoVal {};
template <>
struct ValIO<NoVal> {
    static __device__ __forceinline__ NoVal load(const NoVal*, size_t) { return NoVal(); }
    static __device__ __forceinline__ void store(NoVal*, size_t, NoVal) {}
};

template <typename V>
__global__ void __launch_bounds__(RS_THREADS) rs_downsweep_kernel(const u64* __restrict__ keys_in, u64* __restrict__ keys_out,
                                                                    const V* __restrict__ vals_in, V* __restrict__ vals_out, u32 n,
                                                                    int shift, u32 mask, const u32* __restrict__ table,
                                                                    u32 num_tiles) {
    __shared__ u32 wh[RS_WARPS][RADIX];
    __shared__ u32 digit_base[RADIX];
    __shared__ u32 gbase[RADIX];
    __shared__ u32 scan_sm[RS_THREADS / 32 + 1];
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    u64* skeys = (u64*)dyn_smem;                        // RS_TILE keys
    V* svals = (V*)(dyn_smem + RS_TILE * sizeof(u64));  // RS_TILE values (absent for NoVal)

    const u32 warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < RS_WARPS * RADIX; i += RS_THREADS) (&wh[0][0])[i] = 0;
    __syncthreads();
    const u32 tile = blockIdx.x;
    const size_t tile_base = (size_t)tile * RS_TILE;
    const u32 tile_n = (u32)min((size_t)RS_TILE, (size_t)n - tile_base);
    const size_t base = tile_base + warp * (RS_ITEMS * 32) + lane;

    u64 key[RS_ITEMS];
    V val[RS_ITEMS];
    bool valid[RS_ITEMS];
#pragma unroll
    for (int j = 0; j < RS_ITEMS; ++j) {
        size_t i = base + (size_t)j * 32;
        valid[j] = i < n;
        key[j] = valid[j] ? keys_in[i] : ~0ull;
        if (valid[j]) val[j] = ValIO<V>::load(vals_in, i);
    }
    u32 rank[RS_ITEMS];
    warp_rank_items(key, valid, shift, mask, wh[warp], rank);
    __syncthreads();
    {
        // per-digit exclusive prefix over warps, then over digits
        u32 d = threadIdx.x;
        u32 run = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) {
            u32 c = wh[w][d];
            wh[w][d] = run;
            run += c;
        }
        u32 tot;
        u32 ex = block_excl_scan<u32, RS_THREADS>(run, scan_sm, &tot);
        digit_base[d] = ex;
        gbase[d] = table[(size_t)d * num_tiles + tile] - ex;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < RS_ITEMS; ++j) {
        if (valid[j]) {
            u32 dj = (u32)(key[j] >> shift) & mask;
            u32 p = digit_base[dj] + wh[warp][dj] + rank[j];
            skeys[p] = key[j];
            if (sizeof(V) > 1) svals[p] = val[j];
        }
    }
    __syncthreads();
    for (u32 t = threadIdx.x; t < tile_n; t += RS_THREADS) {
        u64 k = skeys[t];
        u32 d = (u32)(k >> shift) & mask;
        u32 dst = gbase[d] + t;   // mod 2^32: gbase holds (global start - local start)
        keys_out[dst] = k;
        if (sizeof(V) > 1) ValIO<V>::store(vals_out, dst, svals[t]);
    }
}

template <typename V>
int launch_down(const u64* kin, u64* kout, const void* vin, void* vout, u32 n, int shift, u32 mask, const u32* table,
                u32 tiles, cudaStream_t st) {
    const size_t smem = RS_TILE * sizeof(u64) + (sizeof(V) > 1 ? RS_TILE * sizeof(V) : 0);
    CUDA_TRY(cudaFuncSetAttribute(rs_downsweep_kernel<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    rs_downsweep_kernel<V><<<tiles, RS_THREADS, smem, st>>>(kin, kout, (const V*)vin, (V*)vout, n, shift, mask, table, tiles);
    KERNEL_CHECK();
    return CPH_OK;
}

}  // namespace

int radix_sort_u64(u64* ka, u64* kb, void* va, void* vb, int val_bytes, size_t n, int bit_lo, int bit_hi, int* result_in_b,
                   Arena& ar, cudaStream_t st) {
    ARG_CHECK(val_bytes == 0 || val_bytes == 4 || val_bytes == 8);
    ARG_CHECK(bit_lo >= 0 && bit_hi <= 64 && bit_lo <= bit_hi);
    if (n >= ((size_t)1 << 32) - RS_TILE) {
        cph_set_error(__FILE__, __LINE__, "radix_sort_u64: n must be < 2^32 per call");
        return CPH_ERR_TOO_LARGE;
    }
    const u32 tiles = (u32)ceil_div(n > 0 ? n : 1, RS_TILE);
    u32* table = ar.take<u32>((size_t)RADIX * tiles);
    const size_t scan_off = ar.off;   // the scan scratch is re-used by every pass
    {
        Arena probe(nullptr, 0);      // dry run: how much does the scan carve?
        scan_u32(nullptr, (size_t)RADIX * tiles, nullptr, probe, st);
        (void)ar.take<unsigned char>(probe.off);
    }
    if (ar.dry) {
        *result_in_b = 0;
        return CPH_OK;
    }
    ARENA_CHECK(ar);
    int in_b = 0;
    if (n > 0) {
        for (int lo = bit_lo; lo < bit_hi; lo += 8) {
            int bits = bit_hi - lo < 8 ? bit_hi - lo : 8;
            u32 mask = (1u << bits) - 1u;
            u64* kin = in_b ? kb : ka;
            u64* kout = in_b ? ka : kb;
            void* vin = in_b ? vb : va;
            void* vout = in_b ? va : vb;
            rs_upsweep_kernel<<<tiles, RS_THREADS, 0, st>>>(kin, (u32)n, lo, mask, table, tiles);
            KERNEL_CHECK();
            Arena scratch = ar;
            scratch.off = scan_off;
            TRY(scan_u32(table, (size_t)RADIX * tiles, nullptr, scratch, st));
            if (val_bytes == 0) TRY(launch_down<NoVal>(kin, kout, vin, vout, (u32)n, lo, mask, table, tiles, st));
            else if (val_bytes == 4) TRY(launch_down<u32>(kin, kout, vin, vout, (u32)n, lo, mask, table, tiles, st));
            else TRY(launch_down<u64>(kin, kout, vin, vout, (u32)n, lo, mask, table, tiles, st));
            in_b ^= 1;
        }
    }
    *result_in_b = in_b;
    return CPH_OK;
}
