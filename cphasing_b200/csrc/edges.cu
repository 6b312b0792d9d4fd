// Stage 1: contact tables -> hyperedge triplets in raw edge-major order (the CSC view of H), and
// the conversions between that order and the reference's HyperEdges wire order (sorted by row).
//
// Replaces (paths under /root/reference/cphasing/):
//   hypergraph.py:970-1044, 1287-1292   Extractor: 4DN pairs -> two triplets per pair
//   hypergraph.py:1395-1481, 1774-1775  process_pore_c_table: filters + group_by(read, contig)
//   hypergraph.py:1871-1889             generate_edges: read-id offset, sort by contig, casts
// The polars hash group-by becomes: records keyed (read | contig | mapq); when the table is not
// already grouped by read a stable radix sort on the read bits; then two run-local kernels (each
// record inspects only the records of its own read, O(run length), L1-resident) that pick one
// representative per (read, contig), count its records, take the max MAPQ and the distinct-contig
// order of the read; a scan compacts the survivors.  Bytes per record (sorted input): 9 read,
// 8+8 key write/read x2, 10 flags -> HBM bound.
#include "common.cuh"

namespace {

constexpr int TPB = 256;

__device__ __forceinline__ size_t gtid() { return (size_t)blockIdx.x * blockDim.x + threadIdx.x; }

// ------------------------------------------------------------------------------------------------
// Hi-C pairs
// ------------------------------------------------------------------------------------------------
struct PairsIn {
    const u32* c1;
    const u32* c2;
    const int8_t* mapq;
    const u32* p1;
    const u32* p2;
    const i64* len;
    u64 n;
    u32 n_contigs;
    int min_mapq;
    i64 edge_length;
};

__device__ __forceinline__ bool pair_ok(const PairsIn& in, size_t p, u32& a, u32& b, int8_t& q) {
    a = in.c1[p];
    b = in.c2[p];
    q = in.mapq ? in.mapq[p] : (int8_t)0;
    bool ok = a < in.n_contigs && b < in.n_contigs;
    if (in.min_mapq > 0 && in.mapq) ok = ok && (int)q >= in.min_mapq;          // hypergraph.py:970-971
    if (ok && in.edge_length > 0) {                                             // :977-992
        i64 x1 = in.p1[p], x2 = in.p2[p];
        i64 l1 = in.len[a], l2 = in.len[b];
        ok = ((x1 < in.edge_length) || (x1 > l1 - in.edge_length)) && ((x2 < in.edge_length) || (x2 > l2 - in.edge_length));
    }
    return ok;
}

__global__ void __launch_bounds__(TPB) pairs_flag_kernel(PairsIn in, u32* __restrict__ keep) {
    size_t p = gtid();
    if (p >= in.n) return;
    u32 a, b;
    int8_t q;
    keep[p] = pair_ok(in, p, a, b, q) ? 1u : 0u;
}

__global__ void __launch_bounds__(TPB) pairs_emit_kernel(PairsIn in, const u32* __restrict__ pos, u32* __restrict__ rem_row,
                                                         u32* __restrict__ rem_col, u32* __restrict__ rem_count,
                                                         int8_t* __restrict__ rem_mapq) {
    size_t p = gtid();
    if (p >= in.n) return;
    u32 a, b;
    int8_t q;
    if (!pair_ok(in, p, a, b, q)) return;
    u32 col = pos[p];                       // :1031 with_row_count over the surviving pairs
    size_t t = (size_t)col * 2;
    *reinterpret_cast<uint2*>(rem_row + t) = make_uint2(a | CPH_HEAD, b);
    *reinterpret_cast<uint2*>(rem_col + t) = make_uint2(col, col);
    *reinterpret_cast<uint2*>(rem_count + t) = make_uint2(1u, 1u);   // :1287-1292 count = ones
    *reinterpret_cast<char2*>(rem_mapq + t) = make_char2(q, q);
}

// ------------------------------------------------------------------------------------------------
// Pore-C records
// key: [63] invalid | [8+cb, 8+cb+rb) read | [8, 8+cb) contig | [0,8) mapq+128
// ------------------------------------------------------------------------------------------------
struct PorecIn {
    const u32* read;
    const u32* contig;
    const int8_t* mapq;
    const u32* start;
    const u32* end;
    const i64* len;
    u64 n;
    u32 n_contigs;
    int min_mapq;
    i64 edge_length;
};

__global__ void __launch_bounds__(TPB) max_u32_kernel(const u32* __restrict__ x, u64 n, u32* __restrict__ out) {
    u32 m = 0;
    for (size_t i = gtid(); i < n; i += (size_t)gridDim.x * blockDim.x) m = max(m, x[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

__global__ void __launch_bounds__(TPB) porec_key_kernel(PorecIn in, int cb, u64* __restrict__ keys, u32* __restrict__ unsorted_flag) {
    size_t t = gtid();
    if (t >= in.n) return;
    u32 r = in.read[t], c = in.contig[t];
    int8_t q = in.mapq[t];
    bool ok = c < in.n_contigs;
    if (in.min_mapq > 0) ok = ok && (int)q >= in.min_mapq;                      // hypergraph.py:1774-1775
    if (ok && in.edge_length > 0) {                                             // :1449-1456
        i64 st = in.start[t], en = in.end[t];
        i64 pos = st + (en - st) / 2;
        i64 L = in.len[c];
        ok = (pos < in.edge_length) || ((L - pos) < in.edge_length);
    }
    u64 k = ((u64)r << (cb + 8)) | ((u64)(ok ? c : 0u) << 8) | (u64)(u8)((int)q + 128);
    if (!ok) k |= (1ull << 63);
    keys[t] = k;
    if (t > 0 && in.read[t - 1] > r) *unsorted_flag = 1u;   // benign race: everyone writes 1
}

struct KeyView {
    const u64* k;
    int cb;
    u64 n;
    __device__ __forceinline__ bool valid(u64 x) const { return (x >> 63) == 0; }
    __device__ __forceinline__ u64 read(u64 x) const { return (x & ~(1ull << 63)) >> (cb + 8); }
    __device__ __forceinline__ u32 contig(u64 x) const { return (u32)(x >> 8) & ((1u << cb) - 1u); }
    __device__ __forceinline__ int mapq(u64 x) const { return (int)(x & 0xff) - 128; }
};

// rep[t] = 1 iff record t is valid and is the first valid record of its (read, contig)
__global__ void __launch_bounds__(TPB) porec_rep_kernel(KeyView kv, u8* __restrict__ rep) {
    size_t t = gtid();
    if (t >= kv.n) return;
    u64 x = kv.k[t];
    u8 r = 0;
    if (kv.valid(x)) {
        r = 1;
        const u64 rd = kv.read(x);
        const u32 c = kv.contig(x);
        for (size_t u = t; u-- > 0;) {
            u64 y = kv.k[u];
            if (kv.read(y) != rd) break;
            if (kv.valid(y) && kv.contig(y) == c) {
                r = 0;
                break;
            }
        }
    }
    rep[t] = r;
}

// For every representative: count, max mapq, distinct-contig order of its read, head flag.
__global__ void __launch_bounds__(TPB) porec_reduce_kernel(KeyView kv, const u8* __restrict__ rep, u32 min_order, u32 max_order,
                                                           u32* __restrict__ keep, u32* __restrict__ aux, int8_t* __restrict__ mq) {
    size_t t = gtid();
    if (t > kv.n) return;
    if (t == kv.n) {        // sentinel so that pos[t+1] exists after the exclusive scan
        keep[t] = 0;
        return;
    }
    u32 kept = 0;
    if (rep[t]) {
        u64 x = kv.k[t];
        const u64 rd = kv.read(x);
        const u32 c = kv.contig(x);
        u32 cnt = 1, order = 1, before = 0;
        int q = kv.mapq(x);
        for (size_t u = t + 1; u < kv.n; ++u) {
            u64 y = kv.k[u];
            if (kv.read(y) != rd) break;
            order += rep[u];
            if (kv.valid(y) && kv.contig(y) == c) {     // hypergraph.py:1470-1477
                ++cnt;
                q = max(q, kv.mapq(y));
            }
        }
        for (size_t u = t; u-- > 0;) {
            if (kv.read(kv.k[u]) != rd) break;
            before += rep[u];
        }
        order += before;
        if (order >= min_order && order <= max_order) {   // :1461-1468
            kept = 1;
            aux[t] = cnt | (before == 0 ? CPH_HEAD : 0u);
            mq[t] = (int8_t)q;
        }
    }
    keep[t] = kept;
}

__global__ void __launch_bounds__(TPB) porec_emit_kernel(KeyView kv, const u32* __restrict__ pos, const u8* __restrict__ rep,
                                                         const u32* __restrict__ aux, const int8_t* __restrict__ mq,
                                                         u32 read_offset, u32* __restrict__ rem_row, u32* __restrict__ rem_col,
                                                         u32* __restrict__ rem_count, int8_t* __restrict__ rem_mapq) {
    size_t t = gtid();
    if (t >= kv.n || !rep[t]) return;
    u32 p = pos[t];
    if (pos[t + 1] == p) return;   // not kept (pos has n+1 entries)
    u64 x = kv.k[t];
    u32 a = aux[t];
    rem_row[p] = kv.contig(x) | (a & CPH_HEAD);
    rem_col[p] = (u32)kv.read(x) + read_offset;          // generate_edges :1871-1875
    rem_count[p] = a & CPH_ROWM;
    rem_mapq[p] = mq[t];
}

// ------------------------------------------------------------------------------------------------
// wire order <-> edge-major
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB) e2t_key_kernel(const u32* __restrict__ rem_row, const u32* __restrict__ rem_col,
                                                      const u32* __restrict__ rem_count, const int8_t* __restrict__ rem_mapq, u64 n,
                                                      u64* __restrict__ keys, u64* __restrict__ vals) {
    size_t t = gtid();
    if (t >= n) return;
    keys[t] = ((u64)(rem_row[t] & CPH_ROWM) << 32) | rem_col[t];
    vals[t] = (u64)rem_count[t] | ((u64)(u8)(rem_mapq ? rem_mapq[t] : 0) << 32);
}

__global__ void __launch_bounds__(TPB) e2t_out_kernel(const u64* __restrict__ keys, const u64* __restrict__ vals, u64 n,
                                                      i32* __restrict__ row, i32* __restrict__ col, u32* __restrict__ count,
                                                      int8_t* __restrict__ mapq) {
    size_t t = gtid();
    if (t >= n) return;
    u64 k = keys[t], v = vals[t];
    row[t] = (i32)(k >> 32);
    col[t] = (i32)(u32)k;
    count[t] = (u32)v;
    if (mapq) mapq[t] = (int8_t)(u8)(v >> 32);
}

__global__ void __launch_bounds__(TPB) t2e_key_kernel(const i32* __restrict__ row, const i32* __restrict__ col,
                                                      const u32* __restrict__ count, const int8_t* __restrict__ mapq, u64 n,
                                                      u64* __restrict__ keys, u64* __restrict__ vals) {
    size_t t = gtid();
    if (t >= n) return;
    keys[t] = ((u64)(u32)col[t] << 32) | ((u32)row[t] & CPH_ROWM);
    vals[t] = (u64)(count ? count[t] : 1u) | ((u64)(u8)(mapq ? mapq[t] : 0) << 32);
}

__global__ void __launch_bounds__(TPB) t2e_out_kernel(const u64* __restrict__ keys, const u64* __restrict__ vals, u64 n,
                                                      u32* __restrict__ rem_row, u32* __restrict__ rem_col,
                                                      u32* __restrict__ rem_count, int8_t* __restrict__ rem_mapq) {
    size_t t = gtid();
    if (t >= n) return;
    u64 k = keys[t], v = vals[t];
    u32 c = (u32)(k >> 32);
    bool head = (t == 0) || ((u32)(keys[t - 1] >> 32) != c);
    rem_row[t] = (u32)k | (head ? CPH_HEAD : 0u);
    rem_col[t] = c;
    rem_count[t] = (u32)v;
    if (rem_mapq) rem_mapq[t] = (int8_t)(u8)(v >> 32);
}

}  // namespace

// =================================================================================================
extern "C" int cph_pairs_to_edges(const u32* chrom1, const u32* chrom2, const int8_t* mapq, const u32* pos1, const u32* pos2,
                                  const i64* contig_len, u64 n_pairs, u32 n_contigs, int min_mapq, u64 edge_length, u32* rem_row,
                                  u32* rem_col, u32* rem_count, int8_t* rem_mapq, u64* h_counts, void* ws, size_t ws_bytes,
                                  size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(n_pairs < (1ull << 31));
    ARG_CHECK(n_contigs > 0 && n_contigs < CPH_HEAD);
    WS_BEGIN(ar);
    u32* keep = ar.take<u32>(n_pairs + 1);
    u32* total = ar.take<u32>(1);
    PairsIn in{chrom1, chrom2, mapq, pos1, pos2, contig_len, n_pairs, n_contigs, min_mapq, (i64)edge_length};
    if (!ar.dry) {
        ARENA_CHECK(ar);
        ARG_CHECK(n_pairs == 0 || (chrom1 && chrom2));
        ARG_CHECK(rem_row && rem_col && rem_count && rem_mapq && h_counts);
        ARG_CHECK(n_pairs == 0 || edge_length == 0 || (pos1 && pos2 && contig_len));
        if (n_pairs > 0) {
            pairs_flag_kernel<<<grid_for(n_pairs, TPB), TPB, 0, st>>>(in, keep);
            KERNEL_CHECK();
        }
    }
    TRY(scan_u32(keep, n_pairs, total, ar, st));
    WS_QUERY_RETURN(ar);
    u32 h_total = 0;
    if (n_pairs > 0) {
        pairs_emit_kernel<<<grid_for(n_pairs, TPB), TPB, 0, st>>>(in, keep, rem_row, rem_col, rem_count, rem_mapq);
        KERNEL_CHECK();
        CUDA_TRY(cudaMemcpyAsync(&h_total, total, sizeof(u32), cudaMemcpyDeviceToHost, st));
    }
    CUDA_TRY(cudaStreamSynchronize(st));
    h_counts[0] = (u64)h_total * 2;
    h_counts[1] = h_total;
    return CPH_OK;
}

extern "C" int cph_porec_to_edges(const u32* read_idx, const u32* contig_idx, const int8_t* mapq, const u32* start, const u32* end,
                                  const i64* contig_len, u64 n_records, u32 n_contigs, int min_mapq, u32 min_order, u32 max_order,
                                  u64 edge_length, u32 read_offset, int sorted_by_read, u32* rem_row, u32* rem_col, u32* rem_count,
                                  int8_t* rem_mapq, u64* h_counts, void* ws, size_t ws_bytes, size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const u64 N = n_records;
    ARG_CHECK(N < (1ull << 32) - 8192);
    ARG_CHECK(n_contigs > 0 && n_contigs < (1u << 23));
    const int cb = bits_for(n_contigs);   // contig ids < n_contigs <= 2^cb - 1
    WS_BEGIN(ar);
    u64* ka = ar.take<u64>(N);
    u64* kb = ar.take<u64>(N);
    u8* rep = ar.take<u8>(N);
    u32* keep = ar.take<u32>(N + 1);
    u32* aux = ar.take<u32>(N);
    int8_t* mq = ar.take<int8_t>(N);
    u32* scal = ar.take<u32>(4);   // [0] max read  [1] unsorted flag  [2] scan total
    PorecIn in{read_idx, contig_idx, mapq, start, end, contig_len, N, n_contigs, min_mapq, (i64)edge_length};
    u32 h_scal[4] = {0, 0, 0, 0};
    if (!ar.dry) {
        ARENA_CHECK(ar);
        ARG_CHECK(N == 0 || (read_idx && contig_idx && mapq));
        ARG_CHECK(rem_row && rem_col && rem_count && rem_mapq && h_counts);
        ARG_CHECK(N == 0 || edge_length == 0 || (start && end && contig_len));
        CUDA_TRY(cudaMemsetAsync(scal, 0, 4 * sizeof(u32), st));
        if (N > 0) {
            max_u32_kernel<<<148 * 8, TPB, 0, st>>>(read_idx, N, scal);
            KERNEL_CHECK();
            porec_key_kernel<<<grid_for(N, TPB), TPB, 0, st>>>(in, cb, ka, scal + 1);
            KERNEL_CHECK();
            CUDA_TRY(cudaMemcpyAsync(h_scal, scal, 2 * sizeof(u32), cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
        }
    }
    const int rb = bits_for(h_scal[0]);
    if (!ar.dry && 8 + cb + rb > 63) {
        cph_set_error(__FILE__, __LINE__, "read id / contig id bits exceed the 63-bit key");
        return CPH_ERR_TOO_LARGE;
    }
    // dry mode must reserve the sort scratch unconditionally (the table may turn out unsorted)
    int in_b = 0;
    (void)sorted_by_read;   // advisory: the key kernel detects order violations itself
    const bool need_sort = ar.dry || h_scal[1] != 0;
    if (need_sort) TRY(radix_sort_u64(ka, kb, nullptr, nullptr, 0, N, 8 + cb, ar.dry ? 8 + cb + 32 : 8 + cb + rb, &in_b, ar, st));
    const u64* keys = in_b ? kb : ka;
    KeyView kv{keys, cb, N};
    if (!ar.dry && N > 0) {
        porec_rep_kernel<<<grid_for(N, TPB), TPB, 0, st>>>(kv, rep);
        KERNEL_CHECK();
        porec_reduce_kernel<<<grid_for(N + 1, TPB), TPB, 0, st>>>(kv, rep, min_order, max_order, keep, aux, mq);
        KERNEL_CHECK();
    }
    TRY(scan_u32(keep, N > 0 ? N + 1 : 0, scal + 2, ar, st));
    WS_QUERY_RETURN(ar);
    u32 h_total = 0, h_lastcol = 0;
    if (N > 0) {
        porec_emit_kernel<<<grid_for(N, TPB), TPB, 0, st>>>(kv, keep, rep, aux, mq, read_offset, rem_row, rem_col, rem_count,
                                                            rem_mapq);
        KERNEL_CHECK();
        CUDA_TRY(cudaMemcpyAsync(&h_total, scal + 2, sizeof(u32), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        if (h_total > 0) {
            CUDA_TRY(cudaMemcpyAsync(&h_lastcol, rem_col + (h_total - 1), sizeof(u32), cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
        }
    }
    h_counts[0] = h_total;
    h_counts[1] = h_total ? (u64)h_lastcol + 1 : 0;
    return CPH_OK;
}

extern "C" int cph_edges_to_triplets(const u32* rem_row, const u32* rem_col, const u32* rem_count, const int8_t* rem_mapq, u64 nnz,
                                     u32 n_contigs, i32* row, i32* col, u32* count, int8_t* mapq, void* ws, size_t ws_bytes,
                                     size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(nnz < (1ull << 32) - 8192);
    ARG_CHECK(n_contigs > 0 && n_contigs < CPH_HEAD);
    WS_BEGIN(ar);
    u64* ka = ar.take<u64>(nnz);
    u64* kb = ar.take<u64>(nnz);
    u64* va = ar.take<u64>(nnz);
    u64* vb = ar.take<u64>(nnz);
    if (!ar.dry) {
        ARENA_CHECK(ar);
        ARG_CHECK(rem_row && rem_col && rem_count && row && col && count);
        if (nnz > 0) {
            e2t_key_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(rem_row, rem_col, rem_count, rem_mapq, nnz, ka, va);
            KERNEL_CHECK();
        }
    }
    int in_b = 0;
    TRY(radix_sort_u64(ka, kb, va, vb, 8, nnz, 32, 32 + bits_for(n_contigs - 1), &in_b, ar, st));
    WS_QUERY_RETURN(ar);
    if (nnz > 0) {
        e2t_out_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(in_b ? kb : ka, in_b ? vb : va, nnz, row, col, count, mapq);
        KERNEL_CHECK();
    }
    return CPH_OK;
}

extern "C" int cph_triplets_to_edges(const i32* row, const i32* col, const u32* count, const int8_t* mapq, u64 nnz, u32 n_cols,
                                     u32* rem_row, u32* rem_col, u32* rem_count, int8_t* rem_mapq, void* ws, size_t ws_bytes,
                                     size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(nnz < (1ull << 32) - 8192);
    ARG_CHECK(n_cols > 0);
    WS_BEGIN(ar);
    u64* ka = ar.take<u64>(nnz);
    u64* kb = ar.take<u64>(nnz);
    u64* va = ar.take<u64>(nnz);
    u64* vb = ar.take<u64>(nnz);
    if (!ar.dry) {
        ARENA_CHECK(ar);
        ARG_CHECK(row && col && rem_row && rem_col && rem_count);
        if (nnz > 0) {
            t2e_key_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(row, col, count, mapq, nnz, ka, va);
            KERNEL_CHECK();
        }
    }
    int in_b = 0;
    TRY(radix_sort_u64(ka, kb, va, vb, 8, nnz, 32, 32 + bits_for(n_cols - 1), &in_b, ar, st));
    WS_QUERY_RETURN(ar);
    if (nnz > 0) {
        t2e_out_kernel<<<grid_for(nnz, TPB), TPB, 0, st>>>(in_b ? kb : ka, in_b ? vb : va, nnz, rem_row, rem_col, rem_count,
                                                           rem_mapq);
        KERNEL_CHECK();
    }
    return CPH_OK;
}
