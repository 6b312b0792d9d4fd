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
#include "cph.h"
#include <mutex>

namespace {

constexpr int TPB = 256;
constexpr u32 RS_MAX_SMEM_BINS = 49152;      // 192 KB of u32 bins
constexpr u32 RS_PERSIST_MAX_BINS = 24576;   // 2 x u32 per bin: 192 KB
constexpr u32 RS_CHUNK = 32768;              // entries per CTA; count < 2^16 each -> u32 bins cannot overflow

__device__ __forceinline__ size_t gtid() { return (size_t)blockIdx.x * blockDim.x + threadIdx.x; }

__device__ __forceinline__ bool entry_ok(const int8_t* mapq, int min_mapq, size_t t) {
    return mapq == nullptr || (int)mapq[t] >= min_mapq;
}

// per-contig sum of counts; shared-memory privatised histogram, one flush per CTA
__global__ void __launch_bounds__(1024) row_sums_kernel(const u32* __restrict__ rem_row, const u32* __restrict__ rem_count,
                                                       const int8_t* __restrict__ rem_mapq, u64 nnz, u32 n, int min_mapq,
                                                       const u8* __restrict__ row_drop, u32 cmask, int use_smem, u32 chunk,
                                                       u32 bin_limit, unsigned long long* __restrict__ rowsum,
                                                       u32* __restrict__ max_count) {
    extern __shared__ u32 bins[];
    if (use_smem) {
        for (u32 i = threadIdx.x; i < n; i += blockDim.x) bins[i] = 0;
        __syncthreads();
    }
    const size_t lo = (size_t)blockIdx.x * chunk;
    const size_t hi = min((size_t)nnz, lo + chunk);
    u32 mx = 0;
    for (size_t t = lo + threadIdx.x; t < hi; t += blockDim.x) {
        if (!entry_ok(rem_mapq, min_mapq, t)) continue;
        u32 r = rem_row[t] & CPH_ROWM;
        if (r >= n || (row_drop && row_drop[r])) continue;
        u32 c = rem_count[t] & cmask;
        mx = max(mx, c);
        if (use_smem && c < bin_limit) atomicAdd(&bins[r], c);       // chunk * bin_limit < 2^32: a bin cannot overflow
        else atomicAdd(&rowsum[r], (unsigned long long)c);
    }
    if (max_count) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if ((threadIdx.x & 31) == 0 && mx) atomicMax(max_count, mx);
    }
    if (use_smem) {
        __syncthreads();
        for (u32 i = threadIdx.x; i < n; i += blockDim.x) {
            u32 v = bins[i];
            if (v) atomicAdd(&rowsum[i], (unsigned long long)v);
        }
    }
}

// Persistent variant for n <= RS_PERSIST_MAX_BINS: a few CTAs per SM, each streams a contiguous slice with 128-bit loads
// (four entries per load) into shared-memory bins kept as a low u32 word plus a carry word, and flushes every bin ONCE.
// The chunked kernel above flushes n bins per 32 Ki entries: 2.2e7 global atomics onto 4 000 addresses on workload C.
__global__ void __launch_bounds__(512) row_sums_persist_kernel(const u32* __restrict__ rem_row, const u32* __restrict__ rem_count,
                                                               const int8_t* __restrict__ rem_mapq, u64 nnz, u32 n, int min_mapq,
                                                               const u8* __restrict__ row_drop, u32 cmask,
                                                               unsigned long long* __restrict__ rowsum, u32* __restrict__ max_count) {
    extern __shared__ u32 bins[];      // [0, n): low words   [n, 2n): carries
    for (u32 i = threadIdx.x; i < 2 * n; i += blockDim.x) bins[i] = 0;
    __syncthreads();
    const u64 quads = (nnz + 3) / 4;
    const u64 per = (quads + gridDim.x - 1) / gridDim.x;
    const u64 q_lo = (u64)blockIdx.x * per, q_hi = min(quads, q_lo + per);
    u32 mx = 0;
    auto add = [&](u32 rr, u32 cc, int qq) {
        const u32 r = rr & CPH_ROWM;
        if (qq < min_mapq || r >= n || (row_drop && row_drop[r])) return;
        const u32 c = cc & cmask;
        mx = max(mx, c);
        const u32 old = atomicAdd(&bins[r], c);
        if (old + c < old) atomicAdd(&bins[n + r], 1u);
    };
    const bool vec = ((((uintptr_t)rem_row) | ((uintptr_t)rem_count)) & 15) == 0 && (rem_mapq == nullptr || (((uintptr_t)rem_mapq) & 3) == 0);
    for (u64 q = q_lo + threadIdx.x; q < q_hi; q += blockDim.x) {
        const u64 t = q * 4;
        if (vec && t + 4 <= nnz) {
            const uint4 r4 = *reinterpret_cast<const uint4*>(rem_row + t);
            const uint4 c4 = *reinterpret_cast<const uint4*>(rem_count + t);
            int q0 = 127, q1 = 127, q2 = 127, q3 = 127;
            if (rem_mapq) {
                const u32 m4 = *reinterpret_cast<const u32*>(rem_mapq + t);
                q0 = (int8_t)(m4 & 0xff); q1 = (int8_t)((m4 >> 8) & 0xff); q2 = (int8_t)((m4 >> 16) & 0xff); q3 = (int8_t)(m4 >> 24);
            }
            add(r4.x, c4.x, q0); add(r4.y, c4.y, q1); add(r4.z, c4.z, q2); add(r4.w, c4.w, q3);
        } else {
            for (u64 u = t; u < min(nnz, t + 4); ++u) add(rem_row[u], rem_count[u], rem_mapq ? (int)rem_mapq[u] : 127);
        }
    }
    if (max_count) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if ((threadIdx.x & 31) == 0 && mx) atomicMax(max_count, mx);
    }
    __syncthreads();
    for (u32 i = threadIdx.x; i < n; i += blockDim.x) {
        const u64 v = ((u64)bins[n + i] << 32) | bins[i];
        if (v) atomicAdd(&rowsum[i], (unsigned long long)v);
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

// ------------------------------------------------------------------------------------------------
// Fast path (no CSR view requested): one pass of a shared-memory tile kernel, same skeleton as the
// Pore-C tile kernel of edges.cu -- ticketed tiles own the hyperedges that START inside them, the tile
// plus a look-ahead halo is staged in shared memory, per-hyperedge sums run there, a decoupled
// look-back over per-tile totals yields the output offset.  HBM traffic: 9 B read + 8 B written per entry.
// ------------------------------------------------------------------------------------------------
constexpr int IT_THREADS = 512;
constexpr int IT_TILE = 4096;
constexpr int IT_HALO = 1024;
constexpr int IT_WIN = IT_TILE + IT_HALO;
constexpr int IT_WORDS = IT_WIN / 32;
constexpr u64 IT_AGG = 1ull << 62, IT_INCL = 2ull << 62, IT_VAL = (1ull << 62) - 1;

struct IncIn {
    const u32* row;
    const u32* cnt;
    const int8_t* mapq;
    u64 nnz;
    u32 n;
    int min_mapq;
    const i32* row_map;
    u32 col_min_sum;
    u32 cmask;
};

struct ItSmem {
    u32 row[IT_WIN];        // new row (or ROWM if the entry is dropped)
    u32 cnt[IT_WIN];
    u32 res[IT_WIN];        // at a hyperedge start: keep | D << 8
    u32 head[IT_WORDS];
    u32 ok[IT_WORDS];
    u32 opref[IT_WORDS + 1];
    u32 keep[IT_WORDS];
    u32 kpref[IT_WORDS + 1];
    int beg, end, last;
    u32 tile, base, long_k, long_res, has_zero, scan[IT_THREADS / 32 + 1];
    unsigned long long long_end, long_sum;
};

__device__ __forceinline__ void it_word_prefix(const u32* __restrict__ words, u32* __restrict__ pref, int tid) {
    if (tid < 32) {
        constexpr int PER = IT_WORDS / 32;
        u32 loc[PER], sum = 0;
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            loc[k] = sum;
            sum += __popc(words[tid * PER + k]);
        }
        u32 incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            u32 v = __shfl_up_sync(0xffffffffu, incl, o);
            if (tid >= o) incl += v;
        }
        u32 ex = incl - sum;
#pragma unroll
        for (int k = 0; k < PER; ++k) pref[tid * PER + k] = ex + loc[k];
        if (tid == 31) pref[IT_WORDS] = incl;
    }
}
__device__ __forceinline__ u32 it_rank(const u32* __restrict__ words, const u32* __restrict__ pref, int x) {
    int w = x >> 5;
    return w >= IT_WORDS ? pref[IT_WORDS] : pref[w] + __popc(words[w] & ((1u << (x & 31)) - 1u));
}
__device__ __forceinline__ int it_seg_start(const u32* __restrict__ head, int i) {
    int w = i >> 5;
    u32 m = head[w] & (0xffffffffu >> (31 - (i & 31)));
    while (m == 0) m = head[--w];
    return w * 32 + 31 - __clz(m);
}
__device__ __forceinline__ u32 edge_degree(u32 k) {
    // D_e = max(int8(k) - 1, 1) with numpy's int8 wrap-around (algorithms/hypergraph.py:385, 418)
    int d8 = (int)(int8_t)(u8)((k - 1u) & 0xffu);
    return (u32)max(d8, 1);
}
__device__ __forceinline__ u32 inc_new_row(const IncIn& in, size_t t) {   // new row of entry t or ROWM when dropped
    u32 r = in.row[t] & CPH_ROWM;
    bool ok = entry_ok(in.mapq, in.min_mapq, t) && r < in.n;
    i32 nr = ok ? in.row_map[r] : -1;
    return nr >= 0 ? (u32)nr : CPH_ROWM;
}

__global__ void __launch_bounds__(IT_THREADS, 3) incidence_tile_kernel(IncIn in, unsigned long long* __restrict__ tile_state,
                                                                    u32* __restrict__ scal, unsigned long long* __restrict__ heads_total,
                                                                    u32* __restrict__ pem_row, u32* __restrict__ pem_pk) {
    extern __shared__ __align__(16) unsigned char it_raw[];
    ItSmem& S = *reinterpret_cast<ItSmem*>(it_raw);
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid == 0) {
        S.tile = atomicAdd(&scal[3], 1u);
        S.beg = 0x7fffffff;
        S.end = 0x7fffffff;
        S.last = -1;
        S.long_k = 0;
        S.long_sum = 0;
        S.has_zero = 0u;
    }
    __syncthreads();
    const u32 tile = S.tile;
    const size_t lo = (size_t)tile * IT_TILE;
    const size_t N = in.nnz;
    const int own = (int)min((size_t)IT_TILE, N - lo);
    const int wn = (int)min((size_t)IT_WIN, N - lo);
    // stage + heads + ok bits.  The loads of IT_LD window slices are issued together (memory-level parallelism: the
    // kernel is latency-bound, not bandwidth-bound) before any of them is consumed.
    constexpr int IT_LD = 5;
    static_assert(IT_WIN % (IT_THREADS * IT_LD) == 0, "window must split into load batches");
    bool zero_cnt = false;
    for (int i0 = 0; i0 < IT_WIN; i0 += IT_THREADS * IT_LD) {
        u32 lr[IT_LD], lc[IT_LD];
        int lq[IT_LD];
#pragma unroll
        for (int u = 0; u < IT_LD; ++u) {
            const int i = i0 + u * IT_THREADS + tid;
            lr[u] = 0; lc[u] = 0; lq[u] = 127;
            if (i < wn) {
                const size_t t = lo + i;
                lr[u] = in.row[t];
                lc[u] = in.cnt[t];
                if (in.mapq) lq[u] = in.mapq[t];
            }
        }
        i32 lm[IT_LD];
#pragma unroll
        for (int u = 0; u < IT_LD; ++u) {
            const int i = i0 + u * IT_THREADS + tid;
            const u32 r = lr[u] & CPH_ROWM;
            const bool in_ok = i < wn && (in.mapq == nullptr || lq[u] >= in.min_mapq) && r < in.n;
            lm[u] = in_ok ? in.row_map[r] : -1;
        }
#pragma unroll
        for (int u = 0; u < IT_LD; ++u) {
            const int i = i0 + u * IT_THREADS + tid;
            bool head = false, ok = false;
            if (i < wn) {
                const u32 c = lc[u];
                if (in.cmask == 0xffffffffu && c > CPH_CNTM) scal[0] = 1u;      // not representable as float32 count
                const u32 nr = lm[u] >= 0 ? (u32)lm[u] : CPH_ROWM;
                ok = nr != CPH_ROWM;
                S.row[i] = nr;
                S.cnt[i] = c & in.cmask;
                zero_cnt |= ok && (c & in.cmask) == 0u;
                head = (lr[u] & CPH_HEAD) || (lo + i) == 0;
            }
            u32 mh = __ballot_sync(0xffffffffu, head), mo = __ballot_sync(0xffffffffu, ok);
            if (lane == 0) {
                S.head[i >> 5] = mh;
                S.ok[i >> 5] = mo;
            }
        }
    }
    if (zero_cnt) S.has_zero = 1u;      // explicit zero counts (host matrices): the k >= col_min_sum shortcut is off
    __syncthreads();
    // first / last hyperedge start inside the tile and the first one after it, by ONE warp over the head words (shared
    // atomics from every warp on these three words were 12 % of the kernel's instructions); warp 0 meanwhile builds the
    // prefix of the ok bits
    if (tid >= 32 && tid < 64) {
        int b = 0x7fffffff, e = 0x7fffffff, l = -1;
        for (int w = lane; w < IT_WORDS; w += 32) {
            const u32 mh = S.head[w];
            if (mh) {
                const int i = w * 32;
                const u32 inside = (i + 32 <= own) ? mh : (i < own ? (mh & ((1u << (own - i)) - 1u)) : 0u);
                const u32 after = mh & ~inside;
                if (inside) {
                    b = min(b, i + __ffs(inside) - 1);
                    l = max(l, i + 31 - __clz(inside));
                }
                if (after) e = min(e, i + __ffs(after) - 1);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            b = min(b, __shfl_xor_sync(0xffffffffu, b, o));
            e = min(e, __shfl_xor_sync(0xffffffffu, e, o));
            l = max(l, __shfl_xor_sync(0xffffffffu, l, o));
        }
        if (lane == 0) {
            S.beg = b;
            S.end = e;
            S.last = l;
        }
    }
    it_word_prefix(S.ok, S.opref, tid);
    __syncthreads();
    const int beg = S.beg;
    const bool empty = beg == 0x7fffffff;
    bool is_long = false;
    int fend = 0;
    if (!empty) {
        if (S.end != 0x7fffffff) fend = S.end;
        else if (lo + wn == N) fend = wn;
        else {
            is_long = true;
            fend = S.last;
        }
    }
    // per hyperedge (thread at its start): surviving entries k, their summed count -> keep, D_e
    u32 my_heads = 0;
    const bool has_zero = S.has_zero != 0u;
    for (int i0 = 0; i0 < IT_WIN; i0 += IT_THREADS) {
        int i = i0 + tid;
        if (!empty && i >= beg && i < fend && ((S.head[i >> 5] >> (i & 31)) & 1u)) {
            // segment end = next head (or fend)
            int w = i >> 5;
            u32 m = S.head[w] & ~((2u << (i & 31)) - 1u);
            while (m == 0 && ++w < IT_WORDS) m = S.head[w];
            int e = m ? w * 32 + __ffs(m) - 1 : fend;
            e = min(e, fend);
            const u32 k = it_rank(S.ok, S.opref, e) - it_rank(S.ok, S.opref, i);
            // summed count >= col_min_sum (algorithms/hypergraph.py:305-306).  Every count is >= 1 unless the tile saw
            // an explicit zero, so k >= col_min_sum already decides it; only short survivors need the real sum.
            bool edge_keep = k > 0;
            if (edge_keep && (k < in.col_min_sum || has_zero)) {
                u64 colsum = 0;
                for (int j = i; j < e; ++j)
                    if (S.row[j] != CPH_ROWM) colsum += S.cnt[j];
                edge_keep = colsum >= (u64)in.col_min_sum;
            }
            // offset of the first surviving entry (it becomes the head of the packed hyperedge)
            u32 foff = 0;
            if (edge_keep) {
                int fw = i >> 5;
                u32 fm = S.ok[fw] & ~((1u << (i & 31)) - 1u);
                while (fm == 0) fm = S.ok[++fw];
                foff = (u32)(fw * 32 + __ffs(fm) - 1 - i);
            }
            S.res[i] = (edge_keep ? 1u : 0u) | (edge_degree(k) << 8) | (foff << 16);
            my_heads += edge_keep ? 1u : 0u;
        }
    }
    __syncthreads();
    for (int i0 = 0; i0 < IT_WIN; i0 += IT_THREADS) {
        int i = i0 + tid;
        bool keep = false;
        if (!empty && i >= beg && i < fend && S.row[i] != CPH_ROWM) {
            const int s0 = it_seg_start(S.head, i);
            const u32 rs = S.res[s0];
            keep = rs & 1u;
            if (keep) {
                const bool first = (u32)(i - s0) == (rs >> 16);
                S.row[i] |= first ? CPH_HEAD : 0u;
                S.cnt[i] |= ((rs >> 8) & 0xffu) << 24;
            }
        }
        u32 m = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) S.keep[i >> 5] = m;
    }
    __syncthreads();
    it_word_prefix(S.keep, S.kpref, tid);
    // slow path, part 1: the hyperedge that starts at `fend` and runs past the window
    u32 long_kept = 0, long_D = 1;
    size_t ls = 0, le = 0;
    if (is_long) {
        ls = lo + fend;
        if (tid == 0) {
            size_t e = lo + wn;
            while (e < N && !(in.row[e] & CPH_HEAD)) ++e;
            S.long_end = e;
        }
        __syncthreads();
        le = S.long_end;
        u32 k = 0;
        u64 sum = 0;
        for (size_t t = ls + tid; t < le; t += IT_THREADS)
            if (inc_new_row(in, t) != CPH_ROWM) {
                ++k;
                sum += in.cnt[t] & in.cmask;
            }
        k = warp_sum(k);
        sum = warp_sum(sum);
        if (lane == 0 && k) {
            atomicAdd(&S.long_k, k);
            atomicAdd(&S.long_sum, (unsigned long long)sum);
        }
        __syncthreads();
        if (S.long_k > 0 && S.long_sum >= (u64)in.col_min_sum) {
            long_kept = S.long_k;
            long_D = edge_degree(S.long_k);
            if (tid == 0) ++my_heads;
        }
    }
    __syncthreads();
    const u32 fast_total = empty ? 0u : S.kpref[IT_WORDS];
    const u32 total = fast_total + long_kept;
    my_heads = warp_sum(my_heads);
    if (lane == 0 && my_heads) atomicAdd(heads_total, (unsigned long long)my_heads);
    if (tid < 32) {
        u32 base = 0;
        if (tile == 0) {
            if (lane == 0) atomicExch(&tile_state[0], IT_INCL | (u64)total);
        } else {
            if (lane == 0) atomicExch(&tile_state[tile], IT_AGG | (u64)total);
            u64 running = 0;
            long j = (long)tile - 1;
            for (;;) {
                long idx = j - lane;
                u64 st = idx >= 0 ? *((volatile unsigned long long*)&tile_state[idx]) : IT_INCL;
                u32 inval = __ballot_sync(0xffffffffu, (st >> 62) == 0);
                u32 incl = __ballot_sync(0xffffffffu, (st >> 62) == 2);
                u32 first = incl ? (u32)__ffs(incl) - 1u : 32u;
                u32 need = first >= 31u ? 0xffffffffu : ((2u << first) - 1u);
                if (inval & need) continue;
                u64 v = ((u32)lane <= first) ? (st & IT_VAL) : 0ull;
                running += warp_sum(v);
                if (incl) break;
                j -= 32;
            }
            base = (u32)running;
            if (lane == 0) atomicExch(&tile_state[tile], IT_INCL | (running + total));
        }
        if (lane == 0) S.base = base;
    }
    __syncthreads();
    const u32 base = S.base;
    if (!empty) {
        for (int i = beg + tid; i < fend; i += IT_THREADS) {
            u32 w = S.keep[i >> 5];
            if ((w >> (i & 31)) & 1u) {
                u32 p = base + S.kpref[i >> 5] + __popc(w & ((1u << (i & 31)) - 1u));
                pem_row[p] = S.row[i];
                pem_pk[p] = S.cnt[i];
            }
        }
    }
    if (long_kept) {
        u32 carry = 0;
        for (size_t t0 = ls; t0 < le; t0 += IT_THREADS) {
            size_t t = t0 + tid;
            u32 nr = t < le ? inc_new_row(in, t) : CPH_ROWM;
            bool ok = nr != CPH_ROWM;
            u32 tot;
            u32 ex = block_excl_scan<u32, IT_THREADS>(ok ? 1u : 0u, S.scan, &tot);
            if (ok) {
                u32 k = carry + ex;
                u32 p = base + fast_total + k;
                pem_row[p] = nr | (k == 0 ? CPH_HEAD : 0u);
                pem_pk[p] = (long_D << 24) | (in.cnt[t] & in.cmask);
            }
            carry += tot;
        }
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
        u64 c = p & CPH_CNTM;
        double sc = scale[use_high_order ? (p >> 24) : 1u];
        if (ew) sc *= (double)ew[t];
        unsigned long long v = __double2ull_rn((double)(c * c) * sc);
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
    if (n_contigs <= RS_PERSIST_MAX_BINS && nnz >= (1u << 20)) {
        const size_t sm = (size_t)n_contigs * 2 * sizeof(u32);
        static int sms = 0;
        if (!sms) {
            int dev = 0;
            CUDA_TRY(cudaGetDevice(&dev));
            CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        }
        CUDA_TRY(cudaFuncSetAttribute(row_sums_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        int per_sm = (int)((220u * 1024u) / (sm + 1024));
        per_sm = per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm);
        row_sums_persist_kernel<<<sms * per_sm, 512, sm, st>>>(rem_row, rem_count, rem_mapq, nnz, n_contigs, min_mapq, row_drop,
                                                             packed ? CPH_CNTM : 0xffffffffu, (unsigned long long*)rowsum, max_count);
        KERNEL_CHECK();
        return CPH_OK;
    }
    const int use_smem = n_contigs <= RS_MAX_SMEM_BINS;
    const size_t smem = use_smem ? (size_t)n_contigs * sizeof(u32) : 0;
    if (smem > 48 * 1024) CUDA_TRY(cudaFuncSetAttribute(row_sums_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // many bins: a CTA must see enough entries to pay for flushing n bins (8 x larger chunks, 4 x smaller bin limit)
    const u32 chunk = n_contigs > 8192 ? RS_CHUNK * 8 : RS_CHUNK;
    const u32 bin_limit = n_contigs > 8192 ? 16384u : 65536u;
    const int rs_threads = n_contigs > 8192 ? 1024 : TPB;      // 80+ KB of bins per CTA: fill the SM with two big CTAs
    row_sums_kernel<<<grid_for(nnz, chunk), rs_threads, smem, st>>>(rem_row, rem_count, rem_mapq, nnz, n_contigs, min_mapq, row_drop,
                                                            packed ? CPH_CNTM : 0xffffffffu, use_smem, chunk, bin_limit,
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
    // ---- fast path: pem only (no CSR view) ----
    const u32 ntiles = (u32)ceil_div(nnz > 0 ? nnz : 1, (size_t)IT_TILE);
    u32* scal = ar.take<u32>(4);                 // [0] bad count  [3] ticket
    u64* heads_total = ar.take<u64>(1);
    u64* tile_state = ar.take<u64>((size_t)ntiles);
    if (!ar.dry && indptr == nullptr) {
        ARENA_CHECK(ar);
        ARG_CHECK(row_map && pem_row && pem_pk && h_counts);
        ARG_CHECK(nnz == 0 || (rem_row && rem_count));
        h_counts[0] = h_counts[1] = 0;
        if (nnz == 0) return CPH_OK;
        CUDA_TRY(cudaMemsetAsync(scal, 0, 4 * sizeof(u32), st));
        CUDA_TRY(cudaMemsetAsync(heads_total, 0, sizeof(u64), st));
        CUDA_TRY(cudaMemsetAsync(tile_state, 0, (size_t)ntiles * sizeof(u64), st));
        CUDA_TRY(cudaFuncSetAttribute(incidence_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ItSmem)));
        IncIn in{rem_row, rem_count, rem_mapq, nnz, n_contigs, min_mapq, row_map, col_min_sum, cmask};
        incidence_tile_kernel<<<ntiles, IT_THREADS, sizeof(ItSmem), st>>>(in, (unsigned long long*)tile_state, scal,
                                                                         (unsigned long long*)heads_total, pem_row, pem_pk);
        KERNEL_CHECK();
        u32 hs = 0;
        u64 last = 0, heads = 0;
        CUDA_TRY(cudaMemcpyAsync(&hs, scal, sizeof(u32), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaMemcpyAsync(&last, tile_state + (ntiles - 1), sizeof(u64), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaMemcpyAsync(&heads, heads_total, sizeof(u64), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        if (hs) {
            cph_set_error(__FILE__, __LINE__, "an alignment count >= 2^24 cannot be represented exactly (the reference stores counts as float32)");
            return CPH_ERR_TOO_LARGE;
        }
        h_counts[0] = last & IT_VAL;
        h_counts[1] = heads;
        return CPH_OK;
    }
    // ---- general path: also emits the CSR view ----
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

// ---------------------------------------------------------------------------------------------------------------
// cph_incidence_csr_fixed: the CSR view of a small incidence matrix for callers that bring no workspace.
//
// Every other entry point leaves the temporaries to the caller (`ws`).  Some callers have nowhere to get one:
//   * a stage that runs inside somebody else's process (a workflow engine's operator callback, a viewer plug-in) is
//     handed its input and output buffers and nothing more, and the host's memory pool already owns the rest of the GPU,
//     so a cudaMalloc from inside the stage fails or takes memory the host has planned for;
//   * a service that keeps several assemblies in flight on one GPU cannot afford cudaMalloc / cudaFree in the request
//     path at all: both serialise against everything running on the device.
// For them the temporaries live in the array below, which is part of the library image: each GPU that loads libcph.so
// carries these 16 MiB from the load of this module until the process exits, whether the function is ever called or not.
// The function itself allocates nothing and calls nothing that allocates.
//
// `g_fixed_scratch` is a DEVICE array.  On the host its name is only a handle for the runtime's symbol table, not an
// address a kernel could use, so the host code below asks the runtime for the real address -- once per device, with
// that device current -- and hands that pointer on.
// ---------------------------------------------------------------------------------------------------------------
#define CPH_FIXED_SCRATCH_BYTES ((size_t)16 << 20)
static __device__ __align__(256) unsigned char g_fixed_scratch[CPH_FIXED_SCRATCH_BYTES];

namespace {
constexpr int FIXED_MAX_DEVICES = 64;
std::mutex g_fixed_mu;
void* g_fixed_addr[FIXED_MAX_DEVICES];   // device address of g_fixed_scratch per device ordinal, 0 = not asked yet

// device address of the scratch array on the CURRENT device (looked up on the first call per device, cached afterwards)
int fixed_scratch_address(void** out) {
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= FIXED_MAX_DEVICES) {
        cph_set_error(__FILE__, __LINE__, "cph_incidence_csr_fixed: device ordinal beyond the 64 this table holds");
        return CPH_ERR_ARG;
    }
    std::lock_guard<std::mutex> lk(g_fixed_mu);
    if (!g_fixed_addr[dev]) {
        void* p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_fixed_scratch));
        if (!p) {
            cph_set_error(__FILE__, __LINE__, "cph_incidence_csr_fixed: the runtime returned no address for the scratch array");
            return CPH_ERR_CUDA;
        }
        g_fixed_addr[dev] = p;
    }
    *out = g_fixed_addr[dev];
    return CPH_OK;
}
}  // namespace

extern "C" int cph_incidence_csr_fixed(const u32* pem_row, const u32* pem_pk, const u32* pem_col, u64 nnz, u32 n_rows,
                                       i64* indptr, i32* indices, u32* pk, void* stream) {
    if (nnz > CPH_INCIDENCE_CSR_FIXED_MAX_NNZ) {
        char msg[200];
        snprintf(msg, sizeof msg,
                 "cph_incidence_csr_fixed: %llu entries, the built-in scratch holds at most %llu; use cph_incidence_csr with a "
                 "workspace of your own",
                 (unsigned long long)nnz, (unsigned long long)CPH_INCIDENCE_CSR_FIXED_MAX_NNZ);
        cph_set_error(__FILE__, __LINE__, msg);
        return CPH_ERR_TOO_LARGE;
    }
    void* scratch = nullptr;
    TRY(fixed_scratch_address(&scratch));
    // Same kernels as cph_incidence_csr, temporaries carved from the fixed array.  None of them expects zeroes: head
    // flags, sort keys / values and the digit tables are all written by this call before anything reads them, so what
    // the previous call left behind is simply overwritten and nothing is cleared here.
    return cph_incidence_csr(pem_row, pem_pk, pem_col, nnz, n_rows, indptr, indices, pk, scratch, CPH_FIXED_SCRATCH_BYTES, nullptr,
                             stream);
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
