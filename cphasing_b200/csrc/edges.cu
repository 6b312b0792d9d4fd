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
#include "cph.h"
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <emmintrin.h>
#include <mutex>
#include <thread>
#include <vector>

namespace {

constexpr int TPB = 256;

__device__ __forceinline__ size_t gtid() { return (size_t)blockIdx.x * blockDim.x + threadIdx.x; }

// ------------------------------------------------------------------------------------------------
// Pre-filters of the extractors (cph_prefilter of include/cph.h, device view).
//   HCR BED  hypergraph.py:1721-1752 (Pore-C, alignment midpoint), :786-893 (pairs, either end): join_asof(backward) on the
//            BED starts of the same contig = the LAST interval whose start <= pos; the record is "in" iff pos < its end.
//   split    hypergraph.py:1406-1424 (Pore-C: init_idx + midpoint // split_length), :994-1023 (pairs: init_idx + pos //
//            split_length): the contig column then holds codes of the ORIGINAL contig names and the hyperedge rows are
//            the split bins.
// ------------------------------------------------------------------------------------------------
struct PreF {
    const u64* bed_ptr;      // [n_names + 1] or nullptr
    const u32* bed_start;    // sorted by (name, start)
    const u32* bed_end;
    const u32* split_init;   // [n_names] first bin of each original contig (0xffffffff: not in the map) or nullptr
    u64 split_length;
    u32 n_names;             // size of the code space of the contig column(s)
    int hcr_invert;
};

__device__ __forceinline__ bool pf_in_hcr(const PreF& f, u32 name, u32 pos) {
    u64 lo = f.bed_ptr[name], hi = f.bed_ptr[name + 1];      // first interval with start > pos
    const u64 base = lo;
    while (lo < hi) {
        const u64 mid = (lo + hi) >> 1;
        if (f.bed_start[mid] <= pos) lo = mid + 1;
        else hi = mid;
    }
    return lo > base && pos < f.bed_end[lo - 1];
}

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
    PreF pf;
};

__device__ __forceinline__ bool pair_ok(const PairsIn& in, size_t p, u32& a, u32& b, int8_t& q) {
    a = in.c1[p];
    b = in.c2[p];
    q = in.mapq ? in.mapq[p] : (int8_t)0;
    const bool pre = in.pf.bed_ptr != nullptr || in.pf.split_init != nullptr;
    bool ok = pre ? (a < in.pf.n_names && b < in.pf.n_names) : (a < in.n_contigs && b < in.n_contigs);
    if (in.min_mapq > 0 && in.mapq) ok = ok && (int)q >= in.min_mapq;          // hypergraph.py:970-971
    if (ok && pre) {
        const u32 x1 = in.p1[p], x2 = in.p2[p];
        if (in.pf.bed_ptr) {                                                    // :786-893 any end inside an HCR interval
            const bool any = pf_in_hcr(in.pf, a, x1) || pf_in_hcr(in.pf, b, x2);
            ok = in.pf.hcr_invert ? !any : any;
        }
        if (ok && in.pf.split_init) {                                           // :994-1023
            const u32 ia = in.pf.split_init[a], ib = in.pf.split_init[b];
            ok = ia != 0xffffffffu && ib != 0xffffffffu;
            a = ia + (u32)(x1 / in.pf.split_length);
            b = ib + (u32)(x2 / in.pf.split_length);
            ok = ok && a < in.n_contigs && b < in.n_contigs;
        }
    }
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
    // :1287-1292 count = ones; the top byte carries D_e = max(2 - 1, 1) = 1 of the two-entry hyperedge (packed form)
    *reinterpret_cast<uint2*>(rem_count + t) = make_uint2((1u << 24) | 1u, (1u << 24) | 1u);
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
    PreF pf;
};

__global__ void __launch_bounds__(TPB) max_u32_kernel(const u32* __restrict__ x, u64 n, u32* __restrict__ out) {
    u32 m = 0;
    for (size_t i = gtid(); i < n; i += (size_t)gridDim.x * blockDim.x) m = max(m, x[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

constexpr u32 PT_INVALID = 0x00ffffffu;          // contig field of a filtered-out record (n_contigs < 2^23)
__device__ __forceinline__ u32 porec_cvq(const PorecIn& in, size_t t) {   // (contig or INVALID) << 8 | mapq + 128
    u32 c = in.contig[t];
    int8_t q = in.mapq[t];
    const bool pre = in.pf.bed_ptr != nullptr || in.pf.split_init != nullptr;
    bool ok = pre ? c < in.pf.n_names : c < in.n_contigs;
    if (in.min_mapq > 0) ok = ok && (int)q >= in.min_mapq;                      // hypergraph.py:1774-1775
    if (ok && pre) {
        const u32 st = in.start[t], en = in.end[t];
        const u32 mid = st + (en - st) / 2u;                                    // UInt32 arithmetic, as polars does it
        if (in.pf.bed_ptr) {                                                    // :1730-1752
            const bool hit = pf_in_hcr(in.pf, c, mid);
            ok = in.pf.hcr_invert ? !hit : hit;
        }
        if (ok && in.pf.split_init) {                                           // :1406-1424 (no edge filter on this branch)
            const u32 i0 = in.pf.split_init[c];
            ok = i0 != 0xffffffffu;
            c = i0 + (u32)(mid / in.pf.split_length);
            ok = ok && c < in.n_contigs;
        }
    }
    if (ok && in.edge_length > 0 && !in.pf.split_init) {                        // :1449-1456
        i64 st = in.start[t], en = in.end[t];
        i64 pos = st + (en - st) / 2;
        i64 L = in.len[c];
        ok = (pos < in.edge_length) || ((L - pos) < in.edge_length);
    }
    return ((ok ? c : PT_INVALID) << 8) | (u32)(u8)((int)q + 128);
}

__global__ void __launch_bounds__(TPB) porec_key_kernel(PorecIn in, int cb, u64* __restrict__ keys, u32* __restrict__ unsorted_flag) {
    size_t t = gtid();
    if (t >= in.n) return;
    const u32 r = in.read[t];
    const u32 x = porec_cvq(in, t);              // every record filter (MAPQ, HCR, split index, edge) lives there
    const u32 c = x >> 8;
    const bool ok = c != PT_INVALID;
    u64 k = ((u64)r << (cb + 8)) | ((u64)(ok ? c : 0u) << 8) | (u64)(x & 0xffu);
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
// Fast path for tables already grouped by read (what the reference's mappers write): ONE pass of a
// shared-memory tile kernel.  A CTA takes a ticket (so lower tiles are always already running), owns
// the reads that START inside its tile of PT_TILE records, stages the tile plus a PT_HALO look-ahead
// in shared memory (coalesced 9 B/record) and does every per-read step there:
//   heads      bitmask of read starts (ballot)                      -> segment start/end in O(1) by clz/ffs
//   P1         first record of each (read, contig) = representative; the others add their count and
//              max MAPQ to it with shared-memory atomics            (hypergraph.py:1470-1477)
//   P2         distinct-contig order = popcount of representatives in the segment, order filter (:1461-1468)
//   look-back  decoupled look-back over per-tile totals (64-bit status|value words) gives the tile's output
//              offset without a second pass over the input
//   emit       compacted, coalesced stores of (contig|head, read, count, mapq)
// HBM traffic: 9 B per record + 13 B per surviving entry.  The one read per tile that can overflow the
// window is handled by a slow cooperative walk over global memory.
// ------------------------------------------------------------------------------------------------
constexpr int PT_THREADS = 256;
constexpr int PT_TILE = 1792;     // + halo = 2048-record window: 31 KB of shared memory, 6 CTAs of 256 threads per SM
constexpr int PT_HALO = 256;
constexpr int PT_WIN = PT_TILE + PT_HALO;
constexpr int PT_WORDS = PT_WIN / 32;
constexpr u64 PT_AGG = 1ull << 62, PT_INCL = 2ull << 62, PT_VAL = (1ull << 62) - 1;

constexpr int PT_RPT = PT_TILE / PT_THREADS;      // records of the tile per thread: thread t owns the reads that start in its slice
static_assert(PT_TILE % PT_THREADS == 0, "tile must split evenly over the threads");

template <bool WARP>
struct PtSmemT {
    alignas(16) u32 read[PT_WIN];
    alignas(16) u32 cvq[PT_WIN];   // window records, then IN PLACE the distinct-contig list of every read (slot s + d)
    alignas(16) u8 mq[PT_WIN];     // raw MAPQ bytes of the window (bulk-copy staging)
    alignas(8) unsigned long long mbar;
    u16 cnt[PT_WIN];               // record count of the list entry in the same slot (warp kernel: bit 15 = first entry of a hyperedge)
    u16 rn[WARP ? 2 : PT_WIN];     // walk kernel: at a read start, number of entries the read emits (0 = filtered out)
    u16 src[WARP ? 2 : PT_WIN];    // walk kernel: output slot -> window position of the list entry
    u8 dk[WARP ? PT_WIN : 4];      // warp kernel: D_e of the hyperedge the list entry in the same slot belongs to
    u32 head[PT_WORDS + 3];        // + zero padding: the warp kernel reads 96-bit windows of the mask without bounds checks
    u32 hpref[PT_WORDS + 1];
    u32 tile, base, long_order, nheads, scan[PT_THREADS / 32 + 1];
    u32 wstart[PT_THREADS / 32], wcnt[PT_THREADS / 32];   // warp kernel: window position and length of each warp's compacted output
    unsigned long long long_end;
};

// exclusive prefix of popcounts over PT_WORDS words by warp 0; pref[PT_WORDS] = total
__device__ __forceinline__ void pt_word_prefix(const u32* __restrict__ words, u32* __restrict__ pref, int tid) {
    if (tid < 32) {
        constexpr int PER = PT_WORDS / 32;
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
        if (tid == 31) pref[PT_WORDS] = incl;
    }
}
__device__ __forceinline__ int pt_next_head(const u32* __restrict__ head, int i) {   // first head >= i, PT_WIN if none
    if (i >= PT_WIN) return PT_WIN;
    int w = i >> 5;
    u32 m = head[w] & ~((1u << (i & 31)) - 1u);
    while (m == 0) {
        if (++w >= PT_WORDS) return PT_WIN;
        m = head[w];
    }
    return w * 32 + __ffs(m) - 1;
}
__device__ __forceinline__ int pt_prev_head(const u32* __restrict__ head, int i) {   // last head <= i (must exist)
    int w = i >> 5;
    u32 m = head[w] & (0xffffffffu >> (31 - (i & 31)));
    while (m == 0) m = head[--w];
    return w * 32 + 31 - __clz(m);
}
// TMA (bulk async copy) helpers: one elected thread moves the window's three columns global -> shared, completion is
// signalled on an mbarrier (SASS: UBLKCP).  Addresses and sizes must be multiples of 16 bytes.
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, u32 bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, u32 phase) {
    u32 ok;
    do {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok)
                     : "r"(smem_u32(bar)), "r"(phase)
                     : "memory");
    } while (!ok);
}

// D_e = max(int8(k) - 1, 1) with numpy's int8 wrap-around (algorithms/hypergraph.py:385, 418), as incidence.cu edge_degree
__device__ __forceinline__ u32 pt_degree(u32 k) {
    const int d8 = (int)(int8_t)(u8)((k - 1u) & 0xffu);
    return (u32)max(d8, 1);
}

template <bool WARP, int MINB = 6>
__global__ void __launch_bounds__(PT_THREADS, MINB) porec_tile_kernel(PorecIn in, int use_tma, u32 min_order, u32 max_order, u32 read_offset,
                                                                   unsigned long long* __restrict__ tile_state, u32* __restrict__ scal,
                                                                   u32* __restrict__ rem_row, u32* __restrict__ rem_col,
                                                                   u32* __restrict__ rem_count, int8_t* __restrict__ rem_mapq) {
    extern __shared__ __align__(16) unsigned char pt_raw[];
    PtSmemT<WARP>& S = *reinterpret_cast<PtSmemT<WARP>*>(pt_raw);
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid == 0) {
        S.tile = atomicAdd(&scal[3], 1u);        // ticket: every lower tile has already started
        S.long_order = 0;
        S.nheads = 0;
        if (use_tma) mbar_init(&S.mbar, 1);
    }
    __syncthreads();
    const u32 tile = S.tile;
    const size_t lo = (size_t)tile * PT_TILE;
    const size_t N = in.n;
    const int own = (int)min((size_t)PT_TILE, N - lo);           // records of the tile proper
    const int wn = (int)min((size_t)PT_WIN, N - lo);             // records staged
    if (use_tma && wn == PT_WIN) {
        // full window, 16-byte aligned columns, no edge filter: three bulk copies, then an in-place pack of
        // (contig, mapq) -> cvq.  The tile origin lo is a multiple of PT_TILE, so every source stays aligned.
        if (tid == 0) {
            mbar_expect_tx(&S.mbar, PT_WIN * 9u);
            bulk_g2s(S.read, in.read + lo, PT_WIN * 4u, &S.mbar);
            bulk_g2s(S.cvq, in.contig + lo, PT_WIN * 4u, &S.mbar);
            bulk_g2s(S.mq, in.mapq + lo, PT_WIN, &S.mbar);
        }
        mbar_wait(&S.mbar, 0);
        for (int i = tid; i < PT_WIN; i += PT_THREADS) {
            const u32 c = S.cvq[i];
            const int8_t q = (int8_t)S.mq[i];
            bool ok = c < in.n_contigs;
            if (in.min_mapq > 0) ok = ok && (int)q >= in.min_mapq;                  // hypergraph.py:1774-1775
            S.cvq[i] = ((ok ? c : PT_INVALID) << 8) | (u32)(u8)((int)q + 128);
        }
    } else {
        for (int i = tid; i < wn; i += PT_THREADS) {
            size_t t = lo + i;
            S.read[i] = in.read[t];
            S.cvq[i] = porec_cvq(in, t);
        }
    }
    __syncthreads();
    // read starts of the window as a bitmask
    const u32 prev_read = lo > 0 ? in.read[lo - 1] : 0u;
    const bool tail_end = (lo + (size_t)wn == N);                // the table ends inside the window
    if (WARP && tid < 3) S.head[PT_WORDS + tid] = (tid == 0 && tail_end && wn == PT_WIN) ? 1u : 0u;
    for (int i0 = 0; i0 < PT_WIN; i0 += PT_THREADS) {
        int i = i0 + tid;
        bool head = false;
        if (i < wn) {
            u32 r = S.read[i];
            u32 p = i > 0 ? S.read[i - 1] : prev_read;
            head = (lo + i == 0) || (p != r);
            if (i < own && (lo + i) > 0 && p > r) scal[1] = 1u;   // not grouped by read (benign race)
        } else if (WARP && i == wn && tail_end) {
            head = true;        // warp kernel: a virtual read start closes the last read of the table (never counted: is_long
        }                       // needs lo + wn != N, and nr_own only looks below `own` <= wn)
        u32 m = __ballot_sync(0xffffffffu, head);
        if (lane == 0) S.head[i >> 5] = m;
    }
    __syncthreads();
    pt_word_prefix(S.head, S.hpref, tid);
    __syncthreads();
    // a tile OWNS the reads that start inside it (positions < own); the first head at or after `own` ends the last one
    const int nh_total = (int)S.hpref[PT_WORDS];
    const int nr_own = (int)(S.hpref[own >> 5] + __popc(S.head[own >> 5] & ((1u << (own & 31)) - 1u)));
    const bool empty = nr_own == 0;          // a longer read covers the whole tile: its owner emits it
    // the last read of the tile overflows the window (no later head, table not finished): slow path below
    const bool is_long = !empty && nh_total == nr_own && lo + wn != N;
    const int long_s = is_long ? pt_prev_head(S.head, own - 1) : PT_WIN;
    const int fend = long_s;                 // only meaningful when is_long
    // One THREAD per read (hypergraph.py:1461-1477).  Thread t owns the reads that start inside its PT_RPT-record slice
    // of the tile -- their records add up to about PT_RPT whatever the read lengths are, so the pass is balanced by
    // construction.  It walks each read once, keeping the distinct contigs IN PLACE at the front of the read's own span
    // (slot s + d: d-th distinct contig in first-seen order, its record count and its largest MAPQ).  The number of
    // distinct contigs is the read's order; reads outside [min_order, max_order] emit nothing.  No atomics.
    const int sl_lo = tid * PT_RPT, sl_hi = min(sl_lo + PT_RPT, own);
    u32 my_max_read = 0, my_out = 0;
    if constexpr (!WARP) {
    {
        // ONE flat loop over the records of the slice's reads (a nested read / record loop makes the lanes of a warp
        // drift apart at every read boundary).  The first four distinct contigs of the current read live in registers,
        // so the usual order <= 4 never searches the list in shared memory.
        int j = pt_next_head(S.head, sl_lo);
        int jend = j;
        if (j < sl_hi && j != long_s) {
            jend = min(pt_next_head(S.head, sl_hi), wn);              // end of the last read that starts in the slice
            if (is_long && jend > long_s) jend = long_s;              // ... the overflowing read is not ours to walk
        }
        int s = j, nd = 0, lf = 0;
        u32 c0 = PT_INVALID, c1 = PT_INVALID, c2 = PT_INVALID, c3 = PT_INVALID, lc = PT_INVALID;
        auto finish = [&]() {
            const bool keep = (u32)nd >= min_order && (u32)nd <= max_order;
            S.rn[s] = keep ? (u16)nd : (u16)0;
            if (keep) {
                my_out += (u32)nd;
                my_max_read = max(my_max_read, S.read[s]);
            }
        };
        for (; j < jend; ++j) {
            if (j != s && ((S.head[j >> 5] >> (j & 31)) & 1u)) {      // next read of the slice starts here
                finish();
                s = j;
                nd = 0;
                lc = c0 = c1 = c2 = c3 = PT_INVALID;
            }
            const u32 x = S.cvq[j], c = x >> 8;
            if (c == PT_INVALID) continue;
            int f;
            if (c == lc) f = lf;                                      // adjacent duplicate: the likeliest case
            else if (c == c0) f = 0;
            else if (c == c1) f = 1;
            else if (c == c2) f = 2;
            else if (c == c3) f = 3;
            else {
                f = nd < 4 ? nd : 4;
                while (f < nd && (S.cvq[s + f] >> 8) != c) ++f;
            }
            if (f == nd) {
                S.cvq[s + nd] = x;
                S.cnt[s + nd] = 1;
                if (nd == 0) c0 = c;
                else if (nd == 1) c1 = c;
                else if (nd == 2) c2 = c;
                else if (nd == 3) c3 = c;
                ++nd;
            } else {
                S.cnt[s + f] += 1;
                if (x > S.cvq[s + f]) S.cvq[s + f] = x;      // same contig bits, larger MAPQ byte wins
            }
            lc = c;
            lf = f;
        }
        if (jend > s) finish();
    }
    } else {
        // One WARP per chunk of PT_WCH records, one LANE per record (hypergraph.py:1461-1477).  A round takes the next <= 32
        // records that end on a read boundary: the lanes of one (read, contig) group find each other with MATCH.ANY on the key
        // (read start lane << 24 | contig); the lowest lane of a group is its representative (first-seen order), the group's
        // size is the record count, a REDUX.MAX over the group gives the largest MAPQ, the representatives inside the read's
        // lane span are its distinct-contig order.  Kept entries are written back IN PLACE, compacted, at the front of the
        // warp's own part of the window (output never overtakes input), so the copy-out below is a contiguous run per warp.
        constexpr int PT_WCH = PT_TILE / (PT_THREADS / 32);
        const int warp = tid >> 5;
        const int w_lo = warp * PT_WCH, w_hi = min(w_lo + PT_WCH, own);
        int p = w_lo < own ? pt_next_head(S.head, w_lo) : PT_WIN;
        const int p0 = p;
        int o = p;
        // the warp's records: [p0, stop) = whole reads, from the first read start in its chunk to the first one after it
        // (the next warp's p0); the read that overflows the window (long_s) belongs to the slow path
        int stop = p < w_hi ? min(pt_next_head(S.head, w_hi), wn) : p;
        if (long_s >= p && long_s < stop) stop = long_s;
        const u32 lt = lanemask_lt();
        const u32 self = 1u << lane;
        while (p < stop) {
            const int hw = p >> 5;
            const u32 sh = (u32)(p & 31);
            const u32 ha = S.head[hw], hb = S.head[hw + 1], hc = S.head[hw + 2];
            const u32 hm = __funnelshift_r(ha, hb, sh);          // bit l: record p + l starts a read (bit 0 is set)
            u32 em = (hm >> 1) | (__funnelshift_r(hb, hc, sh) << 31);   // bit l: record p + l is the last of its read
            if (stop - p < 32) em &= (1u << (stop - p)) - 1u;
            if (em != 0u) {
                const int L = 31 - __clz(em);                    // the round covers lanes 0..L (whole reads only)
                const bool act = lane <= L;
                const int i = p + lane;
                const u32 x = act ? S.cvq[i] : (PT_INVALID << 8);
                const u32 r = act ? S.read[i] : 0u;
                const int s_l = 31 - __clz(hm & (lt | self));    // lane where my read starts
                const u32 nxt = hm & ~(lt | self);                // read starts after my lane
                const u32 below_next = nxt ? ((nxt & (0u - nxt)) - 1u) : 0xffffffffu;
                const u32 span = below_next & ~((1u << s_l) - 1u);   // lanes of my read
                const u32 c = x >> 8;
                const bool valid = act && c != PT_INVALID;
                const u32 key = valid ? (((u32)s_l << 24) | c) : (0x80000000u | (u32)lane);
                const u32 peers = __match_any_sync(0xffffffffu, key);
                const bool first = valid && (peers & lt) == 0u;
                // largest MAPQ of the group, gathered by its first lane: the other lanes of a group push theirs with one
                // shared-memory atomicMax on the first lane's slot (a REDUX over per-group masks is serialised by the
                // hardware, one group at a time; a shuffle loop over the peers cost 20 % of the kernel's instructions)
                u32 q = x & 0xffu;
                if (__any_sync(0xffffffffu, valid && peers != self)) {
                    u32* qslot = reinterpret_cast<u32*>(S.mq) + (tid & ~31);   // the MAPQ staging bytes are free by now
                    if (first) qslot[lane] = q;
                    __syncwarp();
                    if (valid && !first) atomicMax(&qslot[__ffs(peers) - 1], q);
                    __syncwarp();
                    if (first) q = qslot[lane];
                }
                const u32 rm = __ballot_sync(0xffffffffu, first);
                const u32 nd = __popc(rm & span);
                const bool emit = first && nd >= min_order && nd <= max_order;
                const u32 om = __ballot_sync(0xffffffffu, emit);
                __syncwarp();                                    // every lane holds its record in registers from here on
                if (emit) {
                    const int slot = o + __popc(om & lt);
                    const bool hd = (om & span & lt) == 0u;      // first kept entry of its read = head of the hyperedge
                    S.cvq[slot] = (c << 8) | q;
                    S.read[slot] = r;
                    S.cnt[slot] = (u16)((u32)__popc(peers) | (hd ? 0x8000u : 0u));
                    S.dk[slot] = (u8)pt_degree(nd);
                    my_max_read = max(my_max_read, r);
                }
                __syncwarp();
                o += __popc(om);
                p += L + 1;
            } else {
                // a read of more than 32 records that ends inside the window: 32 records at a time against the list of
                // distinct contigs found so far, kept in place at the front of the read's own span
                int e = min(pt_next_head(S.head, p + 1), stop);
                const u32 r = S.read[p];
                int nd = 0;
                for (int b0 = p; b0 < e; b0 += 32) {
                    const int i = b0 + lane;
                    const bool act = i < e;
                    const u32 x = act ? S.cvq[i] : (PT_INVALID << 8);
                    const u32 c = x >> 8;
                    const bool valid = act && c != PT_INVALID;
                    const u32 key = valid ? c : (0x80000000u | (u32)lane);
                    const u32 peers = __match_any_sync(0xffffffffu, key);
                    const bool first = valid && (peers & lt) == 0u;
                    u32 q = x & 0xffu;
                    {
                        const u32 q0 = q;
                        u32 others = first ? (peers & ~self) : 0u;
                        while (__any_sync(0xffffffffu, others != 0u)) {
                            const int srcl = others ? (__ffs(others) - 1) : lane;
                            const u32 oq = __shfl_sync(0xffffffffu, q0, srcl);
                            q = max(q, oq);
                            others &= others - 1u;
                        }
                    }
                    const u32 gc = (u32)__popc(peers);
                    __syncwarp();
                    int f = -1;
                    if (first)
                        for (int d = 0; d < nd; ++d)
                            if ((S.cvq[p + d] >> 8) == c) {
                                f = d;
                                break;
                            }
                    const u32 nm = __ballot_sync(0xffffffffu, first && f < 0);
                    if (first) {
                        if (f >= 0) {                            // distinct lanes hold distinct contigs: no two touch one slot
                            S.cnt[p + f] = (u16)(S.cnt[p + f] + gc);
                            const u32 old = S.cvq[p + f];
                            if (((c << 8) | q) > old) S.cvq[p + f] = (c << 8) | q;
                        } else {
                            const int d = nd + __popc(nm & lt);
                            S.cvq[p + d] = (c << 8) | q;
                            S.cnt[p + d] = (u16)gc;
                        }
                    }
                    nd += __popc(nm);
                    __syncwarp();
                }
                if ((u32)nd >= min_order && (u32)nd <= max_order) {
                    for (int d0 = 0; d0 < nd; d0 += 32) {        // slide the list down to the output position (o <= p)
                        const int d = d0 + lane;
                        u32 x = 0;
                        u16 cn = 0;
                        if (d < nd) {
                            x = S.cvq[p + d];
                            cn = S.cnt[p + d];
                        }
                        __syncwarp();
                        if (d < nd) {
                            S.cvq[o + d] = x;
                            S.cnt[o + d] = (u16)((cn & 0x7fffu) | (d == 0 ? 0x8000u : 0u));
                            S.dk[o + d] = (u8)pt_degree((u32)nd);
                            S.read[o + d] = r;
                        }
                        __syncwarp();
                    }
                    o += nd;
                    my_max_read = max(my_max_read, r);
                }
                p = e;
            }
        }
        if (lane == 0) {
            S.wstart[warp] = (u32)p0;
            S.wcnt[warp] = (u32)(o - p0);
        }
    }
    // exclusive prefix of the per-thread output counts
    u32 fast_total = 0, my_ex = 0;
    if constexpr (!WARP) {
        my_ex = block_excl_scan<u32, PT_THREADS>(my_out, S.scan, &fast_total);
    } else {
        __syncthreads();
        // exclusive prefix of the warps' output counts up to this thread's own warp, and the tile total
#pragma unroll
        for (int w = 0; w < PT_THREADS / 32; ++w) {
            const u32 c = S.wcnt[w];
            if (w < (tid >> 5)) my_ex += c;
            fast_total += c;
        }
    }
    // slow path, part 1: order of the read that starts at `fend` and runs past the window
    u32 long_kept = 0;
    size_t ls = 0, le = 0;
    u32 rid = 0;
    auto long_rep = [&](size_t t) -> bool {
        u32 c = porec_cvq(in, t) >> 8;
        if (c == PT_INVALID) return false;
        for (size_t u = t; u-- > ls;)
            if ((porec_cvq(in, u) >> 8) == c) return false;
        return true;
    };
    if (is_long) {
        ls = lo + fend;
        rid = S.read[fend];
        if (tid == 0) {
            size_t e = lo + wn;
            while (e < N && in.read[e] == rid) ++e;
            S.long_end = e;
        }
        __syncthreads();
        le = S.long_end;
        u32 nrep = 0;
        for (size_t t = ls + tid; t < le; t += PT_THREADS) nrep += long_rep(t) ? 1u : 0u;
        nrep = warp_sum(nrep);
        if (lane == 0 && nrep) atomicAdd(&S.long_order, nrep);
        __syncthreads();
        const u32 order = S.long_order;
        if (order >= min_order && order <= max_order) {
            long_kept = order;
            my_max_read = max(my_max_read, rid);
        }
    }
    const u32 total = fast_total + long_kept;
    // decoupled look-back (warp 0): exclusive prefix of the tile totals
    if (tid < 32) {
        u32 base = 0;
        // the state word (flag + value in one aligned 64-bit word) is published with a plain volatile store: nothing waits
        // for it here, where an atomicExch would hold warp 0 -- and the seven warps behind the barrier -- for an L2 round trip
        if (tile == 0) {
            if (lane == 0) *((volatile unsigned long long*)&tile_state[0]) = PT_INCL | (u64)total;
        } else {
            if (lane == 0) *((volatile unsigned long long*)&tile_state[tile]) = PT_AGG | (u64)total;
            u64 running = 0;
            long j = (long)tile - 1;
            for (;;) {
                long idx = j - lane;
                u64 st = idx >= 0 ? *((volatile unsigned long long*)&tile_state[idx]) : PT_INCL;
                u32 inval = __ballot_sync(0xffffffffu, (st >> 62) == 0);
                u32 incl = __ballot_sync(0xffffffffu, (st >> 62) == 2);
                u32 first = incl ? (u32)__ffs(incl) - 1u : 32u;
                u32 need = first >= 31u ? 0xffffffffu : ((2u << first) - 1u);
                if (inval & need) {                         // a predecessor has not published yet: re-read
                    __nanosleep(40);                        // ... without taking issue slots from the warps still working
                    continue;
                }
                u64 v = ((u32)lane <= first) ? (st & PT_VAL) : 0ull;
                running += warp_sum(v);
                if (incl) break;
                j -= 32;
            }
            base = (u32)running;
            if (lane == 0) *((volatile unsigned long long*)&tile_state[tile]) = PT_INCL | (running + total);
        }
        if (lane == 0) S.base = base;
    }
    // output slot -> window position, so that the global stores below are coalesced (warps 1.. do this while
    // warp 0 is still looking back)
    if constexpr (!WARP) {
        u32 o = my_ex;
        for (int s = pt_next_head(S.head, sl_lo); s < sl_hi && s != long_s; s = pt_next_head(S.head, s + 1)) {
            const int nd = S.rn[s];
            for (int d = 0; d < nd; ++d) S.src[o + d] = (u16)(s + d);
            o += nd;
        }
    }
    __syncthreads();
    const u32 base = S.base;
    if constexpr (!WARP) {
        for (u32 p = tid; p < fast_total; p += PT_THREADS) {
            const int pos = S.src[p];
            const u32 x = S.cvq[pos];
            const bool first = (S.head[pos >> 5] >> (pos & 31)) & 1u;             // slot s: first entry of the hyperedge
            rem_row[base + p] = (x >> 8) | (first ? CPH_HEAD : 0u);
            rem_col[base + p] = S.read[pos] + read_offset;                      // generate_edges :1871-1875
            rem_count[base + p] = (u32)S.cnt[pos];
            rem_mapq[base + p] = (int8_t)((int)(x & 0xffu) - 128);
        }
    } else {
        // every warp copies its own compacted run out: consecutive lanes, consecutive addresses
        // The count word also carries D_e of the entry's hyperedge in its top byte (D_e << 24 | count = the packed form the
        // clique expansion reads): when the incidence stage has nothing to drop, these arrays ARE the packed incidence.
        const u32 ws = S.wstart[tid >> 5], wc = S.wcnt[tid >> 5];
        u32 nheads = 0;
        for (u32 k = lane; k < wc; k += 32) {
            const u32 x = S.cvq[ws + k];
            const u32 cn = S.cnt[ws + k];
            const size_t p = (size_t)base + my_ex + k;
            rem_row[p] = (x >> 8) | ((cn & 0x8000u) ? CPH_HEAD : 0u);
            rem_col[p] = S.read[ws + k] + read_offset;                          // generate_edges :1871-1875
            rem_count[p] = (cn & 0x7fffu) | ((u32)S.dk[ws + k] << 24);
            rem_mapq[p] = (int8_t)((int)(x & 0xffu) - 128);
            nheads += (cn >> 15) & 1u;
        }
        nheads = warp_sum(nheads);
        if (lane == 0 && nheads) atomicAdd(&S.nheads, nheads);
        __syncthreads();
        if (tid == 0) {
            const u32 nh = S.nheads + (long_kept ? 1u : 0u);
            if (nh) atomicAdd(&scal[2], nh);                             // hyperedges emitted by the whole call
        }
    }
    if (long_kept) {     // slow path, part 2: emit the long read after the tile's fast entries
        u32 carry = 0;
        for (size_t t0 = ls; t0 < le; t0 += PT_THREADS) {
            size_t t = t0 + tid;
            bool rp = t < le && long_rep(t);
            u32 tot;
            u32 ex = block_excl_scan<u32, PT_THREADS>(rp ? 1u : 0u, S.scan, &tot);
            if (rp) {
                u32 c = porec_cvq(in, t) >> 8, cnt = 1;
                int q = in.mapq[t];
                for (size_t u = t + 1; u < le; ++u)
                    if ((porec_cvq(in, u) >> 8) == c) {
                        ++cnt;
                        q = max(q, (int)in.mapq[u]);
                    }
                u32 k = carry + ex;
                u32 p = base + fast_total + k;
                rem_row[p] = c | (k == 0 ? CPH_HEAD : 0u);
                rem_col[p] = rid + read_offset;
                rem_count[p] = WARP ? (cnt | (pt_degree(long_kept) << 24)) : cnt;
                rem_mapq[p] = (int8_t)q;
            }
            carry += tot;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_max_read = max(my_max_read, __shfl_xor_sync(0xffffffffu, my_max_read, o));
    if (lane == 0 && total > 0 && my_max_read) atomicMax(&scal[0], my_max_read);
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
// host cph_prefilter -> device view; false when the combination is not supported
static bool pf_view(const cph_prefilter* pf, PreF* out) {
    *out = PreF{nullptr, nullptr, nullptr, nullptr, 0, 0, 0};
    if (pf == nullptr) return true;
    const bool hcr = pf->bed_ptr != nullptr, split = pf->split_length != 0;
    if (!hcr && !split) return true;
    if (pf->n_names == 0) return false;
    if (hcr && !(pf->bed_start && pf->bed_end)) return false;
    if (split && !pf->split_init) return false;
    out->bed_ptr = hcr ? pf->bed_ptr : nullptr;
    out->bed_start = pf->bed_start;
    out->bed_end = pf->bed_end;
    out->split_init = split ? pf->split_init : nullptr;
    out->split_length = pf->split_length;
    out->n_names = pf->n_names;
    out->hcr_invert = pf->hcr_invert;
    return true;
}

extern "C" int cph_pairs_to_edges(const u32* chrom1, const u32* chrom2, const int8_t* mapq, const u32* pos1, const u32* pos2,
                                  const i64* contig_len, u64 n_pairs, u32 n_contigs, int min_mapq, u64 edge_length, u32* rem_row,
                                  u32* rem_col, u32* rem_count, int8_t* rem_mapq, u64* h_counts, void* ws, size_t ws_bytes,
                                  size_t* h_ws_need, void* stream) {
    return cph_pairs_to_edges_pf(chrom1, chrom2, mapq, pos1, pos2, contig_len, n_pairs, n_contigs, min_mapq, edge_length, nullptr,
                                 rem_row, rem_col, rem_count, rem_mapq, h_counts, ws, ws_bytes, h_ws_need, stream);
}

extern "C" int cph_pairs_to_edges_pf(const u32* chrom1, const u32* chrom2, const int8_t* mapq, const u32* pos1, const u32* pos2,
                                     const i64* contig_len, u64 n_pairs, u32 n_contigs, int min_mapq, u64 edge_length,
                                     const cph_prefilter* prefilter, u32* rem_row, u32* rem_col, u32* rem_count, int8_t* rem_mapq,
                                     u64* h_counts, void* ws, size_t ws_bytes, size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(n_pairs < (1ull << 31));
    ARG_CHECK(n_contigs > 0 && n_contigs < CPH_HEAD);
    PreF pfv;
    ARG_CHECK(pf_view(prefilter, &pfv));
    const bool pre = pfv.bed_ptr != nullptr || pfv.split_init != nullptr;
    ARG_CHECK(!pre || n_pairs == 0 || (pos1 && pos2));
    // the reference joins the contig lengths on the ORIGINAL names while contigsizes holds the split names (:977-992 with
    // split on): nothing survives.  Not reproduced -- refuse the combination instead of guessing.
    ARG_CHECK(!(pfv.split_init && edge_length));
    WS_BEGIN(ar);
    u32* keep = ar.take<u32>(n_pairs + 1);
    u32* total = ar.take<u32>(1);
    PairsIn in{chrom1, chrom2, mapq, pos1, pos2, contig_len, n_pairs, n_contigs, min_mapq, (i64)edge_length, pfv};
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

// which tile kernel runs: the warp-cooperative one (default) or the per-read-slice walk (CPH_POREC_KERNEL=walk; kept for
// A/B measurements -- profiles/)
static int pt_mode() {      // 0: walk   1: warp, 6 CTAs per SM   2 (default): warp, 8 CTAs per SM (32 registers)
    static int m = -1;      // measured on workload C: walk 3.47 ms, warp 3.00 ms, warp8 2.77 ms (profiles/r2_porec_kernels.txt)
    if (m < 0) {
        const char* e = getenv("CPH_POREC_KERNEL");
        m = (e && strcmp(e, "walk") == 0) ? 0 : (e && strcmp(e, "warp6") == 0) ? 1 : 2;
    }
    return m;
}
static cudaError_t pt_prepare() {
    cudaError_t e = cudaFuncSetAttribute(porec_tile_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PtSmemT<false>));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(porec_tile_kernel<true, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PtSmemT<true>));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(porec_tile_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PtSmemT<true>));
}
static void pt_launch(u32 grid, cudaStream_t st, const PorecIn& in, int use_tma, u32 min_order, u32 max_order, u32 read_offset,
                      unsigned long long* tile_state, u32* scal, u32* rem_row, u32* rem_col, u32* rem_count, int8_t* rem_mapq) {
    const int m = pt_mode();
    if (m == 2)
        porec_tile_kernel<true, 8><<<grid, PT_THREADS, sizeof(PtSmemT<true>), st>>>(in, use_tma, min_order, max_order, read_offset,
                                                                                    tile_state, scal, rem_row, rem_col, rem_count, rem_mapq);
    else if (m == 1)
        porec_tile_kernel<true><<<grid, PT_THREADS, sizeof(PtSmemT<true>), st>>>(in, use_tma, min_order, max_order, read_offset, tile_state,
                                                                                 scal, rem_row, rem_col, rem_count, rem_mapq);
    else
        porec_tile_kernel<false><<<grid, PT_THREADS, sizeof(PtSmemT<false>), st>>>(in, use_tma, min_order, max_order, read_offset, tile_state,
                                                                                   scal, rem_row, rem_col, rem_count, rem_mapq);
}

extern "C" int cph_porec_to_edges(const u32* read_idx, const u32* contig_idx, const int8_t* mapq, const u32* start, const u32* end,
                                  const i64* contig_len, u64 n_records, u32 n_contigs, int min_mapq, u32 min_order, u32 max_order,
                                  u64 edge_length, u32 read_offset, int sorted_by_read, u32* rem_row, u32* rem_col, u32* rem_count,
                                  int8_t* rem_mapq, u64* h_counts, void* ws, size_t ws_bytes, size_t* h_ws_need, void* stream) {
    return cph_porec_to_edges_pf(read_idx, contig_idx, mapq, start, end, contig_len, n_records, n_contigs, min_mapq, min_order,
                                 max_order, edge_length, read_offset, sorted_by_read, nullptr, rem_row, rem_col, rem_count, rem_mapq,
                                 h_counts, ws, ws_bytes, h_ws_need, stream);
}

extern "C" int cph_porec_to_edges_pf(const u32* read_idx, const u32* contig_idx, const int8_t* mapq, const u32* start, const u32* end,
                                     const i64* contig_len, u64 n_records, u32 n_contigs, int min_mapq, u32 min_order, u32 max_order,
                                     u64 edge_length, u32 read_offset, int sorted_by_read, const cph_prefilter* prefilter,
                                     u32* rem_row, u32* rem_col, u32* rem_count, int8_t* rem_mapq, u64* h_counts, void* ws,
                                     size_t ws_bytes, size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const u64 N = n_records;
    ARG_CHECK(N < (1ull << 32) - 8192);
    ARG_CHECK(n_contigs > 0 && n_contigs < (1u << 23));
    PreF pfv;
    ARG_CHECK(pf_view(prefilter, &pfv));
    const bool pre = pfv.bed_ptr != nullptr || pfv.split_init != nullptr;
    ARG_CHECK(!pre || N == 0 || (start && end));
    const int cb = bits_for(n_contigs);   // contig ids < n_contigs <= 2^cb - 1
    WS_BEGIN(ar);
    PorecIn in{read_idx, contig_idx, mapq, start, end, contig_len, N, n_contigs, min_mapq, (i64)edge_length, pfv};
    // ---- fast path: table grouped by read (two passes of the shared-memory tile kernel) ----
    const u32 ntiles = (u32)ceil_div(N > 0 ? N : 1, (size_t)PT_TILE);
    u32* scal = ar.take<u32>(4);   // [0] max kept read (fast) / max read (slow)  [1] unsorted flag  [2] scan total  [3] ticket
    u64* tile_state = ar.take<u64>((size_t)ntiles);
    if (!ar.dry) {
        ARENA_CHECK(ar);
        ARG_CHECK(N == 0 || (read_idx && contig_idx && mapq));
        ARG_CHECK(rem_row && rem_col && rem_count && rem_mapq && h_counts);
        ARG_CHECK(N == 0 || edge_length == 0 || (start && end && contig_len));
        h_counts[4] = h_counts[5] = 0;
        if (N == 0) {
            h_counts[0] = h_counts[1] = 0;
            return CPH_OK;
        }
        CUDA_TRY(cudaMemsetAsync(scal, 0, 4 * sizeof(u32), st));
        CUDA_TRY(cudaMemsetAsync(tile_state, 0, (size_t)ntiles * sizeof(u64), st));
        CUDA_TRY(pt_prepare());
        // bulk (TMA) window loads need 16-byte aligned columns; the edge filter reads two more columns -> plain loads
        const int use_tma = edge_length == 0 && !pre && (((uintptr_t)read_idx | (uintptr_t)contig_idx | (uintptr_t)mapq) & 15) == 0;
        pt_launch(ntiles, st, in, use_tma, min_order, max_order, read_offset, (unsigned long long*)tile_state, scal, rem_row, rem_col,
                  rem_count, rem_mapq);
        KERNEL_CHECK();
        u32 hs[4] = {0, 0, 0, 0};
        u64 last = 0;
        CUDA_TRY(cudaMemcpyAsync(hs, scal, 3 * sizeof(u32), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaMemcpyAsync(&last, tile_state + (ntiles - 1), sizeof(u64), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        if (hs[1] == 0) {
            const u64 total = last & ((1ull << 62) - 1);
            h_counts[0] = total;
            h_counts[1] = total ? (u64)hs[0] + read_offset + 1 : 0;
            h_counts[4] = hs[2];                            // hyperedges emitted
            h_counts[5] = pt_mode() != 0 ? 1 : 0;           // the warp kernels pack D_e into the top byte of rem_count
            return CPH_OK;
        }
    }
    // ---- general path: not grouped by read -> key, radix sort by read, run-local kernels ----
    u64* ka = ar.take<u64>(N);
    u64* kb = ar.take<u64>(N);
    u8* rep = ar.take<u8>(N);
    u32* keep = ar.take<u32>(N + 1);
    u32* aux = ar.take<u32>(N);
    int8_t* mq = ar.take<int8_t>(N);
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


namespace {
// ------------------------------------------------------------------------------------------------
// Packed host transfer.  The Pore-C table crosses PCIe at 9 B per record (read u32, contig u32, mapq i8) and PCIe, not the
// kernels, bounds the end-to-end call.  A table grouped by read has small read-id steps and (below 65536 contigs) 16-bit
// contig ids, so host threads re-encode each chunk into pinned staging as
//     delta u8 (read[i] - read[i-1]) | mapq i8 | contig u16 (u32 when n_contigs > 65536) | anchor u32 per 1024 records
// = 4 B per record, and pk_unpack_kernel restores the three device columns bit for bit (anchor + block-local prefix sum of
// the deltas).  A chunk with a read-id step outside 0..255 (or an out-of-range contig id) is copied as it is.  The caller's
// arrays need not be pinned: only the staging ring is.
// ------------------------------------------------------------------------------------------------
constexpr u64 PK_CHUNK = 1u << 20;        // records per packed chunk (default)
constexpr u32 PK_GROUP = 1024;            // records per anchor = per CTA of the unpack kernel
constexpr int PK_NSLOT = 32;              // pinned host slots
constexpr int PK_NDEV = 8;                // device slots

struct PkLayout {
    size_t off_delta, off_mapq, off_contig, off_anchor, bytes;
};
inline PkLayout pk_layout(u64 len, int wide) {
    const size_t L = ceil_div((size_t)len, (size_t)PK_GROUP) * PK_GROUP;
    PkLayout o;
    o.off_delta = 0;
    o.off_mapq = L;
    o.off_contig = 2 * L;
    o.off_anchor = o.off_contig + L * (wide ? 4 : 2);
    o.bytes = o.off_anchor + (L / PK_GROUP) * sizeof(u32);
    o.bytes = (o.bytes + 255) & ~(size_t)255;
    return o;
}

// streaming (non-temporal) stores: the packed bytes are read next by the DMA engine, not by this core, and a normal
// store would first fetch the destination line (a third of the host memory traffic of the packing pass)
inline void pk_stream_copy(unsigned char* __restrict__ dst, const unsigned char* __restrict__ src, u64 n) {    // dst 16 B aligned
    u64 i = 0;
    for (; i + 16 <= n; i += 16) _mm_stream_si128((__m128i*)(dst + i), _mm_loadu_si128((const __m128i*)(src + i)));
    for (; i < n; ++i) dst[i] = src[i];
}

// host: records [lo, hi) -> slot; false when the chunk cannot be packed
bool pk_pack(const u32* __restrict__ hr, const u32* __restrict__ hc, const int8_t* __restrict__ hq, u64 lo, u64 hi, int wide,
             unsigned char* __restrict__ slot) {
    const u64 len = hi - lo;
    const PkLayout L = pk_layout(len, wide);
    unsigned char* __restrict__ d = slot + L.off_delta;
    const u32* __restrict__ r = hr + lo;
    u32 bad = lo > 0 ? r[0] - hr[lo - 1] : 0u;        // order across the chunk border (the anchor carries the value)
    d[0] = 0;
    u64 i = 1;
    for (; i < 16 && i < len; ++i) {
        const u32 dl = r[i] - r[i - 1];
        bad |= dl;
        d[i] = (unsigned char)dl;
    }
    __m128i vb = _mm_setzero_si128();
    for (; i + 16 <= len; i += 16) {
        const __m128i d0 = _mm_sub_epi32(_mm_loadu_si128((const __m128i*)(r + i)), _mm_loadu_si128((const __m128i*)(r + i - 1)));
        const __m128i d1 = _mm_sub_epi32(_mm_loadu_si128((const __m128i*)(r + i + 4)), _mm_loadu_si128((const __m128i*)(r + i + 3)));
        const __m128i d2 = _mm_sub_epi32(_mm_loadu_si128((const __m128i*)(r + i + 8)), _mm_loadu_si128((const __m128i*)(r + i + 7)));
        const __m128i d3 = _mm_sub_epi32(_mm_loadu_si128((const __m128i*)(r + i + 12)), _mm_loadu_si128((const __m128i*)(r + i + 11)));
        vb = _mm_or_si128(vb, _mm_or_si128(_mm_or_si128(d0, d1), _mm_or_si128(d2, d3)));
        // (steps above 255 saturate in the packs; they are caught through vb and the chunk is then not used)
        _mm_stream_si128((__m128i*)(d + i), _mm_packus_epi16(_mm_packs_epi32(d0, d1), _mm_packs_epi32(d2, d3)));
    }
    for (; i < len; ++i) {
        const u32 dl = r[i] - r[i - 1];
        bad |= dl;
        d[i] = (unsigned char)dl;
    }
    alignas(16) u32 lanes[4];
    _mm_store_si128((__m128i*)lanes, vb);
    bad |= lanes[0] | lanes[1] | lanes[2] | lanes[3];
    if (bad >> 8) return false;
    pk_stream_copy(slot + L.off_mapq, (const unsigned char*)(hq + lo), len);
    const u32* __restrict__ c = hc + lo;
    if (wide) pk_stream_copy(slot + L.off_contig, (const unsigned char*)c, len * 4);
    else {
        uint16_t* __restrict__ c16 = (uint16_t*)(slot + L.off_contig);
        __m128i vc = _mm_setzero_si128();
        u32 badc = 0;
        u64 k = 0;
        for (; k + 8 <= len; k += 8) {
            const __m128i a0 = _mm_loadu_si128((const __m128i*)(c + k)), a1 = _mm_loadu_si128((const __m128i*)(c + k + 4));
            vc = _mm_or_si128(vc, _mm_or_si128(a0, a1));
            // low 16 bits of each lane, exactly: sign-extend them so that the signed pack cannot saturate
            _mm_stream_si128((__m128i*)(c16 + k), _mm_packs_epi32(_mm_srai_epi32(_mm_slli_epi32(a0, 16), 16),
                                                                 _mm_srai_epi32(_mm_slli_epi32(a1, 16), 16)));
        }
        for (; k < len; ++k) {
            badc |= c[k];
            c16[k] = (uint16_t)c[k];
        }
        _mm_store_si128((__m128i*)lanes, vc);
        badc |= lanes[0] | lanes[1] | lanes[2] | lanes[3];
        if (badc >> 16) return false;
    }
    u32* a = (u32*)(slot + L.off_anchor);
    for (u64 g = 0; g * PK_GROUP < len; ++g) a[g] = r[g * PK_GROUP];
    _mm_sfence();        // streaming stores are weakly ordered: make them visible before the chunk is announced
    return true;
}

// one CTA per 1024 records, 4 records per thread
template <bool WIDE>
__global__ void __launch_bounds__(256) pk_unpack_kernel(const unsigned char* __restrict__ slot, PkLayout L, u64 len,
                                                        u32* __restrict__ d_read, u32* __restrict__ d_contig,
                                                        int8_t* __restrict__ d_mapq) {
    __shared__ u32 wsum[8];
    const u32 t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const u64 i0 = (u64)blockIdx.x * PK_GROUP + 4 * t;          // multiple of 4; the slot is padded to whole groups
    const uchar4 dl = *(const uchar4*)(slot + L.off_delta + i0);
    const u32 anchor = ((const u32*)(slot + L.off_anchor))[blockIdx.x];
    u32 v0 = t == 0 ? 0u : dl.x, v1 = v0 + dl.y, v2 = v1 + dl.z, v3 = v2 + dl.w;
    u32 x = v3;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u32 y = __shfl_up_sync(0xffffffffu, x, o);
        if ((int)lane >= o) x += y;
    }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    u32 base = anchor + x - v3;
    for (u32 w = 0; w < warp; ++w) base += wsum[w];
    uint4 c;
    if (WIDE) c = *(const uint4*)(slot + L.off_contig + i0 * 4);
    else {
        const ushort4 h = *(const ushort4*)(slot + L.off_contig + i0 * 2);
        c = make_uint4(h.x, h.y, h.z, h.w);
    }
    const char4 q = *(const char4*)(slot + L.off_mapq + i0);
    if (i0 + 4 <= len) {
        *(uint4*)(d_read + i0) = make_uint4(base + v0, base + v1, base + v2, base + v3);
        *(uint4*)(d_contig + i0) = c;
        *(char4*)(d_mapq + i0) = q;
    } else {
        const u32 rv[4] = {base + v0, base + v1, base + v2, base + v3};
        const u32 cv[4] = {c.x, c.y, c.z, c.w};
        const int8_t qv[4] = {q.x, q.y, q.z, q.w};
        for (u32 k = 0; k < 4 && i0 + k < len; ++k) {
            d_read[i0 + k] = rv[k];
            d_contig[i0 + k] = cv[k];
            d_mapq[i0 + k] = qv[k];
        }
    }
}

// pinned staging ring, allocated on first use and kept (cudaHostAlloc of ~130 MB costs tens of ms)
std::mutex g_pk_mutex;
unsigned char* g_pk_host = nullptr;
size_t g_pk_host_bytes = 0;

}  // namespace

extern "C" int cph_release_host_staging() {
    std::lock_guard<std::mutex> lk(g_pk_mutex);
    if (g_pk_host) cudaFreeHost(g_pk_host);
    g_pk_host = nullptr;
    g_pk_host_bytes = 0;
    return CPH_OK;
}

// Host-table variant of the fast path: the three columns live in HOST memory.  Chunks are brought over on a private
// copy stream while porec_tile_kernel already works on the tiles that have landed (its ticket counter and look-back
// state carry across launches), so the group-by hides behind the transfer.  pack_threads = 0: the columns are copied as
// they are (pinned memory for full PCIe speed); > 0: that many host threads re-encode the chunks to 4 B per record
// first (see "Packed host transfer" above; the caller's arrays may be pageable); < 0: two fewer than the machine has
// cores, at most 16.  A launch may only cover tiles whose window ([tile start, +PT_TILE+PT_HALO)) has landed, and must stop
// before a read that is still growing at the end of the landed records (the tile would otherwise see a truncated
// read that overflows the window); the host read column answers that in O(1) per chunk.
// h_counts[3] = bytes that crossed to the device.  h_counts[2] = 1 when the table turned out not to be grouped by read: the staged device columns are complete, the
// outputs are not, and the caller re-runs cph_porec_to_edges on them (general path).
extern "C" int cph_porec_to_edges_host(const u32* h_read_idx, const u32* h_contig_idx, const int8_t* h_mapq, u32* d_read_idx,
                                       u32* d_contig_idx, int8_t* d_mapq, u64 n_records, u32 n_contigs, int min_mapq,
                                       u32 min_order, u32 max_order, u32 read_offset, u64 chunk_records, int pack_threads,
                                       u32* rem_row, u32* rem_col, u32* rem_count, int8_t* rem_mapq, u64* h_counts, void* ws,
                                       size_t ws_bytes, size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const u64 N = n_records;
    ARG_CHECK(N < (1ull << 32) - 8192);
    ARG_CHECK(n_contigs > 0 && n_contigs < (1u << 23));
    int nthreads = pack_threads;
    if (nthreads < 0) {
        const unsigned hw = std::thread::hardware_concurrency();
        nthreads = (int)(hw < 4 ? 1 : hw > 18 ? 16 : hw - 2);       // two cores stay free: this thread feeds the streams
    }
    const bool packed = nthreads > 0;
    const int wide = n_contigs > 65536u;
    u64 pk_chunk = chunk_records ? ceil_div((size_t)chunk_records, (size_t)PK_GROUP) * PK_GROUP : PK_CHUNK;
    if (pk_chunk > 8 * PK_CHUNK) pk_chunk = 8 * PK_CHUNK;
    const PkLayout slot_layout = pk_layout(pk_chunk, wide);
    WS_BEGIN(ar);
    const u32 ntiles = (u32)ceil_div(N > 0 ? N : 1, (size_t)PT_TILE);
    u32* scal = ar.take<u32>(4);
    u64* tile_state = ar.take<u64>((size_t)ntiles);
    unsigned char* dev_ring = packed ? ar.take<unsigned char>(slot_layout.bytes * PK_NDEV) : nullptr;
    WS_QUERY_RETURN(ar);
    ARG_CHECK(h_counts);
    h_counts[0] = h_counts[1] = h_counts[2] = h_counts[3] = h_counts[4] = h_counts[5] = 0;
    if (N == 0) return CPH_OK;
    ARG_CHECK(h_read_idx && h_contig_idx && h_mapq && d_read_idx && d_contig_idx && d_mapq);
    ARG_CHECK(rem_row && rem_col && rem_count && rem_mapq);
    PorecIn in{d_read_idx, d_contig_idx, d_mapq, nullptr, nullptr, nullptr, N, n_contigs, min_mapq, (i64)0,
               PreF{nullptr, nullptr, nullptr, nullptr, 0, 0, 0}};
    u64 chunk = packed ? pk_chunk : (chunk_records ? chunk_records : (u64)(8u << 20));
    if (!packed) chunk = ceil_div(chunk, (size_t)PT_TILE) * PT_TILE;
    const u32 nchunks = (u32)ceil_div(N, chunk);
    const u32 launch_every = packed ? 4u : 1u;          // packed chunks are small: one tile launch per 4 of them

    // ---- host packing threads (packed mode) ----
    std::unique_lock<std::mutex> ring_lock(g_pk_mutex, std::defer_lock);
    struct Shared {
        std::mutex m;
        std::condition_variable cv;
        std::vector<unsigned char> state;       // per chunk: 0 = pending, 1 = packed, 2 = copy as it is
        u64 released = 0;                       // chunks whose host slot may be overwritten
        std::atomic<u32> next{0};
        bool abort = false;
    } sh;
    std::vector<std::thread> workers;
    u32 raw_every = 0;
    if (packed) {
        ring_lock.lock();                       // one packed call at a time owns the pinned ring
        const size_t want = slot_layout.bytes * PK_NSLOT;
        if (g_pk_host_bytes < want) {
            if (g_pk_host) cudaFreeHost(g_pk_host);
            g_pk_host = nullptr;
            g_pk_host_bytes = 0;
            CUDA_TRY(cudaHostAlloc((void**)&g_pk_host, want, cudaHostAllocDefault));
            g_pk_host_bytes = want;
        }
        sh.state.assign(nchunks, 0);
        unsigned char* ring = g_pk_host;
        // Re-encoding is bound by host memory bandwidth (17 B of traffic per record) and leaves PCIe a third idle; a plain
        // copy costs 9 B of host reads per record but 9 B on PCIe.  Sending every k-th chunk as it is (pinned columns only)
        // moves load from the host's memory system to PCIe.  Measured on workload C, ms per end-to-end step without the
        // partition tail: none 39.0, every 6th 38.2, 5th 37.3, 4th 37.9, 3rd 42.8, 2nd 45.8 -- but 41.7 for every 5th in
        // the full default bench run against 39.2-39.4 for none, so it is OFF unless CPH_PACK_RAW_EVERY asks for it.
        {
            cudaPointerAttributes pa;
            bool pinned = true;
            for (const void* q : {(const void*)h_read_idx, (const void*)h_contig_idx, (const void*)h_mapq}) {
                if (cudaPointerGetAttributes(&pa, q) != cudaSuccess || pa.type != cudaMemoryTypeHost) pinned = false;
            }
            (void)cudaGetLastError();
            if (const char* e = getenv("CPH_PACK_RAW_EVERY")) raw_every = pinned ? (u32)atoi(e) : 0u;      // tuning knob
        }
        auto work = [&, ring]() {
            for (;;) {
                const u32 k = sh.next.fetch_add(1);
                if (k >= nchunks) return;
                if (raw_every && k % raw_every == raw_every - 1) {       // this chunk travels as it is (see raw_every)
                    {
                        std::lock_guard<std::mutex> lk(sh.m);
                        sh.state[k] = 2;
                    }
                    sh.cv.notify_all();
                    continue;
                }
                {
                    std::unique_lock<std::mutex> lk(sh.m);
                    sh.cv.wait(lk, [&] { return sh.abort || sh.released + PK_NSLOT > k; });
                    if (sh.abort) return;
                }
                const u64 lo = (u64)k * chunk, hi = lo + chunk < N ? lo + chunk : N;
                const bool ok = pk_pack(h_read_idx, h_contig_idx, h_mapq, lo, hi, wide, ring + (size_t)(k % PK_NSLOT) * slot_layout.bytes);
                {
                    std::lock_guard<std::mutex> lk(sh.m);
                    sh.state[k] = ok ? 1 : 2;
                }
                sh.cv.notify_all();
            }
        };
        const int nt = (u32)nthreads < nchunks ? nthreads : (int)nchunks;
        for (int i = 0; i < nt; ++i) workers.emplace_back(work);
    }

    cudaStream_t cs = nullptr, us = nullptr;
    cudaEvent_t ready = nullptr;
    std::vector<cudaEvent_t> landed(nchunks, nullptr), copied(packed ? nchunks : 0, nullptr);
    int rc = CPH_OK;
    auto cleanup = [&]() {
        if (packed) {
            {
                std::lock_guard<std::mutex> lk(sh.m);
                sh.abort = true;
            }
            sh.cv.notify_all();
            for (std::thread& t : workers) t.join();
            workers.clear();
            if (cs) cudaStreamSynchronize(cs);       // the ring must be idle before the next call may take it
            if (us) cudaStreamSynchronize(us);
        }
        for (cudaEvent_t e : landed)
            if (e) cudaEventDestroy(e);
        for (cudaEvent_t e : copied)
            if (e) cudaEventDestroy(e);
        if (ready) cudaEventDestroy(ready);
        if (cs) cudaStreamDestroy(cs);
        if (us) cudaStreamDestroy(us);
    };
#define HT_TRY(expr)                                                    \
    do {                                                                \
        cudaError_t e__ = (expr);                                       \
        if (e__ != cudaSuccess) {                                       \
            cph_set_error(__FILE__, __LINE__, cudaGetErrorString(e__)); \
            cleanup();                                                  \
            return CPH_ERR_CUDA;                                        \
        }                                                               \
    } while (0)
    HT_TRY(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
    if (packed) HT_TRY(cudaStreamCreateWithFlags(&us, cudaStreamNonBlocking));
    HT_TRY(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming));
    HT_TRY(cudaMemsetAsync(scal, 0, 4 * sizeof(u32), st));
    HT_TRY(cudaMemsetAsync(tile_state, 0, (size_t)ntiles * sizeof(u64), st));
    HT_TRY(pt_prepare());
    // the staging columns may be recycled blocks with work still queued on `st`: copies start after it
    HT_TRY(cudaEventRecord(ready, st));
    HT_TRY(cudaStreamWaitEvent(cs, ready, 0));
    if (packed) HT_TRY(cudaStreamWaitEvent(us, ready, 0));
    const int use_tma = (((uintptr_t)d_read_idx | (uintptr_t)d_contig_idx | (uintptr_t)d_mapq) & 15) == 0;
    u32 t_done = 0, waited = 0;
    u64 rel = 0, h2d_bytes = 0;
    for (u32 c = 0; c < nchunks; ++c) {
        const u64 lo = (u64)c * chunk, hi = lo + chunk < N ? lo + chunk : N, len = hi - lo;
        int how = 2;
        if (packed) {
            for (;;) {       // wait for the chunk to be packed; meanwhile hand finished host slots back to the packers
                bool moved = false;
                while (rel < c) {
                    const cudaError_t q = cudaEventQuery(copied[rel]);
                    if (q == cudaSuccess) {
                        ++rel;
                        moved = true;
                    } else {
                        if (q == cudaErrorNotReady) (void)cudaGetLastError();
                        else HT_TRY(q);
                        break;
                    }
                }
                std::unique_lock<std::mutex> lk(sh.m);
                if (moved) {
                    sh.released = rel;
                    sh.cv.notify_all();
                }
                if (sh.cv.wait_for(lk, std::chrono::microseconds(50), [&] { return sh.state[c] != 0; })) {
                    how = sh.state[c];
                    break;
                }
            }
        }
        HT_TRY(cudaEventCreateWithFlags(&landed[c], cudaEventDisableTiming));
        if (how == 1) {
            const PkLayout L = pk_layout(len, wide);
            unsigned char* dslot = dev_ring + (size_t)(c % PK_NDEV) * slot_layout.bytes;
            if (c >= (u32)PK_NDEV) HT_TRY(cudaStreamWaitEvent(cs, landed[c - PK_NDEV], 0));      // device slot unpacked?
            HT_TRY(cudaMemcpyAsync(dslot, g_pk_host + (size_t)(c % PK_NSLOT) * slot_layout.bytes, L.bytes, cudaMemcpyHostToDevice, cs));
            h2d_bytes += L.bytes;
            HT_TRY(cudaEventCreateWithFlags(&copied[c], cudaEventDisableTiming));
            HT_TRY(cudaEventRecord(copied[c], cs));
            HT_TRY(cudaStreamWaitEvent(us, copied[c], 0));
            const unsigned g = (unsigned)ceil_div((size_t)len, (size_t)PK_GROUP);
            if (wide) pk_unpack_kernel<true><<<g, 256, 0, us>>>(dslot, L, len, d_read_idx + lo, d_contig_idx + lo, d_mapq + lo);
            else pk_unpack_kernel<false><<<g, 256, 0, us>>>(dslot, L, len, d_read_idx + lo, d_contig_idx + lo, d_mapq + lo);
            ++g_cph_launches;
            HT_TRY(cudaGetLastError());
            HT_TRY(cudaEventRecord(landed[c], us));
        } else {
            HT_TRY(cudaMemcpyAsync(d_read_idx + lo, h_read_idx + lo, len * 4, cudaMemcpyHostToDevice, cs));
            HT_TRY(cudaMemcpyAsync(d_contig_idx + lo, h_contig_idx + lo, len * 4, cudaMemcpyHostToDevice, cs));
            HT_TRY(cudaMemcpyAsync(d_mapq + lo, h_mapq + lo, len, cudaMemcpyHostToDevice, cs));
            h2d_bytes += len * 9;
            if (packed) {
                HT_TRY(cudaEventCreateWithFlags(&copied[c], cudaEventDisableTiming));
                HT_TRY(cudaEventRecord(copied[c], cs));
            }
            HT_TRY(cudaEventRecord(landed[c], cs));
        }
        if ((c + 1) % launch_every != 0 && hi != N) continue;
        u32 t_hi;
        if (hi == N) t_hi = ntiles;
        else {
            t_hi = hi >= (u64)PT_HALO ? (u32)((hi - PT_HALO) / PT_TILE) : 0u;
            while (t_hi > t_done && h_read_idx[(u64)t_hi * PT_TILE - 1] == h_read_idx[hi - 1]) --t_hi;
        }
        if (t_hi > t_done) {
            // (packed and unpacked chunks land on different streams: wait for every chunk since the last launch)
            for (; waited <= c; ++waited) HT_TRY(cudaStreamWaitEvent(st, landed[waited], 0));
            pt_launch(t_hi - t_done, st, in, use_tma, min_order, max_order, read_offset, (unsigned long long*)tile_state, scal, rem_row,
                      rem_col, rem_count, rem_mapq);
            ++g_cph_launches;
            HT_TRY(cudaGetLastError());
            t_done = t_hi;
        }
    }
    u32 hs[3] = {0, 0, 0};
    u64 last = 0;
    HT_TRY(cudaMemcpyAsync(hs, scal, 3 * sizeof(u32), cudaMemcpyDeviceToHost, st));
    HT_TRY(cudaMemcpyAsync(&last, tile_state + (ntiles - 1), sizeof(u64), cudaMemcpyDeviceToHost, st));
    HT_TRY(cudaStreamSynchronize(st));
    HT_TRY(cudaStreamSynchronize(cs));
    if (us) HT_TRY(cudaStreamSynchronize(us));
#undef HT_TRY
    cleanup();
    h_counts[3] = h2d_bytes;
    if (hs[1] != 0) {
        h_counts[2] = 1;
        return rc;
    }
    const u64 total = last & ((1ull << 62) - 1);
    h_counts[0] = total;
    h_counts[1] = total ? (u64)hs[0] + read_offset + 1 : 0;
    h_counts[4] = hs[2];
    h_counts[5] = pt_mode() != 0 ? 1 : 0;
    return rc;
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
