// Stage 2: clique expansion  A = H diag(d) H^T  on the packed edge-major form.
//
// Replaces HyperGraph.clique_expansion_init (/root/reference/cphasing/algorithms/hypergraph.py:358-485):
// scipy's two-pass single-thread csr_matmat on float32 (after a CSC->CSR transpose of the scaled H)
// becomes one pass over the hyperedge runs.  Every entry a of a hyperedge owns the pairs (a, b) with
// b after it in the run; the term round(c_a c_b 2^F / D_e) is added with one 64-bit integer RED to
// the packed upper triangle of a dense accumulator (L2 resident for n <= 5.6k, HBM beyond).  Integer
// accumulation makes the result independent of the order of the atomics (deterministic) and exact
// for the integer contact counts (use_high_order = 0).  The diagonal never sees an atomic: it is a
// CSR row reduction (cph_diag_terms).  A 32x32-tiled epilogue mirrors the triangle through shared
// memory, thresholds (algorithms/hypergraph.py:453-455), optionally normalises (:590-602) and
// emits CSR with sorted indices.
//
// Roofline: HBM.  Algorithmic bytes per call: 8 nnz(H) read + 12 nnz(A) + 16 (n+1) written
// (SURVEY.md 8d); the work figure is sum_e k_e (k_e - 1) / 2 REDs.
#include <cstdlib>
#include "common.cuh"
#include <cmath>
#include <cstdlib>
#include <vector>

namespace {

constexpr int TPB = 256;
constexpr int TILE = 32;

__device__ __forceinline__ size_t gtid() { return (size_t)blockIdx.x * blockDim.x + threadIdx.x; }

__device__ __forceinline__ u64 tri_index(u64 i, u64 j, u64 n) {   // i <= j
    return (i * (2 * n - i + 1)) / 2 + (j - i);
}

__device__ __forceinline__ void fill_scale(double* scale, int frac_bits) {
    // scale[D] = 2^F / D ; D = 0 unused
    for (int d = threadIdx.x; d < 128; d += blockDim.x) scale[d] = d ? ldexp(1.0, frac_bits) / (double)d : 0.0;
}

// warp per CSR row: rowsq[i] = sum c^2
__global__ void __launch_bounds__(TPB) row_sq_kernel(const i64* __restrict__ indptr, const u32* __restrict__ pk, u32 n,
                                                     u64* __restrict__ rowsq) {
    const u32 row = (u32)(gtid() >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    u64 s = 0;
    for (i64 t = indptr[row] + lane, e = indptr[row + 1]; t < e; t += 32) {
        u64 c = pk[t] & CPH_CNTM;
        s += c * c;
    }
    s = warp_sum(s);
    if (lane == 0) rowsq[row] = s;
}

// warp per CSR row: diag_fx[i] = sum round(c^2 w 2^F / D)
__global__ void __launch_bounds__(TPB) diag_kernel(const i64* __restrict__ indptr, const u32* __restrict__ pk,
                                                   const float* __restrict__ ew, u32 n, int use_high_order, int frac_bits,
                                                   u64* __restrict__ diag) {
    __shared__ double scale[128];
    fill_scale(scale, frac_bits);
    __syncthreads();
    const u32 row = (u32)(gtid() >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    u64 s = 0;
    for (i64 t = indptr[row] + lane, e = indptr[row + 1]; t < e; t += 32) {
        u32 p = pk[t];
        u64 c = p & CPH_CNTM;
        double sc = scale[use_high_order ? (p >> 24) : 1u];
        if (ew) sc *= (double)ew[t];
        s += __double2ull_rn((double)(c * c) * sc);
    }
    s = warp_sum(s);
    if (lane == 0) diag[row] = s;
}

// thread per pem entry; owns the pairs with the later entries of its hyperedge
__global__ void __launch_bounds__(TPB) expand_kernel(const u32* __restrict__ pem_row, const u32* __restrict__ pem_pk,
                                                     const float* __restrict__ ew, u64 nnz, u32 n, int use_high_order, int frac_bits,
                                                     unsigned long long* __restrict__ acc) {
    __shared__ double scale[128];
    fill_scale(scale, frac_bits);
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t a = gtid(); a < nnz; a += stride) {
        size_t b = a + 1;
        if (b >= nnz) break;
        u32 rb = pem_row[b];
        if (rb & CPH_HEAD) continue;          // last entry of its hyperedge: owns nothing
        const u32 pa = pem_pk[a];
        const u64 i = pem_row[a] & CPH_ROWM;
        double sc = scale[use_high_order ? (pa >> 24) : 1u];
        if (ew) sc *= (double)ew[a];
        const u64 ca = pa & CPH_CNTM;
        for (;;) {
            const u64 j = rb & CPH_ROWM;
            const u64 cb = pem_pk[b] & CPH_CNTM;
            unsigned long long v = __double2ull_rn((double)(ca * cb) * sc);
            u64 lo = i < j ? i : j, hi = i < j ? j : i;
            if (i == j) v += v;   // duplicated (row, col): both orders land on the diagonal
            atomicAdd(&acc[tri_index(lo, hi, n)], v);
            if (++b >= nnz) break;
            rb = pem_row[b];
            if (rb & CPH_HEAD) break;
        }
    }
}

// Fused form of diag_hist_kernel + expand_kernel: ONE pass over the pem.  Thread per entry (grid-stride, persistent
// CTAs): the entry adds its own c^2 term to the diagonal (shared-memory u64 bins flushed once per CTA when n fits,
// global REDs otherwise) and owns the pairs with the later entries of its hyperedge.  The loads of the NEXT grid-stride
// iteration are issued before the pair loop of the current one, so the DRAM latency of the coalesced pem stream hides
// behind the RED traffic; the triangle index stays in 32-bit arithmetic when n(n+1) < 2^32.
constexpr int XT = 512;
template <bool SMEM_DIAG, bool IDX32>
__global__ void __launch_bounds__(XT, 3) expand_diag_kernel(const u32* __restrict__ pem_row, const u32* __restrict__ pem_pk,
                                                            const float* __restrict__ ew, u64 nnz, u32 n, int use_high_order,
                                                            int frac_bits, unsigned long long* __restrict__ acc,
                                                            unsigned long long* __restrict__ diag) {
    extern __shared__ unsigned long long xbins[];
    __shared__ double scale[128];
    fill_scale(scale, frac_bits);
    if (SMEM_DIAG)
        for (u32 i = threadIdx.x; i < n; i += XT) xbins[i] = 0ull;
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * XT;
    size_t a = gtid();
    u32 ra = 0, pa = 0, rb = CPH_HEAD;
    float wa = 1.0f;
    if (a < nnz) {
        ra = pem_row[a];
        pa = pem_pk[a];
        rb = a + 1 < nnz ? pem_row[a + 1] : CPH_HEAD;
        if (ew) wa = ew[a];
    }
    const u32 two_n1 = 2u * n + 1u;
    while (a < nnz) {
        const size_t an = a + stride;
        u32 ra2 = 0, pa2 = 0, rb2 = CPH_HEAD;
        float wa2 = 1.0f;
        if (an < nnz) {     // prefetch the next iteration's entry
            ra2 = pem_row[an];
            pa2 = pem_pk[an];
            rb2 = an + 1 < nnz ? pem_row[an + 1] : CPH_HEAD;
            if (ew) wa2 = ew[an];
        }
        const u32 i = ra & CPH_ROWM;
        double sc = scale[use_high_order ? (pa >> 24) : 1u];
        if (ew) sc *= (double)wa;
        const u64 ca = pa & CPH_CNTM;
        // integer product first (exact, < 2^48), then ONE rounding: every kernel that forms a term does it this way
        const unsigned long long dv = __double2ull_rn((double)(ca * ca) * sc);
        if (SMEM_DIAG) atomicAdd(&xbins[i], dv);
        else atomicAdd(&diag[i], dv);
        size_t b = a + 1;
        while (!(rb & CPH_HEAD)) {
            const u32 j = rb & CPH_ROWM;
            const u64 cb = pem_pk[b] & CPH_CNTM;
            unsigned long long v = __double2ull_rn((double)(ca * cb) * sc);
            const u32 lo = i < j ? i : j, hi = i < j ? j : i;
            if (i == j) v += v;   // duplicated (row, col): both orders land on the diagonal
            if (IDX32) atomicAdd(&acc[(lo * (two_n1 - lo)) / 2u + (hi - lo)], v);
            else atomicAdd(&acc[tri_index(lo, hi, n)], v);
            if (++b >= nnz) break;
            rb = pem_row[b];
        }
        a = an;
        ra = ra2;
        pa = pa2;
        rb = rb2;
        wa = wa2;
    }
    if (SMEM_DIAG) {
        __syncthreads();
        for (u32 i = threadIdx.x; i < n; i += XT) {
            const unsigned long long v = xbins[i];
            if (v) atomicAdd(&diag[i], v);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Binned expansion for accumulators that do not fit in L2 (n > ~5.6k; config D: n = 20k, 1.6 GB triangle).
// A RED into an HBM-resident accumulator is a random 32-byte read-modify-write in DRAM.  Here the pairs are first
// written out grouped by accumulator BLOCK (contig-block pair (I, J), Bs x Bs cells <= 48 MB), then applied block after
// block, so that while one block is being updated its cells stay in L2.  The accumulator layout does not change -- only
// the order of the (order-free, exact) integer additions does, so the result is bit-identical to the direct kernel.
//   bin_count_kernel : CTA per chunk of XB_CHUNK entries; pairs per block -> cnt[bin * grid + cta]; diagonal terms
//   (scan of cnt: exclusive offsets, bin-major, so every CTA knows where its records of every bin go -- no global atomics)
//   bin_fill_kernel  : same enumeration, writes one 8-byte record per pair
//   bin_apply_kernel : streams the records bin by bin, one RED each
// record = [63:41] cell index inside the block (li * Bs + lj) | [40:34] D_e | [33:0] c_a * c_b ; pairs whose count product
// needs more than 34 bits are added directly.  Scratch: 8 B per pair.
// ------------------------------------------------------------------------------------------------
constexpr int XB_T = 256;
constexpr u32 XB_CHUNK = 8192;
constexpr u32 XB_BS_MAX = 2508;           // 2508^2 x 8 B = 50 MB; the record's cell field (23 bits) allows 2896
constexpr u32 XB_MAXNB = 46;              // n <= 115368 at the default block size
constexpr int XB_MAXBINS = XB_MAXNB * (XB_MAXNB + 1) / 2;   // 1081
constexpr u64 XB_PROD_MAX = (1ull << 34) - 1;

struct BinGeom {
    u32 n, bs, nb, nbins;
};
__device__ __forceinline__ u32 bin_of(const BinGeom& g, u32 I, u32 J) { return I * (2u * g.nb - I + 1u) / 2u + (J - I); }   // I <= J

// enumerate the pairs owned by the entries [lo, hi) of this CTA: f(i, j, ca, cb, D) with i = row of the owning entry
template <class F>
__device__ __forceinline__ void xb_for_pairs(const u32* __restrict__ pem_row, const u32* __restrict__ pem_pk, u64 nnz, size_t lo,
                                             size_t hi, F f) {
    for (size_t a = lo + threadIdx.x; a < hi; a += XB_T) {
        size_t b = a + 1;
        if (b >= nnz) break;
        u32 rb = pem_row[b];
        if (rb & CPH_HEAD) continue;
        const u32 pa = pem_pk[a];
        const u32 i = pem_row[a] & CPH_ROWM;
        for (;;) {
            f(i, rb & CPH_ROWM, (u64)(pa & CPH_CNTM), (u64)(pem_pk[b] & CPH_CNTM), pa >> 24);
            if (++b >= nnz) break;
            rb = pem_row[b];
            if (rb & CPH_HEAD) break;
        }
    }
}

__global__ void __launch_bounds__(XB_T) bin_count_kernel(const u32* __restrict__ pem_row, const u32* __restrict__ pem_pk, u64 nnz,
                                                         BinGeom g, int use_high_order, int frac_bits, u64* __restrict__ cnt,
                                                         unsigned long long* __restrict__ diag) {
    __shared__ u32 c[XB_MAXBINS];
    __shared__ double scale[128];
    fill_scale(scale, frac_bits);
    for (int k = threadIdx.x; k < XB_MAXBINS; k += XB_T) c[k] = 0;
    __syncthreads();
    const size_t lo = (size_t)blockIdx.x * XB_CHUNK, hi = min((size_t)nnz, lo + XB_CHUNK);
    for (size_t a = lo + threadIdx.x; a < hi; a += XB_T) {       // diagonal terms (global REDs: n is large here)
        const u32 pa = pem_pk[a];
        const u64 ca = pa & CPH_CNTM;
        atomicAdd(&diag[pem_row[a] & CPH_ROWM], __double2ull_rn((double)(ca * ca) * scale[use_high_order ? (pa >> 24) : 1u]));
    }
    xb_for_pairs(pem_row, pem_pk, nnz, lo, hi, [&](u32 i, u32 j, u64 ca, u64 cb, u32) {
        if (ca * cb > XB_PROD_MAX) return;
        const u32 l = i < j ? i : j, h = i < j ? j : i;
        atomicAdd(&c[bin_of(g, l / g.bs, h / g.bs)], 1u);
    });
    __syncthreads();
    for (u32 k = threadIdx.x; k < g.nbins; k += XB_T) cnt[(size_t)k * gridDim.x + blockIdx.x] = c[k];
}

__global__ void __launch_bounds__(XB_T) bin_fill_kernel(const u32* __restrict__ pem_row, const u32* __restrict__ pem_pk, u64 nnz,
                                                        BinGeom g, int use_high_order, int frac_bits, const u64* __restrict__ off,
                                                        u64* __restrict__ rec, unsigned long long* __restrict__ acc) {
    __shared__ u32 c[XB_MAXBINS];
    __shared__ u64 base[XB_MAXBINS];
    __shared__ double scale[128];
    fill_scale(scale, frac_bits);
    for (u32 k = threadIdx.x; k < XB_MAXBINS; k += XB_T) {
        c[k] = 0;
        base[k] = k < g.nbins ? off[(size_t)k * gridDim.x + blockIdx.x] : 0;
    }
    __syncthreads();
    const size_t lo = (size_t)blockIdx.x * XB_CHUNK, hi = min((size_t)nnz, lo + XB_CHUNK);
    xb_for_pairs(pem_row, pem_pk, nnz, lo, hi, [&](u32 i, u32 j, u64 ca, u64 cb, u32 D) {
        const u32 l = i < j ? i : j, h = i < j ? j : i;
        const u64 prod = ca * cb;
        if (!use_high_order) D = 1;
        if (prod > XB_PROD_MAX) {       // does not fit the record: straight to the accumulator
            unsigned long long v = __double2ull_rn((double)prod * scale[D]);
            if (i == j) v += v;
            atomicAdd(&acc[tri_index(l, h, g.n)], v);
            return;
        }
        const u32 I = l / g.bs, J = h / g.bs;
        const u32 k = bin_of(g, I, J);
        const u64 cell = (u64)(l - I * g.bs) * g.bs + (h - J * g.bs);
        __stcs(&rec[base[k] + atomicAdd(&c[k], 1u)], (cell << 41) | ((u64)D << 34) | prod);    // streaming: keep L2 for acc
    });
}

__global__ void __launch_bounds__(XB_T) bin_apply_kernel(const u64* __restrict__ rec, const u64* __restrict__ off, u32 grid_fill,
                                                         BinGeom g, int frac_bits, unsigned long long* __restrict__ acc) {
    __shared__ double scale[128];
    fill_scale(scale, frac_bits);
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * XB_T;
    u32 k = 0;
    for (u32 I = 0; I < g.nb; ++I)
        for (u32 J = I; J < g.nb; ++J, ++k) {
            const u64 beg = off[(size_t)k * grid_fill], end = off[(size_t)(k + 1) * grid_fill];   // off has nbins*grid + 1 entries
            for (u64 p = beg + gtid(); p < end; p += stride) {
                const u64 r = __ldcs(&rec[p]);          // evict-first: the block being updated must stay in L2
                const u64 cell = r >> 41;
                const u32 li = (u32)(cell / g.bs), lj = (u32)(cell % g.bs);
                unsigned long long v = __double2ull_rn((double)(r & XB_PROD_MAX) * scale[(r >> 34) & 127u]);
                const u64 i = (u64)I * g.bs + li, j = (u64)J * g.bs + lj;
                if (i == j) v += v;
                atomicAdd(&acc[tri_index(i, j, g.n)], v);
            }
        }
}

// ------------------------------------------------------------------------------------------------
// Swept expansion (round 2) for accumulators beyond L2.  What the measurements said (profiles/red_ceiling.txt,
// profiles/r2a_bench_C.json): scattered 64-bit REDs run at ~190 G/s into an L2-resident region and at ~21 G/s into HBM;
// every form that MATERIALISES the pairs (binned: 13.5 ms, banded pair records: 23 ms on the workload-D shard) pays 16 B of
// scattered traffic per pair.  This form materialises nothing:
//   sw_prep_kernel : ONE pass that sorts every hyperedge run by contig (stable rank inside the run, runs <= 64 entries)
//                    and packs the entry to 4 bytes:  head(1) | D_e(7) | count(24 - rb) | row(rb),  rb = bits(n - 1).
//                    With sorted runs the pairs (a, b > a) owned by entry a all lie in row(a) of the upper triangle.
//   sw_band_kernel : the packed triangle is cut into row bands of <= ~64 MB; for one band [r0, r1) the kernel streams the
//                    4-byte entries (coalesced, evict-first), keeps the entries whose row is in the band in a shared-memory
//                    queue, and walks their run suffix out of the shared-memory tile -- one RED per pair into a band that
//                    was zeroed just before and stays L2 resident.  The diagonal terms of the band's rows go to
//                    shared-memory bins (two 32-bit atomics with carry) and are flushed once per CTA.
// DRAM traffic: 12 B per entry for the preparation + 4 B per entry and band.  Integer sums are order-free: acc / diag come
// out bit-identical to the direct kernel.  Entries whose count does not fit the field, or runs longer than 64, raise a
// device flag and the caller falls back to the direct kernel.
// ------------------------------------------------------------------------------------------------
constexpr int SW_T = 256;
constexpr int SW_PER = 8;
constexpr u32 SW_TILE = SW_T * SW_PER;        // 2048 entries per tile
constexpr u32 SW_HALO = 64;                   // longest run
constexpr u32 SW_BINS_MAX = 7168;             // band rows whose diagonal lives in shared-memory bins (56 KB)

// A tile owns the runs that START in it (a run crossing the end of the tile is finished out of the halo, a run that started
// in the previous tile is left to that tile), so its output range [first head in the tile, first head at or after the end
// of the tile) is contiguous: the sorted entries are staged in shared memory and leave with coalesced stores.
__global__ void __launch_bounds__(SW_T) sw_prep_kernel(const u32* __restrict__ pem_row, const u32* __restrict__ pem_pk, u64 nnz,
                                                       u32 rb, int use_high_order, u32* __restrict__ pe4, u32* __restrict__ flag) {
    constexpr u32 L = SW_TILE + SW_HALO;
    constexpr int NQ = (L + SW_T - 1) / SW_T;
    constexpr int NW = NQ * SW_T / 32;
    __shared__ u32 keys[L];                  // (row << 12) | position in the tile: distinct, ordered by (row, position)
    __shared__ u32 outs[L];
    __shared__ u32 hm[NW + 2];               // head bits; positions past the input count as heads
    const u64 t0 = (u64)blockIdx.x * SW_TILE;
    if (threadIdx.x < 2) hm[NW + threadIdx.x] = 0xffffffffu;
    const u32 nv = (u32)min((u64)L, nnz - t0);          // valid positions of the window (32-bit arithmetic from here on)
    const u32* __restrict__ prow = pem_row + t0;
    const u32* __restrict__ ppk = pem_pk + t0;
    const u32 lane = threadIdx.x & 31;
    u32 pkr[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) {           // all global loads of the tile are in flight together
        const u32 x = q * SW_T + threadIdx.x;
        const bool ok = x < nv;
        u32 v = ok ? prow[x] : CPH_HEAD;
        pkr[q] = ok ? ppk[x] : 0u;
        if (q == 0 && x == 0 && blockIdx.x == 0) v |= CPH_HEAD;
        if (x < L) keys[x] = ((v & CPH_ROWM) << 12) | x;
        const u32 mh = __ballot_sync(0xffffffffu, (v & CPH_HEAD) != 0u);
        if (lane == 0) hm[x >> 5] = mh;
    }
    __syncthreads();
    const u32 cmax = 1u << (24 - rb);
    u32 bad = 0;
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        const u32 x = q * SW_T + threadIdx.x;
        // start of my run: last head at or before x, at most 63 back (two words back at most)
        const u32 wi = x >> 5, bi = x & 31u;
        const u32 m0 = x < nv ? hm[wi] & (0xffffffffu >> (31u - bi)) : 0u;
        const u32 m1 = wi >= 1u ? hm[wi - 1u] : 0u, m2 = wi >= 2u ? hm[wi - 2u] : 0u;
        const u32 s = m0 ? wi * 32u + 31u - (u32)__clz(m0)
                         : (m1 ? wi * 32u - 1u - (u32)__clz(m1) : (m2 ? wi * 32u - 33u - (u32)__clz(m2) : 0xffffffffu));
        // end of my run: first head after x (positions past the window are heads)
        const u32 a0 = bi == 31u ? 0u : hm[wi] & (0xffffffffu << (bi + 1u));
        const u32 a1 = hm[wi + 1u], a2 = hm[wi + 2u];
        const u32 e = a0 ? wi * 32u + (u32)__ffs(a0) - 1u
                         : (a1 ? wi * 32u + 31u + (u32)__ffs(a1) : (a2 ? wi * 32u + 63u + (u32)__ffs(a2) : 0xffffffffu));
        const bool mine = x < nv && s != 0xffffffffu && s < SW_TILE;       // runs that start here (see above)
        if (x < nv && ((s == 0xffffffffu && x >= SW_HALO) || (s != 0xffffffffu && (x - s >= SW_HALO || e - s > SW_HALO)))) bad = 1u;
        if (mine && e - s <= SW_HALO) {
            const u32 ka = keys[x];
            u32 rank = 0;
            for (u32 b = s; b < e; ++b) rank += keys[b] < ka ? 1u : 0u;
            const u32 pk = pkr[q];
            const u32 c = pk & CPH_CNTM;
            if (c >= cmax) bad = 1u;
            const u32 D = use_high_order ? (pk >> 24) & 127u : 1u;
            outs[s + rank] = (rank == 0 ? CPH_HEAD : 0u) | (D << 24) | (c << rb) | (ka >> 12);
        }
    }
    if (bad) atomicOr(flag, 1u);
    __syncthreads();
    // my output range: [first head of the tile, first head at or after its end)
    const u32 h0 = hm[0], h1 = hm[1], e0 = hm[SW_TILE / 32], e1 = hm[SW_TILE / 32 + 1];
    const u32 lo = h0 ? (u32)__ffs(h0) - 1u : (h1 ? 32u + (u32)__ffs(h1) - 1u : L);
    u32 hi = e0 ? SW_TILE + (u32)__ffs(e0) - 1u : (e1 ? SW_TILE + 32u + (u32)__ffs(e1) - 1u : L);
    if (lo >= SW_TILE) return;               // no run starts here (only when a run is too long: flagged)
    hi = min(hi, nv);
    u32* __restrict__ po = pe4 + t0;
    for (u32 x = lo + threadIdx.x; x < hi; x += SW_T) po[x] = outs[x];
}

// Band pass, warp-autonomous: every warp streams its own 512-entry tiles (4 x uint4 per lane + a 64-entry halo), keeps the
// NEXT tile's loads in flight in registers while it works, and only needs __syncwarp; the CTA meets twice (bins).
constexpr int SWB_T = 512;
constexpr int SWB_WARPS = SWB_T / 32;
constexpr u32 SWB_WT = 512;                   // entries per warp tile
constexpr u32 SWB_EMPTY = 0xffffffffu;        // never a real entry (D_e <= 126)

__device__ __forceinline__ uint4 sw_load4(const u32* __restrict__ pe4, u64 g, u64 nnz) {
    uint4 x;
    if (g + 3 < nnz) {
        x = __ldcs(reinterpret_cast<const uint4*>(pe4 + g));       // evict-first: the band must stay in L2
    } else {
        x.x = g < nnz ? pe4[g] : SWB_EMPTY;
        x.y = g + 1 < nnz ? pe4[g + 1] : SWB_EMPTY;
        x.z = g + 2 < nnz ? pe4[g + 2] : SWB_EMPTY;
        x.w = SWB_EMPTY;
    }
    return x;
}

template <bool IDX32>
__global__ void __launch_bounds__(SWB_T, 2) sw_band_kernel(const u32* __restrict__ pe4, u64 nnz, u32 n, u32 rb, u32 r0, u32 r1,
                                                           int frac_bits, int smem_bins, unsigned long long* __restrict__ acc,
                                                           unsigned long long* __restrict__ diag) {
    extern __shared__ __align__(16) u32 sw_bins[];            // [2 * (r1 - r0)] lo / hi words of the diagonal terms
    __shared__ __align__(16) u32 wtiles[SWB_WARPS][SWB_WT + SW_HALO];
    __shared__ double scale[128];
    fill_scale(scale, frac_bits);
    const u32 nrows = r1 - r0;
    if (smem_bins)
        for (u32 x = threadIdx.x; x < 2 * nrows; x += SWB_T) sw_bins[x] = 0u;
    __syncthreads();
    const u32 rmask = (1u << rb) - 1u, cmask = (1u << (24 - rb)) - 1u;
    const u32 lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    u32* __restrict__ wt = wtiles[w];
    const u64 ntiles = (nnz + SWB_WT - 1) / SWB_WT;
    const u64 nwarps = (u64)gridDim.x * SWB_WARPS;
    u64 t = (u64)blockIdx.x * SWB_WARPS + w;
    uint4 r[4], h = make_uint4(SWB_EMPTY, SWB_EMPTY, SWB_EMPTY, SWB_EMPTY);
    if (t < ntiles) {
#pragma unroll
        for (int k = 0; k < 4; ++k) r[k] = sw_load4(pe4, t * SWB_WT + 4ull * (k * 32 + lane), nnz);
        if (lane < SW_HALO / 4) h = sw_load4(pe4, t * SWB_WT + SWB_WT + 4ull * lane, nnz);
    }
    for (; t < ntiles; t += nwarps) {
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 4; ++k) reinterpret_cast<uint4*>(wt)[k * 32 + lane] = r[k];
        if (lane < SW_HALO / 4) reinterpret_cast<uint4*>(wt)[128 + lane] = h;
        __syncwarp();
        // which of my 16 entries lie in the band
        u32 m = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const u32 ee[4] = {r[k].x, r[k].y, r[k].z, r[k].w};
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if ((u32)((ee[c] & rmask) - r0) < nrows && ee[c] != SWB_EMPTY) m |= 1u << (k * 4 + c);
        }
        const u64 tn = t + nwarps;
        if (tn < ntiles) {                    // next tile in flight while this one is worked on
#pragma unroll
            for (int k = 0; k < 4; ++k) r[k] = sw_load4(pe4, tn * SWB_WT + 4ull * (k * 32 + lane), nnz);
            if (lane < SW_HALO / 4) h = sw_load4(pe4, tn * SWB_WT + SWB_WT + 4ull * lane, nnz);
        }
        while (m) {
            const u32 bit = (u32)__ffs(m) - 1u;
            m &= m - 1u;
            u32 x = ((bit >> 2) * 32u + lane) * 4u + (bit & 3u);
            const u32 e = wt[x];
            const u32 i = e & rmask;
            const u64 ca = (e >> rb) & cmask;
            const double sc = scale[(e >> 24) & 127u];
            const unsigned long long dv = __double2ull_rn((double)(ca * ca) * sc);
            if (smem_bins) {
                const u32 lo = (u32)dv, hi = (u32)(dv >> 32);
                const u32 old = atomicAdd(&sw_bins[2 * (i - r0)], lo);
                const u32 carry = (old + lo < old) ? 1u : 0u;
                if (hi + carry) atomicAdd(&sw_bins[2 * (i - r0) + 1], hi + carry);
            } else {
                atomicAdd(&diag[i], dv);
            }
            unsigned long long* __restrict__ arow;
            if (IDX32) arow = acc + ((i * (2u * n + 1u - i)) / 2u - i);
            else arow = acc + (((u64)i * (2ull * n + 1ull - i)) / 2ull - i);
            u32 eb = wt[++x];
            while (!(eb & CPH_HEAD)) {
                const u32 j = eb & rmask;
                unsigned long long v = __double2ull_rn((double)(ca * (u64)((eb >> rb) & cmask)) * sc);
                if (j == i) v += v;     // duplicated (row, col): both orders land on the diagonal cell
                atomicAdd(arow + j, v);
                eb = wt[++x];
            }
        }
    }
    if (smem_bins) {
        __syncthreads();
        for (u32 x = threadIdx.x; x < nrows; x += SWB_T) {
            const unsigned long long v = (unsigned long long)sw_bins[2 * x] | ((unsigned long long)sw_bins[2 * x + 1] << 32);
            if (v) atomicAdd(&diag[r0 + x], v);
        }
    }
}

struct Epi {
    const u64* acc;
    const u64* diag;
    u32 n;
    u32 nt;
    double inv_scale;
    float min_weight;
};

// loads the (I, J) tile (I <= J) of the symmetric matrix into shared memory as doubles
__device__ __forceinline__ void load_tile(const Epi& p, u32 I, u32 J, double (*t)[TILE + 1]) {
    const u32 warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (u32 r = warp; r < TILE; r += TPB / 32) {
        u64 i = (u64)I * TILE + r, j = (u64)J * TILE + lane;
        double v = 0.0;
        if (i < p.n && j < p.n) {
            u64 raw = (i <= j) ? p.acc[tri_index(i, j, p.n)] : p.acc[tri_index(j, i, p.n)];
            if (i == j) raw += p.diag[i];
            v = (double)raw * p.inv_scale;
        }
        t[r][lane] = v;
    }
}

__device__ __forceinline__ bool keep_entry(double v, float min_weight) {
    // algorithms/hypergraph.py:453-455  A.data[A.data < min_weight] = 0 ; eliminate_zeros  (float32 compare)
    if (v == 0.0) return false;
    if (min_weight > 0.0f && (float)v < min_weight) return false;
    return true;
}

__global__ void __launch_bounds__(TPB) epi_count_kernel(Epi p, u64* __restrict__ cnt) {
    const u32 I = blockIdx.y, J = blockIdx.x;
    if (I > J) return;
    __shared__ double t[TILE][TILE + 1];
    load_tile(p, I, J, t);
    __syncthreads();
    const u32 warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (u32 r = warp; r < TILE; r += TPB / 32) {
        u32 i = I * TILE + r;
        u32 m = __ballot_sync(0xffffffffu, keep_entry(t[r][lane], p.min_weight));
        if (lane == 0 && i < p.n) cnt[(u64)i * p.nt + J] = __popc(m);
        if (I != J) {
            u32 jr = J * TILE + r;   // transposed: row jr of A, columns of tile I
            u32 m2 = __ballot_sync(0xffffffffu, keep_entry(t[lane][r], p.min_weight));
            if (lane == 0 && jr < p.n) cnt[(u64)jr * p.nt + I] = __popc(m2);
        }
    }
}

__global__ void __launch_bounds__(TPB) epi_emit_kernel(Epi p, const u64* __restrict__ off, const double* __restrict__ node_w,
                                                       i32* __restrict__ indices, float* __restrict__ d32, double* __restrict__ d64) {
    const u32 I = blockIdx.y, J = blockIdx.x;
    if (I > J) return;
    __shared__ double t[TILE][TILE + 1];
    load_tile(p, I, J, t);
    __syncthreads();
    const u32 warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const u32 lt = lanemask_lt();
    for (u32 r = warp; r < TILE; r += TPB / 32) {
        {
            u32 i = I * TILE + r, j = J * TILE + lane;
            double v = t[r][lane];
            bool k = keep_entry(v, p.min_weight);
            u32 m = __ballot_sync(0xffffffffu, k);
            if (k) {
                u64 pos = off[(u64)i * p.nt + J] + __popc(m & lt);
                indices[pos] = (i32)j;
                float nw = node_w ? (float)(node_w[i] * node_w[j]) : 1.0f;   // hyperpartition.py:568 NW float32
                if (d32) d32[pos] = node_w ? (float)v * nw : (float)v;
                if (d64) d64[pos] = node_w ? v * (double)nw : v;
            }
        }
        if (I != J) {
            u32 i = J * TILE + r, j = I * TILE + lane;
            double v = t[lane][r];
            bool k = keep_entry(v, p.min_weight);
            u32 m = __ballot_sync(0xffffffffu, k);
            if (k) {
                u64 pos = off[(u64)i * p.nt + I] + __popc(m & lt);
                indices[pos] = (i32)j;
                float nw = node_w ? (float)(node_w[i] * node_w[j]) : 1.0f;
                if (d32) d32[pos] = node_w ? (float)v * nw : (float)v;
                if (d64) d64[pos] = node_w ? v * (double)nw : v;
            }
        }
    }
}

__global__ void __launch_bounds__(TPB) epi_indptr_kernel(const u64* __restrict__ off, u32 n, u32 nt, i64* __restrict__ indptr) {
    size_t i = gtid();
    if (i <= n) indptr[i] = (i64)off[(u64)i * nt];   // off has n*nt + 1 entries; entry n*nt is the total
}

}  // namespace

extern "C" int cph_row_sq_sums(const i64* indptr, const u32* pk, u32 n, u64* rowsq, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(indptr && rowsq && n > 0);
    row_sq_kernel<<<grid_for((size_t)n * 32, TPB), TPB, 0, st>>>(indptr, pk, n, rowsq);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_diag_terms(const i64* indptr, const u32* pk, const float* csr_entry_w, u32 n, int use_high_order, int frac_bits,
                              u64* diag_fx, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(indptr && diag_fx && n > 0);
    ARG_CHECK(frac_bits >= 0 && frac_bits <= 52);
    diag_kernel<<<grid_for((size_t)n * 32, TPB), TPB, 0, st>>>(indptr, pk, csr_entry_w, n, use_high_order, frac_bits, diag_fx);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_expand_accumulate(const u32* pem_row, const u32* pem_pk, const float* entry_w, u64 nnz, u32 n, int use_high_order,
                                     int frac_bits, int zero_first, u64* acc, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(acc && n > 0);
    ARG_CHECK(frac_bits >= 0 && frac_bits <= 52);
    const size_t tri = (size_t)n * ((size_t)n + 1) / 2;
    if (zero_first) CUDA_TRY(cudaMemsetAsync(acc, 0, tri * sizeof(u64), st));
    if (nnz == 0) return CPH_OK;
    ARG_CHECK(pem_row && pem_pk);
    // persistent-ish grid: 148 SMs x 8 resident CTAs of 256 threads
    size_t want = ceil_div(nnz, TPB);
    unsigned grid = (unsigned)(want < 148 * 8 ? want : 148 * 8);
    expand_kernel<<<grid, TPB, 0, st>>>(pem_row, pem_pk, entry_w, nnz, n, use_high_order, frac_bits, (unsigned long long*)acc);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_expand_accumulate_diag(const u32* pem_row, const u32* pem_pk, const float* entry_w, u64 nnz, u32 n,
                                          int use_high_order, int frac_bits, int zero_first, u64* acc, u64* diag_fx, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(acc && diag_fx && n > 0);
    ARG_CHECK(frac_bits >= 0 && frac_bits <= 52);
    const size_t tri = (size_t)n * ((size_t)n + 1) / 2;
    if (zero_first) CUDA_TRY(cudaMemsetAsync(acc, 0, tri * sizeof(u64), st));
    CUDA_TRY(cudaMemsetAsync(diag_fx, 0, (size_t)n * sizeof(u64), st));
    if (nnz == 0) return CPH_OK;
    ARG_CHECK(pem_row && pem_pk);
    const bool smem_diag = n <= 4096;               // 32 KB of u64 bins, 3 CTAs of 512 threads per SM
    const bool idx32 = n <= 46340;                  // n (n + 1) < 2^31
    const size_t smem = smem_diag ? (size_t)n * sizeof(u64) : 0;
    size_t want = ceil_div(nnz, (size_t)XT);
    unsigned grid = (unsigned)(want < 148 * 3 ? want : 148 * 3);
    unsigned long long* A = (unsigned long long*)acc;
    unsigned long long* Dg = (unsigned long long*)diag_fx;
#define XLAUNCH(S, I)                                                                                                          \
    do {                                                                                                                       \
        if (smem > 48 * 1024)                                                                                                  \
            CUDA_TRY(cudaFuncSetAttribute(expand_diag_kernel<S, I>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   \
        expand_diag_kernel<S, I><<<grid, XT, smem, st>>>(pem_row, pem_pk, entry_w, nnz, n, use_high_order, frac_bits, A, Dg);  \
    } while (0)
    if (smem_diag) XLAUNCH(true, true);
    else if (idx32) XLAUNCH(false, true);
    else XLAUNCH(false, false);
#undef XLAUNCH
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_expand_accumulate_binned(int phase, const u32* pem_row, const u32* pem_pk, u64 nnz, u32 n, int use_high_order,
                                            int frac_bits, int zero_first, u64* acc, u64* diag_fx, u64* pair_records,
                                            u64* h_pairs, void* ws, size_t ws_bytes, size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(n > 0);
    ARG_CHECK(frac_bits >= 0 && frac_bits <= 52);
    // blocks of at most 50 MB (bs x bs u64 cells).  Measured on the workload-D shard (n = 20k, 371 M pairs; scripts/
    // diag_binned.py): block size 600 / 887 / 1774 / 2508 -> fill 7.9 / 7.7 / 6.7 / 5.0 ms, apply 6.0 / 6.5 / 5.3 / 5.4 ms:
    // smaller blocks do not make the apply pass faster (it is bound by the rate of scattered 64-bit REDs, ~70 G/s, not
    // by DRAM) and make the fill pass slower (more write streams per CTA).
    u32 bs_max = XB_BS_MAX;
    if (const char* e = getenv("CPH_XB_BS")) {          // tuning knob for scripts/diag_binned.py
        const long v = atol(e);
        if (v >= 64 && v <= 65536) bs_max = (u32)v;
    }
    u32 nb = (u32)ceil_div((size_t)n, (size_t)bs_max);
    ARG_CHECK(nb <= XB_MAXNB);
    BinGeom g{n, (u32)ceil_div((size_t)n, (size_t)nb), nb, nb * (nb + 1) / 2};
    const unsigned grid = grid_for(nnz, XB_CHUNK);
    WS_BEGIN(ar);
    const size_t cells = (size_t)g.nbins * grid + 1;
    u64* cnt = ar.take<u64>(cells);
    u64* total = ar.take<u64>(1);
    if (!ar.dry && phase == 1) {
        ARENA_CHECK(ar);
        ARG_CHECK(acc && diag_fx && h_pairs && (nnz == 0 || (pem_row && pem_pk)));
        const size_t tri = (size_t)n * ((size_t)n + 1) / 2;
        if (zero_first) CUDA_TRY(cudaMemsetAsync(acc, 0, tri * sizeof(u64), st));
        CUDA_TRY(cudaMemsetAsync(diag_fx, 0, (size_t)n * sizeof(u64), st));
        CUDA_TRY(cudaMemsetAsync(cnt, 0, cells * sizeof(u64), st));
        if (nnz > 0) {
            bin_count_kernel<<<grid, XB_T, 0, st>>>(pem_row, pem_pk, nnz, g, use_high_order, frac_bits, cnt,
                                                    (unsigned long long*)diag_fx);
            KERNEL_CHECK();
        }
    }
    // (the scan carves its scratch in every mode so that both phases see the same workspace layout)
    if (ar.dry || phase == 1) TRY(scan_u64(cnt, cells, total, ar, st));
    WS_QUERY_RETURN(ar);
    if (phase == 1) {       // phase 1 ends here: the caller sizes pair_records from *h_pairs and comes back with phase 2
        CUDA_TRY(cudaMemcpyAsync(h_pairs, total, sizeof(u64), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        return CPH_OK;
    }
    ARG_CHECK(phase >= 2 && phase <= 4 && acc);          // 3 / 4: fill only / apply only (timing the two halves apart)
    if (nnz == 0) return CPH_OK;
    ARG_CHECK(pem_row && pem_pk && pair_records);
    if (phase != 4) {
        bin_fill_kernel<<<grid, XB_T, 0, st>>>(pem_row, pem_pk, nnz, g, use_high_order, frac_bits, cnt, pair_records,
                                               (unsigned long long*)acc);
        KERNEL_CHECK();
    }
    if (phase != 3) {
        bin_apply_kernel<<<148 * 8, XB_T, 0, st>>>(pair_records, cnt, grid, g, frac_bits, (unsigned long long*)acc);
        KERNEL_CHECK();
    }
    return CPH_OK;
}

// Swept form, step 1: sorted + packed entries (4 B each) for cph_expand_sweep_band.  *flag (device u32, zeroed here) becomes
// non-zero when an entry does not fit the packed form (count >= 2^(24 - bits(n-1)) or a run longer than 64).
extern "C" int cph_expand_sweep_prep(const u32* pem_row, const u32* pem_pk, u64 nnz, u32 n, int use_high_order, u32* pe4, u32* flag,
                                     void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(n > 0 && flag && (nnz == 0 || (pem_row && pem_pk && pe4)));
    const u32 rb = (u32)bits_for(n > 1 ? n - 1 : 1);
    if (rb > 20) {
        cph_set_error(__FILE__, __LINE__, "swept expansion: more than 2^20 contigs (use the binned or the direct form)");
        return CPH_ERR_TOO_LARGE;
    }
    CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(u32), st));
    if (nnz == 0) return CPH_OK;
    sw_prep_kernel<<<grid_for(nnz, SW_TILE), SW_T, 0, st>>>(pem_row, pem_pk, nnz, rb, use_high_order, pe4, flag);
    KERNEL_CHECK();
    return CPH_OK;
}

// Swept form, step 2: one row band [row_lo, row_hi) of the packed triangle.  Zeroes the band's cells and diag_fx[row_lo:row_hi)
// first when zero_first != 0, then adds every pair whose smaller contig lies in the band and the diagonal terms of its rows.
extern "C" int cph_expand_sweep_band(const u32* pe4, u64 nnz, u32 n, u32 row_lo, u32 row_hi, int frac_bits, int zero_first, u64* acc,
                                     u64* diag_fx, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(acc && diag_fx && n > 0 && row_lo < row_hi && row_hi <= n);
    ARG_CHECK(frac_bits >= 0 && frac_bits <= 52);
    const u32 rb = (u32)bits_for(n > 1 ? n - 1 : 1);
    ARG_CHECK(rb <= 20);
    auto tri0 = [&](u64 i) { return (i * (2ull * n - i + 1ull)) / 2ull; };
    if (zero_first) {
        CUDA_TRY(cudaMemsetAsync(acc + tri0(row_lo), 0, (tri0(row_hi) - tri0(row_lo)) * sizeof(u64), st));
        CUDA_TRY(cudaMemsetAsync(diag_fx + row_lo, 0, (size_t)(row_hi - row_lo) * sizeof(u64), st));
    }
    if (nnz == 0) return CPH_OK;
    ARG_CHECK(pe4);
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        CUDA_TRY(cudaGetDevice(&dev));
        CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    }
    const u32 nrows = row_hi - row_lo;
    const int smem_bins = nrows <= SW_BINS_MAX;
    const size_t smem = smem_bins ? (size_t)nrows * 8 : 0;
    const int per_sm = smem + 39 * 1024 <= 100 * 1024 ? 2 : 1;
    const u64 nwt = ceil_div(nnz, (size_t)SWB_WT * SWB_WARPS);
    const unsigned grid = (unsigned)(nwt < (u64)sms * per_sm ? nwt : (u64)sms * per_sm);
    unsigned long long* A = (unsigned long long*)acc;
    unsigned long long* Dg = (unsigned long long*)diag_fx;
    if (n <= 65535) {
        if (smem > 8 * 1024)
            CUDA_TRY(cudaFuncSetAttribute(sw_band_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        sw_band_kernel<true><<<grid, SWB_T, smem, st>>>(pe4, nnz, n, rb, row_lo, row_hi, frac_bits, smem_bins, A, Dg);
    } else {
        if (smem > 8 * 1024)
            CUDA_TRY(cudaFuncSetAttribute(sw_band_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        sw_band_kernel<false><<<grid, SWB_T, smem, st>>>(pe4, nnz, n, rb, row_lo, row_hi, frac_bits, smem_bins, A, Dg);
    }
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_expand_count(const u64* acc, const u64* diag_fx, u32 n, int frac_bits, float min_weight, u64* h_nnz, void* ws,
                                size_t ws_bytes, size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(n > 0);
    const u32 nt = (u32)ceil_div(n, TILE);
    ARG_CHECK(nt <= 65535);
    WS_BEGIN(ar);
    const size_t cells = (size_t)n * nt + 1;
    u64* cnt = ar.take<u64>(cells);
    u64* total = ar.take<u64>(1);
    if (!ar.dry) {
        ARENA_CHECK(ar);
        ARG_CHECK(acc && diag_fx && h_nnz);
        CUDA_TRY(cudaMemsetAsync(cnt, 0, cells * sizeof(u64), st));
        Epi p{acc, diag_fx, n, nt, ldexp(1.0, -frac_bits), min_weight};
        epi_count_kernel<<<dim3(nt, nt), TPB, 0, st>>>(p, cnt);
        KERNEL_CHECK();
    }
    TRY(scan_u64(cnt, cells, total, ar, st));
    WS_QUERY_RETURN(ar);
    CUDA_TRY(cudaMemcpyAsync(h_nnz, total, sizeof(u64), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return CPH_OK;
}

extern "C" int cph_expand_emit(const u64* acc, const u64* diag_fx, u32 n, int frac_bits, float min_weight, const double* node_w,
                               const void* ws, i64* indptr, i32* indices, float* data32, double* data64, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    ARG_CHECK(acc && diag_fx && ws && indptr && n > 0);
    const u32 nt = (u32)ceil_div(n, TILE);
    const u64* off = (const u64*)ws;   // first carve of cph_expand_count
    Epi p{acc, diag_fx, n, nt, ldexp(1.0, -frac_bits), min_weight};
    epi_emit_kernel<<<dim3(nt, nt), TPB, 0, st>>>(p, off, node_w, indices, data32, data64);
    KERNEL_CHECK();
    epi_indptr_kernel<<<grid_for((size_t)n + 1, TPB), TPB, 0, st>>>(off, n, nt, indptr);
    KERNEL_CHECK();
    return CPH_OK;
}
