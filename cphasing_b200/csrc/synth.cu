// Device twin of cphasing_b200/synth.py: seeded, counter-based synthetic contact tables of the shapes
// BASELINE.json names, generated straight into HBM so that bench.py never has to
// build 10^8-10^9 records on the host.  Same hash (splitmix64 finaliser), same draw slots and the
// same laws as the numpy generator; libm differences (log1p, floor of a product) can flip a draw that
// lands within an ulp of a boundary, so the two are statistically, not byte-for-byte, identical --
// parity tests use the numpy tables, the bench uses these and hands the CPU baseline a slice of them.
// This is workload generation (bench infrastructure), not part of the reference's path.
#include "common.cuh"

namespace {

constexpr int TPB = 256;
constexpr u64 PHI = 0x9E3779B97F4A7C15ull;
constexpr u64 KMUL = 0xD1B54A32D192ED03ull;
constexpr int SLOT_ORDER = 0, SLOT_GROUP = 1, SLOT_MONO = 8, SLOT_REC = 1024, MAX_MULT = 8, MAX_ORDER = 50;

struct Genome {
    u32 n, G, ploidy;
    const double* cum;     // [n+1] cumulative length over `order`
    const double* gcum;    // [G+1] cumulative group length
    const i32* gstart;     // [G+1]
    const i32* order;      // [n]
    const i64* length;     // [n]
};

__device__ __forceinline__ u64 mix64(u64 x) {
    u64 z = x + PHI;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ u64 edge_hash(u64 seed, u64 e) { return mix64(seed ^ ((e + 1) * PHI)); }
__device__ __forceinline__ double draw(u64 h, u32 j) { return (double)(mix64(h + (u64)j * KMUL) >> 11) * (1.0 / 9007199254740992.0); }

// largest p in [lo, hi] with a[p] <= x  (searchsorted(side="right") - 1)
__device__ __forceinline__ i32 upper_idx(const double* __restrict__ a, i32 lo, i32 hi, double x) {
    while (lo < hi) {
        i32 mid = (lo + hi + 1) >> 1;
        if (a[mid] <= x) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

__device__ __forceinline__ u32 order_of(u64 h, int law) {
    double u = draw(h, SLOT_ORDER);
    int k;
    if (law == 0) {          // 2 + Poisson(2)
        double p = 0.1353352832366127, c = p;   // e^-2
        int x = 0;
        while (u >= c && x < 47) {
            ++x;
            p *= 2.0 / x;
            c += p;
        }
        k = 2 + x;
    } else if (law == 1) {   // 2 + floor(Exp(mean 8))
        k = 2 + (int)floor(-8.0 * log1p(-u));
    } else {
        k = 2;
    }
    return (u32)min(k, MAX_ORDER);
}

__device__ __forceinline__ u32 home_group(const Genome& g, u64 h) {
    double x = draw(h, SLOT_GROUP) * g.gcum[g.G];
    i32 p = upper_idx(g.gcum, 0, (i32)g.G, x);
    return (u32)min(max(p, 0), (i32)g.G - 1);
}

__device__ __forceinline__ u32 monomer_contig(const Genome& g, u64 h, u32 s, u32 grp) {
    double uc = draw(h, SLOT_MONO + 8 * s + 0), uh = draw(h, SLOT_MONO + 8 * s + 1), up = draw(h, SLOT_MONO + 8 * s + 2);
    u32 chrom = grp / g.ploidy, hap = grp % g.ploidy;
    u32 ohap = g.ploidy > 1 ? (hap + 1 + (u32)floor(uh * (g.ploidy - 1))) % g.ploidy : hap;
    u32 tgt = uc < 0.85 ? grp : chrom * g.ploidy + ohap;
    i32 lo = g.gstart[tgt], hi = g.gstart[tgt + 1];
    if (uc >= 0.95) {
        lo = 0;
        hi = (i32)g.n;
    }
    double x = g.cum[lo] + up * (g.cum[hi] - g.cum[lo]);
    i32 p = upper_idx(g.cum, 0, (i32)g.n, x);
    p = min(max(p, lo), hi - 1);
    return (u32)g.order[p];
}

__device__ __forceinline__ u32 mult_of(u64 h, u32 s) {
    double um = draw(h, SLOT_MONO + 8 * s + 3);
    int mu = 1 + (int)floor(log1p(-um) / -1.2039728043259361);   // log(0.3)
    return (u32)min(max(mu, 1), MAX_MULT);
}

__device__ __forceinline__ int8_t mapq_of(double u) {
    int q = u < 0.1 ? 0 : (u < 0.3 ? 1 + (int)floor((u - 0.1) / 0.2 * 59.0) : 60);
    return (int8_t)min(q, 60);
}

__global__ void __launch_bounds__(TPB) porec_count_kernel(u64 seed, u64 first, u64 m, int law, u64* __restrict__ nrec) {
    size_t e = (size_t)blockIdx.x * TPB + threadIdx.x;
    if (e > m) return;
    if (e == m) {
        nrec[e] = 0;
        return;
    }
    u64 h = edge_hash(seed, first + e);
    u32 k = order_of(h, law), tot = 0;
    for (u32 s = 0; s < k; ++s) tot += mult_of(h, s);
    nrec[e] = tot;
}

__global__ void __launch_bounds__(TPB) porec_fill_kernel(Genome g, u64 seed, u64 first, u64 m, int law, const u64* __restrict__ off,
                                                         u32* __restrict__ read_idx, u32* __restrict__ contig_idx,
                                                         u32* __restrict__ start, u32* __restrict__ end,
                                                         int8_t* __restrict__ mapq) {
    size_t e = (size_t)blockIdx.x * TPB + threadIdx.x;
    if (e >= m) return;
    u64 h = edge_hash(seed, first + e);
    u32 k = order_of(h, law);
    u32 grp = home_group(g, h);
    u64 o = off[e];
    for (u32 s = 0; s < k; ++s) {
        u32 ctg = monomer_contig(g, h, s, grp);
        u32 mu = mult_of(h, s);
        i64 L = g.length[ctg];
        i64 st = (i64)floor(draw(h, SLOT_MONO + 8 * s + 4) * (double)L);
        i64 span = 100 + (i64)floor(draw(h, SLOT_MONO + 8 * s + 5) * 4900.0);
        i64 en = min(st + span, L);
        for (u32 r = 0; r < mu; ++r, ++o) {
            read_idx[o] = (u32)(first + e);
            contig_idx[o] = ctg;
            if (start) start[o] = (u32)st;
            if (end) end[o] = (u32)en;
            mapq[o] = mapq_of(draw(h, SLOT_REC + 16 * s + r));
        }
    }
}

__global__ void __launch_bounds__(TPB) pairs_fill_kernel(Genome g, u64 seed, u64 first, u64 m, double intra, u32* __restrict__ c1o,
                                                         u32* __restrict__ c2o, u32* __restrict__ p1o, u32* __restrict__ p2o,
                                                         int8_t* __restrict__ mapq) {
    size_t e = (size_t)blockIdx.x * TPB + threadIdx.x;
    if (e >= m) return;
    u64 h = edge_hash(seed, first + e);
    u32 grp = home_group(g, h);
    u32 c1 = monomer_contig(g, h, 0, grp), c2 = monomer_contig(g, h, 1, grp);
    if (draw(h, SLOT_MONO + 3) < intra) c2 = c1;
    c1o[e] = c1;
    c2o[e] = c2;
    if (p1o) p1o[e] = (u32)floor(draw(h, SLOT_MONO + 4) * (double)g.length[c1]);
    if (p2o) p2o[e] = (u32)floor(draw(h, SLOT_MONO + 8 + 4) * (double)g.length[c2]);
    mapq[e] = mapq_of(draw(h, SLOT_REC));
}

}  // namespace

// nrec_off: u64[m+1]; on return holds the exclusive prefix of records per concatemer; *h_total = N
extern "C" int cph_synth_porec_count(u64 seed, u64 first, u64 m, int law, u64* nrec_off, u64* h_total, void* ws, size_t ws_bytes,
                                     size_t* h_ws_need, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    WS_BEGIN(ar);
    u64* total = ar.take<u64>(1);
    if (!ar.dry) {
        ARENA_CHECK(ar);
        ARG_CHECK(nrec_off && h_total);
        porec_count_kernel<<<grid_for(m + 1, TPB), TPB, 0, st>>>(seed, first, m, law, nrec_off);
        KERNEL_CHECK();
    }
    TRY(scan_u64(nrec_off, m + 1, total, ar, st));
    WS_QUERY_RETURN(ar);
    CUDA_TRY(cudaMemcpyAsync(h_total, total, sizeof(u64), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return CPH_OK;
}

extern "C" int cph_synth_porec_fill(u64 seed, u64 first, u64 m, int law, u32 n, u32 G, u32 ploidy, const double* cum,
                                    const double* gcum, const i32* gstart, const i32* order, const i64* length, const u64* nrec_off,
                                    u32* read_idx, u32* contig_idx, u32* start, u32* end, int8_t* mapq, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (m == 0) return CPH_OK;
    ARG_CHECK(cum && gcum && gstart && order && length && nrec_off && read_idx && contig_idx && mapq);
    ARG_CHECK(first + m < (1ull << 32));
    Genome g{n, G, ploidy, cum, gcum, gstart, order, length};
    porec_fill_kernel<<<grid_for(m, TPB), TPB, 0, st>>>(g, seed, first, m, law, nrec_off, read_idx, contig_idx, start, end, mapq);
    KERNEL_CHECK();
    return CPH_OK;
}

extern "C" int cph_synth_pairs(u64 seed, u64 first, u64 m, double intra, u32 n, u32 G, u32 ploidy, const double* cum,
                               const double* gcum, const i32* gstart, const i32* order, const i64* length, u32* chrom1, u32* chrom2,
                               u32* pos1, u32* pos2, int8_t* mapq, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (m == 0) return CPH_OK;
    ARG_CHECK(cum && gcum && gstart && order && length && chrom1 && chrom2 && mapq);
    Genome g{n, G, ploidy, cum, gcum, gstart, order, length};
    pairs_fill_kernel<<<grid_for(m, TPB), TPB, 0, st>>>(g, seed, first, m, intra, chrom1, chrom2, pos1, pos2, mapq);
    KERNEL_CHECK();
    return CPH_OK;
}
