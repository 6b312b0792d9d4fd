// Louvain / Leiden modularity clustering on the device (sm_100a) -- replaces the igraph calls of IRMM
// (cphasing/algorithms/hypergraph.py:1239-1252: community_multilevel / community_leiden on the graph built
// from triu(A) with the diagonal removed, :1157-1244).
//
// Design (B200-first, deterministic):
//   * every weight is converted once to signed 64-bit FIXED POINT (scale 2^s chosen from nnz * max|w| so that
//     the total 2m stays below 2^61).  Vertex strengths, community totals, neighbour-community sums and the
//     internal-weight sum are then exact integers: atomics commute, so results do not depend on thread order.
//     Gains are evaluated in double from those exact integers with one fixed expression.
//   * local moving: ONE WARP PER VERTEX.  The warp sums its row's weights per neighbouring community in a private
//     dense int64 strip (L2-resident scratch, one strip per resident warp), evaluates
//     gain(c) = w(v,c) - gamma k_v tot_c / 2m for every neighbouring community, picks the best with ties to the
//     LOWEST community id (shuffle arg-max), and resets the strip.  Vertices are visited in `n_phases`
//     interleaved phases per sweep; inside a phase all proposals read one snapshot and are committed together
//     (integer atomics on tot / size), so a sweep is Gauss-Seidel across phases and Jacobi inside one.  Swaps of
//     two singletons are excluded by the lower-id rule (Lu, Halappanavar & Kalyanaraman 2015).  After each sweep
//     the exact modularity is evaluated; the best labelling seen is kept (a synchronous sweep may lose quality).
//   * Leiden refinement (Traag et al. 2019, Alg. 4, greedy limit): rounds of propose / grant inside each community;
//     only vertices that are alone move, only into well-connected sub-communities with non-negative gain; a
//     move into a singleton whose own vertex moves in the same round is retried next round, which keeps every
//     sub-community connected.  E(S, C-S) is recomputed exactly (integer) every round.
//   * aggregation: one warp per super-vertex sums its members' rows into the same dense strip and emits the
//     non-zero columns in increasing order (no atomics on the output, no n^2 matrix).
// Level graphs are "ragged CSR": row v occupies [beg[v], end[v]) of idx / w.
#include "common.cuh"
#include "cph.h"
#include <math.h>
#include <string.h>
#include <stdlib.h>
#include <cooperative_groups.h>

namespace {

typedef unsigned long long ull;

struct LGraph {
    int n;
    const i64* beg;
    const i64* end;
    const i32* idx;
    const i64* w;
};

constexpr int CM_THREADS = 256;
constexpr int CM_WPB = CM_THREADS / 32;
constexpr double CM_PG_SCALE = 1.0 / 16777216.0;      // predicted gains are summed as int64 of gain * 2^-24

__device__ __forceinline__ i64 ld_cg(const i64* p) { return __ldcg((const long long*)p); }
__device__ __forceinline__ void st_cg(i64* p, i64 v) { __stcg((long long*)p, (long long)v); }
__device__ __forceinline__ void add64(i64* p, i64 v) { atomicAdd((ull*)p, (ull)v); }

// penalty term gamma * a * b / 2m, ONE expression used everywhere (determinism)
__device__ __forceinline__ double pen(double gamma, i64 a, i64 b, double m2) { return gamma * (double)a * (double)b / m2; }

// ---------------------------------------------------------------------------------------------------------------
// level-0 preparation
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(CM_THREADS) absmax_kernel(const i64* __restrict__ indptr, const i32* __restrict__ indices,
                                                            const float* __restrict__ d32, const double* __restrict__ d64,
                                                            int n, ull* __restrict__ out_bits) {
    int v = blockIdx.x * CM_WPB + (threadIdx.x >> 5);
    if (v >= n) return;
    double mx = 0;
    for (i64 t = indptr[v] + lane_id(); t < indptr[v + 1]; t += 32) {
        if (indices[t] == v) continue;
        double x = d64 ? d64[t] : (double)d32[t];
        x = fabs(x);
        if (x == x && x > mx && !isinf(x)) mx = x;
    }
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane_id() == 0 && mx > 0) atomicMax(out_bits, (ull)__double_as_longlong(mx));
}

// w0[t] = round(A_t * 2^s) (0 on the diagonal, for non-finite values), k[v] = row sum, m2 += k[v]
__global__ void __launch_bounds__(CM_THREADS) to_fixed_kernel(const i64* __restrict__ indptr, const i32* __restrict__ indices,
                                                              const float* __restrict__ d32, const double* __restrict__ d64,
                                                              int n, double scale, i64* __restrict__ w0, i64* __restrict__ k,
                                                              i64* __restrict__ m2) {
    int v = blockIdx.x * CM_WPB + (threadIdx.x >> 5);
    if (v >= n) return;
    i64 s = 0;
    for (i64 t = indptr[v] + lane_id(); t < indptr[v + 1]; t += 32) {
        double x = d64 ? d64[t] : (double)d32[t];
        i64 q = 0;
        if (indices[t] != v && x == x && !isinf(x)) q = llrint(x * scale);
        w0[t] = q;
        s += q;
    }
    s = warp_sum(s);
    if (lane_id() == 0) {
        k[v] = s;
        add64(m2, s);
    }
}

// strengths of an aggregated level (self loops included)
__global__ void __launch_bounds__(CM_THREADS) strength_kernel(LGraph g, i64* __restrict__ k) {
    int v = blockIdx.x * CM_WPB + (threadIdx.x >> 5);
    if (v >= g.n) return;
    i64 s = 0;
    for (i64 t = g.beg[v] + lane_id(); t < g.end[v]; t += 32) s += g.w[t];
    s = warp_sum(s);
    if (lane_id() == 0) k[v] = s;
}

__global__ void iota_kernel(i32* a, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = i;
}

// tot[c] = sum of k over members, csize[c] = members (both zeroed by the caller)
__global__ void comm_totals_kernel(const i32* __restrict__ comm, const i64* __restrict__ k, int n, i64* __restrict__ tot,
                                   i32* __restrict__ csize) {
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    int c = comm[v];
    add64(&tot[c], k[v]);
    atomicAdd(&csize[c], 1);
}

// ---------------------------------------------------------------------------------------------------------------
// local moving
// ---------------------------------------------------------------------------------------------------------------
struct Best {
    double gain;
    int c;
};
__device__ __forceinline__ Best warp_best(Best b) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double g = __shfl_xor_sync(0xffffffffu, b.gain, o);
        int c = __shfl_xor_sync(0xffffffffu, b.c, o);
        if (c >= 0 && (b.c < 0 || g > b.gain || (g == b.gain && c < b.c))) {
            b.gain = g;
            b.c = c;
        }
    }
    return b;
}

// group of G threads (a warp, or the whole CTA for high-degree levels) that works on one vertex at a time
template <int G>
__device__ __forceinline__ void group_sync() {
    if (G == 32) __syncwarp();
    else __syncthreads();
}
template <int G>
__device__ __forceinline__ Best group_best(Best b, Best* sm) {
    b = warp_best(b);
    if (G == 32) return b;
    const int w = threadIdx.x >> 5;
    if (lane_id() == 0) sm[w] = b;
    __syncthreads();
    Best r = sm[0];
#pragma unroll
    for (int q = 1; q < G / 32; ++q) {
        Best o = sm[q];
        if (o.c >= 0 && (r.c < 0 || o.gain > r.gain || (o.gain == r.gain && o.c < r.c))) r = o;
    }
    __syncthreads();
    return r;
}

// proposals for the vertices v = first + phase + i * n_phases (< last).  prop[v] = chosen community (or comm[v]).
// `strip` = one private row of g.n int64 zeros per group (left zero again on exit).
// one phase of proposals by this CTA's groups (a warp, or the whole CTA on dense level-0 graphs, per vertex)
template <int G>
__device__ __forceinline__ void propose_phase(const LGraph& g, int phase, int n_phases, int first, int last,
                                              const i32* __restrict__ comm, const i64* __restrict__ tot,
                                              const i32* __restrict__ csize, const i64* __restrict__ k, double gamma, double m2,
                                              i64* __restrict__ strip, i32* __restrict__ prop, i64* __restrict__ pgain,
                                              Best* sm_best) {
    constexpr int GPB = CM_THREADS / G;                       // groups per block
    const int n_groups = gridDim.x * GPB;
    const int gi = blockIdx.x * GPB + threadIdx.x / G;
    const u32 lane = threadIdx.x % G;
    i64* sc = strip + (size_t)gi * g.n;
    for (i64 v = (i64)first + phase + (i64)gi * n_phases; v < last; v += (i64)n_groups * n_phases) {
        const int old = comm[v];
        const i64 kv = k[v];
        const i64 b = g.beg[v], e = g.end[v];
        for (i64 t = b + lane; t < e; t += G) {
            int u = g.idx[t];
            i64 w = g.w[t];
            if (u != v && w != 0) add64(&sc[comm[u]], w);
        }
        group_sync<G>();
        const double stay = (double)ld_cg(&sc[old]) - pen(gamma, kv, tot[old] - kv, m2);
        const bool old_single = csize[old] == 1;
        Best best;
        best.gain = 0;
        best.c = -1;
        for (i64 t = b + lane; t < e; t += G) {
            int u = g.idx[t];
            i64 w = g.w[t];
            if (u == v || w == 0) continue;
            int c = comm[u];
            if (c == old) continue;
            // two singletons that decide in the SAME phase could swap for ever: there only the higher id moves.  A singleton
            // neighbour of another phase keeps still while v decides, so the sequential rule applies unchanged
            // (at n_phases = n this is the oracle's sweep).
            if (old_single && c > old && csize[c] == 1 && (u - (int)v) % n_phases == 0) continue;
            double gain = (double)ld_cg(&sc[c]) - pen(gamma, kv, tot[c], m2);
            if (best.c < 0 || gain > best.gain || (gain == best.gain && c < best.c)) {
                best.gain = gain;
                best.c = c;
            }
        }
        best = group_best<G>(best, sm_best);
        group_sync<G>();
        for (i64 t = b + lane; t < e; t += G) {
            int u = g.idx[t];
            if (u != v) st_cg(&sc[comm[u]], 0);
        }
        group_sync<G>();
        if (lane == 0) {
            const bool mv = best.c >= 0 && best.gain > stay;
            prop[v] = mv ? best.c : old;
            // predicted modularity gain of this move on the phase's snapshot, fixed point (exact integer sums: the
            // sweep's predicted total does not depend on the order of the atomics)
            pgain[v] = mv ? llrint((best.gain - stay) * CM_PG_SCALE) : 0;
        }
    }
}

// proposals for the vertices v = first + phase + i * n_phases (< last).  prop[v] = chosen community (or comm[v]).
// `strip` = one private row of g.n int64 zeros per group (left zero again on exit).
template <int G>
__global__ void __launch_bounds__(CM_THREADS) move_propose_kernel(LGraph g, int phase, int n_phases, int first, int last,
                                                                  const i32* __restrict__ comm, const i64* __restrict__ tot,
                                                                  const i32* __restrict__ csize, const i64* __restrict__ k,
                                                                  double gamma, double m2, i64* __restrict__ strip,
                                                                  i32* __restrict__ prop, i64* __restrict__ pgain) {
    __shared__ Best sm_best[CM_WPB];
    propose_phase<G>(g, phase, n_phases, first, last, comm, tot, csize, k, gamma, m2, strip, prop, pgain, sm_best);
}

// A whole sweep in ONE cooperative launch: for every phase, proposals -> grid barrier -> commit -> grid barrier.  With the
// adaptive phase count the first sweeps of a level run up to 128 phases of a few dozen vertices each; as separate launches
// (propose + commit per phase) that is 256 launch latencies per sweep and most of a clustering's time.
template <int G>
__global__ void __launch_bounds__(CM_THREADS) sweep_kernel(LGraph g, int n_phases, i32* __restrict__ comm, i64* __restrict__ tot,
                                                           i32* __restrict__ csize, const i64* __restrict__ k, double gamma,
                                                           double m2, i64* __restrict__ strip, i32* __restrict__ prop,
                                                           i64* __restrict__ pgain, ull* __restrict__ moves,
                                                           ull* __restrict__ pred) {
    __shared__ Best sm_best[CM_WPB];
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    const i64 gt = (i64)blockIdx.x * blockDim.x + threadIdx.x, nt = (i64)gridDim.x * blockDim.x;
    u32 moved = 0;
    i64 gsum = 0;
    for (int ph = 0; ph < n_phases; ++ph) {
        propose_phase<G>(g, ph, n_phases, 0, g.n, comm, tot, csize, k, gamma, m2, strip, prop, pgain, sm_best);
        grid.sync();
        for (i64 v = ph + gt * n_phases; v < g.n; v += nt * n_phases) {
            const int p = prop[v], old = comm[v];
            if (p != old) {
                const i64 kv = k[v];
                add64(&tot[p], kv);
                add64(&tot[old], -kv);
                atomicAdd(&csize[p], 1);
                atomicSub(&csize[old], 1);
                comm[v] = p;
                ++moved;
                gsum += pgain[v];
            }
        }
        grid.sync();
    }
    moved = warp_sum(moved);
    gsum = warp_sum(gsum);
    if (lane_id() == 0 && moved) {
        atomicAdd(moves, (ull)moved);
        atomicAdd(pred, (ull)gsum);
    }
}

__global__ void move_commit_kernel(int phase, int n_phases, int first, int last, i32* __restrict__ comm,
                                   const i32* __restrict__ prop, const i64* __restrict__ pgain, const i64* __restrict__ k,
                                   i64* __restrict__ tot, i32* __restrict__ csize, ull* __restrict__ moves,
                                   ull* __restrict__ pred) {
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    i64 v = (i64)first + phase + i * n_phases;
    bool mv = false;
    i64 gs = 0;
    if (v < last) {
        int p = prop[v], old = comm[v];
        if (p != old) {
            i64 kv = k[v];
            add64(&tot[p], kv);
            add64(&tot[old], -kv);
            atomicAdd(&csize[p], 1);
            atomicSub(&csize[old], 1);
            comm[v] = p;
            mv = true;
            gs = pgain[v];
        }
    }
    u32 m = __ballot_sync(0xffffffffu, mv);
    gs = warp_sum(gs);
    if (lane_id() == 0 && m) {
        atomicAdd(moves, (ull)__popc(m));
        atomicAdd(pred, (ull)gs);
    }
}

// internal weight: in_sum += sum over stored entries whose end points share a community (self loops included)
__global__ void __launch_bounds__(CM_THREADS) internal_kernel(LGraph g, const i32* __restrict__ comm, i64* __restrict__ in_sum) {
    int v = blockIdx.x * CM_WPB + (threadIdx.x >> 5);
    if (v >= g.n) return;
    int c = comm[v];
    i64 s = 0;
    for (i64 t = g.beg[v] + lane_id(); t < g.end[v]; t += 32)
        if (comm[g.idx[t]] == c) s += g.w[t];
    s = warp_sum(s);
    if (lane_id() == 0 && s) add64(in_sum, s);
}

// sum_c (tot_c / 2m)^2 in a fixed order (one block): deterministic double
__global__ void __launch_bounds__(1024) sumsq_kernel(const i64* __restrict__ tot, int n, double m2, double* __restrict__ out) {
    __shared__ double sm[1024];
    double s = 0;
    for (int c = threadIdx.x; c < n; c += 1024) {
        double x = (double)tot[c] / m2;
        s += x * x;
    }
    sm[threadIdx.x] = s;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = sm[0];
}

// ---------------------------------------------------------------------------------------------------------------
// Leiden refinement
// ---------------------------------------------------------------------------------------------------------------
__global__ void refine_init_kernel(const i64* __restrict__ k, int n, i32* __restrict__ refined, i32* __restrict__ ssize,
                                   i64* __restrict__ stot) {
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    refined[v] = v;
    ssize[v] = 1;
    stot[v] = k[v];
}

// ext[s] = E(S, C - S): weight between sub-community s and the rest of its community (ext zeroed by the caller)
__global__ void __launch_bounds__(CM_THREADS) refine_ext_kernel(LGraph g, const i32* __restrict__ comm,
                                                                const i32* __restrict__ refined, i64* __restrict__ ext) {
    int v = blockIdx.x * CM_WPB + (threadIdx.x >> 5);
    if (v >= g.n) return;
    const int c = comm[v], s = refined[v];
    i64 x = 0;
    for (i64 t = g.beg[v] + lane_id(); t < g.end[v]; t += 32) {
        int u = g.idx[t];
        if (u != v && comm[u] == c && refined[u] != s) x += g.w[t];
    }
    x = warp_sum(x);
    if (lane_id() == 0 && x) add64(&ext[s], x);
}

// target[v] = sub-community a lone, well-connected vertex wants to join (-1: none / not eligible)
__global__ void __launch_bounds__(CM_THREADS) refine_propose_kernel(LGraph g, const i32* __restrict__ comm,
                                                                    const i64* __restrict__ ctot, const i64* __restrict__ k,
                                                                    const i32* __restrict__ refined, const i32* __restrict__ ssize,
                                                                    const i64* __restrict__ stot, const i64* __restrict__ ext,
                                                                    double gamma, double m2, i64* __restrict__ strip,
                                                                    i32* __restrict__ target) {
    const int n_warps = gridDim.x * CM_WPB;
    const int gw = blockIdx.x * CM_WPB + (threadIdx.x >> 5);
    const u32 lane = lane_id();
    i64* sc = strip + (size_t)gw * g.n;
    for (int v = gw; v < g.n; v += n_warps) {
        int tgt = -1;
        const bool alone = refined[v] == v && ssize[v] == 1;
        const int c = comm[v];
        const i64 kv = k[v], cT = ctot[c];
        if (alone && !((double)ext[v] < pen(gamma, kv, cT - kv, m2))) {
            const i64 b = g.beg[v], e = g.end[v];
            for (i64 t = b + lane; t < e; t += 32) {
                int u = g.idx[t];
                i64 w = g.w[t];
                if (u != v && w != 0 && comm[u] == c) add64(&sc[refined[u]], w);
            }
            __syncwarp();
            Best best;
            best.gain = 0;
            best.c = -1;
            for (i64 t = b + lane; t < e; t += 32) {
                int u = g.idx[t];
                i64 w = g.w[t];
                if (u == v || w == 0 || comm[u] != c) continue;
                int s = refined[u];
                if (s > v && ssize[s] == 1) continue;                                   // two lone vertices: higher id moves
                i64 st = stot[s];
                if ((double)ext[s] < pen(gamma, st, cT - st, m2)) continue;           // S not well connected
                double gain = (double)ld_cg(&sc[s]) - pen(gamma, kv, st, m2);
                if (gain < 0) continue;
                if (best.c < 0 || gain > best.gain || (gain == best.gain && s < best.c)) {
                    best.gain = gain;
                    best.c = s;
                }
            }
            best = warp_best(best);
            __syncwarp();
            for (i64 t = b + lane; t < e; t += 32) {
                int u = g.idx[t];
                if (u != v && comm[u] == c) st_cg(&sc[refined[u]], 0);
            }
            __syncwarp();
            tgt = best.c;
        }
        if (lane == 0) target[v] = tgt;
    }
}

// a move v -> s is granted unless vertex s itself moves this round (target[] is read-only here)
__global__ void refine_apply_kernel(const i32* __restrict__ target, const i64* __restrict__ k, int n, i32* __restrict__ refined,
                                    i32* __restrict__ ssize, i64* __restrict__ stot, ull* __restrict__ moves) {
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    bool mv = false;
    if (v < n) {
        int s = target[v];
        if (s >= 0 && target[s] < 0) {
            refined[v] = s;
            add64(&stot[s], k[v]);
            atomicAdd(&ssize[s], 1);
            ssize[v] = 0;
            stot[v] = 0;
            mv = true;
        }
    }
    u32 m = __ballot_sync(0xffffffffu, mv);
    if (lane_id() == 0 && m) atomicAdd(moves, (ull)__popc(m));
}

// ---------------------------------------------------------------------------------------------------------------
// label compaction and aggregation
// ---------------------------------------------------------------------------------------------------------------
__global__ void flag_labels_kernel(const i32* __restrict__ lab, int n, u32* __restrict__ flag) {
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) flag[lab[v]] = 1;
}

// one block: in-place exclusive scan of a[0..n), total -> *total
__global__ void __launch_bounds__(1024) block_scan_u32_kernel(u32* __restrict__ a, int n, u32* __restrict__ total) {
    __shared__ u32 sm[1024 / 32 + 1];
    u32 carry = 0;
    for (int s = 0; s < n; s += 1024) {
        int i = s + threadIdx.x;
        u32 v = i < n ? a[i] : 0u, tot;
        u32 ex = block_excl_scan<u32, 1024>(v, sm, &tot);
        if (i < n) a[i] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) *total = carry;
}

__global__ void apply_map_kernel(i32* __restrict__ lab, int n, const u32* __restrict__ map) {
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) lab[v] = (i32)map[lab[v]];
}

// per super-vertex: member count and summed row length
__global__ void agg_count_kernel(LGraph g, const i32* __restrict__ lab, u32* __restrict__ mcnt, i64* __restrict__ degsum) {
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= g.n) return;
    int c = lab[v];
    atomicAdd(&mcnt[c], 1u);
    add64(&degsum[c], g.end[v] - g.beg[v]);
}

// one block: mbeg = exclusive scan of mcnt (ng + 1 entries), off = exclusive scan of min(ng, degsum)
__global__ void __launch_bounds__(1024) agg_offsets_kernel(const u32* __restrict__ mcnt, const i64* __restrict__ degsum, int ng,
                                                           u32* __restrict__ mbeg, i64* __restrict__ off) {
    __shared__ i64 sm[1024 / 32 + 1];
    i64 c1 = 0, c2 = 0;
    for (int s = 0; s <= ng; s += 1024) {
        int i = s + threadIdx.x;
        i64 a = i < ng ? (i64)mcnt[i] : 0, tot;
        i64 ex = block_excl_scan<i64, 1024>(a, sm, &tot);
        if (i <= ng) mbeg[i] = (u32)(c1 + ex);
        c1 += tot;
        i64 b = i < ng ? (degsum[i] < (i64)ng ? degsum[i] : (i64)ng) : 0;
        ex = block_excl_scan<i64, 1024>(b, sm, &tot);
        if (i <= ng) off[i] = c2 + ex;
        c2 += tot;
    }
}

__global__ void agg_scatter_kernel(const i32* __restrict__ lab, int n, const u32* __restrict__ mbeg, u32* __restrict__ cursor,
                                   i32* __restrict__ members) {
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    int c = lab[v];
    members[mbeg[c] + atomicAdd(&cursor[c], 1u)] = v;
}

// one warp per super-vertex: dense-strip accumulation of the members' rows, emission in increasing column order
__global__ void __launch_bounds__(CM_THREADS) agg_rows_kernel(LGraph g, const i32* __restrict__ lab, int ng,
                                                              const u32* __restrict__ mbeg, const i32* __restrict__ members,
                                                              const i64* __restrict__ off, i64* __restrict__ strip, int stride,
                                                              i64* __restrict__ nbeg, i64* __restrict__ nend,
                                                              i32* __restrict__ nidx, i64* __restrict__ nw) {
    const int n_warps = gridDim.x * CM_WPB;
    const int gw = blockIdx.x * CM_WPB + (threadIdx.x >> 5);
    const u32 lane = lane_id();
    i64* sc = strip + (size_t)gw * stride;
    for (int c = gw; c < ng; c += n_warps) {
        for (u32 q = mbeg[c]; q < mbeg[c + 1]; ++q) {
            int v = members[q];
            for (i64 t = g.beg[v] + lane; t < g.end[v]; t += 32) {
                i64 w = g.w[t];
                if (w != 0) add64(&sc[lab[g.idx[t]]], w);
            }
        }
        __syncwarp();
        i64 o = off[c];
        const i64 o0 = o;
        for (int j0 = 0; j0 < ng; j0 += 32) {
            int j = j0 + lane;
            i64 x = j < ng ? ld_cg(&sc[j]) : 0;
            u32 m = __ballot_sync(0xffffffffu, x != 0);
            if (x != 0) {
                i64 p = o + __popc(m & lanemask_lt());
                nidx[p] = j;
                nw[p] = x;
                st_cg(&sc[j], 0);
            }
            o += __popc(m);
        }
        __syncwarp();
        if (lane == 0) {
            nbeg[c] = o0;
            nend[c] = o;
        }
    }
}

// Aggregation onto FEW super-vertices (ng <= AGG_DENSE_MAX): a warp per super-vertex would leave the machine empty (at
// n = 20k, ng ~ 100 that was 8 ms per launch, the largest item of a clustering), so the members' rows are shared out over
// the whole grid instead.  CTA b takes a slice of `members` (vertices sorted by super-vertex), sums the rows of one
// super-vertex at a time into a shared-memory strip of ng int64 (entries that stay inside the super-vertex -- most of
// them -- go to a register) and flushes the strip into the dense ng x ng matrix M with one RED per non-zero.  M lives in
// the (all-zero) strip scratch; agg_emit_kernel turns its rows into the CSR of the next level and zeroes them again.
constexpr int AGG_DENSE_MAX = 4096;

__global__ void __launch_bounds__(CM_THREADS) agg_dense_kernel(LGraph g, const i32* __restrict__ lab, int ng,
                                                               const i32* __restrict__ members, int per_cta, i64* __restrict__ M) {
    extern __shared__ i64 agg_sm[];
    for (int j = threadIdx.x; j < ng; j += CM_THREADS) agg_sm[j] = 0;
    __syncthreads();
    const int q0 = blockIdx.x * per_cta, q1 = min(g.n, q0 + per_cta);
    if (q0 >= q1) return;
    int cur = lab[members[q0]];
    i64 own = 0;
    for (int q = q0; q <= q1; ++q) {
        const int v = q < q1 ? members[q] : -1;
        const int c = q < q1 ? lab[v] : -1;          // the same for every thread of the CTA
        if (c != cur) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) own += __shfl_xor_sync(0xffffffffu, own, o);
            if (lane_id() == 0 && own != 0) atomicAdd((ull*)&agg_sm[cur], (ull)own);
            __syncthreads();
            for (int j = threadIdx.x; j < ng; j += CM_THREADS) {
                const i64 x = agg_sm[j];
                if (x != 0) {
                    add64(&M[(size_t)cur * ng + j], x);
                    agg_sm[j] = 0;
                }
            }
            __syncthreads();
            own = 0;
            cur = c;
        }
        if (q == q1) break;
        for (i64 t = g.beg[v] + threadIdx.x; t < g.end[v]; t += CM_THREADS) {
            const i64 w = g.w[t];
            if (w != 0) {
                const int l = lab[g.idx[t]];
                if (l == c) own += w;
                else atomicAdd((ull*)&agg_sm[l], (ull)w);
            }
        }
    }
}

// one warp per super-vertex: row c of M -> CSR row in increasing column order; M is left all zero
__global__ void __launch_bounds__(CM_THREADS) agg_emit_kernel(int ng, const i64* __restrict__ off, i64* __restrict__ M,
                                                              i64* __restrict__ nbeg, i64* __restrict__ nend,
                                                              i32* __restrict__ nidx, i64* __restrict__ nw) {
    const int n_warps = gridDim.x * CM_WPB;
    const u32 lane = lane_id();
    for (int c = blockIdx.x * CM_WPB + (threadIdx.x >> 5); c < ng; c += n_warps) {
        i64* row = M + (size_t)c * ng;
        i64 o = off[c];
        const i64 o0 = o;
        for (int j0 = 0; j0 < ng; j0 += 32) {
            const int j = j0 + lane;
            const i64 x = j < ng ? ld_cg(&row[j]) : 0;
            const u32 m = __ballot_sync(0xffffffffu, x != 0);
            if (x != 0) {
                const i64 p = o + __popc(m & lanemask_lt());
                nidx[p] = j;
                nw[p] = x;
                st_cg(&row[j], 0);
            }
            o += __popc(m);
        }
        if (lane == 0) {
            nbeg[c] = o0;
            nend[c] = o;
        }
    }
}

__global__ void gather_kernel(const i32* __restrict__ src, const i32* index, int n, i32* dst) {   // dst may alias index
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) dst[v] = src[index[v]];
}

// next-level community of each super-vertex: init[lab[v]] = comm[v] (all members agree)
__global__ void scatter_comm_kernel(const i32* __restrict__ lab, const i32* __restrict__ comm, int n, i32* __restrict__ init) {
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) init[lab[v]] = comm[v];
}

// ---------------------------------------------------------------------------------------------------------------
// host orchestration
// ---------------------------------------------------------------------------------------------------------------
struct Scalars {   // one 64-byte device block
    i64 m2;
    ull moves;
    i64 in_sum;
    double sumsq;
    ull absmax_bits;
    u32 total;
    u32 pad;
    ull pred;          // predicted gain of the sweep's moves (int64, CM_PG_SCALE)
};

struct State {
    cudaStream_t st;
    int n0;
    i64 nnz0;
    double gamma, m2;
    int n_phases, max_sweeps, pmax, pdiv;
    double eta_min;      // adaptive phases: a sweep whose moves realise less than this share of their predicted gain is redone
    int coop_blocks;     // > 0: co-resident CTAs for the one-launch sweep (sweep_kernel); 0: one launch per phase
    int blocks;          // persistent grid of the warp-per-vertex kernels; strip has blocks * CM_WPB rows of n0
    Scalars* sc;
    i64 *w0, *gw[2], *gbeg[2], *gend[2];
    i32* gidx[2];
    i64 *k, *tot, *stot, *ext, *degsum, *off, *strip;
    i64* pgain;
    i32* snap;
    i32 *csize, *comm, *best, *best_it, *prop, *refined, *ssize, *target, *lab, *members, *top, *member, *init;
    u32 *flag, *mcnt, *mbeg, *cursor;
    // statistics
    int levels, sweeps, rounds, retries;
    bool trace;          // CPH_COMMUNITY_TRACE=1: one stderr line per sweep
    bool noisy;          // adaptive phases: a sweep fell short of its predicted gain (see local_moving)
};

inline unsigned tgrid(int n) { return grid_for((size_t)n, 256); }
inline unsigned wgrid(int n) { return grid_for((size_t)n, CM_WPB); }

int carve(State& S, Arena& ar) {
    const size_t n = (size_t)S.n0, nz = (size_t)(S.nnz0 > 0 ? S.nnz0 : 1);
    S.sc = ar.take<Scalars>(1);
    S.w0 = ar.take<i64>(nz);
    for (int i = 0; i < 2; ++i) {
        S.gidx[i] = ar.take<i32>(nz);
        S.gw[i] = ar.take<i64>(nz);
        S.gbeg[i] = ar.take<i64>(n + 1);
        S.gend[i] = ar.take<i64>(n + 1);
    }
    S.k = ar.take<i64>(n + 1);
    S.tot = ar.take<i64>(n + 1);
    S.stot = ar.take<i64>(n + 1);
    S.ext = ar.take<i64>(n + 1);
    S.degsum = ar.take<i64>(n + 1);
    S.off = ar.take<i64>(n + 2);
    S.csize = ar.take<i32>(n + 1);
    S.comm = ar.take<i32>(n + 1);
    S.best = ar.take<i32>(n + 1);
    S.best_it = ar.take<i32>(n + 1);
    S.prop = ar.take<i32>(n + 1);
    S.pgain = ar.take<i64>(n + 1);
    S.snap = ar.take<i32>(n + 1);
    S.refined = ar.take<i32>(n + 1);
    S.ssize = ar.take<i32>(n + 1);
    S.target = ar.take<i32>(n + 1);
    S.lab = ar.take<i32>(n + 1);
    S.members = ar.take<i32>(n + 1);
    S.top = ar.take<i32>(n + 1);
    S.member = ar.take<i32>(n + 1);
    S.init = ar.take<i32>(n + 1);
    S.flag = ar.take<u32>(n + 2);
    S.mcnt = ar.take<u32>(n + 2);
    S.mbeg = ar.take<u32>(n + 2);
    S.cursor = ar.take<u32>(n + 2);
    // strips: at most 1 GiB, at most 4 CTAs per SM, at least one CTA
    size_t per_block = (size_t)CM_WPB * n * sizeof(i64);
    size_t blocks = per_block ? ((size_t)1 << 30) / per_block : 1;
    if (blocks > 148 * 4) blocks = 148 * 4;
    if (blocks > ceil_div(n, CM_WPB)) blocks = ceil_div(n, CM_WPB);
    if (blocks < 1) blocks = 1;
    S.blocks = (int)blocks;
    S.strip = ar.take<i64>(blocks * CM_WPB * n);
    return CPH_OK;
}

template <typename T>
int fetch(State& S, const T* dev, T* host) {
    CUDA_TRY(cudaMemcpyAsync(host, dev, sizeof(T), cudaMemcpyDeviceToHost, S.st));
    CUDA_TRY(cudaStreamSynchronize(S.st));
    return CPH_OK;
}

int totals(State& S, const LGraph& g, const i32* comm) {
    CUDA_TRY(cudaMemsetAsync(S.tot, 0, sizeof(i64) * (size_t)g.n, S.st));
    CUDA_TRY(cudaMemsetAsync(S.csize, 0, sizeof(i32) * (size_t)g.n, S.st));
    comm_totals_kernel<<<tgrid(g.n), 256, 0, S.st>>>(comm, S.k, g.n, S.tot, S.csize);
    KERNEL_CHECK();
    return CPH_OK;
}

// exact modularity of (g, comm) given tot; synchronises
int quality(State& S, const LGraph& g, const i32* comm, double* Q) {
    CUDA_TRY(cudaMemsetAsync(&S.sc->in_sum, 0, sizeof(i64), S.st));
    internal_kernel<<<wgrid(g.n), CM_THREADS, 0, S.st>>>(g, comm, &S.sc->in_sum);
    KERNEL_CHECK();
    sumsq_kernel<<<1, 1024, 0, S.st>>>(S.tot, g.n, S.m2, &S.sc->sumsq);
    KERNEL_CHECK();
    Scalars h;
    TRY(fetch(S, S.sc, &h));
    *Q = (double)h.in_sum / S.m2 - S.gamma * h.sumsq;
    return CPH_OK;
}

int local_moving(State& S, const LGraph& g, bool* moved) {
    // rows of >= 128 entries on average (the dense level-0 graph of a clique expansion): a CTA per vertex keeps the
    // machine full where a warp per vertex would leave it latency bound
    const bool cta_mode = g.w == S.w0 && S.nnz0 / (g.n > 0 ? g.n : 1) >= 128;
    TRY(totals(S, g, S.comm));
    double bestQ;
    TRY(quality(S, g, S.comm, &bestQ));
    CUDA_TRY(cudaMemcpyAsync(S.best, S.comm, sizeof(i32) * (size_t)g.n, cudaMemcpyDeviceToDevice, S.st));
    bool improved = false, dirty = false;
    int stall = 0;
    // Phases per sweep.  A fixed count when the caller gave one.  Otherwise adaptive, from what the sweeps themselves show:
    // every proposal is decided on the snapshot of its phase and carries its PREDICTED modularity gain; the sweep's exact
    // gain is known afterwards.  When many vertices move together their decisions interfere and the sweep realises only
    // part of what it predicted -- that is when few phases end 3-7 % of |Q| below the sequential sweep on noisy graphs
    // (profiles/r2j_community_samephase.txt), while on graphs with a clear structure 8 phases lose nothing
    // (profiles/r2s_community_eta.txt).  So a clustering starts at 8 phases; the first sweep that realises less than
    // eta_min = 95 % of its prediction is UNDONE and marks the graph as noisy: from then on
    // P = clamp(moves of the previous sweep / 32, 8, 128), the rule that stays within 0.7 % of the sequential oracle on 30
    // noisy cases (profiles/r2k_community_adaptive.txt) at 2-3 x the time, which clean graphs now do not pay.
    const bool adaptive = S.n_phases <= 0;
    auto by_moves = [&](ull mv) {
        ull want = (mv + (ull)S.pdiv - 1) / (ull)S.pdiv;
        return (int)(want < 8 ? 8 : (want > (ull)S.pmax ? (ull)S.pmax : want));
    };
    // a level of at most pmax vertices (the aggregated levels, where one bad merge of two communities costs most) is swept
    // one vertex per phase: the sequential sweep, at no cost worth mentioning
    const bool tiny = adaptive && g.n <= S.pmax;
    int Pcur = adaptive ? (tiny ? g.n : (S.noisy ? by_moves((ull)g.n) : 8)) : S.n_phases;
    double Qcur = bestQ;
    for (int sweep = 0; sweep < S.max_sweeps; ++sweep) {
        int P = Pcur;
        if (P > g.n) P = g.n > 0 ? g.n : 1;
        if (adaptive && !S.noisy && !tiny)
            CUDA_TRY(cudaMemcpyAsync(S.snap, S.comm, sizeof(i32) * (size_t)g.n, cudaMemcpyDeviceToDevice, S.st));
        CUDA_TRY(cudaMemsetAsync(&S.sc->pred, 0, sizeof(ull), S.st));
        CUDA_TRY(cudaMemsetAsync(&S.sc->moves, 0, sizeof(ull), S.st));
        if (S.coop_blocks > 0) {
            // one cooperative launch for the whole sweep; the grid is what is co-resident and what the strips allow
            const int per = cta_mode ? 1 : CM_WPB;
            int cnt = (g.n + P - 1) / P;
            int blocks = (int)ceil_div((size_t)cnt, (size_t)per);
            if (blocks > S.coop_blocks) blocks = S.coop_blocks;
            if (blocks < 1) blocks = 1;
            LGraph gg = g;
            int Pn = P;
            i32* comm = S.comm;
            i64* tot = S.tot;
            i32* csize = S.csize;
            const i64* kk = S.k;
            double gamma = S.gamma, m2 = S.m2;
            i64* strip = S.strip;
            i32* prop = S.prop;
            i64* pgain = S.pgain;
            ull* mv = &S.sc->moves;
            ull* pred = &S.sc->pred;
            void* args[] = {&gg, &Pn, &comm, &tot, &csize, &kk, &gamma, &m2, &strip, &prop, &pgain, &mv, &pred};
            const void* fn = cta_mode ? (const void*)sweep_kernel<CM_THREADS> : (const void*)sweep_kernel<32>;
            CUDA_TRY(cudaLaunchCooperativeKernel(fn, dim3(blocks), dim3(CM_THREADS), args, 0, S.st));
            KERNEL_CHECK();
        } else
        for (int ph = 0; ph < P; ++ph) {
            int cnt = (g.n - ph + P - 1) / P;
            if (cnt <= 0) continue;
            if (cta_mode) {          // one CTA per vertex: S.blocks * CM_WPB strips are available
                int blocks = cnt < S.blocks * CM_WPB ? cnt : S.blocks * CM_WPB;
                move_propose_kernel<CM_THREADS><<<blocks, CM_THREADS, 0, S.st>>>(g, ph, P, 0, g.n, S.comm, S.tot, S.csize, S.k,
                                                                                 S.gamma, S.m2, S.strip, S.prop, S.pgain);
            } else {
                int blocks = (int)ceil_div((size_t)cnt, CM_WPB);
                if (blocks > S.blocks) blocks = S.blocks;
                move_propose_kernel<32><<<blocks, CM_THREADS, 0, S.st>>>(g, ph, P, 0, g.n, S.comm, S.tot, S.csize, S.k, S.gamma,
                                                                         S.m2, S.strip, S.prop, S.pgain);
            }
            KERNEL_CHECK();
            move_commit_kernel<<<tgrid(cnt), 256, 0, S.st>>>(ph, P, 0, g.n, S.comm, S.prop, S.pgain, S.k, S.tot, S.csize,
                                                             &S.sc->moves, &S.sc->pred);
            KERNEL_CHECK();
        }
        ++S.sweeps;
        ull mv;
        TRY(fetch(S, &S.sc->moves, &mv));
        if (mv == 0) break;
        double Q;
        TRY(quality(S, g, S.comm, &Q));
        if (adaptive) {
            ull pr;
            TRY(fetch(S, &S.sc->pred, &pr));
            const double dq_pred = 2.0 * ((double)(i64)pr / CM_PG_SCALE) / S.m2;
            const double eta = dq_pred > 1e-15 ? (Q - Qcur) / dq_pred : 1.0;
            if (S.trace) fprintf(stderr, "[community] n=%d sweep=%d P=%d moves=%llu Q=%.6f eta=%.4f noisy=%d\n", g.n, sweep, P, mv, Q, eta, (int)S.noisy);
            if (!S.noisy && !tiny && eta < S.eta_min) {
                // too much interference: undo the sweep, repeat it with the phase count its number of moves asks for
                S.noisy = true;
                ++S.retries;
                const int Pn = by_moves(mv);
                if (Pn > Pcur) {
                    CUDA_TRY(cudaMemcpyAsync(S.comm, S.snap, sizeof(i32) * (size_t)g.n, cudaMemcpyDeviceToDevice, S.st));
                    TRY(totals(S, g, S.comm));
                    Pcur = Pn;
                    --sweep;
                    continue;
                }
            }
            if (S.noisy && !tiny) Pcur = by_moves(mv);
        }
        Qcur = Q;
        if (Q > bestQ + 1e-14 * fabs(bestQ)) {
            bestQ = Q;
            CUDA_TRY(cudaMemcpyAsync(S.best, S.comm, sizeof(i32) * (size_t)g.n, cudaMemcpyDeviceToDevice, S.st));
            improved = true;
            dirty = false;
            stall = 0;
        } else {
            dirty = true;
            if (++stall >= 2) break;
        }
    }
    if (dirty) {
        CUDA_TRY(cudaMemcpyAsync(S.comm, S.best, sizeof(i32) * (size_t)g.n, cudaMemcpyDeviceToDevice, S.st));
        TRY(totals(S, g, S.comm));
    }
    *moved = improved;
    return CPH_OK;
}

int refine(State& S, const LGraph& g) {
    refine_init_kernel<<<tgrid(g.n), 256, 0, S.st>>>(S.k, g.n, S.refined, S.ssize, S.stot);
    KERNEL_CHECK();
    int blocks = (int)ceil_div((size_t)g.n, CM_WPB);
    if (blocks > S.blocks) blocks = S.blocks;
    for (int round = 0; round < 4 * 64; ++round) {
        CUDA_TRY(cudaMemsetAsync(S.ext, 0, sizeof(i64) * (size_t)g.n, S.st));
        refine_ext_kernel<<<wgrid(g.n), CM_THREADS, 0, S.st>>>(g, S.comm, S.refined, S.ext);
        KERNEL_CHECK();
        refine_propose_kernel<<<blocks, CM_THREADS, 0, S.st>>>(g, S.comm, S.tot, S.k, S.refined, S.ssize, S.stot, S.ext, S.gamma,
                                                               S.m2, S.strip, S.target);
        KERNEL_CHECK();
        CUDA_TRY(cudaMemsetAsync(&S.sc->moves, 0, sizeof(ull), S.st));
        refine_apply_kernel<<<tgrid(g.n), 256, 0, S.st>>>(S.target, S.k, g.n, S.refined, S.ssize, S.stot, &S.sc->moves);
        KERNEL_CHECK();
        ++S.rounds;
        ull mv;
        TRY(fetch(S, &S.sc->moves, &mv));
        if (mv == 0) break;
    }
    return CPH_OK;
}

// relabel lab[0..n) (values in [0, range)) to 0..ng-1 in increasing label order; synchronises
int compact_labels(State& S, i32* lab, int n, int range, int* ng) {
    CUDA_TRY(cudaMemsetAsync(S.flag, 0, sizeof(u32) * ((size_t)range + 1), S.st));
    flag_labels_kernel<<<tgrid(n), 256, 0, S.st>>>(lab, n, S.flag);
    KERNEL_CHECK();
    block_scan_u32_kernel<<<1, 1024, 0, S.st>>>(S.flag, range, &S.sc->total);
    KERNEL_CHECK();
    apply_map_kernel<<<tgrid(n), 256, 0, S.st>>>(lab, n, S.flag);
    KERNEL_CHECK();
    u32 t;
    TRY(fetch(S, &S.sc->total, &t));
    *ng = (int)t;
    return CPH_OK;
}

int aggregate(State& S, const LGraph& g, const i32* lab, int ng, int dst, LGraph* out) {
    CUDA_TRY(cudaMemsetAsync(S.mcnt, 0, sizeof(u32) * ((size_t)ng + 1), S.st));
    CUDA_TRY(cudaMemsetAsync(S.cursor, 0, sizeof(u32) * ((size_t)ng + 1), S.st));
    CUDA_TRY(cudaMemsetAsync(S.degsum, 0, sizeof(i64) * ((size_t)ng + 1), S.st));
    agg_count_kernel<<<tgrid(g.n), 256, 0, S.st>>>(g, lab, S.mcnt, S.degsum);
    KERNEL_CHECK();
    agg_offsets_kernel<<<1, 1024, 0, S.st>>>(S.mcnt, S.degsum, ng, S.mbeg, S.off);
    KERNEL_CHECK();
    agg_scatter_kernel<<<tgrid(g.n), 256, 0, S.st>>>(lab, g.n, S.mbeg, S.cursor, S.members);
    KERNEL_CHECK();
    int blocks = (int)ceil_div((size_t)ng, CM_WPB);
    if (blocks > S.blocks) blocks = S.blocks;
    if (ng <= AGG_DENSE_MAX && (size_t)ng * ng <= (size_t)S.blocks * CM_WPB * S.n0) {
        const int per_cta = (int)ceil_div((size_t)g.n, (size_t)148 * 8);
        agg_dense_kernel<<<(unsigned)ceil_div((size_t)g.n, (size_t)per_cta), CM_THREADS, sizeof(i64) * (size_t)ng, S.st>>>(
            g, lab, ng, S.members, per_cta, S.strip);
        KERNEL_CHECK();
        agg_emit_kernel<<<blocks, CM_THREADS, 0, S.st>>>(ng, S.off, S.strip, S.gbeg[dst], S.gend[dst], S.gidx[dst], S.gw[dst]);
        KERNEL_CHECK();
    } else {
        agg_rows_kernel<<<blocks, CM_THREADS, 0, S.st>>>(g, lab, ng, S.mbeg, S.members, S.off, S.strip, S.n0, S.gbeg[dst],
                                                         S.gend[dst], S.gidx[dst], S.gw[dst]);
        KERNEL_CHECK();
    }
    out->n = ng;
    out->beg = S.gbeg[dst];
    out->end = S.gend[dst];
    out->idx = S.gidx[dst];
    out->w = S.gw[dst];
    return CPH_OK;
}

// fixed-point level-0 graph from CSR(A): fills w0, k, m2.  *empty = 1 when the graph has no usable weight.
int prepare(State& S, const i64* indptr, const i32* indices, const float* d32, const double* d64, LGraph* g0, int* empty) {
    const int n = S.n0;
    CUDA_TRY(cudaMemsetAsync(S.sc, 0, sizeof(Scalars), S.st));
    absmax_kernel<<<wgrid(n), CM_THREADS, 0, S.st>>>(indptr, indices, d32, d64, n, &S.sc->absmax_bits);
    KERNEL_CHECK();
    ull bits;
    TRY(fetch(S, &S.sc->absmax_bits, &bits));
    double mx;
    memcpy(&mx, &bits, sizeof(double));
    *empty = 1;
    if (!(mx > 0)) return CPH_OK;
    int e;
    frexp(mx * (double)(S.nnz0 > 0 ? S.nnz0 : 1), &e);     // sum |w| < 2^e
    int s = 60 - e;
    if (s > 52) s = 52;
    to_fixed_kernel<<<wgrid(n), CM_THREADS, 0, S.st>>>(indptr, indices, d32, d64, n, ldexp(1.0, s), S.w0, S.k, &S.sc->m2);
    KERNEL_CHECK();
    i64 m2;
    TRY(fetch(S, &S.sc->m2, &m2));
    S.m2 = (double)m2;
    g0->n = n;
    g0->beg = indptr;
    g0->end = indptr + 1;
    g0->idx = indices;
    g0->w = S.w0;
    *empty = (m2 == 0);
    return CPH_OK;
}

int renumber_first_appearance(State& S, const i32* dev_in, i32* dev_out, u32* n_groups) {
    const size_t n = (size_t)S.n0;
    i32* h = (i32*)malloc(sizeof(i32) * (n + 1) * 2);
    if (!h) {
        cph_set_error(__FILE__, __LINE__, "host allocation failed");
        return CPH_ERR_ARG;
    }
    i32* map = h + n + 1;
    cudaError_t er = cudaMemcpyAsync(h, dev_in, sizeof(i32) * n, cudaMemcpyDeviceToHost, S.st);
    if (er == cudaSuccess) er = cudaStreamSynchronize(S.st);
    if (er != cudaSuccess) {
        free(h);
        CUDA_TRY(er);
    }
    for (size_t i = 0; i <= n; ++i) map[i] = -1;
    i32 gcount = 0;
    for (size_t v = 0; v < n; ++v) {
        i32 c = h[v];
        if (c < 0 || (size_t)c > n) c = (i32)n;
        if (map[c] < 0) map[c] = gcount++;
        h[v] = map[c];
    }
    er = cudaMemcpyAsync(dev_out, h, sizeof(i32) * n, cudaMemcpyHostToDevice, S.st);
    if (er == cudaSuccess) er = cudaStreamSynchronize(S.st);
    free(h);
    CUDA_TRY(er);
    *n_groups = (u32)gcount;
    return CPH_OK;
}

}  // namespace

extern "C" {

int cph_community(const int64_t* indptr, const int32_t* indices, const float* data32, const double* data64, uint32_t n,
                  uint64_t nnz, double resolution, int method, int n_iterations, int n_phases, int max_sweeps,
                  const int32_t* init_membership, int32_t* membership, double* h_quality, uint32_t* h_stats, void* ws,
                  size_t ws_bytes, size_t* h_ws_need, cph_stream_t stream) {
    ARG_CHECK(n > 0 && n < (1u << 30));
    ARG_CHECK(method == 0 || method == 1);
    State S;
    memset(&S, 0, sizeof(S));
    S.st = (cudaStream_t)stream;
    S.n0 = (int)n;
    S.nnz0 = (i64)nnz;
    S.gamma = resolution;
    S.n_phases = n_phases;              // <= 0: adaptive (local_moving)
    S.pmax = 128;
    S.pdiv = 32;
    if (const char* e = getenv("CPH_COMMUNITY_PMAX")) S.pmax = atoi(e) > 0 ? atoi(e) : S.pmax;     // experiment knobs
    if (const char* e = getenv("CPH_COMMUNITY_PDIV")) S.pdiv = atoi(e) > 0 ? atoi(e) : S.pdiv;
    S.eta_min = 0.95;
    S.trace = getenv("CPH_COMMUNITY_TRACE") != nullptr;
    if (const char* e = getenv("CPH_COMMUNITY_ETA")) S.eta_min = atof(e);
    S.max_sweeps = max_sweeps > 0 ? max_sweeps : 64;
    WS_BEGIN(ar);
    TRY(carve(S, ar));
    WS_QUERY_RETURN(ar);
    ARG_CHECK(indptr && membership && (nnz == 0 || (indices && (data32 || data64))));
    {   // co-resident grid of the one-launch sweep (CPH_COMMUNITY_SWEEP=launch: one launch per phase, for A/B timing)
        static int coop_max = -1;
        if (coop_max < 0) {
            int dev = 0, sms = 0, ok = 0, nb_cta = 0, nb_warp = 0;
            CUDA_TRY(cudaGetDevice(&dev));
            CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
            CUDA_TRY(cudaDeviceGetAttribute(&ok, cudaDevAttrCooperativeLaunch, dev));
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb_cta, sweep_kernel<CM_THREADS>, CM_THREADS, 0));
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb_warp, sweep_kernel<32>, CM_THREADS, 0));
            int nb = nb_cta < nb_warp ? nb_cta : nb_warp;
            if (nb > 2) nb = 2;                 // a grid barrier costs more the more CTAs take part
            coop_max = ok ? sms * nb : 0;
        }
        const char* e = getenv("CPH_COMMUNITY_SWEEP");
        S.coop_blocks = (e && strcmp(e, "launch") == 0) ? 0 : (coop_max < S.blocks ? coop_max : S.blocks);
    }
    if (n_iterations < 1) n_iterations = 1;
    CUDA_TRY(cudaMemsetAsync(S.strip, 0, sizeof(i64) * (size_t)S.blocks * CM_WPB * n, S.st));

    // initial membership (compacted) or singletons
    if (init_membership) {
        CUDA_TRY(cudaMemcpyAsync(S.member, init_membership, sizeof(i32) * n, cudaMemcpyDeviceToDevice, S.st));
    } else {
        iota_kernel<<<tgrid((int)n), 256, 0, S.st>>>(S.member, (int)n);
        KERNEL_CHECK();
    }
    LGraph g0;
    int empty = 0;
    TRY(prepare(S, indptr, indices, data32, data64, &g0, &empty));
    u32 groups = 0;
    if (empty) {
        TRY(renumber_first_appearance(S, S.member, membership, &groups));
        if (h_quality) *h_quality = 0.0;
        if (h_stats) {
            h_stats[0] = groups;
            h_stats[1] = h_stats[2] = h_stats[3] = 0;
        }
        return CPH_OK;
    }
    if (init_membership) {
        int nc;
        TRY(compact_labels(S, S.member, (int)n, (int)n, &nc));
    }

    double bestQ_it = 0.0;
    for (int it = 0; it < n_iterations; ++it) {
        LGraph g = g0;
        strength_kernel<<<wgrid(g.n), CM_THREADS, 0, S.st>>>(g, S.k);
        KERNEL_CHECK();
        iota_kernel<<<tgrid((int)n), 256, 0, S.st>>>(S.top, (int)n);
        KERNEL_CHECK();
        CUDA_TRY(cudaMemcpyAsync(S.comm, S.member, sizeof(i32) * n, cudaMemcpyDeviceToDevice, S.st));
        for (int level = 0; level < 64; ++level) {
            bool moved = false;
            TRY(local_moving(S, g, &moved));
            ++S.levels;
            int ng = 0;
            if (method == 1) {
                TRY(refine(S, g));
                CUDA_TRY(cudaMemcpyAsync(S.lab, S.refined, sizeof(i32) * (size_t)g.n, cudaMemcpyDeviceToDevice, S.st));
                TRY(compact_labels(S, S.lab, g.n, g.n, &ng));
                scatter_comm_kernel<<<tgrid(g.n), 256, 0, S.st>>>(S.lab, S.comm, g.n, S.init);
                KERNEL_CHECK();
            } else {
                CUDA_TRY(cudaMemcpyAsync(S.lab, S.comm, sizeof(i32) * (size_t)g.n, cudaMemcpyDeviceToDevice, S.st));
                TRY(compact_labels(S, S.lab, g.n, g.n, &ng));
            }
            gather_kernel<<<tgrid((int)n), 256, 0, S.st>>>(S.comm, S.top, (int)n, S.member);
            KERNEL_CHECK();
            bool done = (!moved && level > 0) || ng == g.n;
            if (method == 0 && !moved) done = true;
            if (done) break;
            LGraph a;
            TRY(aggregate(S, g, S.lab, ng, level & 1, &a));
            gather_kernel<<<tgrid((int)n), 256, 0, S.st>>>(S.lab, S.top, (int)n, S.top);
            KERNEL_CHECK();
            if (method == 1) {
                CUDA_TRY(cudaMemcpyAsync(S.comm, S.init, sizeof(i32) * (size_t)ng, cudaMemcpyDeviceToDevice, S.st));
                int nc;
                TRY(compact_labels(S, S.comm, ng, g.n, &nc));
            } else {
                iota_kernel<<<tgrid(ng), 256, 0, S.st>>>(S.comm, ng);
                KERNEL_CHECK();
            }
            g = a;
            strength_kernel<<<wgrid(g.n), CM_THREADS, 0, S.st>>>(g, S.k);
            KERNEL_CHECK();
        }
        // the next iteration starts from this partition: labels must be < n and compact
        int nc;
        TRY(compact_labels(S, S.member, (int)n, (int)n, &nc));
        if (n_iterations > 1) {     // keep the best partition over the iterations; stop when one brings nothing
            strength_kernel<<<wgrid(g0.n), CM_THREADS, 0, S.st>>>(g0, S.k);
            KERNEL_CHECK();
            TRY(totals(S, g0, S.member));
            double Qi;
            TRY(quality(S, g0, S.member, &Qi));
            if (it == 0 || Qi > bestQ_it + 1e-12 * fabs(bestQ_it)) {
                const bool first = it == 0;
                const double gain = first ? 1.0 : Qi - bestQ_it;
                bestQ_it = Qi;
                CUDA_TRY(cudaMemcpyAsync(S.best_it, S.member, sizeof(i32) * n, cudaMemcpyDeviceToDevice, S.st));
                if (!first && gain < 1e-7 * fabs(Qi)) break;
            } else {
                CUDA_TRY(cudaMemcpyAsync(S.member, S.best_it, sizeof(i32) * n, cudaMemcpyDeviceToDevice, S.st));
                break;
            }
        }
    }

    // final quality on the level-0 graph
    strength_kernel<<<wgrid(g0.n), CM_THREADS, 0, S.st>>>(g0, S.k);
    KERNEL_CHECK();
    TRY(totals(S, g0, S.member));
    double Q;
    TRY(quality(S, g0, S.member, &Q));
    TRY(renumber_first_appearance(S, S.member, membership, &groups));
    if (h_quality) *h_quality = Q;
    if (h_stats) {
        h_stats[0] = groups;
        h_stats[1] = (u32)S.levels;
        h_stats[2] = (u32)S.sweeps;
        h_stats[3] = (u32)S.rounds;
    }
    return CPH_OK;
}

// modularity of a given membership (labels in [0, n)) with the same fixed-point arithmetic
int cph_modularity(const int64_t* indptr, const int32_t* indices, const float* data32, const double* data64, uint32_t n,
                   uint64_t nnz, double resolution, const int32_t* membership, double* h_quality, void* ws, size_t ws_bytes,
                   size_t* h_ws_need, cph_stream_t stream) {
    ARG_CHECK(n > 0 && n < (1u << 30));
    State S;
    memset(&S, 0, sizeof(S));
    S.st = (cudaStream_t)stream;
    S.n0 = (int)n;
    S.nnz0 = (i64)nnz;
    S.gamma = resolution;
    WS_BEGIN(ar);
    S.sc = ar.take<Scalars>(1);
    S.w0 = ar.take<i64>(nnz > 0 ? nnz : 1);
    S.k = ar.take<i64>((size_t)n + 1);
    S.tot = ar.take<i64>((size_t)n + 1);
    S.csize = ar.take<i32>((size_t)n + 1);
    S.member = ar.take<i32>((size_t)n + 1);
    WS_QUERY_RETURN(ar);
    ARG_CHECK(indptr && membership && h_quality && (nnz == 0 || (indices && (data32 || data64))));
    LGraph g0;
    int empty = 0;
    TRY(prepare(S, indptr, indices, data32, data64, &g0, &empty));
    *h_quality = 0.0;
    if (empty) return CPH_OK;
    TRY(totals(S, g0, membership));
    TRY(quality(S, g0, membership, h_quality));
    return CPH_OK;
}

}  // extern "C"
