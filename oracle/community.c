/* CPU oracle for the clustering stage -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * The reference hands the graph to python-igraph 0.11.9 / igraph C 0.10.16 (not in the reference tree, not
 * installable here): cphasing/algorithms/hypergraph.py:1239-1252
 *     community_multilevel(weights, resolution)                           (Louvain, Blondel et al. 2008)
 *     community_leiden(weights, resolution_parameter, objective_function="modularity")  (Traag et al. 2019)
 * This file restates the PUBLISHED algorithms sequentially in double precision:
 *   - Louvain: repeated local moving (vertices in index order, best gain, ties -> lowest community id) until no
 *     vertex moves, then aggregation, until a level moves nothing.
 *   - Leiden: local moving, then refinement inside each community (singletons only, well-connectedness tests
 *     of the paper, greedy choice = the theta -> 0 limit of the randomised rule), aggregation by the refined
 *     partition with the unrefined one as the initial assignment; n_iterations passes.
 * Quality: Q = sum_c [ in_c / 2m - gamma (tot_c / 2m)^2 ]  (python-igraph's resolution convention for
 * objective_function="modularity": resolution / 2m, node weights = strength).
 * PARITY UNPINNED against igraph itself (RNG-driven vertex order there); pinned only against networkx's
 * independent modularity / Louvain implementations in tests/test_oracle_community.py.
 *
 * Graph input: symmetric CSR (both directions stored), self loops allowed (an entry (v,v) of weight w adds w
 * to k_v and to in_c once).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

typedef struct {
    int n;
    int64_t* ptr;
    int* idx;
    double* w;
} graph_t;

static void graph_free(graph_t* g) {
    free(g->ptr);
    free(g->idx);
    free(g->w);
}

static double total_weight(const graph_t* g) {
    double s = 0;
    for (int64_t t = 0; t < g->ptr[g->n]; ++t) s += g->w[t];
    return s; /* = 2m */
}

/* one local-moving phase; comm (in/out), returns number of moves */
static long local_moving(const graph_t* g, int* comm, double gamma, double m2, const int* bound /* or NULL */) {
    int n = g->n;
    double* k = (double*)calloc(n, sizeof(double));
    double* tot = (double*)calloc(n, sizeof(double));
    double* acc = (double*)calloc(n, sizeof(double));
    int* touched = (int*)malloc(sizeof(int) * (size_t)n);
    long* stamp = (long*)calloc(n, sizeof(long));
    long clock = 0;
    for (int v = 0; v < n; ++v) {
        for (int64_t t = g->ptr[v]; t < g->ptr[v + 1]; ++t) k[v] += g->w[t];
    }
    for (int v = 0; v < n; ++v) tot[comm[v]] += k[v];
    long moves_total = 0;
    for (int sweep = 0; sweep < 1000; ++sweep) {
        long moves = 0;
        for (int v = 0; v < n; ++v) {
            int nt = 0, old = comm[v];
            ++clock;
            for (int64_t t = g->ptr[v]; t < g->ptr[v + 1]; ++t) {
                int u = g->idx[t];
                if (u == v || g->w[t] == 0) continue;
                if (bound && bound[u] != bound[v]) continue;
                int c = comm[u];
                if (stamp[c] != clock) {
                    stamp[c] = clock;
                    if (c != old) touched[nt++] = c;
                }
                acc[c] += g->w[t];
            }
            double stay = acc[old] - gamma * k[v] * (tot[old] - k[v]) / m2;
            double best = stay;
            int bc = old;
            for (int q = 0; q < nt; ++q) {
                int c = touched[q];
                double gain = acc[c] - gamma * k[v] * tot[c] / m2;
                if (gain > best || (gain == best && bc != old && c < bc)) {
                    best = gain;
                    bc = c;
                }
            }
            for (int q = 0; q < nt; ++q) acc[touched[q]] = 0;
            acc[old] = 0;
            if (bc != old) {
                tot[old] -= k[v];
                tot[bc] += k[v];
                comm[v] = bc;
                ++moves;
            }
        }
        moves_total += moves;
        if (moves == 0) break;
    }
    free(k); free(tot); free(acc); free(touched); free(stamp);
    return moves_total;
}

/* renumber labels to 0..g-1 in order of first appearance; returns g */
static int renumber(int* lab, int n) {
    int mx = 0;
    for (int v = 0; v < n; ++v) mx = lab[v] > mx ? lab[v] : mx;
    int* map = (int*)malloc(sizeof(int) * ((size_t)mx + 1));
    for (int i = 0; i <= mx; ++i) map[i] = -1;
    int g = 0;
    for (int v = 0; v < n; ++v) {
        if (map[lab[v]] < 0) map[lab[v]] = g++;
        lab[v] = map[lab[v]];
    }
    free(map);
    return g;
}

/* aggregate g by labels (0..ng-1) into a dense-accumulated CSR */
static graph_t aggregate(const graph_t* g, const int* lab, int ng) {
    graph_t a;
    a.n = ng;
    double* dense = (double*)calloc((size_t)ng * ng, sizeof(double));
    for (int v = 0; v < g->n; ++v)
        for (int64_t t = g->ptr[v]; t < g->ptr[v + 1]; ++t) dense[(size_t)lab[v] * ng + lab[g->idx[t]]] += g->w[t];
    int64_t nnz = 0;
    for (size_t i = 0; i < (size_t)ng * ng; ++i) nnz += dense[i] != 0;
    a.ptr = (int64_t*)malloc(sizeof(int64_t) * ((size_t)ng + 1));
    a.idx = (int*)malloc(sizeof(int) * (size_t)(nnz > 0 ? nnz : 1));
    a.w = (double*)malloc(sizeof(double) * (size_t)(nnz > 0 ? nnz : 1));
    int64_t p = 0;
    for (int i = 0; i < ng; ++i) {
        a.ptr[i] = p;
        for (int j = 0; j < ng; ++j)
            if (dense[(size_t)i * ng + j] != 0) {
                a.idx[p] = j;
                a.w[p++] = dense[(size_t)i * ng + j];
            }
    }
    a.ptr[ng] = p;
    free(dense);
    return a;
}

/* Leiden refinement (Traag 2019, Alg. 4, greedy limit): inside each community of `comm`, start from
 * singletons; a vertex that is still a singleton and is well connected to its community joins the
 * well-connected sub-community with the largest non-negative gain.  refined (out): sub-community label = a
 * vertex id. */
static void refine(const graph_t* g, const int* comm, double gamma, double m2, int* refined) {
    int n = g->n;
    double* k = (double*)calloc(n, sizeof(double));
    double* ctot = (double*)calloc(n, sizeof(double));   /* community totals */
    double* stot = (double*)calloc(n, sizeof(double));   /* sub-community totals */
    double* ext = (double*)calloc(n, sizeof(double));    /* E(S, C - S) */
    int* ssize = (int*)malloc(sizeof(int) * (size_t)n);
    double* acc = (double*)calloc(n, sizeof(double));
    int* touched = (int*)malloc(sizeof(int) * (size_t)n);
    long* stamp = (long*)calloc(n, sizeof(long));
    long clock = 0;
    for (int v = 0; v < n; ++v) {
        for (int64_t t = g->ptr[v]; t < g->ptr[v + 1]; ++t) {
            k[v] += g->w[t];
            int u = g->idx[t];
            if (u != v && comm[u] == comm[v]) ext[v] += g->w[t];
        }
        refined[v] = v;
        ssize[v] = 1;
    }
    for (int v = 0; v < n; ++v) {
        ctot[comm[v]] += k[v];
        stot[v] = k[v];
    }
    for (int v = 0; v < n; ++v) {
        if (ssize[refined[v]] != 1) continue;                              /* only singletons move */
        double cT = ctot[comm[v]];
        if (ext[v] < gamma * k[v] * (cT - k[v]) / m2) continue;            /* v not well connected */
        int nt = 0;
        ++clock;
        for (int64_t t = g->ptr[v]; t < g->ptr[v + 1]; ++t) {
            int u = g->idx[t];
            if (u == v || comm[u] != comm[v] || g->w[t] == 0) continue;
            int s = refined[u];
            if (stamp[s] != clock) {
                stamp[s] = clock;
                touched[nt++] = s;
            }
            acc[s] += g->w[t];
        }
        double best = -1;  /* staying a singleton has gain 0: only non-negative gains are eligible */
        int bs = -1;
        for (int q = 0; q < nt; ++q) {
            int s = touched[q];
            if (ext[s] < gamma * stot[s] * (cT - stot[s]) / m2) continue;  /* S not well connected */
            double gain = acc[s] - gamma * k[v] * stot[s] / m2;
            if (gain >= 0 && (bs < 0 || gain > best || (gain == best && s < bs))) {
                best = gain;
                bs = s;
            }
        }
        if (bs >= 0) {
            ext[bs] = ext[bs] + ext[v] - 2 * acc[bs];
            stot[bs] += k[v];
            ssize[bs] += 1;
            ssize[v] = 0;
            refined[v] = bs;
        }
        for (int q = 0; q < nt; ++q) acc[touched[q]] = 0;
    }
    free(k); free(ctot); free(stot); free(ext); free(ssize); free(acc); free(touched); free(stamp);
}

static graph_t graph_copy_in(int n, const int64_t* ptr, const int* idx, const double* w) {
    graph_t g;
    g.n = n;
    int64_t nnz = ptr[n];
    g.ptr = (int64_t*)malloc(sizeof(int64_t) * ((size_t)n + 1));
    g.idx = (int*)malloc(sizeof(int) * (size_t)(nnz > 0 ? nnz : 1));
    g.w = (double*)malloc(sizeof(double) * (size_t)(nnz > 0 ? nnz : 1));
    memcpy(g.ptr, ptr, sizeof(int64_t) * ((size_t)n + 1));
    memcpy(g.idx, idx, sizeof(int) * (size_t)nnz);
    memcpy(g.w, w, sizeof(double) * (size_t)nnz);
    return g;
}

double oracle_modularity(int n, const int64_t* ptr, const int* idx, const double* w, const int* member, double gamma) {
    double m2 = 0;
    double* tot = (double*)calloc(n, sizeof(double));
    double* in = (double*)calloc(n, sizeof(double));
    for (int v = 0; v < n; ++v)
        for (int64_t t = ptr[v]; t < ptr[v + 1]; ++t) {
            m2 += w[t];
            tot[member[v]] += w[t];
            if (member[idx[t]] == member[v]) in[member[v]] += w[t];
        }
    double q = 0;
    if (m2 != 0)
        for (int c = 0; c < n; ++c) q += in[c] / m2 - gamma * (tot[c] / m2) * (tot[c] / m2);
    free(tot); free(in);
    return q;
}

/* method: 0 = Louvain, 1 = Leiden.  member: in = initial membership (or all -1 for singletons), out = result
 * renumbered 0..g-1 by first appearance.  Returns the number of communities. */
int oracle_community(int n, const int64_t* ptr, const int* idx, const double* w, double gamma, int method, int n_iterations,
                     int* member) {
    graph_t g0 = graph_copy_in(n, ptr, idx, w);
    const double m2 = total_weight(&g0);
    int has_init = n > 0 && member[0] >= 0;
    if (!has_init)
        for (int v = 0; v < n; ++v) member[v] = v;
    if (m2 == 0) {
        graph_free(&g0);
        return renumber(member, n);
    }
    if (method == 0) n_iterations = 1;
    for (int it = 0; it < n_iterations; ++it) {
        graph_t g = graph_copy_in(g0.n, g0.ptr, g0.idx, g0.w);
        int* top = (int*)malloc(sizeof(int) * (size_t)n);      /* level vertex of each original vertex */
        int* comm = (int*)malloc(sizeof(int) * (size_t)n);     /* community of each level vertex */
        for (int v = 0; v < n; ++v) {
            top[v] = v;
            comm[v] = member[v];
        }
        renumber(comm, n);
        for (int level = 0; level < 64; ++level) {
            long moves = local_moving(&g, comm, gamma, m2, NULL);
            int* lab = (int*)malloc(sizeof(int) * (size_t)g.n);
            int* init = NULL;
            int ng;
            if (method == 1) {
                refine(&g, comm, gamma, m2, lab);
                ng = renumber(lab, g.n);
                init = (int*)malloc(sizeof(int) * (size_t)ng);
                for (int v = 0; v < g.n; ++v) init[lab[v]] = comm[v];
            } else {
                memcpy(lab, comm, sizeof(int) * (size_t)g.n);
                ng = renumber(lab, g.n);
            }
            for (int v = 0; v < n; ++v) member[v] = comm[top[v]];
            int done = (moves == 0 && level > 0) || ng == g.n;
            if (method == 0 && moves == 0) done = 1;
            if (done) {
                free(lab);
                free(init);
                break;
            }
            graph_t a = aggregate(&g, lab, ng);
            for (int v = 0; v < n; ++v) top[v] = lab[top[v]];
            graph_free(&g);
            g = a;
            for (int c = 0; c < ng; ++c) comm[c] = init ? init[c] : c;
            if (init) renumber(comm, ng);
            free(lab);
            free(init);
        }
        graph_free(&g);
        free(top);
        free(comm);
    }
    graph_free(&g0);
    return renumber(member, n);
}
