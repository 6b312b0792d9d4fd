// Shared device/host helpers for libcph (sm_100a).  No torch types, no third-party headers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <stdio.h>

typedef uint8_t u8;
typedef uint16_t u16;
typedef uint32_t u32;
typedef uint64_t u64;
typedef int32_t i32;
typedef int64_t i64;

// status codes (mirrored in include/cph.h)
#define CPH_OK 0
#define CPH_ERR_CUDA 1
#define CPH_ERR_ARG 2
#define CPH_ERR_WORKSPACE 3
#define CPH_ERR_TOO_LARGE 4
#define CPH_ERR_NOT_CONVERGED 5

void cph_set_error(const char* file, int line, const char* what);

#define CUDA_TRY(expr)                                                              \
    do {                                                                            \
        cudaError_t e__ = (expr);                                                   \
        if (e__ != cudaSuccess) {                                                   \
            cph_set_error(__FILE__, __LINE__, cudaGetErrorString(e__));             \
            return CPH_ERR_CUDA;                                                    \
        }                                                                           \
    } while (0)

extern unsigned long long g_cph_launches;   // api.cu; host-side counter behind cph_launch_count()
#define KERNEL_CHECK()                                                              \
    do {                                                                            \
        ++g_cph_launches;                                                           \
        CUDA_TRY(cudaGetLastError());                                               \
    } while (0)

#define ARG_CHECK(cond)                                                             \
    do {                                                                            \
        if (!(cond)) {                                                              \
            cph_set_error(__FILE__, __LINE__, "bad argument: " #cond);              \
            return CPH_ERR_ARG;                                                     \
        }                                                                           \
    } while (0)

#define TRY(expr)                                                                   \
    do {                                                                            \
        int s__ = (expr);                                                           \
        if (s__ != CPH_OK) return s__;                                              \
    } while (0)

static inline size_t ceil_div(size_t a, size_t b) { return (a + b - 1) / b; }
static inline int bits_for(uint64_t max_value) {   // number of bits needed to hold max_value (>= 1)
    int b = 1;
    while (b < 64 && (max_value >> b) != 0) ++b;
    return b;
}
static inline unsigned grid_for(size_t n, int threads) { return (unsigned)ceil_div(n > 0 ? n : 1, (size_t)threads); }

// workspace-query protocol shared by every entry point that takes (ws, ws_bytes, h_ws_need)
#define WS_BEGIN(ar_name)                                                           \
    Arena ar_name(ws, ws ? ws_bytes : 0);                                           \
    if (ws == nullptr) ar_name.dry = true;
#define WS_QUERY_RETURN(ar_name)                                                    \
    do {                                                                            \
        if ((ar_name).dry) {                                                        \
            if (h_ws_need) *h_ws_need = (ar_name).off + 256;                        \
            return CPH_OK;                                                          \
        }                                                                           \
        ARENA_CHECK(ar_name);                                                       \
    } while (0)

#define CPH_HEAD 0x80000000u
#define CPH_ROWM 0x7fffffffu
#define CPH_CNTM 0x00ffffffu

// ---------------------------------------------------------------------------------------------
// Workspace arena: every entry point carves its temporaries out of one caller-owned buffer
// (a torch tensor on the Python side).  The *_workspace() query runs the same carve sequence in
// dry mode, so the two can never disagree.
// ---------------------------------------------------------------------------------------------
struct Arena {
    char* base;
    size_t cap;
    size_t off;
    bool dry;
    bool ok;
    Arena(void* b, size_t c) : base((char*)b), cap(c), off(0), dry(b == nullptr && c == 0), ok(true) {}
    template <typename T>
    T* take(size_t count) {
        size_t bytes = (count * sizeof(T) + 255) & ~(size_t)255;
        size_t o = off;
        off += bytes;
        if (dry) return (T*)nullptr;
        if (off > cap) {
            ok = false;
            return (T*)nullptr;
        }
        return (T*)(base + o);
    }
};

#define ARENA_CHECK(ar)                                                             \
    do {                                                                            \
        if (!(ar).dry && !(ar).ok) {                                                \
            cph_set_error(__FILE__, __LINE__, "workspace too small");               \
            return CPH_ERR_WORKSPACE;                                               \
        }                                                                           \
    } while (0)

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ u32 lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ u32 lanemask_lt() {
    u32 m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// block-wide exclusive scan of one value per thread; THREADS must be a multiple of 32, <= 1024.
// `sm` needs THREADS/32 + 1 entries.  Returns the exclusive prefix; *total gets the block sum.
template <typename T, int THREADS>
__device__ __forceinline__ T block_excl_scan(T v, T* sm, T* total) {
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        T t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (u32)o) incl += t;
    }
    if (lane == 31) sm[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        constexpr int NW = THREADS / 32;
        T w = (lane < NW) ? sm[lane] : (T)0;
        T wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            T t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= (u32)o) wi += t;
        }
        if (lane < NW) sm[lane] = wi - w;   // exclusive warp offsets
        if (lane == 31) sm[NW] = wi;        // block total (lanes >= NW contribute 0)
    }
    __syncthreads();
    T res = sm[warp] + (incl - v);
    *total = sm[THREADS / 32];
    __syncthreads();   // sm reusable afterwards
    return res;
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
#endif

// ---------------------------------------------------------------------------------------------
// internal primitives (scan.cu / radix.cu)
// ---------------------------------------------------------------------------------------------
// in-place exclusive scan of n values; if total_out != nullptr the grand total is written there
// (device pointer).  Carves its scratch from `ar`.
int scan_u32(u32* data, size_t n, u32* total_out, Arena& ar, cudaStream_t st);
int scan_u64(u64* data, size_t n, u64* total_out, Arena& ar, cudaStream_t st);

// LSD radix sort of u64 keys over bits [bit_lo, bit_hi).  Stable.  `a` holds the input, `b` is a
// same-sized scratch; *result_in_b tells where the sorted keys ended up.  VAL_BYTES in {0,4,8}.
int radix_sort_u64(u64* ka, u64* kb, void* va, void* vb, int val_bytes, size_t n, int bit_lo, int bit_hi,
                   int* result_in_b, Arena& ar, cudaStream_t st);
