// Microbenchmarks behind the clique-expansion design: what one B200 sustains for
//   (a) scattered 64/32-bit RED.ADD into an L2-resident (64 MB) and an HBM-resident (1.6 GB) buffer,
//   (b) shared-memory atomics at scattered addresses (u32 ATOMS.ADD, u64 atomicAdd, u64 as 2 x u32 + carry),
//   (c) scattered 32-byte sector gathers (the edge-run fetch of a row-block / Gustavson expansion),
//   (d) bulk shared->global reductions (cp.reduce.async.bulk .add.u64: the strip flush of a row-block kernel).
// Every figure is operations (or bytes) per second over CUDA-event time, best of 3 after a warm-up launch.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/atomics atomics.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

typedef unsigned long long u64;
typedef unsigned int u32;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ u32 mix(u32 x) {   // cheap integer hash (one multiply-xorshift round per draw)
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

// ---- (a) global REDs ----
template <typename T>
__global__ void __launch_bounds__(512) red_global(T* buf, u64 cells, int per_thread, u32 hot_cells, u32 hot_permille) {
    u32 s = (blockIdx.x * 512u + threadIdx.x) * 2654435761u + 12345u;
    for (int k = 0; k < per_thread; ++k) {
        s = mix(s + k);
        u64 c;
        if (hot_cells && (s % 1000u) < hot_permille) c = (u64)(mix(s ^ 0x9e3779b9u) % hot_cells) * 47ull % cells;
        else c = ((u64)s * (u64)mix(s ^ 0x85ebca6bu)) % cells;
        atomicAdd(&buf[c], (T)(s | 1u));
    }
}

// ---- (b) shared-memory atomics ----
// mode 0: u32 atomicAdd; 1: u64 atomicAdd; 2: u64 as two u32 atomics with carry; 3: non-atomic u32 load/add/store (racy: rate only)
template <int MODE>
__global__ void __launch_bounds__(512) smem_atomics(u64* out, int per_thread, u32 cells) {
    extern __shared__ u32 sm[];
    for (u32 i = threadIdx.x; i < cells * (MODE == 1 || MODE == 2 ? 2u : 1u); i += 512) sm[i] = 0;
    __syncthreads();
    u32 s = (blockIdx.x * 512u + threadIdx.x) * 2654435761u + 777u;
    u64* sm64 = (u64*)sm;
    for (int k = 0; k < per_thread; ++k) {
        s = mix(s + k);
        const u32 c = s % cells;
        const u64 v = ((u64)s << 8) | 1u;
        if (MODE == 0) atomicAdd(&sm[c], (u32)v);
        else if (MODE == 1) atomicAdd(&sm64[c], v);
        else if (MODE == 2) {
            const u32 lo = (u32)v, hi = (u32)(v >> 32);
            const u32 old = atomicAdd(&sm[2 * c], lo);
            const u32 carry = (old + lo) < old ? 1u : 0u;
            if (hi + carry) atomicAdd(&sm[2 * c + 1], hi + carry);
        } else sm[c] += (u32)v;
    }
    __syncthreads();
    u64 acc = 0;
    for (u32 i = threadIdx.x; i < cells; i += 512) acc += sm[i];
    if (acc == 0x1234567ull) out[0] = acc;
}

// ---- (c) scattered 32-byte sector gathers ----
__global__ void __launch_bounds__(512) gather32(const uint4* __restrict__ buf, u64 sectors, int per_thread, u64* out) {
    u32 s = (blockIdx.x * 512u + threadIdx.x) * 2654435761u + 99u;
    u32 acc = 0;
    for (int k = 0; k < per_thread; k += 4) {
        u64 a[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            s = mix(s + k + u);
            a[u] = (((u64)s * (u64)mix(s ^ 0x85ebca6bu)) % sectors) * 2;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint4 x = buf[a[u]], y = buf[a[u] + 1];
            acc += x.x ^ y.w;
        }
    }
    if (acc == 0x12345u) out[0] = acc;
}

// ---- (d) bulk reduce shared -> global ----
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(128) bulk_reduce(u64* buf, u64 cells, u32 strip_cells, int reps) {
    extern __shared__ __align__(128) u64 strip[];
    for (u32 i = threadIdx.x; i < strip_cells; i += 128) strip[i] = i + 1;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        u32 s = blockIdx.x * 2654435761u + 5u;
        const u64 slots = cells / strip_cells;
        for (int k = 0; k < reps; ++k) {
            s = mix(s + k);
            u64* dst = buf + (u64)(s % slots) * strip_cells;
            asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.u64 [%0], [%1], %2;" ::"l"(dst),
                         "r"(smem_u32(strip)), "r"(strip_cells * 8u)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            if ((k & 7) == 7) asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

template <class F>
static float best_ms(F f) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    f();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        CK(cudaEventRecord(a));
        f();
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main() {
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, 0));
    printf("# %s, %d SMs, L2 %d MB\n", p.name, p.multiProcessorCount, p.l2CacheSize >> 20);
    const u64 big = 1600ull << 20;
    void* buf;
    CK(cudaMalloc(&buf, big));
    CK(cudaMemset(buf, 0, big));
    u64* out;
    CK(cudaMalloc(&out, 64));
    const int sms = p.multiProcessorCount;

    printf("\n## (a) scattered global RED.ADD (no return), uniform random cells\n");
    printf("%-10s %-8s %-10s %-10s %12s %10s\n", "type", "buffer", "ctas/SM", "hot", "G RED/s", "ms");
    for (u64 mb : {32ull, 64ull, 256ull, 1600ull}) {
        for (int cps : {2, 4}) {
            const int per = 2048;
            const double ops = (double)sms * cps * 512 * per;
            float ms = best_ms([&] { red_global<u64><<<sms * cps, 512>>>((u64*)buf, (mb << 20) / 8, per, 0, 0); });
            printf("%-10s %-8llu %-10d %-10s %12.1f %10.3f\n", "u64", mb, cps, "-", ops / ms / 1e6, ms);
        }
    }
    for (u64 mb : {32ull, 800ull}) {
        const int per = 2048, cps = 4;
        const double ops = (double)sms * cps * 512 * per;
        float ms = best_ms([&] { red_global<u32><<<sms * cps, 512>>>((u32*)buf, (mb << 20) / 4, per, 0, 0); });
        printf("%-10s %-8llu %-10d %-10s %12.1f %10.3f\n", "u32", mb, cps, "-", ops / ms / 1e6, ms);
    }
    // workload-C-like skew: 72 % of the updates into 168k hot cells of the 64 MB triangle
    {
        const int per = 2048, cps = 4;
        const double ops = (double)sms * cps * 512 * per;
        float ms = best_ms([&] { red_global<u64><<<sms * cps, 512>>>((u64*)buf, (64ull << 20) / 8, per, 168000, 720); });
        printf("%-10s %-8d %-10d %-10s %12.1f %10.3f\n", "u64", 64, cps, "72%/168k", ops / ms / 1e6, ms);
        ms = best_ms([&] { red_global<u64><<<sms * cps, 512>>>((u64*)buf, (1600ull << 20) / 8, per, 2400000, 720); });
        printf("%-10s %-8d %-10d %-10s %12.1f %10.3f\n", "u64", 1600, cps, "72%/2.4M", ops / ms / 1e6, ms);
    }

    printf("\n## (b) shared-memory atomics, scattered cells, 512 threads per CTA\n");
    printf("%-28s %-8s %-8s %12s %12s\n", "mode", "cells", "ctas/SM", "G op/s", "op/clk/SM");
    const double clk = p.clockRate * 1e3;
    auto smem_run = [&](int mode, u32 cells, int cps, const char* nm) {
        const int per = 4096;
        const size_t bytes = (size_t)cells * ((mode == 1 || mode == 2) ? 8 : 4);
        const double ops = (double)sms * cps * 512 * per;
        float ms = 0;
        if (mode == 0) { CK(cudaFuncSetAttribute(smem_atomics<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            ms = best_ms([&] { smem_atomics<0><<<sms * cps, 512, bytes>>>(out, per, cells); }); }
        if (mode == 1) { CK(cudaFuncSetAttribute(smem_atomics<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            ms = best_ms([&] { smem_atomics<1><<<sms * cps, 512, bytes>>>(out, per, cells); }); }
        if (mode == 2) { CK(cudaFuncSetAttribute(smem_atomics<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            ms = best_ms([&] { smem_atomics<2><<<sms * cps, 512, bytes>>>(out, per, cells); }); }
        if (mode == 3) { CK(cudaFuncSetAttribute(smem_atomics<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            ms = best_ms([&] { smem_atomics<3><<<sms * cps, 512, bytes>>>(out, per, cells); }); }
        printf("%-28s %-8u %-8d %12.1f %12.2f\n", nm, cells, cps, ops / ms / 1e6, ops / (ms * 1e-3) / clk / sms);
    };
    for (int cps : {1, 2, 4}) {
        smem_run(0, 8192, cps, "u32 atomicAdd");
        smem_run(1, 4096, cps, "u64 atomicAdd");
        smem_run(2, 4096, cps, "u64 = 2 x u32 + carry");
        smem_run(3, 8192, cps, "u32 ld/add/st (non-atomic)");
    }
    smem_run(0, 40960, 1, "u32 atomicAdd (160 KB)");
    smem_run(2, 20480, 1, "u64 = 2 x u32 + carry (160 KB)");

    printf("\n## (c) scattered 32-byte sector gathers (2 x LDG.128)\n");
    printf("%-8s %-8s %12s %12s %10s\n", "buffer", "ctas/SM", "G sector/s", "GB/s", "ms");
    for (u64 mb : {64ull, 1440ull}) {
        for (int cps : {2, 4}) {
            const int per = 1024;
            const double ops = (double)sms * cps * 512 * per;
            float ms = best_ms([&] { gather32<<<sms * cps, 512>>>((const uint4*)buf, (mb << 20) / 32, per, out); });
            printf("%-8llu %-8d %12.1f %12.1f %10.3f\n", mb, cps, ops / ms / 1e6, ops * 32 / ms / 1e6, ms);
        }
    }

    printf("\n## (d) cp.reduce.async.bulk .add.u64 shared -> global, one strip per op, random strip slots\n");
    printf("%-8s %-10s %-8s %12s %12s\n", "buffer", "strip B", "ctas/SM", "GB/s", "G cell/s");
    CK(cudaFuncSetAttribute(bulk_reduce, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    for (u64 mb : {64ull, 1600ull}) {
        for (u32 strip : {512u, 4096u}) {            // cells
            for (int cps : {2, 8}) {
                const int reps = strip == 512 ? 512 : 128;
                const double cells_done = (double)sms * cps * reps * strip;
                float ms = best_ms([&] { bulk_reduce<<<sms * cps, 128, strip * 8>>>((u64*)buf, (mb << 20) / 8, strip, reps); });
                printf("%-8llu %-10u %-8d %12.1f %12.1f\n", mb, strip * 8, cps, cells_done * 8 / ms / 1e6, cells_done / ms / 1e6);
            }
        }
    }
    return 0;
}
