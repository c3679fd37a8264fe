// random_access.cu -- what a hash-table probe can cost on this GPU: independent random accesses over a footprint of F
// bytes (8-byte loads, 32-byte loads, 64-bit CAS), as G accesses / s, against the footprint (TLB reach, L2) and the
// number of independent accesses per thread.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/random_access tools/random_access.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint64_t mix(uint64_t x) {
    x ^= x >> 33; x *= 0xFF51AFD7ED558CCDull; x ^= x >> 33; x *= 0xC4CEB9FE1A85EC53ull; x ^= x >> 29;
    return x;
}

// MODE 0: 8-byte ld.cg   1: 32-byte ld (v4.b64)   2: atomicCAS(0 -> 0) 64-bit   3: prefetch.global.L2 only
// WIN > 0: thread blocks walk the footprint window by window (all accesses of a grid-wide time slice fall into one
// window of WIN bytes): what a radix partition of the probes by table region would give
template <int MODE, int U>
__global__ void __launch_bounds__(256) k_rand(const uint64_t* __restrict__ tab, uint64_t words, int iters, uint64_t win_words, uint64_t* sink) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t acc = 0;
    const uint64_t nwin = win_words ? words / win_words : 1;
    for (int it = 0; it < iters; ++it) {
        uint64_t idx[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            uint64_t h = mix(t * 0x9E3779B97F4A7C15ull + (uint64_t)(it * U + u));
            if (win_words) {
                const uint64_t w = ((uint64_t)it * nwin / iters);  // the window of this time slice
                idx[u] = w * win_words + (h % win_words);
            } else {
                idx[u] = h % words;
            }
            if (MODE == 1) idx[u] &= ~3ull;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (MODE == 0) {
                acc += __ldcg(tab + idx[u]);
            } else if (MODE == 1) {
                uint64_t a, b, c, d;
                asm volatile("ld.global.v4.b64 {%0, %1, %2, %3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(tab + idx[u]));
                acc += a ^ b ^ c ^ d;
            } else if (MODE == 2) {
                acc += atomicCAS((unsigned long long*)(tab + idx[u]), 0xFFFFFFFFFFFFFFFFull, 1ull);
            } else if (MODE == 3) {
                asm volatile("prefetch.global.L2 [%0];" ::"l"(tab + idx[u]));
            } else if (MODE == 4) {  // what a hash-table insert does: load the line, then CAS into it
                const uint64_t v = __ldcg(tab + idx[u]);
                acc += atomicCAS((unsigned long long*)(tab + idx[u]), v + 0xFFFFFFFFull, 1ull);
            } else if (MODE == 5) {  // load, then a plain 8-byte store into the line
                const uint64_t v = __ldcg(tab + idx[u]);
                if (v == 0x12345ull) __stcg((unsigned long long*)(tab + idx[u]) + 1, 0ull);
                else __stcg((unsigned long long*)(tab + (idx[u] ^ 1)), 0ull);
            } else if (MODE == 6) {  // fire-and-forget atomic (RED)
                atomicMin((unsigned long long*)(tab + idx[u]), 0ull);
            } else if (MODE == 7) {  // load 2 x 32 bytes (an aligned pair), CAS, 8-byte + 16-byte store: the state-table insert
                uint64_t a, b, c, d, e, f, g, h;
                const uint64_t* q = tab + (idx[u] & ~7ull);
                asm volatile("ld.global.cg.v4.b64 {%0, %1, %2, %3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(q));
                asm volatile("ld.global.cg.v4.b64 {%0, %1, %2, %3}, [%4];" : "=l"(e), "=l"(f), "=l"(g), "=l"(h) : "l"(q + 4));
                const uint64_t old = atomicCAS((unsigned long long*)q, a + (b ^ c ^ d ^ e ^ f ^ g ^ h) + 0xFFFFFFFFull, 1ull);
                if (old != 0x12345ull) {
                    __stcg((unsigned long long*)q + 1, 0ull);
                    asm volatile("st.global.cg.v2.b64 [%0], {%1, %2};" ::"l"(q + 2), "l"(0ull), "l"(0ull) : "memory");
                }
            } else if (MODE >= 8 && MODE <= 13) {
                // variants of the insert: aligned 64-byte pair load, then
                //   8: CAS64 only   9: CAS64 + one full 32-byte store   10: one full 32-byte store, no CAS
                //   11: CAS64 + 16-byte store of the upper half   12: CAS128 only   13: CAS128 + 16-byte store of the upper half
                uint64_t a, b, c, d, e, f, g, h;
                uint64_t* q = const_cast<uint64_t*>(tab) + (idx[u] & ~7ull);
                asm volatile("ld.global.cg.v4.b64 {%0, %1, %2, %3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(q));
                asm volatile("ld.global.cg.v4.b64 {%0, %1, %2, %3}, [%4];" : "=l"(e), "=l"(f), "=l"(g), "=l"(h) : "l"(q + 4));
                const uint64_t x = a + (b ^ c ^ d ^ e ^ f ^ g ^ h);  // 0 on the zeroed table
                uint64_t old = 0;
                if (MODE == 8 || MODE == 9 || MODE == 11) old = atomicCAS((unsigned long long*)q, x + 0xFFFFFFFFull, 1ull);
                if (MODE == 12 || MODE == 13) {
                    uint64_t o0, o1;
                    asm volatile("{\n\t.reg .b128 c, s, o;\n\tmov.b128 c, {%2, %3};\n\tmov.b128 s, {%4, %5};\n\t"
                                 "atom.global.cas.b128 o, [%6], c, s;\n\tmov.b128 {%0, %1}, o;\n\t}"
                                 : "=l"(o0), "=l"(o1) : "l"(x + 0xFFFFFFFFull), "l"(x), "l"(1ull), "l"(1ull), "l"(q) : "memory");
                    old = o0 ^ o1;
                }
                if (old != 0x12345ull) {
                    if (MODE == 9 || MODE == 10)
                        asm volatile("st.global.cg.v4.b64 [%0], {%1, %2, %3, %4};" ::"l"(q), "l"(x), "l"(x), "l"(x), "l"(x) : "memory");
                    if (MODE == 11 || MODE == 13) asm volatile("st.global.cg.v2.b64 [%0], {%1, %2};" ::"l"(q + 2), "l"(x), "l"(x) : "memory");
                }
                acc += old;
            }
        }
    }
    if (acc == 0x123456789ull) *sink = acc;
}

template <int MODE, int U>
static void run(const char* name, const uint64_t* tab, uint64_t bytes, uint64_t win, uint64_t* sink, int ctas_per_sm) {
    const int iters = 64 / U * 4;
    const int grid = 148 * ctas_per_sm;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    k_rand<MODE, U><<<grid, 256>>>(tab, bytes / 8, iters, win / 8, sink);
    cudaEventRecord(a);
    const int reps = 4;
    for (int r = 0; r < reps; ++r) k_rand<MODE, U><<<grid, 256>>>(tab, bytes / 8, iters, win / 8, sink);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    const double n = (double)reps * grid * 256.0 * iters * U;
    printf("{\"op\": \"%s\", \"footprint_gb\": %.3f, \"window_mb\": %.1f, \"per_thread\": %d, \"ctas_per_sm\": %d, \"g_access_per_s\": %.2f}\n", name,
           bytes / 1e9, win / 1e6, U, ctas_per_sm, n / (ms * 1e-3) / 1e9);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); exit(1); }
}

int main(int argc, char** argv) {
    if (argc > 1) {  // L2 fetch granularity hint (32 / 64 / 128 bytes)
        cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(argv[1]));
        size_t g = 0;
        cudaDeviceGetLimit(&g, cudaLimitMaxL2FetchGranularity);
        printf("# cudaLimitMaxL2FetchGranularity requested %s -> %zu (%s)\n", argv[1], g, cudaGetErrorString(e));
    }
    uint64_t* sink;
    cudaMalloc(&sink, 8);
    const uint64_t sizes[] = {64ull << 20, 1ull << 30, 8ull << 30, 32ull << 30};
    uint64_t* tab = nullptr;
    if (cudaMalloc(&tab, sizes[3]) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMemset(tab, 0, sizes[3]);
    for (uint64_t bytes : sizes) {
        run<0, 1>("ld8", tab, bytes, 0, sink, 8);
        run<0, 4>("ld8", tab, bytes, 0, sink, 8);
        run<1, 1>("ld32", tab, bytes, 0, sink, 8);
        run<1, 4>("ld32", tab, bytes, 0, sink, 8);
        run<2, 1>("cas64", tab, bytes, 0, sink, 8);
        run<2, 4>("cas64", tab, bytes, 0, sink, 8);
        run<3, 4>("prefetch_l2", tab, bytes, 0, sink, 8);
        run<4, 1>("ld8_then_cas", tab, bytes, 0, sink, 8);
        run<4, 4>("ld8_then_cas", tab, bytes, 0, sink, 8);
        run<5, 1>("ld8_then_st8", tab, bytes, 0, sink, 8);
        run<5, 4>("ld8_then_st8", tab, bytes, 0, sink, 8);
        run<6, 1>("red_min64", tab, bytes, 0, sink, 8);
        run<6, 4>("red_min64", tab, bytes, 0, sink, 8);
        run<7, 1>("ld64_cas_st24", tab, bytes, 0, sink, 8);
        run<7, 4>("ld64_cas_st24", tab, bytes, 0, sink, 8);
        run<8, 1>("ld64_cas64", tab, bytes, 0, sink, 8);
        run<9, 1>("ld64_cas64_st32", tab, bytes, 0, sink, 8);
        run<10, 1>("ld64_st32", tab, bytes, 0, sink, 8);
        run<11, 1>("ld64_cas64_st16", tab, bytes, 0, sink, 8);
        run<12, 1>("ld64_cas128", tab, bytes, 0, sink, 8);
        run<13, 1>("ld64_cas128_st16", tab, bytes, 0, sink, 8);
    }
    // windows over the 32 GB footprint
    const uint64_t wins[] = {32ull << 20, 128ull << 20, 256ull << 20, 1ull << 30};
    for (uint64_t w : wins) {
        run<0, 4>("ld8", tab, sizes[3], w, sink, 8);
        run<2, 4>("cas64", tab, sizes[3], w, sink, 8);
    }
    return 0;
}
