// copy_ceiling.cu -- what does pure data movement cost for one ACEnv.step launch (2^20 envs, m=36)?
//
// Standalone microbenchmark (no torch): the memory behaviour of the env-step kernel WITHOUT the row
// arithmetic, to separate "HBM / launch-structure ceiling" from "instruction issue" in the roofline
// discussion.  Variants:
//   ring_ldg   persistent TMA ring over 72-B rows (in place), aux arrays through per-thread LDG/STG
//              (= round-1 kernel structure: actions 1 B, steps 4 B r/w, step_out 8 B w + 2 B r, reward 4 B, status 1 B)
//   ring_tma   same ring, aux arrays (action 1 B r, meta 4 B r/w, record 8 B w) through the bulk copies too
//   plain      grid-wide 16-byte LDG/STG copy of the same bytes (no shared memory)
// each with / without programmatic dependent launch, K launches in one CUDA graph over 4 rotating batches.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o copy_ceiling copy_ceiling.cu && ./copy_ceiling
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "../ac-solver-caltech_b200/csrc/acx_tma.cuh"

using namespace acx;

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e = (x);                                                          \
        if (e != cudaSuccess) {                                                       \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
            exit(1);                                                                  \
        }                                                                             \
    } while (0)

struct Bufs {
    uint8_t* state;
    uint8_t* actions;
    int32_t* steps;
    uint8_t* rec;
    float* rew;
    uint8_t* status;
    uint32_t* meta;
};

constexpr int ROWB = 72;

template <int TR, int S, bool WAIT>
__global__ void __launch_bounds__(TR) k_ring_ldg(Bufs b, int64_t ntiles) {
    constexpr uint32_t TILEB = TR * ROWB;
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)S * TILEB);
    const int tid = threadIdx.x;
    const uint32_t smem0 = smem_u32(smem), bar0 = smem_u32(bars);
    if (tid == 0) {
        for (int s = 0; s < S; ++s) mbar_init(bar0 + 8 * s, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (WAIT) asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int64_t first = blockIdx.x, stride = gridDim.x;
    if (tid == 0) {
        for (int s = 0; s < S; ++s) {
            const int64_t t = first + (int64_t)s * stride;
            if (t < ntiles) {
                mbar_expect_tx(bar0 + 8 * s, TILEB);
                bulk_load(smem0 + s * TILEB, b.state + t * TILEB, TILEB, bar0 + 8 * s);
            }
        }
    }
    uint32_t s = 0, parity = 0, sprev = S - 1;
    bool first_it = true;
    for (int64_t t = first; t < ntiles; t += stride) {
        const int64_t row = t * TR + tid;
        const int move = b.actions[row];
        const int32_t st = b.steps[row];
        const uchar2 ln = *reinterpret_cast<const uchar2*>(b.rec + 8 * row + 6);
        mbar_wait(bar0 + 8 * s, parity);
        uint32_t* srow = reinterpret_cast<uint32_t*>(smem + s * TILEB + tid * ROWB);
        srow[move & 7] ^= 1u;  // touch the row
        b.rew[row] = (float)move;
        b.steps[row] = st + 1;
        reinterpret_cast<uint2*>(b.rec)[row] = make_uint2((uint32_t)move, ((uint32_t)ln.x << 16) | ((uint32_t)ln.y << 24));
        b.status[row] = 0;
        fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            bulk_store(b.state + t * TILEB, smem0 + s * TILEB, TILEB);
            bulk_commit();
            if (!first_it) {
                bulk_wait_read<1>();
                const int64_t tn = t + (int64_t)(S - 1) * stride;
                if (tn < ntiles) {
                    mbar_expect_tx(bar0 + 8 * sprev, TILEB);
                    bulk_load(smem0 + sprev * TILEB, b.state + tn * TILEB, TILEB, bar0 + 8 * sprev);
                }
            }
        }
        first_it = false;
        sprev = s;
        if (++s == S) {
            s = 0;
            parity ^= 1u;
        }
    }
    if (tid == 0) bulk_wait_all<0>();
}

// aux arrays through the ring as well: stage = [state TR*72 | meta TR*4 | actions TR | rec TR*8]
template <int TR, int S, bool WAIT>
__global__ void __launch_bounds__(TR) k_ring_tma(Bufs b, int64_t ntiles) {
    constexpr uint32_t TILEB = TR * ROWB, METAB = TR * 4, ACTB = TR, RECB = TR * 8;
    constexpr uint32_t STAGEB = TILEB + METAB + ACTB + RECB;
    static_assert(STAGEB % 16 == 0, "");
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)S * STAGEB);
    const int tid = threadIdx.x;
    const uint32_t smem0 = smem_u32(smem), bar0 = smem_u32(bars);
    if (tid == 0) {
        for (int s = 0; s < S; ++s) mbar_init(bar0 + 8 * s, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (WAIT) asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int64_t first = blockIdx.x, stride = gridDim.x;
    auto issue = [&](int st, int64_t t) {
        const uint32_t base = smem0 + st * STAGEB, bar = bar0 + 8 * st;
        mbar_expect_tx(bar, TILEB + METAB + ACTB);
        bulk_load(base, b.state + t * TILEB, TILEB, bar);
        bulk_load(base + TILEB, reinterpret_cast<uint8_t*>(b.meta) + t * METAB, METAB, bar);
        bulk_load(base + TILEB + METAB, b.actions + t * ACTB, ACTB, bar);
    };
    if (tid == 0) {
        for (int s = 0; s < S; ++s) {
            const int64_t t = first + (int64_t)s * stride;
            if (t < ntiles) issue(s, t);
        }
    }
    uint32_t s = 0, parity = 0, sprev = S - 1;
    bool first_it = true;
    for (int64_t t = first; t < ntiles; t += stride) {
        mbar_wait(bar0 + 8 * s, parity);
        uint8_t* stage = smem + s * STAGEB;
        uint32_t* srow = reinterpret_cast<uint32_t*>(stage + tid * ROWB);
        uint32_t* meta = reinterpret_cast<uint32_t*>(stage + TILEB);
        const int move = stage[TILEB + METAB + tid];
        srow[move & 7] ^= 1u;
        const uint32_t mt = meta[tid];
        meta[tid] = mt + 1;
        reinterpret_cast<uint2*>(stage + TILEB + METAB + ACTB)[tid] = make_uint2((uint32_t)move, mt);
        fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            const uint32_t base = smem0 + s * STAGEB;
            bulk_store(b.state + t * TILEB, base, TILEB);
            bulk_store(reinterpret_cast<uint8_t*>(b.meta) + t * METAB, base + TILEB, METAB);
            bulk_store(b.rec + t * RECB, base + TILEB + METAB + ACTB, RECB);
            bulk_commit();
            if (!first_it) {
                bulk_wait_read<1>();
                const int64_t tn = t + (int64_t)(S - 1) * stride;
                if (tn < ntiles) issue(sprev, tn);
            }
        }
        first_it = false;
        sprev = s;
        if (++s == S) {
            s = 0;
            parity ^= 1u;
        }
    }
    if (tid == 0) bulk_wait_all<0>();
}

template <bool WAIT>
__global__ void __launch_bounds__(256) k_plain(Bufs b, int64_t n) {
    if (WAIT) asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int64_t nv = n * ROWB / 16;
    uint4* p = reinterpret_cast<uint4*>(b.state);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = tid; i < nv; i += nt) {
        uint4 v = p[i];
        v.x ^= 1u;
        p[i] = v;
    }
    for (int64_t i = tid; i < n; i += nt) {
        const uint32_t mt = b.meta[i];
        b.meta[i] = mt + 1;
        reinterpret_cast<uint2*>(b.rec)[i] = make_uint2(mt, b.actions[i]);
    }
}

template <typename F>
static float time_graph(F launch, int K, cudaStream_t st) {
    cudaGraph_t g;
    cudaGraphExec_t ge;
    CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal));
    for (int k = 0; k < K; ++k) launch(k);
    CK(cudaStreamEndCapture(st, &g));
    CK(cudaGraphInstantiate(&ge, g, 0));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int w = 0; w < 2; ++w) CK(cudaGraphLaunch(ge, st));
    CK(cudaStreamSynchronize(st));
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CK(cudaEventRecord(e0, st));
        CK(cudaGraphLaunch(ge, st));
        CK(cudaEventRecord(e1, st));
        CK(cudaStreamSynchronize(st));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    CK(cudaGraphExecDestroy(ge));
    CK(cudaGraphDestroy(g));
    return best * 1e3f / K;  // us per launch
}

template <typename Kern, typename... Args>
static void launch_pdl(Kern kern, dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool pdl, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    CK(cudaLaunchKernelEx(&cfg, kern, args...));
}

template <int TR, int S, bool TMA_AUX, bool WAIT>
static void run_ring(const char* name, std::vector<Bufs>& bufs, int64_t n, int K, cudaStream_t st, bool pdl, int ctas_cap) {
    constexpr size_t smem_ldg = (size_t)S * TR * ROWB + 8 * S;
    constexpr size_t smem_tma = (size_t)S * (TR * ROWB + TR * 13) + 8 * S;
    const size_t smem = TMA_AUX ? smem_tma : smem_ldg;
    auto kern = TMA_AUX ? k_ring_tma<TR, S, WAIT> : k_ring_ldg<TR, S, WAIT>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, TR, smem));
    if (ctas_cap > 0 && occ > ctas_cap) occ = ctas_cap;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int64_t ntiles = n / TR;
    const int grid = (int)std::min<int64_t>(ntiles, (int64_t)sms * occ);
    const float us = time_graph([&](int k) { launch_pdl(kern, dim3(grid), dim3(TR), smem, st, pdl, bufs[k % bufs.size()], ntiles); }, K, st);
    const double alg = 153.0 * n;
    printf("%-10s TR=%3d S=%d ctas/SM=%2d grid=%5d pdl=%d wait=%d : %7.2f us/launch  alg %.0f GB/s (frac of 6538.9: %.3f)\n", name, TR, S, occ,
           grid, (int)pdl, (int)WAIT, us, alg / us * 1e-3, alg / us * 1e-3 / 6538.9);
}

int main(int argc, char** argv) {
    const int64_t n = 1 << 20;
    const int K = argc > 1 ? atoi(argv[1]) : 20;
    const int NB = 4;
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    std::vector<Bufs> bufs(NB);
    for (auto& b : bufs) {
        CK(cudaMalloc(&b.state, n * ROWB));
        CK(cudaMalloc(&b.actions, n));
        CK(cudaMalloc(&b.steps, n * 4));
        CK(cudaMalloc(&b.rec, n * 8));
        CK(cudaMalloc(&b.rew, n * 4));
        CK(cudaMalloc(&b.status, n));
        CK(cudaMalloc(&b.meta, n * 4));
        CK(cudaMemset(b.state, 1, n * ROWB));
        CK(cudaMemset(b.actions, 3, n));
        CK(cudaMemset(b.steps, 0, n * 4));
        CK(cudaMemset(b.rec, 0, n * 8));
        CK(cudaMemset(b.meta, 0, n * 4));
    }
    printf("one launch = 2^20 rows of 72 B in place + aux arrays; K=%d launches per graph, %d rotating batches\n", K, NB);
    // memcpy reference: D2D copy of the state bytes (read + write = 151 MB)
    {
        uint8_t* tmp;
        CK(cudaMalloc(&tmp, n * ROWB));
        const float us = time_graph([&](int k) { CK(cudaMemcpyAsync(tmp, bufs[k % NB].state, n * ROWB, cudaMemcpyDeviceToDevice, st)); }, K, st);
        printf("cudaMemcpyAsync D2D of 75.5 MB: %7.2f us  (%.0f GB/s read+write)\n", us, 2.0 * n * ROWB / us * 1e-3);
        CK(cudaFree(tmp));
    }
    for (int pdl = 0; pdl < 2; ++pdl) {
        run_ring<128, 3, false, true>("ring_ldg", bufs, n, K, st, pdl, 0);
        run_ring<128, 3, true, true>("ring_tma", bufs, n, K, st, pdl, 0);
        run_ring<128, 4, true, true>("ring_tma", bufs, n, K, st, pdl, 0);
        run_ring<128, 2, true, true>("ring_tma", bufs, n, K, st, pdl, 0);
        run_ring<64, 4, true, true>("ring_tma", bufs, n, K, st, pdl, 0);
        run_ring<64, 3, true, true>("ring_tma", bufs, n, K, st, pdl, 0);
        run_ring<256, 3, true, true>("ring_tma", bufs, n, K, st, pdl, 0);
        run_ring<256, 2, true, true>("ring_tma", bufs, n, K, st, pdl, 0);
    }
    // upper bound for a per-CTA flag dependency instead of griddepcontrol.wait
    run_ring<128, 3, true, false>("ring_tma", bufs, n, K, st, true, 0);
    run_ring<128, 3, false, false>("ring_ldg", bufs, n, K, st, true, 0);
    run_ring<64, 4, true, false>("ring_tma", bufs, n, K, st, true, 0);
    run_ring<256, 2, true, false>("ring_tma", bufs, n, K, st, true, 0);
    for (int cap : {1, 2, 4, 6}) run_ring<128, 3, true, true>("ring_tma", bufs, n, K, st, true, cap);
    for (int cap : {1, 2, 4}) run_ring<128, 3, true, false>("ring_tma", bufs, n, K, st, true, cap);
    {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
        for (int pdl = 0; pdl < 2; ++pdl)
            for (int mult : {4, 8, 16}) {
                const float us = time_graph([&](int k) { launch_pdl(k_plain<true>, dim3(sms * mult), dim3(256), 0, st, pdl, bufs[k % NB], n); }, K, st);
                printf("plain      grid=%d*%d pdl=%d : %7.2f us/launch  alg %.0f GB/s (frac %.3f)\n", sms, mult, pdl, us, 153.0 * n / us * 1e-3,
                       153.0 * n / us * 1e-3 / 6538.9);
            }
    }
    return 0;
}
