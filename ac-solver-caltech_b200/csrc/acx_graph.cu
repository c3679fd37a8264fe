// acx_graph.cu -- C ABI of the graph engine (acx_graph.cuh): every per-element functor runs as one kernel
// launch over a grid sized in whole waves of the 148 SMs; the two stream compactions of a step (new nodes in
// discovery order, edges in emission order) are single-pass cub scans.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/acx.h"
#include "acx_common.h"
#include "acx_graph.cuh"
#include "acx_ms.cuh"

#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

namespace acx {

template <class F>
__global__ void __launch_bounds__(256) k_graph_apply(int64_t n, const F f) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) f(i);
}

template <class F>
__global__ void k_graph_blocks(const F f) {
    extern __shared__ uint4 g_smem[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(g_smem);
    f.phase1((int64_t)blockIdx.x, (int)threadIdx.x, smem);
    __syncthreads();
    f.phase2((int64_t)blockIdx.x, (int)threadIdx.x, smem);
}

struct FlagToCount {
    __host__ __device__ __forceinline__ int32_t operator()(const uint8_t& f) const { return (int32_t)f; }
};

static int64_t scan_bytes(int64_t n) {
    cub::TransformInputIterator<int32_t, FlagToCount, const uint8_t*> in(nullptr, FlagToCount());
    size_t need = 0;
    if (cub::DeviceScan::ExclusiveSum(nullptr, need, in, (int32_t*)nullptr, (int)n) != cudaSuccess) return -1;
    return (int64_t)need;
}

struct GpuExec {
    cudaStream_t stream;
    int sms;
    cudaError_t err = cudaSuccess;
    template <class F>
    void run(int64_t n, const F& f) {
        if (n <= 0) return;
        int64_t grid = (n + 255) / 256;
        const int64_t wave = (int64_t)sms * 8;  // 8 CTAs of 256 threads per SM
        if (grid > wave) grid = wave * ((grid + wave - 1) / wave > 4 ? 4 : (grid + wave - 1) / wave);
        k_graph_apply<F><<<(unsigned)grid, 256, 0, stream>>>(n, f);
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess && err == cudaSuccess) err = e;
    }
    template <class F>
    void run_shuffled(int64_t n, const F& f) { run(n, f); }
    template <class F>
    void run_blocks(int64_t blocks, int threads, size_t smem, const F& f) {
        if (blocks <= 0) return;
        cudaError_t e = cudaSuccess;
        if (smem > 48 * 1024) e = cudaFuncSetAttribute(k_graph_blocks<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) {
            k_graph_blocks<F><<<(unsigned)blocks, threads, smem, stream>>>(f);
            e = cudaGetLastError();
        }
        if (e != cudaSuccess && err == cudaSuccess) err = e;
    }
    int scan(const uint8_t* flag, int32_t* excl, int64_t n, void* tmp, int64_t bytes) {
        cub::TransformInputIterator<int32_t, FlagToCount, const uint8_t*> in(flag, FlagToCount());
        size_t tb = (size_t)bytes;
        const cudaError_t e = cub::DeviceScan::ExclusiveSum(tmp, tb, in, excl, (int)n, stream);
        if (e != cudaSuccess) return acx_fail(ACX_ERR_CUDA, "acx_graph_step: cub::DeviceScan: %s", cudaGetErrorString(e));
        return 0;
    }
};

static int make_exec(GpuExec& ex, acx_stream_t stream) {
    int dev = 0, sms = 148;
    ACX_CUDA(cudaGetDevice(&dev));
    ACX_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    ex.stream = (cudaStream_t)stream;
    ex.sms = sms;
    return ACX_OK;
}

static bool width_ok(int w) { return w == 2 || w == 4 || w == 8; }

#define ACX_GRAPH_DISPATCH(w, CALL)      \
    switch (w) {                         \
        case 2: { constexpr int W = 2; CALL; } break; \
        case 4: { constexpr int W = 4; CALL; } break; \
        default: { constexpr int W = 8; CALL; } break; \
    }

}  // namespace acx

using namespace acx;

extern "C" int64_t acx_graph_scan_bytes(int64_t n_candidates) {
    if (n_candidates < 0 || n_candidates > 0x7FFFFFFFll) return acx_fail(ACX_ERR_ARG, "acx_graph_scan_bytes: need 0 <= n <= 2^31 - 1");
    const int64_t b = scan_bytes(n_candidates);
    if (b < 0) return acx_fail(ACX_ERR_CUDA, "acx_graph_scan_bytes: cub size query failed");
    return b > 0 ? b : 1;
}

extern "C" int acx_graph_step(const acx_graph_step_t* step, acx_stream_t stream) {
    if (!step) return acx_fail(ACX_ERR_ARG, "acx_graph_step: null step");
    const acx_graph_step_t& a = *step;
    if (!width_ok(a.width_words)) return acx_fail(ACX_ERR_ARG, "acx_graph_step: width_words must be 2, 4 or 8");
    const int nm = a.classic ? 14 : 12;
    if (a.n_parents < 0 || a.parent_first < 0 || a.n_nodes < 0 || a.parent_first + a.n_parents > a.n_nodes)
        return acx_fail(ACX_ERR_ARG, "acx_graph_step: parents must be committed nodes");
    const int64_t C = a.n_parents * nm;
    if (C > 0x7FFFFFFFll) return acx_fail(ACX_ERR_ARG, "acx_graph_step: at most 2^31 - 1 candidates per step");
    if (a.tcap < 2 || (a.tcap & (a.tcap - 1)) || a.tcap < 2 * (a.n_nodes + C))
        return acx_fail(ACX_ERR_ARG, "acx_graph_step: table capacity must be a power of two >= 2 * (n_nodes + candidates)");
    if (a.n_nodes + C > (int64_t)G_VAL) return acx_fail(ACX_ERR_ARG, "acx_graph_step: node ids are 47 bits");
    if (C == 0) return ACX_OK;
    if (!a.nodes || !a.table || !a.cand || !a.ref || !a.csize || !a.flag || !a.excl || !a.scan_tmp || !a.node_size || !a.counts)
        return acx_fail(ACX_ERR_ARG, "acx_graph_step: null pointer");
    if (a.want_edges && (!a.edges || !a.edge_filt)) return acx_fail(ACX_ERR_ARG, "acx_graph_step: want_edges needs edges / edge_filt");
    if (((uintptr_t)a.nodes | (uintptr_t)a.cand) & 31u) return acx_fail(ACX_ERR_ARG, "acx_graph_step: nodes / cand must be 32-byte aligned");
    if (a.n_parents > 0x7FFFFFFFll * 32) return acx_fail(ACX_ERR_ARG, "acx_graph_step: too many parents for one launch");
    if (a.scan_tmp_bytes < scan_bytes(C)) return acx_fail(ACX_ERR_ARG, "acx_graph_step: scan scratch too small");
    GpuExec ex;
    if (int rc = make_exec(ex, stream)) return rc;
    int rc = 0;
    ACX_GRAPH_DISPATCH(a.width_words, rc = (graph_step<W, GpuExec>(ex, a)));
    if (rc) return rc;
    ACX_CUDA(ex.err);
    return ACX_OK;
}

extern "C" int acx_graph_pack(const int8_t* rows, int64_t n, int m, const int32_t* inst, int width_words, uint32_t* nodes_out,
                              uint8_t* ok_out, acx_stream_t stream) {
    if (n < 0 || m < 1 || !width_ok(width_words)) return acx_fail(ACX_ERR_ARG, "acx_graph_pack: need n >= 0, m >= 1, width_words in {2,4,8}");
    if (n == 0) return ACX_OK;
    if (!rows || !nodes_out || !ok_out || ((uintptr_t)nodes_out & 31u)) return acx_fail(ACX_ERR_ARG, "acx_graph_pack: null or unaligned pointer");
    GpuExec ex;
    if (int rc = make_exec(ex, stream)) return rc;
    ACX_GRAPH_DISPATCH(width_words, ex.run(n, GPackRows<W>{rows, m, inst, nodes_out, ok_out}));
    ACX_CUDA(ex.err);
    return ACX_OK;
}

extern "C" int acx_graph_unpack(const uint32_t* nodes, int64_t n, int width_words, int m, int8_t* rows_out, int32_t* inst_out,
                                acx_stream_t stream) {
    if (n < 0 || m < 1 || !width_ok(width_words)) return acx_fail(ACX_ERR_ARG, "acx_graph_unpack: need n >= 0, m >= 1, width_words in {2,4,8}");
    if (n == 0) return ACX_OK;
    if (!nodes || !rows_out || ((uintptr_t)nodes & 31u)) return acx_fail(ACX_ERR_ARG, "acx_graph_unpack: null or unaligned pointer");
    GpuExec ex;
    if (int rc = make_exec(ex, stream)) return rc;
    ACX_GRAPH_DISPATCH(width_words, ex.run(n, GUnpackRows<W>{nodes, m, rows_out, inst_out}));
    ACX_CUDA(ex.err);
    return ACX_OK;
}

extern "C" int acx_graph_rehash(const uint32_t* nodes, int64_t n_nodes, int width_words, uint64_t* table, int64_t tcap,
                                acx_stream_t stream) {
    if (n_nodes < 0 || !width_ok(width_words) || tcap < 2 || (tcap & (tcap - 1)) || tcap < 2 * n_nodes)
        return acx_fail(ACX_ERR_ARG, "acx_graph_rehash: need a power-of-two table of at least 2 * n_nodes slots");
    if (n_nodes == 0) return ACX_OK;
    if (!nodes || !table || ((uintptr_t)nodes & 31u)) return acx_fail(ACX_ERR_ARG, "acx_graph_rehash: null or unaligned pointer");
    GpuExec ex;
    if (int rc = make_exec(ex, stream)) return rc;
    ACX_GRAPH_DISPATCH(width_words, ex.run(n_nodes, GRehash<W>{nodes, table, tcap}));
    ACX_CUDA(ex.err);
    return ACX_OK;
}

extern "C" int acx_graph_count_instances(const uint32_t* nodes, int64_t n_nodes, int width_words, int64_t* counts, int64_t n_inst,
                                         acx_stream_t stream) {
    if (n_nodes < 0 || n_inst < 1 || !width_ok(width_words)) return acx_fail(ACX_ERR_ARG, "acx_graph_count_instances: bad argument");
    if (n_nodes == 0) return ACX_OK;
    if (!nodes || !counts || ((uintptr_t)nodes & 31u)) return acx_fail(ACX_ERR_ARG, "acx_graph_count_instances: null or unaligned pointer");
    GpuExec ex;
    if (int rc = make_exec(ex, stream)) return rc;
    ACX_GRAPH_DISPATCH(width_words, ex.run(n_nodes, GCountInst<W>{nodes, counts, n_inst}));
    ACX_CUDA(ex.err);
    return ACX_OK;
}

extern "C" int acx_ms_enumerate(int w_len, int64_t first, int64_t count, uint64_t* key_out, uint64_t* rel_out, acx_stream_t stream) {
    if (w_len < 1 || w_len > 15 || first < 0 || count < 0 || first + count > (1ll << (2 * w_len)))
        return acx_fail(ACX_ERR_ARG, "acx_ms_enumerate: need 1 <= w_len <= 15 and [first, first + count) inside [0, 4^w_len)");
    if (count == 0) return ACX_OK;
    if (!key_out || !rel_out) return acx_fail(ACX_ERR_ARG, "acx_ms_enumerate: null pointer");
    GpuExec ex;
    if (int rc = make_exec(ex, stream)) return rc;
    ex.run(count, MsEnumerate{w_len, first, key_out, rel_out});
    ACX_CUDA(ex.err);
    return ACX_OK;
}
