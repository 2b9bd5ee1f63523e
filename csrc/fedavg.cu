// FedAvg aggregation kernels and the peer-memory runtime they need.
//
// The reference's server averages the clients' state_dicts key by key in Python after shipping them through
// pickle + gzip + TCP (server.py:67-79 aggregate_models; transport server.py:29-65, client1.py:228-336).  Here every
// client keeps its parameters in one contiguous arena in peer-mappable GPU memory and the aggregation is ONE kernel
// per GPU that reads the peers' arenas over NVLink/NVSwitch, accumulates in fp32 in client order (bit parity with the
// reference's `+=` loop), divides by N, and stores the result -- plus the bf16 compute copy of the GEMM weights --
// straight into every client's buffers.  No NCCL on this path.
//
//   two-shot (default): rank r reduces shard r from all N arenas and pushes the averaged shard to all N arenas
//                       (in place; ingress (N-1)/N*S + egress (N-1)/N*S per GPU, both link directions busy)
//   one-shot          : every rank reduces the whole arena from all peers into its own separate output buffer
//   local             : all sources reachable from one GPU, no cross-GPU handshake (server-side socket fallback,
//                       single-process tests)
//
// Cross-GPU handshake: per-rank signal pad in peer memory, flags written with st.release.sys and polled with
// ld.acquire.sys; bounded spins set an error flag instead of hanging, and a rank whose start handshake timed out
// leaves every arena untouched.  The host reads the flag after the round and clears it (dcad users: pad word 33).
#include <string.h>

#include "common.cuh"
#include "../include/dcad_b200.h"

namespace dcad {

constexpr int FED_MAX = DCAD_FED_MAX_CLIENTS;
constexpr int PAD_READY = 0;     // [0,16)  : ready flags, slot k written by rank k
constexpr int PAD_DONE = 16;     // [16,32) : done flags
constexpr int PAD_COUNTER = 32;  // local block-completion counter
constexpr int PAD_ERROR = 33;    // set to 1 when a bounded spin expired

struct FedArgs {
    void* arena[FED_MAX];
    void* out[FED_MAX];
    void* mirror[FED_MAX];
    uint32_t* pad[FED_MAX];
    float weight[FED_MAX];
    int use_weights;
    void* dst[FED_MAX];  // two-shot destinations (defaults to `arena`); may be a superset of the sources
    int n_dst;           // number of ranks / destinations / pads
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ bool spin_until_ge(const uint32_t* p, uint32_t target) {
    for (uint32_t i = 0; i < (1u << 24); ++i) {
        if (static_cast<int32_t>(ld_acquire_sys(p) - target) >= 0) return true;
        __nanosleep(64);
    }
    return false;
}

// Start handshake of a round (all threads of every block call it).  Block 0 publishes this rank's ready flag to every
// peer; every block waits until every peer's flag has reached `epoch`.  Returns false -- for the whole block, after
// raising the error word -- when a peer did not show up within the bounded spin.
__device__ __forceinline__ bool start_barrier(const FedArgs& a, uint32_t* my_pad, int rank, int nr, uint32_t epoch) {
    __shared__ int s_timed_out;
    if (threadIdx.x == 0) s_timed_out = 0;
    __syncthreads();
    if (blockIdx.x == 0 && threadIdx.x < nr) st_release_sys(a.pad[threadIdx.x] + PAD_READY + rank, epoch);
    if (threadIdx.x < nr) {
        if (!spin_until_ge(my_pad + PAD_READY + threadIdx.x, epoch)) {
            atomicExch(my_pad + PAD_ERROR, 1u);
            s_timed_out = 1;
        }
    }
    __syncthreads();
    return s_timed_out == 0;
}

// End handshake: the last block of this rank to finish tells every peer that this rank's stores are visible, then
// waits for theirs.  Every block takes part in the block counter (also one that skipped its work after a time-out, so
// the counter is back at zero for the next round), but a rank whose error word is set does not signal completion:
// its peers then time out too and every client of the round reports the failure instead of trusting a partial result.
__device__ __forceinline__ void end_barrier(const FedArgs& a, uint32_t* my_pad, int rank, int nr, uint32_t epoch) {
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t prev = atomicAdd(my_pad + PAD_COUNTER, 1u);
        if (prev == gridDim.x - 1) {
            my_pad[PAD_COUNTER] = 0;
            __threadfence_system();
            if (atomicAdd(my_pad + PAD_ERROR, 0u) == 0u) {
                for (int k = 0; k < nr; ++k) st_release_sys(a.pad[k] + PAD_DONE + rank, epoch);
                for (int k = 0; k < nr; ++k)
                    if (!spin_until_ge(my_pad + PAD_DONE + k, epoch)) atomicExch(my_pad + PAD_ERROR, 1u);
            }
        }
    }
}

template <bool IN_BF16>
struct VecIO;
template <>
struct VecIO<false> {  // fp32 arena, 4 elements per 16-byte access
    static constexpr int ELEMS = 4;
    __device__ static void load(const void* base, int64_t vec, float (&v)[4]) {
        const float4 t = *(reinterpret_cast<const float4*>(base) + vec);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    __device__ static void store(void* base, int64_t vec, const float (&v)[4]) {
        *(reinterpret_cast<float4*>(base) + vec) = make_float4(v[0], v[1], v[2], v[3]);
    }
};
template <>
struct VecIO<true> {  // bf16 arena, 8 elements per 16-byte access
    static constexpr int ELEMS = 8;
    __device__ static void load(const void* base, int64_t vec, float (&v)[8]) {
        const uint4 t = *(reinterpret_cast<const uint4*>(base) + vec);
        const float2 a = unpack_bf16x2(t.x), b = unpack_bf16x2(t.y), c = unpack_bf16x2(t.z), d = unpack_bf16x2(t.w);
        v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
    }
    __device__ static void store(void* base, int64_t vec, const float (&v)[8]) {
        uint4 o;
        o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
        o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
        *(reinterpret_cast<uint4*>(base) + vec) = o;
    }
};

// Average U vector positions (16 bytes each) across n clients, in client order.  All U*n loads are issued before
// any arithmetic so that many peer reads per thread are in flight: NVLink round trips are ~2-3 us, and
// 770 GB/s x 3 us = 2.3 MB must be outstanding per GPU to keep the links busy.
template <bool IN_BF16, int U>
__device__ __forceinline__ void reduce_vecs(const FedArgs& a, int n, const int64_t (&vec)[U], const bool (&ok)[U],
                                            float (&acc)[U][VecIO<IN_BF16>::ELEMS]) {
    constexpr int E = VecIO<IN_BF16>::ELEMS;
    constexpr int KB = (IN_BF16 ? 8 : 16) / U;  // clients per batch: at most 64 staged floats per thread
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
        for (int j = 0; j < E; ++j) acc[u][j] = 0.f;
    for (int k0 = 0; k0 < n; k0 += KB) {
        float t[U][KB][E];
#pragma unroll
        for (int kk = 0; kk < KB; ++kk)
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (k0 + kk < n && ok[u]) VecIO<IN_BF16>::load(a.arena[k0 + kk], vec[u], t[u][kk]);
#pragma unroll
        for (int kk = 0; kk < KB; ++kk) {
            if (k0 + kk < n) {
                const float w = a.use_weights ? a.weight[k0 + kk] : 1.0f;
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    if (!ok[u]) continue;
                    if (a.use_weights) {
#pragma unroll
                        for (int j = 0; j < E; ++j)
                            acc[u][j] = (k0 + kk == 0) ? w * t[u][kk][j] : fmaf(w, t[u][kk][j], acc[u][j]);
                    } else {
#pragma unroll
                        for (int j = 0; j < E; ++j)
                            acc[u][j] = (k0 + kk == 0) ? t[u][kk][j] : acc[u][j] + t[u][kk][j];
                    }
                }
            }
        }
    }
    if (!a.use_weights) {
        const float nf = static_cast<float>(n);
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int j = 0; j < E; ++j) acc[u][j] = __fdiv_rn(acc[u][j], nf);  // reference: base[key] /= len(models)
    }
}

template <bool IN_BF16>
__device__ __forceinline__ void store_mirror(void* mirror, int64_t elem0, int64_t bf16_begin, int64_t bf16_end,
                                             const float (&v)[VecIO<IN_BF16>::ELEMS]) {
    if constexpr (!IN_BF16) {
        if (mirror && elem0 >= bf16_begin && elem0 < bf16_end) {
            uint2 o;
            o.x = pack_bf16x2(v[0], v[1]);
            o.y = pack_bf16x2(v[2], v[3]);
            *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(mirror) + (elem0 - bf16_begin)) = o;
        }
    }
}

constexpr int FED_U = 4;  // vector positions per thread per iteration

// ---- local: sources readable from this GPU, single destination ----
template <bool IN_BF16>
__global__ void __launch_bounds__(512)
fedavg_local_kernel(const FedArgs a, int n, int64_t num_vecs, int64_t bf16_begin, int64_t bf16_end) {
    constexpr int E = VecIO<IN_BF16>::ELEMS;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t v0 = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; v0 < num_vecs; v0 += stride * FED_U) {
        int64_t vec[FED_U];
        bool ok[FED_U];
#pragma unroll
        for (int u = 0; u < FED_U; ++u) { vec[u] = v0 + u * stride; ok[u] = vec[u] < num_vecs; }
        float acc[FED_U][E];
        reduce_vecs<IN_BF16, FED_U>(a, n, vec, ok, acc);
#pragma unroll
        for (int u = 0; u < FED_U; ++u) {
            if (!ok[u]) continue;
            VecIO<IN_BF16>::store(a.out[0], vec[u], acc[u]);
            store_mirror<IN_BF16>(a.mirror[0], vec[u] * E, bf16_begin, bf16_end, acc[u]);
        }
    }
}

// ---- peer-to-peer: one launch per rank, kernels on different GPUs handshake through the signal pads ----
template <bool IN_BF16>
__global__ void __launch_bounds__(512)
fedavg_p2p_kernel(const FedArgs a, int rank, int n, int64_t num_vecs, int mode, int64_t bf16_begin, int64_t bf16_end,
                  uint32_t epoch) {
    // n = number of SOURCE arenas that are averaged; a.n_dst = number of ranks (kernels) taking part: they share the
    // shards, the handshake and -- two-shot -- all receive the result.  Normally the two coincide; a data-parallel
    // client group contributes one source but several ranks.
    constexpr int E = VecIO<IN_BF16>::ELEMS;
    const int nr = a.n_dst;
    uint32_t* my_pad = a.pad[rank];
    // start barrier: every peer's arena is final (their launch is stream-ordered after their local training).  A peer
    // that never arrives (dead client, hung process) must not cost the others their models: the block that times out
    // raises the error word and NO block of this rank that saw the time-out reads or writes a single parameter.
    if (!start_barrier(a, my_pad, rank, nr, epoch)) num_vecs = 0;

    int64_t lo = 0, hi = num_vecs;
    if (mode == DCAD_FED_TWO_SHOT) {
        lo = num_vecs * rank / nr;
        hi = num_vecs * (rank + 1) / nr;
    }
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t v0 = lo + static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; v0 < hi; v0 += stride * FED_U) {
        int64_t vec[FED_U];
        bool ok[FED_U];
#pragma unroll
        for (int u = 0; u < FED_U; ++u) { vec[u] = v0 + u * stride; ok[u] = vec[u] < hi; }
        float acc[FED_U][E];
        reduce_vecs<IN_BF16, FED_U>(a, n, vec, ok, acc);
#pragma unroll
        for (int u = 0; u < FED_U; ++u) {
            if (!ok[u]) continue;
            if (mode == DCAD_FED_TWO_SHOT) {
                for (int k = 0; k < nr; ++k) {
                    VecIO<IN_BF16>::store(a.dst[k], vec[u], acc[u]);
                    store_mirror<IN_BF16>(a.mirror[k], vec[u] * E, bf16_begin, bf16_end, acc[u]);
                }
            } else {
                VecIO<IN_BF16>::store(a.out[rank], vec[u], acc[u]);
                store_mirror<IN_BF16>(a.mirror[rank], vec[u] * E, bf16_begin, bf16_end, acc[u]);
            }
        }
    }
    // end barrier: the last block of this rank tells every peer its pushes are visible, then waits for theirs
    __threadfence_system();
    __syncthreads();
    end_barrier(a, my_pad, rank, nr, epoch);
}

// ---- NVLS: in-switch reduction over a multicast mapping of all clients' arenas ----
// multimem.ld_reduce pulls element i already summed over all N GPUs by the NVSwitch (ingress S/N per GPU instead of
// (N-1)/N*S), multimem.st writes the averaged value (and its bf16 copy) into all N arenas with one store.  The
// summation order inside the switch is unspecified, so this path is fp32-tolerance, not bit, equal to the reference
// loop; it cannot apply per-client weights.
__device__ __forceinline__ float4 multimem_ld_reduce_add_f32x4(const void* mc_ptr) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc_ptr) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st_f32x4(void* mc_ptr, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 ::"l"(mc_ptr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_b32x2(void* mc_ptr, uint32_t a, uint32_t b) {
    asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};"
                 ::"l"(mc_ptr), "f"(__uint_as_float(a)), "f"(__uint_as_float(b)) : "memory");
}

__global__ void __launch_bounds__(512)
fedavg_nvls_kernel(const FedArgs a, float* mc_arena, __nv_bfloat16* mc_mirror, int rank, int n, int64_t num_vecs,
                   int64_t bf16_begin, int64_t bf16_end, uint32_t epoch) {
    uint32_t* my_pad = a.pad[rank];
    if (!start_barrier(a, my_pad, rank, n, epoch)) num_vecs = 0;  // (see fedavg_p2p_kernel)
    const int64_t lo = num_vecs * rank / n, hi = num_vecs * (rank + 1) / n;
    const float inv_n = 1.0f / static_cast<float>(n);
    const float nf = static_cast<float>(n);
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    constexpr int NU = 4;  // 16-byte switch-reduced loads in flight per thread (8 measured no faster: 0.690 vs 0.659 ms at 8 GPUs)
    for (int64_t v0 = lo + static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; v0 < hi; v0 += stride * NU) {
        float4 acc[NU];
#pragma unroll
        for (int u = 0; u < NU; ++u)
            if (v0 + u * stride < hi) acc[u] = multimem_ld_reduce_add_f32x4(mc_arena + (v0 + u * stride) * 4);
#pragma unroll
        for (int u = 0; u < NU; ++u) {
            const int64_t vec = v0 + u * stride;
            if (vec >= hi) continue;
            float4 r = acc[u];
            r.x = __fdiv_rn(r.x, nf); r.y = __fdiv_rn(r.y, nf); r.z = __fdiv_rn(r.z, nf); r.w = __fdiv_rn(r.w, nf);
            multimem_st_f32x4(mc_arena + vec * 4, r);
            const int64_t e0 = vec * 4;
            if (mc_mirror && e0 >= bf16_begin && e0 < bf16_end)
                multimem_st_b32x2(mc_mirror + (e0 - bf16_begin), pack_bf16x2(r.x, r.y), pack_bf16x2(r.z, r.w));
        }
    }
    (void)inv_n;
    __threadfence_system();
    __syncthreads();
    end_barrier(a, my_pad, rank, n, epoch);
}

static int fill_args(FedArgs& fa, const void* const* arenas, const void* const* outs, const void* const* mirrors,
                     const void* const* pads, const float* weights, int n) {
    if (n < 1 || n > FED_MAX) return DCAD_ERR_TOO_LARGE;
    for (int k = 0; k < FED_MAX; ++k) {
        fa.arena[k] = k < n ? const_cast<void*>(arenas[k]) : nullptr;
        fa.out[k] = (outs && k < n) ? const_cast<void*>(outs[k]) : nullptr;
        fa.mirror[k] = (mirrors && k < n) ? const_cast<void*>(mirrors[k]) : nullptr;
        fa.pad[k] = (pads && k < n) ? reinterpret_cast<uint32_t*>(const_cast<void*>(pads[k])) : nullptr;
        fa.weight[k] = (weights && k < n) ? weights[k] : 0.f;
    }
    fa.use_weights = weights ? 1 : 0;
    for (int k = 0; k < FED_MAX; ++k) fa.dst[k] = fa.arena[k];
    fa.n_dst = n;
    return 0;
}

}  // namespace dcad

using namespace dcad;

extern "C" int dcad_fedavg_local(const void* const* srcs, void* dst, void* dst_bf16, int n_clients, int64_t count,
                                 int in_bf16, const float* weights, int64_t bf16_begin, int64_t bf16_end,
                                 int max_ctas, void* stream) {
    if (count <= 0) return 0;
    const int E = in_bf16 ? 8 : 4;
    if (count % E || (bf16_begin & 3) || (bf16_end & 3)) return DCAD_ERR_BAD_ALIGNMENT;
    FedArgs fa;
    int rc = fill_args(fa, srcs, nullptr, nullptr, nullptr, weights, n_clients);
    if (rc) return rc;
    fa.out[0] = dst;
    fa.mirror[0] = dst_bf16;
    const int64_t num_vecs = count / E;
    int64_t blocks = (num_vecs + 511) / 512;
    const int64_t cap = max_ctas > 0 ? max_ctas : 148 * 2;
    if (blocks > cap) blocks = cap;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (in_bf16) fedavg_local_kernel<true><<<(int)blocks, 512, 0, s>>>(fa, n_clients, num_vecs, bf16_begin, bf16_end);
    else fedavg_local_kernel<false><<<(int)blocks, 512, 0, s>>>(fa, n_clients, num_vecs, bf16_begin, bf16_end);
    return (int)cudaGetLastError();
}

extern "C" int dcad_fedavg_p2p(const void* const* arenas, const void* const* outs, const void* const* mirrors,
                               const void* const* pads, const float* weights, int rank, int n_clients, int64_t count,
                               int in_bf16, int mode, int64_t bf16_begin, int64_t bf16_end, uint32_t epoch,
                               int max_ctas, void* stream) {
    if (count <= 0) return 0;
    const int E = in_bf16 ? 8 : 4;
    if (count % E || (bf16_begin & 3) || (bf16_end & 3)) return DCAD_ERR_BAD_ALIGNMENT;
    if (rank < 0 || rank >= n_clients || pads == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (mode == DCAD_FED_ONE_SHOT && outs == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    FedArgs fa;
    int rc = fill_args(fa, arenas, outs, mirrors, pads, weights, n_clients);
    if (rc) return rc;
    const int64_t num_vecs = count / E;
    const int64_t work = mode == DCAD_FED_TWO_SHOT ? (num_vecs + n_clients - 1) / n_clients : num_vecs;
    int64_t blocks = (work + 511) / 512;
    const int64_t cap = max_ctas > 0 ? max_ctas : 148;  // one co-resident wave: the end barrier needs all blocks to run
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (in_bf16)
        fedavg_p2p_kernel<true><<<(int)blocks, 512, 0, s>>>(fa, rank, n_clients, num_vecs, mode, bf16_begin, bf16_end, epoch);
    else
        fedavg_p2p_kernel<false><<<(int)blocks, 512, 0, s>>>(fa, rank, n_clients, num_vecs, mode, bf16_begin, bf16_end, epoch);
    return (int)cudaGetLastError();
}

// Two-shot FedAvg where the averaged SOURCES (one arena per client) are a subset of the participating RANKS: every
// rank reduces 1/n_ranks of the arena from the n_src sources and pushes it to all n_ranks destinations.  Used when a
// client spans several GPUs (intra-client data parallelism): one representative arena per client is averaged, all
// GPUs of all clients receive the result and its bf16 copy.  All arrays are HOST arrays.
extern "C" int dcad_fedavg_p2p_grouped(const void* const* srcs, int n_src, const void* const* dsts,
                                       const void* const* mirrors, const void* const* pads, int n_ranks,
                                       const float* weights, int rank, int64_t count, int in_bf16, int64_t bf16_begin,
                                       int64_t bf16_end, uint32_t epoch, int max_ctas, void* stream) {
    if (count <= 0) return 0;
    const int E = in_bf16 ? 8 : 4;
    if (count % E || (bf16_begin & 3) || (bf16_end & 3)) return DCAD_ERR_BAD_ALIGNMENT;
    if (n_src < 1 || n_src > FED_MAX || n_ranks < 1 || n_ranks > FED_MAX) return DCAD_ERR_TOO_LARGE;
    if (rank < 0 || rank >= n_ranks || pads == nullptr || dsts == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    FedArgs fa;
    int rc = fill_args(fa, srcs, nullptr, nullptr, nullptr, weights, n_src);
    if (rc) return rc;
    for (int k = 0; k < FED_MAX; ++k) {
        fa.dst[k] = k < n_ranks ? const_cast<void*>(dsts[k]) : nullptr;
        fa.mirror[k] = (mirrors && k < n_ranks) ? const_cast<void*>(mirrors[k]) : nullptr;
        fa.pad[k] = k < n_ranks ? reinterpret_cast<uint32_t*>(const_cast<void*>(pads[k])) : nullptr;
    }
    fa.n_dst = n_ranks;
    const int64_t num_vecs = count / E;
    int64_t blocks = ((num_vecs + n_ranks - 1) / n_ranks + 511) / 512;
    const int64_t cap = max_ctas > 0 ? max_ctas : 148;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (in_bf16)
        fedavg_p2p_kernel<true><<<(int)blocks, 512, 0, s>>>(fa, rank, n_src, num_vecs, DCAD_FED_TWO_SHOT, bf16_begin, bf16_end, epoch);
    else
        fedavg_p2p_kernel<false><<<(int)blocks, 512, 0, s>>>(fa, rank, n_src, num_vecs, DCAD_FED_TWO_SHOT, bf16_begin, bf16_end, epoch);
    return (int)cudaGetLastError();
}

extern "C" int dcad_fedavg_nvls(void* mc_arena, void* mc_mirror, const void* const* pads, int rank, int n_clients,
                                int64_t count, int64_t bf16_begin, int64_t bf16_end, uint32_t epoch, int max_ctas,
                                void* stream) {
    if (count <= 0) return 0;
    if (count % 4 || (bf16_begin & 3) || (bf16_end & 3)) return DCAD_ERR_BAD_ALIGNMENT;
    if (rank < 0 || rank >= n_clients || pads == nullptr || mc_arena == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    FedArgs fa;
    const void* dummy[DCAD_FED_MAX_CLIENTS] = {};
    int rc = fill_args(fa, dummy, nullptr, nullptr, pads, nullptr, n_clients);
    if (rc) return rc;
    const int64_t num_vecs = count / 4;
    int64_t blocks = ((num_vecs + n_clients - 1) / n_clients + 511) / 512;
    const int64_t cap = max_ctas > 0 ? max_ctas : 148;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    fedavg_nvls_kernel<<<(int)blocks, 512, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        fa, reinterpret_cast<float*>(mc_arena), reinterpret_cast<__nv_bfloat16*>(mc_mirror), rank, n_clients, num_vecs,
        bf16_begin, bf16_end, epoch);
    return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// peer-memory runtime (these DO allocate / map memory; they are set-up calls, not hot-path calls)
// ------------------------------------------------------------------------------------------------
extern "C" int dcad_p2p_alloc(void** ptr, size_t bytes) {
    DCAD_CUDA_TRY(cudaMalloc(ptr, bytes));
    DCAD_CUDA_TRY(cudaMemset(*ptr, 0, bytes));
    return 0;
}
extern "C" int dcad_p2p_free(void* ptr) { return (int)cudaFree(ptr); }
extern "C" int dcad_p2p_ipc_get(void* ptr, void* handle64) {
    cudaIpcMemHandle_t h;
    DCAD_CUDA_TRY(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle64, &h, sizeof(h));
    return 0;
}
extern "C" int dcad_p2p_ipc_open(const void* handle64, void** ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    DCAD_CUDA_TRY(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
extern "C" int dcad_p2p_ipc_close(void* ptr) { return (int)cudaIpcCloseMemHandle(ptr); }
extern "C" int dcad_p2p_enable_peer_access(int peer_device) {
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
    return (int)e;
}
extern "C" int dcad_p2p_can_access(int device, int peer, int* out) {
    return (int)cudaDeviceCanAccessPeer(out, device, peer);
}
