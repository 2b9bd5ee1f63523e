// Flat Adam over the contiguous parameter arena: one launch updates all 102 tensors (66.4 M parameters), emits
// the bf16 compute copy of the GEMM weights in the same pass and clears the gradients that are accumulated with
// atomics, so optimizer.zero_grad() costs nothing extra.
//
// Replaces torch.optim.Adam(lr=2e-5).step() + optimizer.zero_grad() of the reference (client1.py:106,110,380) with
// identical math (betas (0.9, 0.999), eps 1e-8, no weight decay, bias-corrected, non-amsgrad).
// The step counter lives in device memory (rng_state[1]) so the launch is replayable inside a CUDA graph.
#include "common.cuh"
#include "../include/dcad_b200.h"

namespace dcad {

__global__ void __launch_bounds__(256)
adam_flat_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                 __nv_bfloat16* __restrict__ p_bf16, int64_t n, int64_t bf16_begin, int64_t bf16_end,
                 int64_t zero_begin, int64_t zero_end, float lr, float beta1, float beta2, float eps,
                 const uint64_t* __restrict__ state, const uint8_t* __restrict__ row_flags, int64_t rows_begin,
                 int64_t rows_end, int row_len, unsigned int* __restrict__ advance_ticket) {
    __shared__ float s_c[2];
    pdl_trigger();
    pdl_wait();
    if (threadIdx.x == 0) {
        const double t = static_cast<double>(state ? state[1] + 1 : 1);
        const double bc1 = 1.0 - pow(static_cast<double>(beta1), t);
        const double bc2 = 1.0 - pow(static_cast<double>(beta2), t);
        s_c[0] = static_cast<float>(static_cast<double>(lr) / bc1);  // step size
        s_c[1] = static_cast<float>(1.0 / sqrt(bc2));                // 1/sqrt(bias_correction2)
    }
    __syncthreads();
    const float step_size = s_c[0], inv_sqrt_bc2 = s_c[1];
    const float omb1 = 1.0f - beta1, omb2 = 1.0f - beta2;
    auto update = [&](int64_t i) {
        float4 pp = *reinterpret_cast<const float4*>(p + i);
        const float4 gg = *reinterpret_cast<const float4*>(g + i);
        float4 mm = *reinterpret_cast<const float4*>(m + i);
        float4 vv = *reinterpret_cast<const float4*>(v + i);
#define DCAD_ADAM_ONE(c)                                                   \
        mm.c = beta1 * mm.c + omb1 * gg.c;                                 \
        vv.c = beta2 * vv.c + omb2 * gg.c * gg.c;                          \
        pp.c -= step_size * (mm.c / (sqrtf(vv.c) * inv_sqrt_bc2 + eps));
        DCAD_ADAM_ONE(x) DCAD_ADAM_ONE(y) DCAD_ADAM_ONE(z) DCAD_ADAM_ONE(w)
#undef DCAD_ADAM_ONE
        *reinterpret_cast<float4*>(p + i) = pp;
        *reinterpret_cast<float4*>(m + i) = mm;
        *reinterpret_cast<float4*>(v + i) = vv;
        if (i >= zero_begin && i < zero_end) *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p_bf16 && i >= bf16_begin && i < bf16_end) {
            uint2 o;
            o.x = pack_bf16x2(pp.x, pp.y);
            o.y = pack_bf16x2(pp.z, pp.w);
            *reinterpret_cast<uint2*>(p_bf16 + (i - bf16_begin)) = o;
        }
    };
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x * 4;
    const int64_t first = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
    if (row_flags) {
        // Embedding rows that never received a gradient have g = m = v = 0: the update is the identity and the row is
        // skipped.  The dense pieces around the table go through the grid-stride loop; the table itself is walked by
        // rows, so an untouched row costs one flag read, not one per vector.  (Testing the flag per 4-element
        // vector cost 25 us for a fully skipped table -- 11.5 M warp instructions, ncu -- and a blocked mapping with
        // the flags staged in shared memory was slower still: the dense part then sits in ~50 blocks.)
        const int64_t table = rows_end - rows_begin;
        const int64_t n_dense = n - table;
        for (int64_t j = first; j < n_dense; j += stride) update(j < rows_begin ? j : j + table);
        // Groups of 8 rows per block: one warp reads the 8 flags (one load + ballot), then the block's warps share the
        // 128-element chunks of the LIVE rows, so a touched row is updated by several warps at once instead of by one
        // warp in row_len/128 dependent iterations.
        const uint32_t n_rows = static_cast<uint32_t>(table / row_len);
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
        const int chunks = (row_len + 127) >> 7;
        __shared__ uint32_t s_mask;
        for (uint32_t g0 = blockIdx.x * 8u; g0 < n_rows; g0 += gridDim.x * 8u) {
            if (warp == 0) {
                const uint32_t r = g0 + lane;
                const uint32_t mk = __ballot_sync(0xffffffffu, lane < 8 && r < n_rows && row_flags[r] != 0);
                if (lane == 0) s_mask = mk;
            }
            __syncthreads();
            const uint32_t mk = s_mask;
            const int units = __popc(mk) * chunks;
            for (int u = warp; u < units; u += n_warps) {
                const int k = u / chunks, c = (u - k * chunks) * 128 + lane * 4;
                uint32_t t = mk;
                for (int j = 0; j < k; ++j) t &= t - 1;  // drop the k lowest live rows
                const uint32_t r = g0 + (__ffs(t) - 1);
                if (c < row_len) update(rows_begin + static_cast<int64_t>(r) * row_len + c);
            }
            __syncthreads();  // s_mask is rewritten by the next group
        }
    } else {
        for (int64_t i = first; i < n; i += stride) update(i);
    }
    if (advance_ticket) {
        // Last launch of the step: the block that finishes last advances the step counter (every block read it at its
        // start, and none can read it after this point), which saves the separate one-thread launch.
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(advance_ticket, 1u) == gridDim.x - 1) {
                *advance_ticket = 0u;
                const_cast<uint64_t*>(state)[1] += 1;
            }
        }
    }
}

// Sharded variant for intra-client data parallelism: this rank owns elements [begin, end) of the arena.  Gradients
// were summed over the group (wgrad TMA reduce-add / reduce-scatter), so they are scaled by grad_scale = 1/G here.
// The updated values are pushed to every peer of the group over NVLink in the same pass: the bf16 GEMM weights into
// all peers' mirrors, and (for the region that kernels read in fp32: embeddings, LayerNorm, biases, classifier) the
// fp32 values into all peers' parameter arenas -- the parameter all-gather, fused into the optimizer.
struct AdamPeers {
    float* p[16];
    __nv_bfloat16* bf16[16];
    int n;
    int self;
};

__global__ void __launch_bounds__(256)
adam_dp_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
               int64_t begin, int64_t end, int64_t bf16_begin, int64_t bf16_end, int64_t push32_begin,
               int64_t push32_end, float lr, float beta1, float beta2, float eps, float grad_scale,
               const uint64_t* __restrict__ state, const AdamPeers peers) {
    __shared__ float s_c[2];
    if (threadIdx.x == 0) {
        const double t = static_cast<double>(state ? state[1] + 1 : 1);
        const double bc1 = 1.0 - pow(static_cast<double>(beta1), t);
        const double bc2 = 1.0 - pow(static_cast<double>(beta2), t);
        s_c[0] = static_cast<float>(static_cast<double>(lr) / bc1);
        s_c[1] = static_cast<float>(1.0 / sqrt(bc2));
    }
    __syncthreads();
    const float step_size = s_c[0], inv_sqrt_bc2 = s_c[1];
    const float omb1 = 1.0f - beta1, omb2 = 1.0f - beta2;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x * 4;
    for (int64_t i = begin + (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i < end; i += stride) {
        float4 pp = *reinterpret_cast<const float4*>(p + i);
        float4 gg = *reinterpret_cast<const float4*>(g + i);
        float4 mm = *reinterpret_cast<const float4*>(m + i);
        float4 vv = *reinterpret_cast<const float4*>(v + i);
        gg.x *= grad_scale; gg.y *= grad_scale; gg.z *= grad_scale; gg.w *= grad_scale;
#define DCAD_ADAM_ONE(c)                                                   \
        mm.c = beta1 * mm.c + omb1 * gg.c;                                 \
        vv.c = beta2 * vv.c + omb2 * gg.c * gg.c;                          \
        pp.c -= step_size * (mm.c / (sqrtf(vv.c) * inv_sqrt_bc2 + eps));
        DCAD_ADAM_ONE(x) DCAD_ADAM_ONE(y) DCAD_ADAM_ONE(z) DCAD_ADAM_ONE(w)
#undef DCAD_ADAM_ONE
        *reinterpret_cast<float4*>(p + i) = pp;
        *reinterpret_cast<float4*>(m + i) = mm;
        *reinterpret_cast<float4*>(v + i) = vv;
        *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i >= push32_begin && i < push32_end) {
            for (int k = 0; k < peers.n; ++k)
                if (k != peers.self) *reinterpret_cast<float4*>(peers.p[k] + i) = pp;
        }
        if (i >= bf16_begin && i < bf16_end) {
            uint2 o;
            o.x = pack_bf16x2(pp.x, pp.y);
            o.y = pack_bf16x2(pp.z, pp.w);
            for (int k = 0; k < peers.n; ++k) *reinterpret_cast<uint2*>(peers.bf16[k] + (i - bf16_begin)) = o;
        }
    }
}

// Row-sparse sharded Adam for the word-embedding table of a data-parallel client (rows [row_lo, row_hi) are owned by
// this rank).  A vocabulary row that no GPU of the group has ever touched has g = m = v = 0 on every GPU: its update is
// the identity and it moves no bytes.  For a touched row the owner sums the G locally accumulated gradient rows
// (peer reads over NVLink), applies Adam (fp32 moments live with the owner only) and stores the new row into every
// GPU's parameter arena -- reduce-scatter, optimizer and all-gather of the table in one pass over the rows that
// matter.  Flow sentences use a few thousand of the 30,522 rows, so the table's share of the per-step NVLink traffic
// drops from 2 x (G-1)/G x 94 MB to a few MB.  flags[k][row] != 0: GPU k has touched the row since the optimizer
// state was reset (written by embed_ln_bwd into peer-mapped memory).  One warp per row.
struct AdamRowPeers {
    float* p[8];             // parameter arenas, word-table base
    const float* g[8];       // gradient arenas, word-table base (locally accumulated, this step)
    const uint8_t* flags[8];
    int n;
};

__global__ void __launch_bounds__(256)
adam_dp_rows_kernel(float* __restrict__ m, float* __restrict__ v, float* __restrict__ g_own, const AdamRowPeers peers,
                    int row_lo, int row_hi, int row_len, float lr, float beta1, float beta2, float eps, float grad_scale,
                    const uint64_t* __restrict__ state) {
    __shared__ float s_c[2];
    if (threadIdx.x == 0) {
        const double t = static_cast<double>(state ? state[1] + 1 : 1);
        const double bc1 = 1.0 - pow(static_cast<double>(beta1), t);
        const double bc2 = 1.0 - pow(static_cast<double>(beta2), t);
        s_c[0] = static_cast<float>(static_cast<double>(lr) / bc1);
        s_c[1] = static_cast<float>(1.0 / sqrt(bc2));
    }
    __syncthreads();
    const float step_size = s_c[0], inv_sqrt_bc2 = s_c[1];
    const float omb1 = 1.0f - beta1, omb2 = 1.0f - beta2;
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int row = row_lo + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5); row < row_hi; row += warps) {
        uint32_t f = (lane < peers.n) ? peers.flags[lane][row] : 0u;
        if (__ballot_sync(0xffffffffu, f != 0u) == 0u) continue;
        const size_t base = static_cast<size_t>(row) * row_len;
        // two 16-byte chunks per lane and pass; all peer loads of a pass are issued before any arithmetic (an NVLink
        // round trip is ~2 us: the row costs a few round trips, not n_peers x chunks of them)
        for (int c0 = lane * 4; c0 < row_len; c0 += 256) {
            float4 t[2][8];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int c = c0 + h * 128;
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (k < peers.n && c < row_len) t[h][k] = *reinterpret_cast<const float4*>(peers.g[k] + base + c);
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int c = c0 + h * 128;
                if (c >= row_len) continue;
                float4 gg = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (k < peers.n) { gg.x += t[h][k].x; gg.y += t[h][k].y; gg.z += t[h][k].z; gg.w += t[h][k].w; }
                gg.x *= grad_scale; gg.y *= grad_scale; gg.z *= grad_scale; gg.w *= grad_scale;
                float4 pp = *reinterpret_cast<const float4*>(peers.p[0] + base + c);  // every replica holds the same value
                float4 mm = *reinterpret_cast<const float4*>(m + base + c);
                float4 vv = *reinterpret_cast<const float4*>(v + base + c);
#define DCAD_ADAM_ONE(x_)                                                         \
                mm.x_ = beta1 * mm.x_ + omb1 * gg.x_;                             \
                vv.x_ = beta2 * vv.x_ + omb2 * gg.x_ * gg.x_;                     \
                pp.x_ -= step_size * (mm.x_ / (sqrtf(vv.x_) * inv_sqrt_bc2 + eps));
                DCAD_ADAM_ONE(x) DCAD_ADAM_ONE(y) DCAD_ADAM_ONE(z) DCAD_ADAM_ONE(w)
#undef DCAD_ADAM_ONE
                *reinterpret_cast<float4*>(m + base + c) = mm;
                *reinterpret_cast<float4*>(v + base + c) = vv;
                *reinterpret_cast<float4*>(g_own + base + c) = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int k = 0; k < peers.n; ++k) *reinterpret_cast<float4*>(peers.p[k] + base + c) = pp;
            }
        }
    }
}

// One-block cross-GPU barrier over the group's signal pads (words [48, 64) of each pad, slot k written by rank k).
// The epoch is derived from the device-side step counter so the launch is replayable inside a CUDA graph.
__global__ void xgpu_barrier_kernel(uint32_t* const* pads_dev_unused, const uint64_t* state, uint32_t phase, int rank,
                                    int n, uint32_t* p0, uint32_t* p1, uint32_t* p2, uint32_t* p3, uint32_t* p4,
                                    uint32_t* p5, uint32_t* p6, uint32_t* p7) {
    (void)pads_dev_unused;
    uint32_t* pads[8] = {p0, p1, p2, p3, p4, p5, p6, p7};
    const uint32_t epoch = static_cast<uint32_t>(state[1]) * 8u + phase + 1u;
    __threadfence_system();
    if (threadIdx.x < n) {
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pads[threadIdx.x] + 48 + rank), "r"(epoch) : "memory");
        const uint32_t* mine = pads[rank] + 48 + threadIdx.x;
        uint32_t v = 0;
        for (uint32_t it = 0; it < (1u << 26); ++it) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
            if (static_cast<int32_t>(v - epoch) >= 0) break;
            __nanosleep(32);
        }
        if (static_cast<int32_t>(v - epoch) < 0) pads[rank][33] = 1;  // error word, same as the FedAvg kernels
    }
}

__global__ void step_advance_kernel(uint64_t* state) {
    pdl_trigger();
    pdl_wait();
    if (threadIdx.x == 0 && blockIdx.x == 0) state[1] += 1;
}

}  // namespace dcad

using namespace dcad;

// Grid cap of the flat Adam kernel in 256-thread blocks per SM.  The kernel runs on the side stream underneath the
// backward pass; a finer grid lets its blocks interleave with the GEMM CTAs of the chain.  Measured in the training
// step (batch 16, 96 executed positions, 200 steps): 4 -> 1.307 ms, 8 -> 1.281, 16 -> 1.259, 32 -> 1.251,
// 48 -> 1.255, 64 -> 1.263.
static int g_adam_blocks_per_sm = 32;
extern "C" int dcad_set_adam_blocks_per_sm(int blocks) {
    const int prev = g_adam_blocks_per_sm;
    if (blocks >= 1 && blocks <= 64) g_adam_blocks_per_sm = blocks;
    return prev;
}

// n, bf16_begin/end, zero_begin/end are element offsets into the arena and must be multiples of 4.
extern "C" int dcad_adam_step_rows(float* p, float* g, float* m, float* v, void* p_bf16, int64_t n,
                                   int64_t bf16_begin, int64_t bf16_end, int64_t zero_begin, int64_t zero_end,
                                   float lr, float beta1, float beta2, float eps, const uint64_t* state,
                                   const uint8_t* row_flags, int64_t rows_begin, int64_t rows_end, int row_len,
                                   unsigned int* advance_ticket, void* stream) {
    if (n <= 0) return 0;
    if (advance_ticket && state == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (row_flags && (row_len <= 0 || (row_len & 3) || (rows_begin & 3))) return DCAD_ERR_BAD_ALIGNMENT;
    if (row_flags && (rows_begin < 0 || rows_end > n || rows_end < rows_begin || (rows_end - rows_begin) % row_len != 0 ||
                      (rows_end - rows_begin) / row_len >= (int64_t(1) << 31)))
        return DCAD_ERR_BAD_ARGUMENT;  // the table is walked by whole rows
    if ((n | bf16_begin | bf16_end | zero_begin | zero_end) & 3) return DCAD_ERR_BAD_ALIGNMENT;
    if ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
         reinterpret_cast<uintptr_t>(v)) & 15u)
        return DCAD_ERR_BAD_ALIGNMENT;
    int64_t blocks = (n / 4 + 255) / 256;
    const int64_t cap = 148 * static_cast<int64_t>(g_adam_blocks_per_sm);
    if (blocks > cap) blocks = cap;
    return (int)launch_pdl(adam_flat_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0,
                           reinterpret_cast<cudaStream_t>(stream), p, g, m, v, reinterpret_cast<__nv_bfloat16*>(p_bf16),
                           n, bf16_begin, bf16_end, zero_begin, zero_end, lr, beta1, beta2, eps, state, row_flags,
                           rows_begin, rows_end, row_len, advance_ticket);
}

extern "C" int dcad_adam_step(float* p, float* g, float* m, float* v, void* p_bf16, int64_t n, int64_t bf16_begin,
                              int64_t bf16_end, int64_t zero_begin, int64_t zero_end, float lr, float beta1,
                              float beta2, float eps, const uint64_t* state, void* stream) {
    return dcad_adam_step_rows(p, g, m, v, p_bf16, n, bf16_begin, bf16_end, zero_begin, zero_end, lr, beta1, beta2,
                               eps, state, nullptr, 0, 0, 1, nullptr, stream);
}

extern "C" int dcad_step_advance(uint64_t* state, void* stream) {
    return (int)launch_pdl(step_advance_kernel, dim3(1), dim3(32), 0, reinterpret_cast<cudaStream_t>(stream), state);
}

// Sharded Adam + fused parameter all-gather (see adam_dp_kernel).  peer_p / peer_bf16 are HOST arrays of n_peers
// device pointers (this rank's own buffers included at index `self`).
extern "C" int dcad_adam_step_dp(float* p, float* g, float* m, float* v, int64_t begin, int64_t end, int64_t bf16_begin,
                                 int64_t bf16_end, int64_t push32_begin, int64_t push32_end, float lr, float beta1,
                                 float beta2, float eps, float grad_scale, const uint64_t* state,
                                 const void* const* peer_p, const void* const* peer_bf16, int n_peers, int self,
                                 void* stream) {
    if (end <= begin) return 0;
    if ((begin | end | bf16_begin | bf16_end | push32_begin | push32_end) & 3) return DCAD_ERR_BAD_ALIGNMENT;
    if (n_peers < 1 || n_peers > 16 || self < 0 || self >= n_peers) return DCAD_ERR_BAD_ARGUMENT;
    AdamPeers peers;
    for (int k = 0; k < 16; ++k) {
        peers.p[k] = (k < n_peers && peer_p) ? reinterpret_cast<float*>(const_cast<void*>(peer_p[k])) : nullptr;
        peers.bf16[k] = (k < n_peers && peer_bf16) ? reinterpret_cast<__nv_bfloat16*>(const_cast<void*>(peer_bf16[k])) : nullptr;
    }
    peers.n = n_peers;
    peers.self = self;
    if (!peer_p) { push32_begin = 0; push32_end = 0; }
    if (!peer_bf16) { bf16_begin = 0; bf16_end = 0; }
    int64_t blocks = ((end - begin) / 4 + 255) / 256;
    const int64_t cap = 148 * 16;
    if (blocks > cap) blocks = cap;
    adam_dp_kernel<<<static_cast<int>(blocks), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        p, g, m, v, begin, end, bf16_begin, bf16_end, push32_begin, push32_end, lr, beta1, beta2, eps, grad_scale, state,
        peers);
    return (int)cudaGetLastError();
}

// Row-sparse sharded Adam of the word-embedding table (see adam_dp_rows_kernel).  All pointer arrays are HOST arrays of
// n device pointers: p / g point at the word table inside each GPU's parameter / gradient arena, flags at each GPU's
// touched-row bytes.  peers.p[0] must be THIS rank's own table (it is the replica the old value is read from); m, v,
// g_own are this rank's moment / gradient tables.  Rows [row_lo, row_hi) are updated; row_len % 4 == 0.
extern "C" int dcad_adam_step_dp_rows(float* m, float* v, float* g_own, void* const* p_ptrs, const void* const* g_ptrs,
                                      const void* const* flag_ptrs, int n, int row_lo, int row_hi, int row_len, float lr,
                                      float beta1, float beta2, float eps, float grad_scale, const uint64_t* state,
                                      void* stream) {
    if (row_hi <= row_lo) return 0;
    if (n < 1 || n > 8 || (row_len & 3)) return DCAD_ERR_BAD_ARGUMENT;
    AdamRowPeers peers;
    for (int k = 0; k < 8; ++k) {
        peers.p[k] = k < n ? reinterpret_cast<float*>(p_ptrs[k]) : nullptr;
        peers.g[k] = k < n ? reinterpret_cast<const float*>(g_ptrs[k]) : nullptr;
        peers.flags[k] = k < n ? reinterpret_cast<const uint8_t*>(flag_ptrs[k]) : nullptr;
    }
    peers.n = n;
    const int rows = row_hi - row_lo;
    int blocks = (rows + 7) / 8;  // 8 warps (rows) per block
    if (blocks > 148 * 8) blocks = 148 * 8;
    adam_dp_rows_kernel<<<blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(m, v, g_own, peers, row_lo, row_hi,
                                                                                      row_len, lr, beta1, beta2, eps,
                                                                                      grad_scale, state);
    return (int)cudaGetLastError();
}

extern "C" int dcad_xgpu_barrier(const void* const* pads, int rank, int n, const uint64_t* state, uint32_t phase,
                                 void* stream) {
    if (n < 1 || n > 8 || rank < 0 || rank >= n || phase > 7) return DCAD_ERR_BAD_ARGUMENT;
    uint32_t* p[8];
    for (int k = 0; k < 8; ++k) p[k] = k < n ? reinterpret_cast<uint32_t*>(const_cast<void*>(pads[k])) : nullptr;
    xgpu_barrier_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(nullptr, state, phase, rank, n, p[0], p[1],
                                                                                p[2], p[3], p[4], p[5], p[6], p[7]);
    return (int)cudaGetLastError();
}
