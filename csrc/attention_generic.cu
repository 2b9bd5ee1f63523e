// Multi-head self-attention for sequences LONGER than one 128-row tile (128 < S <= 512, head_dim 64): the fallback
// behind dcad_attn_fwd / dcad_attn_bwd.
//
// The reference pads / truncates every flow sentence to max_len = 128 (client1.py:27,38-45) and the tcgen05 kernels of
// attention.cu are specialised for exactly that case (whole sequence = one tile).  max_len is a parameter of the
// reference's Dataset, though, and DistilBERT has 512 positions, so longer sequences must not be an error.  This file
// covers them with plain CUDA-core kernels: correctness first, no tensor cores, not tuned -- a flow sentence is ~84
// tokens, so nothing on the benchmarked path comes here.
//
// Same math as attention.cu / HF's eager attention (transformers modeling_distilbert.py:126-151): scores = QK^T / 8,
// key-padding mask from the per-sample valid length, softmax, dropout on the probabilities, context = P V.  The
// dropout mask is a pure function of (seed, step, stream, element), so the two backward kernels regenerate it.
//   forward          grid (B*H, S/16): 16 query rows per CTA, probabilities staged in shared memory
//   backward, dQ     grid (B*H, S/16): recomputes P and dS for 16 query rows, dQ = dS K
//   backward, dK dV  grid (B*H, S/8) : recomputes P and dS for 8 KEYS against every query, dV = P^T dO, dK = dS^T Q
// Splitting the backward by output (queries / keys) instead of using atomics keeps it deterministic and needs no
// fp32 scratch.  Query rows past the sample's length are padding: context and gradients are written as zeros.
//
// Layout as in attention.cu: qkv [B*S, 3*H*64] bf16 (Q | K | V column blocks), ctx / dctx [B*S, H*64] bf16,
// lse2 [B*H*S] fp32 = log2-domain logsumexp of the scaled scores.
#include "common.cuh"
#include "../include/dcad_b200.h"

namespace dcad {

constexpr int AG_D = 64;          // head dim
constexpr int AG_R = 16;          // query rows per CTA (forward, dQ)
constexpr int AG_RK = 8;          // keys per CTA (dK dV): two [8][512] fp32 arrays stay under 48 KB of static smem
constexpr int AG_MAXS = 512;      // learned position table of DistilBERT
constexpr int AG_THREADS = 256;
constexpr float AG_LOG2E = 1.4426950408889634f;

__device__ __forceinline__ uint64_t ag_seed(const uint64_t* rng) {
    return rng ? rng[0] + rng[1] * 0xD1B54A32D192ED03ULL : 0ULL;
}
// probability (b*H+h, q, k) is kept with probability 1 - p
__device__ __forceinline__ bool ag_keep(uint64_t seed, uint32_t stream, int bh, int q, int k, int S, uint32_t thresh) {
    return rng_u32(seed, stream, static_cast<uint32_t>((bh * S + q) * S + k)) >= thresh;
}
// dot product of two 64-element bf16 vectors (16-byte aligned), fp32 accumulation
__device__ __forceinline__ float dot64(const __nv_bfloat16* a, const __nv_bfloat16* b) {
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < AG_D / 8; ++i) {
        const uint4 ua = __ldg(reinterpret_cast<const uint4*>(a) + i);
        const uint4 ub = __ldg(reinterpret_cast<const uint4*>(b) + i);
        const float2 a0 = unpack_bf16x2(ua.x), a1 = unpack_bf16x2(ua.y), a2 = unpack_bf16x2(ua.z), a3 = unpack_bf16x2(ua.w);
        const float2 b0 = unpack_bf16x2(ub.x), b1 = unpack_bf16x2(ub.y), b2 = unpack_bf16x2(ub.z), b3 = unpack_bf16x2(ub.w);
        acc += a0.x * b0.x + a0.y * b0.y + a1.x * b1.x + a1.y * b1.y + a2.x * b2.x + a2.y * b2.y + a3.x * b3.x + a3.y * b3.y;
    }
    return acc;
}

struct AgPtrs {
    const __nv_bfloat16* q;   // row 0 of this (batch, head): element (s, d) at q[s * ld + d]
    const __nv_bfloat16* k;
    const __nv_bfloat16* v;
    int ld;
};
__device__ __forceinline__ AgPtrs ag_ptrs(const __nv_bfloat16* qkv, int b, int h, int S, int H) {
    const int HD = H * AG_D;
    const __nv_bfloat16* base = qkv + static_cast<size_t>(b) * S * 3 * HD + h * AG_D;
    return AgPtrs{base, base + HD, base + 2 * HD, 3 * HD};
}

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(AG_THREADS)
attn_gen_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ ctx, float* __restrict__ lse2_out,
                    const int32_t* __restrict__ lens, int S, int H, float scale, uint32_t thresh, float keep_scale,
                    const uint64_t* __restrict__ rng, uint32_t stream_id) {
    pdl_trigger();
    pdl_wait();
    __shared__ float s_p[AG_R][AG_MAXS];   // scores, then (dropped, normalised) probabilities
    const int bh = blockIdx.x, b = bh / H, h = bh % H;
    const int q0 = blockIdx.y * AG_R;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int len = lens ? min(lens[b], S) : S;
    const AgPtrs P = ag_ptrs(qkv, b, h, S, H);
    const uint64_t seed = ag_seed(rng);
    // scores: thread -> key(s), all 16 query rows
    for (int k = tid; k < len; k += AG_THREADS) {
#pragma unroll 4
        for (int r = 0; r < AG_R; ++r) {
            const int q = q0 + r;
            s_p[r][k] = (q < len) ? dot64(P.q + static_cast<size_t>(q) * P.ld, P.k + static_cast<size_t>(k) * P.ld) * scale
                                  : 0.f;
        }
    }
    __syncthreads();
    // softmax: warp w owns rows 2w, 2w+1
    for (int r = warp * 2; r < warp * 2 + 2; ++r) {
        const int q = q0 + r;
        if (q >= len || q >= S) {
            for (int k = lane; k < len; k += 32) s_p[r][k] = 0.f;
            if (lane == 0 && q < S) lse2_out[static_cast<size_t>(bh) * S + q] = 0.f;
            continue;
        }
        float mx = -INFINITY;
        for (int k = lane; k < len; k += 32) mx = fmaxf(mx, s_p[r][k]);
        mx = warp_max(mx);
        float sum = 0.f;
        for (int k = lane; k < len; k += 32) sum += fast_exp2((s_p[r][k] - mx) * AG_LOG2E);
        sum = warp_sum(sum);
        const float l2 = mx * AG_LOG2E + log2f(sum);
        if (lane == 0) lse2_out[static_cast<size_t>(bh) * S + q] = l2;
        for (int k = lane; k < len; k += 32) {
            float p = fast_exp2(s_p[r][k] * AG_LOG2E - l2);
            if (thresh) p = ag_keep(seed, stream_id, bh, q, k, S, thresh) ? p * keep_scale : 0.f;
            s_p[r][k] = p;
        }
    }
    __syncthreads();
    // context: thread -> (row r, 4 output dims)
    const int r = tid >> 4, d4 = (tid & 15) * 4;
    const int q = q0 + r;
    if (q < S) {
        float o0 = 0.f, o1 = 0.f, o2 = 0.f, o3 = 0.f;
        if (q < len) {
            for (int k = 0; k < len; ++k) {
                const float p = s_p[r][k];
                const uint2 u = __ldg(reinterpret_cast<const uint2*>(P.v + static_cast<size_t>(k) * P.ld + d4));
                const float2 v01 = unpack_bf16x2(u.x), v23 = unpack_bf16x2(u.y);
                o0 = fmaf(p, v01.x, o0); o1 = fmaf(p, v01.y, o1); o2 = fmaf(p, v23.x, o2); o3 = fmaf(p, v23.y, o3);
            }
        }
        uint2 o;
        o.x = pack_bf16x2(o0, o1); o.y = pack_bf16x2(o2, o3);
        *reinterpret_cast<uint2*>(ctx + (static_cast<size_t>(b) * S + q) * (H * AG_D) + h * AG_D + d4) = o;
    }
}

// ------------------------------------------------------------------------------------------------
// backward.  For a (query q, key k) pair: p = exp2(s * log2e - lse2[q]); with dropout p_d = keep ? p / (1 - pd) : 0 and
// dp_d = keep ? (dO[q] . V[k]) / (1 - pd) : 0;  delta[q] = dO[q] . O[q];  dS = p * (dp_d - delta) * scale.
// ------------------------------------------------------------------------------------------------
struct AgPair { float pd, ds; };
__device__ __forceinline__ AgPair ag_pair(const AgPtrs& P, const __nv_bfloat16* dO_q, float l2, float delta, int bh,
                                          int q, int k, int S, float scale, uint64_t seed, uint32_t stream_id,
                                          uint32_t thresh, float keep_scale) {
    const float s = dot64(P.q + static_cast<size_t>(q) * P.ld, P.k + static_cast<size_t>(k) * P.ld) * scale;
    const float p = fast_exp2(s * AG_LOG2E - l2);
    float dp = dot64(dO_q, P.v + static_cast<size_t>(k) * P.ld);
    float pd = p;
    if (thresh) {
        const bool keep = ag_keep(seed, stream_id, bh, q, k, S, thresh);
        pd = keep ? p * keep_scale : 0.f;
        dp = keep ? dp * keep_scale : 0.f;
    }
    return AgPair{pd, p * (dp - delta) * scale};
}

// dQ for 16 query rows
__global__ void __launch_bounds__(AG_THREADS)
attn_gen_bwd_dq_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ ctx,
                       const __nv_bfloat16* __restrict__ dctx, const float* __restrict__ lse2,
                       __nv_bfloat16* __restrict__ dqkv, const int32_t* __restrict__ lens, int S, int H, float scale,
                       uint32_t thresh, float keep_scale, const uint64_t* __restrict__ rng, uint32_t stream_id) {
    pdl_trigger();
    pdl_wait();
    __shared__ float s_ds[AG_R][AG_MAXS];
    __shared__ float s_delta[AG_R];
    const int bh = blockIdx.x, b = bh / H, h = bh % H;
    const int q0 = blockIdx.y * AG_R;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int len = lens ? min(lens[b], S) : S;
    const int HD = H * AG_D;
    const AgPtrs P = ag_ptrs(qkv, b, h, S, H);
    const uint64_t seed = ag_seed(rng);
    const __nv_bfloat16* dO = dctx + static_cast<size_t>(b) * S * HD + h * AG_D;
    const __nv_bfloat16* O = ctx + static_cast<size_t>(b) * S * HD + h * AG_D;
    for (int r = warp * 2; r < warp * 2 + 2; ++r) {   // delta[r] = dO[q] . O[q]: one lane pair of values each
        const int q = q0 + r;
        float part = 0.f;
        if (q < len) {
            const float2 a = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(dO + static_cast<size_t>(q) * HD + lane * 2));
            const float2 o = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(O + static_cast<size_t>(q) * HD + lane * 2));
            part = a.x * o.x + a.y * o.y;
        }
        part = warp_sum(part);
        if (lane == 0) s_delta[r] = part;
    }
    __syncthreads();
    for (int k = tid; k < len; k += AG_THREADS) {
#pragma unroll 2
        for (int r = 0; r < AG_R; ++r) {
            const int q = q0 + r;
            float ds = 0.f;
            if (q < len)
                ds = ag_pair(P, dO + static_cast<size_t>(q) * HD, lse2[static_cast<size_t>(bh) * S + q], s_delta[r], bh, q,
                             k, S, scale, seed, stream_id, thresh, keep_scale).ds;
            s_ds[r][k] = ds;
        }
    }
    __syncthreads();
    const int r = tid >> 4, d4 = (tid & 15) * 4;
    const int q = q0 + r;
    if (q < S) {
        float o0 = 0.f, o1 = 0.f, o2 = 0.f, o3 = 0.f;
        if (q < len) {
            for (int k = 0; k < len; ++k) {
                const float ds = s_ds[r][k];
                const uint2 u = __ldg(reinterpret_cast<const uint2*>(P.k + static_cast<size_t>(k) * P.ld + d4));
                const float2 k01 = unpack_bf16x2(u.x), k23 = unpack_bf16x2(u.y);
                o0 = fmaf(ds, k01.x, o0); o1 = fmaf(ds, k01.y, o1); o2 = fmaf(ds, k23.x, o2); o3 = fmaf(ds, k23.y, o3);
            }
        }
        uint2 o;
        o.x = pack_bf16x2(o0, o1); o.y = pack_bf16x2(o2, o3);
        *reinterpret_cast<uint2*>(dqkv + (static_cast<size_t>(b) * S + q) * 3 * HD + h * AG_D + d4) = o;
    }
}

// dK and dV for 8 keys
__global__ void __launch_bounds__(AG_THREADS)
attn_gen_bwd_dkv_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ ctx,
                        const __nv_bfloat16* __restrict__ dctx, const float* __restrict__ lse2,
                        __nv_bfloat16* __restrict__ dqkv, const int32_t* __restrict__ lens, int S, int H, float scale,
                        uint32_t thresh, float keep_scale, const uint64_t* __restrict__ rng, uint32_t stream_id) {
    pdl_trigger();
    pdl_wait();
    __shared__ float s_pd[AG_RK][AG_MAXS];   // [key][query]
    __shared__ float s_ds[AG_RK][AG_MAXS];
    const int bh = blockIdx.x, b = bh / H, h = bh % H;
    const int k0 = blockIdx.y * AG_RK;
    const int tid = threadIdx.x;
    const int len = lens ? min(lens[b], S) : S;
    const int HD = H * AG_D;
    const AgPtrs P = ag_ptrs(qkv, b, h, S, H);
    const uint64_t seed = ag_seed(rng);
    const __nv_bfloat16* dO = dctx + static_cast<size_t>(b) * S * HD + h * AG_D;
    const __nv_bfloat16* O = ctx + static_cast<size_t>(b) * S * HD + h * AG_D;
    for (int q = tid; q < len; q += AG_THREADS) {     // thread -> query, all 8 keys of this CTA
        const __nv_bfloat16* dO_q = dO + static_cast<size_t>(q) * HD;
        const float delta = dot64(dO_q, O + static_cast<size_t>(q) * HD);
        const float l2 = lse2[static_cast<size_t>(bh) * S + q];
#pragma unroll 2
        for (int r = 0; r < AG_RK; ++r) {
            const int k = k0 + r;
            AgPair pr{0.f, 0.f};
            if (k < len) pr = ag_pair(P, dO_q, l2, delta, bh, q, k, S, scale, seed, stream_id, thresh, keep_scale);
            s_pd[r][q] = pr.pd;
            s_ds[r][q] = pr.ds;
        }
    }
    __syncthreads();
    const int r = tid >> 5, d2 = (tid & 31) * 2;     // thread -> (key r, 2 output dims)
    const int k = k0 + r;
    if (k < S) {
        float v0 = 0.f, v1 = 0.f, g0 = 0.f, g1 = 0.f;
        if (k < len) {
            for (int q = 0; q < len; ++q) {
                const float pd = s_pd[r][q], ds = s_ds[r][q];
                const float2 d01 = unpack_bf16x2(__ldg(reinterpret_cast<const uint32_t*>(dO + static_cast<size_t>(q) * HD + d2)));
                const float2 q01 = unpack_bf16x2(__ldg(reinterpret_cast<const uint32_t*>(P.q + static_cast<size_t>(q) * P.ld + d2)));
                v0 = fmaf(pd, d01.x, v0); v1 = fmaf(pd, d01.y, v1);
                g0 = fmaf(ds, q01.x, g0); g1 = fmaf(ds, q01.y, g1);
            }
        }
        __nv_bfloat16* row = dqkv + (static_cast<size_t>(b) * S + k) * 3 * HD + h * AG_D + d2;
        *reinterpret_cast<uint32_t*>(row + HD) = pack_bf16x2(g0, g1);          // dK
        *reinterpret_cast<uint32_t*>(row + 2 * HD) = pack_bf16x2(v0, v1);      // dV
    }
}

static inline uint32_t ag_drop_thresh(float p) {
    if (p <= 0.f) return 0u;
    double t = static_cast<double>(p) * 4294967296.0;
    if (t > 4294967295.0) t = 4294967295.0;
    return static_cast<uint32_t>(t);
}

// called by dcad_attn_fwd / dcad_attn_bwd (attention.cu) for 128 < S <= 512
int attn_generic_fwd(const void* qkv, void* ctx, float* lse2, const int32_t* lens, int B, int S, int H, float p_drop,
                     const uint64_t* rng, uint32_t stream_id, cudaStream_t stream) {
    if (S > AG_MAXS) return DCAD_ERR_BAD_SHAPE;
    const float ks = p_drop > 0.f ? 1.0f / (1.0f - p_drop) : 1.0f;
    return (int)launch_pdl(attn_gen_fwd_kernel, dim3(B * H, (S + AG_R - 1) / AG_R), dim3(AG_THREADS), 0, stream,
                           reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<__nv_bfloat16*>(ctx), lse2, lens,
                           S, H, 0.125f, ag_drop_thresh(p_drop), ks, rng, stream_id);
}

int attn_generic_bwd(const void* qkv, const void* ctx, const void* dctx, const float* lse2, void* dqkv,
                     const int32_t* lens, int B, int S, int H, float p_drop, const uint64_t* rng, uint32_t stream_id,
                     cudaStream_t stream) {
    if (S > AG_MAXS) return DCAD_ERR_BAD_SHAPE;
    const float ks = p_drop > 0.f ? 1.0f / (1.0f - p_drop) : 1.0f;
    const dim3 grid(B * H, (S + AG_R - 1) / AG_R);
    cudaError_t e = launch_pdl(attn_gen_bwd_dq_kernel, grid, dim3(AG_THREADS), 0, stream,
                               reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<const __nv_bfloat16*>(ctx),
                               reinterpret_cast<const __nv_bfloat16*>(dctx), lse2, reinterpret_cast<__nv_bfloat16*>(dqkv),
                               lens, S, H, 0.125f, ag_drop_thresh(p_drop), ks, rng, stream_id);
    if (e != cudaSuccess) return (int)e;
    return (int)launch_pdl(attn_gen_bwd_dkv_kernel, dim3(B * H, (S + AG_RK - 1) / AG_RK), dim3(AG_THREADS), 0, stream,
                           reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<const __nv_bfloat16*>(ctx),
                           reinterpret_cast<const __nv_bfloat16*>(dctx), lse2, reinterpret_cast<__nv_bfloat16*>(dqkv),
                           lens, S, H, 0.125f, ag_drop_thresh(p_drop), ks, rng, stream_id);
}

}  // namespace dcad
