// Fused multi-head self-attention for sequences of up to 128 tokens (the reference truncates/pads every flow
// sentence to max_len=128, client1.py:27,38-45), head_dim 64: one CTA per (batch, head), the whole sequence is a
// single 128-row tile, so there is no online-softmax loop and the backward needs no atomics.
//
// forward : S = Q K^T (tcgen05, TMEM) -> masked softmax + dropout in registers -> P (bf16, smem) -> O = P V (tcgen05)
// backward: S = Q K^T, dP = dO V^T -> P, dS in registers -> dV = P^T dO, dK = dS^T Q, dQ = dS K (three more MMAs)
//
// Replaces HF DistilBERT's attention core (transformers modeling_distilbert.py:126-151 eager /
// F.scaled_dot_product_attention) reached from reference client1.py:61, incl. attention dropout 0.1 and the
// key-padding mask (here: per-sample valid length instead of a [B,1,S,S] mask tensor).
//
// Layout: qkv is [B*S, 3*H*64] bf16 (Q | K | V column blocks, head-major inside each), ctx is [B*S, H*64].
#include <mutex>

#include "common.cuh"
#include "../include/dcad_b200.h"

namespace dcad {

int make_tensor_map_bf16_3d(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t ld1,
                            uint64_t ld2, uint32_t b0, uint32_t b1);
// attention_generic.cu: CUDA-core fallback for 128 < S <= 512
int attn_generic_fwd(const void* qkv, void* ctx, float* lse2, const int32_t* lens, int B, int S, int H, float p_drop,
                     const uint64_t* rng, uint32_t stream_id, cudaStream_t stream);
int attn_generic_bwd(const void* qkv, const void* ctx, const void* dctx, const float* lse2, void* dqkv,
                     const int32_t* lens, int B, int S, int H, float p_drop, const uint64_t* rng, uint32_t stream_id,
                     cudaStream_t stream);

constexpr int AT_S = 128;        // tile rows (max sequence length)
constexpr int AT_D = 64;         // head dim
constexpr int TILE_BYTES = AT_S * AT_D * 2;  // 16 KB: [128][64] bf16, 128-byte rows, SW128
constexpr float LOG2E = 1.4426950408889634f;

// Attention-probability dropout draws ONE 32-bit hash per PAIR of adjacent keys and compares its two 16-bit halves
// with the top 16 bits of the threshold (drop probability quantised to 1/65536: 0.1 -> 0.09999): the softmax loops
// are ALU-bound and the hash was a fifth of their instructions.  Forward and backward use the same pairing.
// j = position inside the (even-aligned) run of keys starting at first_idx; h carries the pair's hash from the even
// element to the odd one (the loops are fully unrolled, so this is straight-line code)
__device__ __forceinline__ bool attn_keep(uint64_t seed, uint32_t stream, uint32_t first_idx, int j, uint32_t th16,
                                          uint32_t& h) {
    if ((j & 1) == 0) h = rng_u32(seed, stream, (first_idx >> 1) + (j >> 1));
    return ((h >> ((j & 1) * 16)) & 0xFFFFu) >= th16;
}

// A 32-key chunk of a row is FULL (all keys valid), PARTIAL or EMPTY with respect to the sample's length.  The class is
// the same for a whole warp (the length is per sample, the chunk per warp), so the softmax loops are specialised on it:
// no per-element length compare for full chunks, no exp / hash / multiply at all for empty ones.  A typical flow
// sentence (84 tokens of 128) has 2 full, 1 partial and 1 empty chunk per row.
enum { CHUNK_FULL = 0, CHUNK_PARTIAL = 1, CHUNK_EMPTY = 2 };
__device__ __forceinline__ int chunk_mode(int first_key, int len) {
    return first_key + 32 <= len ? CHUNK_FULL : (first_key >= len ? CHUNK_EMPTY : CHUNK_PARTIAL);
}

template <int MODE>
__device__ __forceinline__ float fwd_chunk_max(const uint32_t (&r)[32], int first_key, int len, float mx) {
    if constexpr (MODE == CHUNK_EMPTY) return mx;
#pragma unroll
    for (int j = 0; j < 32; ++j)
        if (MODE == CHUNK_FULL || first_key + j < len) mx = fmaxf(mx, __uint_as_float(r[j]));
    return mx;
}

// probabilities (dropped, not yet normalised) of one chunk into pv; returns the chunk's contribution to the row sum
template <int MODE>
__device__ __forceinline__ float fwd_chunk_exp(const uint32_t (&r)[32], int first_key, int len, float sc2, float m2,
                                               uint64_t seed, uint32_t stream_id, uint32_t drop_first, uint32_t thresh,
                                               uint32_t th16, float keep_scale, float (&pv)[32]) {
    float sum = 0.f;
    if constexpr (MODE == CHUNK_EMPTY) {
#pragma unroll
        for (int j = 0; j < 32; ++j) pv[j] = 0.f;
        return 0.f;
    }
    uint32_t hpair = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        float e = (MODE == CHUNK_FULL || first_key + j < len) ? fast_exp2(__uint_as_float(r[j]) * sc2 - m2) : 0.f;
        sum += e;
        if (thresh) e = attn_keep(seed, stream_id, drop_first, j, th16, hpair) ? e * keep_scale : 0.f;
        pv[j] = e;
    }
    return sum;
}

// backward: recomputed probabilities (dropped) into pd and dS = P o (dP_dropped - delta) * scale into ds
template <int MODE>
__device__ __forceinline__ void bwd_chunk(const uint32_t (&rs)[32], const uint32_t (&rp)[32], int first_key, int len,
                                          bool row_ok, float sc2, float l2, float delta, float scale, uint64_t seed,
                                          uint32_t stream_id, uint32_t drop_first, uint32_t thresh, uint32_t th16,
                                          float keep_scale, float (&pd)[32], float (&ds)[32]) {
    if constexpr (MODE == CHUNK_EMPTY) {
#pragma unroll
        for (int j = 0; j < 32; ++j) { pd[j] = 0.f; ds[j] = 0.f; }
        return;
    }
    uint32_t hpair = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const bool valid = (MODE == CHUNK_FULL || first_key + j < len) && row_ok;
        const float pr = valid ? fast_exp2(__uint_as_float(rs[j]) * sc2 - l2) : 0.f;
        float dp = __uint_as_float(rp[j]);
        float pk = pr;
        if (thresh) {
            const bool keep = attn_keep(seed, stream_id, drop_first, j, th16, hpair);
            pk = keep ? pr * keep_scale : 0.f;
            dp = keep ? dp * keep_scale : 0.f;
        }
        pd[j] = pk;
        ds[j] = valid ? pr * (dp - delta) * scale : 0.f;
    }
}

__device__ __forceinline__ uint64_t attn_seed(const uint64_t* rng) {
    return rng ? rng[0] + rng[1] * 0xD1B54A32D192ED03ULL : 0ULL;
}

// write 8 consecutive bf16 (keys j0..j0+7 of row `row`) into a K-major SW128 [128][128] tile made of two
// [128][64] slabs
__device__ __forceinline__ void st_tile8(uint8_t* tile, int row, int j0, const float* v) {
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
    uint8_t* p = tile + (j0 >> 6) * TILE_BYTES + sw128_offset(row, (j0 & 63) >> 3);
    *reinterpret_cast<uint4*>(p) = o;
}

// ------------------------------------------------------------------------------------------------
// forward.  256 threads: thread (row = tid & 127, hsel = tid >> 7) owns query `row` and keys [64*hsel, 64*hsel+64).
// Two threads per row double the warps per scheduler: the softmax phase is a long dependent ALU chain per thread and
// was latency-bound with one warp per scheduler.
// ------------------------------------------------------------------------------------------------
constexpr int AT_THREADS = 256;

__global__ void __launch_bounds__(AT_THREADS, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, __nv_bfloat16* __restrict__ ctx,
                float* __restrict__ lse2_out, const int32_t* __restrict__ lens, int S, int H, float scale,
                uint32_t thresh, float keep_scale, const uint64_t* __restrict__ rng, uint32_t stream_id) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bar_load, bar_mma1, bar_mma2;
    __shared__ uint32_t tmem_base_smem;
    __shared__ float s_max[2][AT_S], s_sum[2][AT_S];

    const int bh = blockIdx.x, b = bh / H, h = bh % H;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int row = tid & (AT_S - 1), hsel = tid >> 7;
    uint8_t* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw));
    uint8_t* sQ = smem; uint8_t* sK = smem + TILE_BYTES; uint8_t* sV = smem + 2 * TILE_BYTES;
    uint8_t* sP = smem + 3 * TILE_BYTES;  // 2 slabs: keys 0..63 | 64..127
    const int HD = H * AT_D;

    if (tid == 0) {
        tma_prefetch_desc(&tm_qkv);
        mbar_init(smem_u32(&bar_load), 1);
        mbar_init(smem_u32(&bar_mma1), 1);
        mbar_init(smem_u32(&bar_mma2), 1);
        fence_barrier_init();
    }
    if (warp == 0) { tmem_alloc(smem_u32(&tmem_base_smem), 256); tmem_relinquish(); }
    pdl_trigger();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_smem;
    pdl_wait();  // prologue above overlaps the previous kernel's tail
    const uint32_t tS = tmem, tO = tmem + 128;

    if (tid == 0) {
        const uint32_t bar = smem_u32(&bar_load);
        mbar_arrive_expect_tx(bar, 3 * TILE_BYTES);
        tma_load_3d(smem_u32(sQ), &tm_qkv, bar, h * AT_D, 0, b);
        tma_load_3d(smem_u32(sK), &tm_qkv, bar, HD + h * AT_D, 0, b);
        tma_load_3d(smem_u32(sV), &tm_qkv, bar, 2 * HD + h * AT_D, 0, b);
    }
    mbar_wait(smem_u32(&bar_load), 0);
    if (tid == 0) {
        tc_fence_after();
        constexpr uint32_t idesc = make_idesc_bf16(128, 128, kMajorK, kMajorK);
#pragma unroll
        for (int k = 0; k < AT_D / 16; ++k)
            umma_bf16_ss(tS, make_smem_desc_sw128(smem_u32(sQ) + k * 32, 16, 1024),
                         make_smem_desc_sw128(smem_u32(sK) + k * 32, 16, 1024), idesc, k != 0);
        umma_commit(smem_u32(&bar_mma1));
    }
    const int len = min(lens ? lens[b] : S, S);
    const uint64_t seed = attn_seed(rng);
    const uint32_t th16 = thresh >> 16;
    const float sc2 = scale * LOG2E;
    mbar_wait(smem_u32(&bar_mma1), 0);
    tc_fence_after();

    const uint32_t t_lane = (static_cast<uint32_t>(warp & 3) * 32u) << 16;
    const int key0 = hsel * 64;
    // the two 32-key chunks of this thread's half stay in registers for both passes
    uint32_t r0[32], r1[32];
    tmem_ld_32x32b_x32(tS + t_lane + key0, r0);
    tmem_ld_32x32b_x32(tS + t_lane + key0 + 32, r1);
    tmem_ld_wait();
    // Query rows at or past the sample's length are padding: nothing downstream reads their context vector (they are
    // masked as keys and only the CLS row reaches the head), so a warp whose 32 rows are all padding skips the softmax
    // and writes zeros.
    const bool dead_rows = (warp & 3) * 32 >= len;                                 // warp-uniform
    const int mode0 = dead_rows ? CHUNK_EMPTY : chunk_mode(key0, len);
    const int mode1 = dead_rows ? CHUNK_EMPTY : chunk_mode(key0 + 32, len);
    float mx = -INFINITY;
    if (mode0 == CHUNK_FULL) mx = fwd_chunk_max<CHUNK_FULL>(r0, key0, len, mx);
    else if (mode0 == CHUNK_PARTIAL) mx = fwd_chunk_max<CHUNK_PARTIAL>(r0, key0, len, mx);
    if (mode1 == CHUNK_FULL) mx = fwd_chunk_max<CHUNK_FULL>(r1, key0 + 32, len, mx);
    else if (mode1 == CHUNK_PARTIAL) mx = fwd_chunk_max<CHUNK_PARTIAL>(r1, key0 + 32, len, mx);
    s_max[hsel][row] = mx;
    __syncthreads();
    const float m2 = fmaxf(s_max[0][row], s_max[1][row]) * sc2;  // key 0 is always valid -> finite
    float sum = 0.f;
    const uint32_t drop_base = (static_cast<uint32_t>(bh) * AT_S + row) * AT_S;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        float pv[32];
        const int fk = key0 + c * 32, mode = c == 0 ? mode0 : mode1;
        const uint32_t (&r)[32] = c == 0 ? r0 : r1;
        if (mode == CHUNK_FULL)
            sum += fwd_chunk_exp<CHUNK_FULL>(r, fk, len, sc2, m2, seed, stream_id, drop_base + fk, thresh, th16, keep_scale, pv);
        else if (mode == CHUNK_PARTIAL)
            sum += fwd_chunk_exp<CHUNK_PARTIAL>(r, fk, len, sc2, m2, seed, stream_id, drop_base + fk, thresh, th16, keep_scale, pv);
        else
            sum += fwd_chunk_exp<CHUNK_EMPTY>(r, fk, len, sc2, m2, seed, stream_id, drop_base + fk, thresh, th16, keep_scale, pv);
#pragma unroll
        for (int j = 0; j < 32; j += 8) st_tile8(sP, row, fk + j, pv + j);
    }
    s_sum[hsel][row] = sum;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
        tc_fence_after();
        constexpr uint32_t idesc = make_idesc_bf16(128, 64, kMajorK, kMajorMN);
#pragma unroll
        for (int k = 0; k < AT_S / 16; ++k)
            umma_bf16_ss(tO, make_smem_desc_sw128(smem_u32(sP) + (k >> 2) * TILE_BYTES + (k & 3) * 32, 16, 1024),
                         make_smem_desc_sw128(smem_u32(sV) + k * 2048, TILE_BYTES, 1024), idesc, k != 0);
        umma_commit(smem_u32(&bar_mma2));
    }
    const float total = s_sum[0][row] + s_sum[1][row];
    if (hsel == 0 && lse2_out && row < S) lse2_out[static_cast<size_t>(bh) * AT_S + row] = m2 + log2f(total);
    const float inv = dead_rows ? 0.f : 1.0f / total;
    mbar_wait(smem_u32(&bar_mma2), 0);
    tc_fence_after();
    {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tO + t_lane + hsel * 32, r);  // this thread stores 32 of the 64 head-dim columns
        tmem_ld_wait();
        if (row < S) {
            __nv_bfloat16* dst = ctx + (static_cast<size_t>(b) * S + row) * HD + h * AT_D + hsel * 32;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
                uint4 o;
                o.x = pack_bf16x2(__uint_as_float(r[j]) * inv, __uint_as_float(r[j + 1]) * inv);
                o.y = pack_bf16x2(__uint_as_float(r[j + 2]) * inv, __uint_as_float(r[j + 3]) * inv);
                o.z = pack_bf16x2(__uint_as_float(r[j + 4]) * inv, __uint_as_float(r[j + 5]) * inv);
                o.w = pack_bf16x2(__uint_as_float(r[j + 6]) * inv, __uint_as_float(r[j + 7]) * inv);
                *reinterpret_cast<uint4*>(dst + j) = o;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 256); }
}

// ------------------------------------------------------------------------------------------------
// backward (same thread mapping as forward)
// ------------------------------------------------------------------------------------------------
// Two CTAs are resident per SM (B*H = 192 CTAs run as ONE wave on 148 SMs): 112 KB of tiles + a few barrier words of
// dynamic shared memory and 256 TMEM columns each.  To get there, P's first slab reuses V's tile (V is dead once
// dP = dO V^T has been issued) and dV / dK / dQ reuse the TMEM columns of S / dP (dead once every thread has pulled
// its chunks into registers).  No static shared memory: the dynamic buffer is 1024-byte aligned by declaration.
extern __shared__ __align__(1024) uint8_t smem_attn_bwd[];

__global__ void __launch_bounds__(AT_THREADS, 2)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_dctx,
                const __nv_bfloat16* __restrict__ ctx, const __nv_bfloat16* __restrict__ dctx,
                const float* __restrict__ lse2, __nv_bfloat16* __restrict__ dqkv, const int32_t* __restrict__ lens,
                int S, int H, float scale, uint32_t thresh, float keep_scale, const uint64_t* __restrict__ rng,
                uint32_t stream_id, float* __restrict__ dbias) {
    uint8_t* smem = smem_attn_bwd;
    if ((smem_u32(smem) & 1023u) != 0u) __trap();  // the swizzled tiles need the declared alignment
    const int bh = blockIdx.x, b = bh / H, h = bh % H;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int row = tid & (AT_S - 1), hsel = tid >> 7;
    uint8_t* sQ = smem; uint8_t* sK = smem + TILE_BYTES; uint8_t* sdO = smem + 2 * TILE_BYTES;
    uint8_t* sV = smem + 3 * TILE_BYTES;
    uint8_t* sP = sV;                      // 2 slabs: P (dropped, normalised); slab 0 overwrites V
    uint8_t* sdS = smem + 5 * TILE_BYTES;  // 2 slabs: dS (scaled)
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 7 * TILE_BYTES);
    uint64_t& bar_load = bars[0];
    uint64_t& bar_mma1 = bars[1];
    uint64_t& bar_mma2 = bars[2];
    uint32_t& tmem_base_smem = *reinterpret_cast<uint32_t*>(bars + 3);
    const int HD = H * AT_D;

    if (tid == 0) {
        tma_prefetch_desc(&tm_qkv);
        tma_prefetch_desc(&tm_dctx);
        mbar_init(smem_u32(&bar_load), 1);
        mbar_init(smem_u32(&bar_mma1), 1);
        mbar_init(smem_u32(&bar_mma2), 1);
        fence_barrier_init();
    }
    if (warp == 0) { tmem_alloc(smem_u32(&tmem_base_smem), 256); tmem_relinquish(); }
    pdl_trigger();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_smem;
    pdl_wait();  // prologue above overlaps the previous kernel's tail
    const uint32_t tS = tmem, tdP = tmem + 128;                       // phase 1
    const uint32_t tdV = tmem, tdK = tmem + 64, tdQ = tmem + 128;     // phase 2 (reuses the columns of S / dP)

    if (tid == 0) {
        const uint32_t bar = smem_u32(&bar_load);
        mbar_arrive_expect_tx(bar, 4 * TILE_BYTES);
        tma_load_3d(smem_u32(sQ), &tm_qkv, bar, h * AT_D, 0, b);
        tma_load_3d(smem_u32(sK), &tm_qkv, bar, HD + h * AT_D, 0, b);
        tma_load_3d(smem_u32(sV), &tm_qkv, bar, 2 * HD + h * AT_D, 0, b);
        tma_load_3d(smem_u32(sdO), &tm_dctx, bar, h * AT_D, 0, b);
    }
    // delta = rowsum(dO * O), from global while the TMA loads fly (both threads of a row compute it)
    float delta = 0.f;
    if (row < S) {
        const uint4* po = reinterpret_cast<const uint4*>(ctx + (static_cast<size_t>(b) * S + row) * HD + h * AT_D);
        const uint4* pd = reinterpret_cast<const uint4*>(dctx + (static_cast<size_t>(b) * S + row) * HD + h * AT_D);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint4 o = __ldg(po + i), d = __ldg(pd + i);
            const float2 o0 = unpack_bf16x2(o.x), o1 = unpack_bf16x2(o.y), o2 = unpack_bf16x2(o.z), o3 = unpack_bf16x2(o.w);
            const float2 d0 = unpack_bf16x2(d.x), d1 = unpack_bf16x2(d.y), d2 = unpack_bf16x2(d.z), d3 = unpack_bf16x2(d.w);
            delta += o0.x * d0.x + o0.y * d0.y + o1.x * d1.x + o1.y * d1.y + o2.x * d2.x + o2.y * d2.y +
                     o3.x * d3.x + o3.y * d3.y;
        }
    }
    mbar_wait(smem_u32(&bar_load), 0);
    if (tid == 0) {
        tc_fence_after();
        constexpr uint32_t idesc = make_idesc_bf16(128, 128, kMajorK, kMajorK);
#pragma unroll
        for (int k = 0; k < AT_D / 16; ++k)  // S = Q K^T
            umma_bf16_ss(tS, make_smem_desc_sw128(smem_u32(sQ) + k * 32, 16, 1024),
                         make_smem_desc_sw128(smem_u32(sK) + k * 32, 16, 1024), idesc, k != 0);
#pragma unroll
        for (int k = 0; k < AT_D / 16; ++k)  // dP = dO V^T
            umma_bf16_ss(tdP, make_smem_desc_sw128(smem_u32(sdO) + k * 32, 16, 1024),
                         make_smem_desc_sw128(smem_u32(sV) + k * 32, 16, 1024), idesc, k != 0);
        umma_commit(smem_u32(&bar_mma1));
    }
    const int len = min(lens ? lens[b] : S, S);
    const uint64_t seed = attn_seed(rng);
    const uint32_t th16 = thresh >> 16;
    const float sc2 = scale * LOG2E;
    const float l2 = (row < S) ? lse2[static_cast<size_t>(bh) * AT_S + row] : 0.f;
    const uint32_t t_lane = (static_cast<uint32_t>(warp & 3) * 32u) << 16;
    const uint32_t drop_base = (static_cast<uint32_t>(bh) * AT_S + row) * AT_S;
    const int key0 = hsel * 64;
    mbar_wait(smem_u32(&bar_mma1), 0);
    tc_fence_after();
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
        // padded query rows have dO = 0 and receive dQ = 0: their P / dS rows are simply zero (see the forward kernel)
        const int fk = key0 + c * 32;
        const int mode = (warp & 3) * 32 >= len ? CHUNK_EMPTY : chunk_mode(fk, len);  // warp-uniform
        float pd[32], ds[32];
        if (mode == CHUNK_EMPTY) {  // nothing to read or compute: P and dS of padded keys are zero
#pragma unroll
            for (int j = 0; j < 32; ++j) { pd[j] = 0.f; ds[j] = 0.f; }
        } else {
            uint32_t rs[32], rp[32];
            tmem_ld_32x32b_x32(tS + t_lane + fk, rs);
            tmem_ld_32x32b_x32(tdP + t_lane + fk, rp);
            tmem_ld_wait();
            if (mode == CHUNK_FULL)
                bwd_chunk<CHUNK_FULL>(rs, rp, fk, len, row < S, sc2, l2, delta, scale, seed, stream_id, drop_base + fk,
                                      thresh, th16, keep_scale, pd, ds);
            else
                bwd_chunk<CHUNK_PARTIAL>(rs, rp, fk, len, row < S, sc2, l2, delta, scale, seed, stream_id,
                                         drop_base + fk, thresh, th16, keep_scale, pd, ds);
        }
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
            st_tile8(sP, row, key0 + c * 32 + j, pd + j);
            st_tile8(sdS, row, key0 + c * 32 + j, ds + j);
        }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
        tc_fence_after();
        constexpr uint32_t idesc_t = make_idesc_bf16(128, 64, kMajorMN, kMajorMN);
        constexpr uint32_t idesc_n = make_idesc_bf16(128, 64, kMajorK, kMajorMN);
#pragma unroll
        for (int k = 0; k < AT_S / 16; ++k)  // dV[key, d] = sum_q P[q, key] dO[q, d]
            umma_bf16_ss(tdV, make_smem_desc_sw128(smem_u32(sP) + k * 2048, TILE_BYTES, 1024),
                         make_smem_desc_sw128(smem_u32(sdO) + k * 2048, TILE_BYTES, 1024), idesc_t, k != 0);
#pragma unroll
        for (int k = 0; k < AT_S / 16; ++k)  // dK[key, d] = sum_q dS[q, key] Q[q, d]
            umma_bf16_ss(tdK, make_smem_desc_sw128(smem_u32(sdS) + k * 2048, TILE_BYTES, 1024),
                         make_smem_desc_sw128(smem_u32(sQ) + k * 2048, TILE_BYTES, 1024), idesc_t, k != 0);
#pragma unroll
        for (int k = 0; k < AT_S / 16; ++k)  // dQ[q, d] = sum_key dS[q, key] K[key, d]
            umma_bf16_ss(tdQ, make_smem_desc_sw128(smem_u32(sdS) + (k >> 2) * TILE_BYTES + (k & 3) * 32, 16, 1024),
                         make_smem_desc_sw128(smem_u32(sK) + k * 2048, TILE_BYTES, 1024), idesc_n, k != 0);
        umma_commit(smem_u32(&bar_mma2));
    }
    mbar_wait(smem_u32(&bar_mma2), 0);
    tc_fence_after();
    // rows of dQ are queries, rows of dK/dV are keys: thread (row, hsel) writes 32 head-dim columns of token `row`
#pragma unroll 1
    for (int which = 0; which < 3; ++which) {
        const uint32_t tsrc = which == 0 ? tdQ : (which == 1 ? tdK : tdV);
        uint32_t r[32];
        tmem_ld_32x32b_x32(tsrc + t_lane + hsel * 32, r);
        tmem_ld_wait();
        if (row < S) {
            __nv_bfloat16* dst =
                dqkv + (static_cast<size_t>(b) * S + row) * (3 * HD) + which * HD + h * AT_D + hsel * 32;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
                uint4 o;
                o.x = pack_bf16x2(__uint_as_float(r[j]), __uint_as_float(r[j + 1]));
                o.y = pack_bf16x2(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
                o.z = pack_bf16x2(__uint_as_float(r[j + 4]), __uint_as_float(r[j + 5]));
                o.w = pack_bf16x2(__uint_as_float(r[j + 6]), __uint_as_float(r[j + 7]));
                *reinterpret_cast<uint4*>(dst + j) = o;
            }
        }
        if (dbias) {
            // gradient of the fused QKV bias = column sums of dqkv: this warp's 32 tokens x 32 columns are reduced
            // with a shuffle butterfly and leave as one fp32 atomic per column (no separate pass over dqkv)
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = row < S ? __uint_as_float(r[j]) : 0.f;
            const uint32_t lane = static_cast<uint32_t>(tid) & 31u;
            warp_colsum<32>(v, lane);
            atomicAdd(dbias + which * HD + h * AT_D + hsel * 32 + lane, v[0]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 256); }
}

static inline uint32_t drop_thresh_attn(float p) {
    if (p <= 0.f) return 0u;
    double t = static_cast<double>(p) * 4294967296.0;
    if (t > 4294967295.0) t = 4294967295.0;
    return static_cast<uint32_t>(t);
}

}  // namespace dcad

using namespace dcad;

// qkv: [B*S, 3*H*64] bf16; ctx: [B*S, H*64] bf16; lse2: [B*H*max(S,128)] fp32 (log2-domain logsumexp, for backward).
// S <= 128: the tcgen05 kernels of this file; 128 < S <= 512: the CUDA-core fallback of attention_generic.cu.
extern "C" int dcad_attn_fwd(const void* qkv, void* ctx, float* lse2, const int32_t* lens, int B, int S, int H,
                             float p_drop, const uint64_t* rng, uint32_t stream_id, void* stream) {
    if (B <= 0) return 0;
    if (S <= 0 || S > 512 || H <= 0) return DCAD_ERR_BAD_SHAPE;
    if (p_drop > 0.f && rng == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (S > AT_S)
        return attn_generic_fwd(qkv, ctx, lse2, lens, B, S, H, p_drop, rng, stream_id, reinterpret_cast<cudaStream_t>(stream));
    const int HD = H * AT_D;
    CUtensorMap tm;
    int rc = make_tensor_map_bf16_3d(&tm, qkv, 3 * HD, S, B, 3 * HD, static_cast<uint64_t>(S) * 3 * HD, AT_D, AT_S);
    if (rc) return rc;
    constexpr int SMEM = 5 * TILE_BYTES + 1024;
    static DeviceOnce once;
    DCAD_CUDA_TRY(ensure_dynamic_smem(attn_fwd_kernel, SMEM, once));
    const float ks = p_drop > 0.f ? 1.0f / (1.0f - p_drop) : 1.0f;
    return (int)launch_pdl(attn_fwd_kernel, dim3(B * H), dim3(AT_THREADS), SMEM, reinterpret_cast<cudaStream_t>(stream),
                           tm, reinterpret_cast<__nv_bfloat16*>(ctx), lse2, lens, S, H, 0.125f,
                           drop_thresh_attn(p_drop), ks, rng, stream_id);
}

// dctx: [B*S, H*64] bf16 gradient of ctx; dqkv: [B*S, 3*H*64] bf16 output
extern "C" int dcad_attn_bwd(const void* qkv, const void* ctx, const void* dctx, const float* lse2, void* dqkv,
                             const int32_t* lens, int B, int S, int H, float p_drop, const uint64_t* rng,
                             uint32_t stream_id, float* dbias, void* stream) {
    if (B <= 0) return 0;
    if (S <= 0 || S > 512 || H <= 0) return DCAD_ERR_BAD_SHAPE;
    if (p_drop > 0.f && rng == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (S > AT_S) {
        int rc = attn_generic_bwd(qkv, ctx, dctx, lse2, dqkv, lens, B, S, H, p_drop, rng, stream_id,
                                  reinterpret_cast<cudaStream_t>(stream));
        if (rc || dbias == nullptr) return rc;
        return dcad_colsum_bf16(dqkv, dbias, B * S, 3 * H * AT_D, 3 * H * AT_D, stream);  // fused-bias contract: += colsum
    }
    const int HD = H * AT_D;
    CUtensorMap tq, td;
    int rc = make_tensor_map_bf16_3d(&tq, qkv, 3 * HD, S, B, 3 * HD, static_cast<uint64_t>(S) * 3 * HD, AT_D, AT_S);
    if (rc) return rc;
    rc = make_tensor_map_bf16_3d(&td, dctx, HD, S, B, HD, static_cast<uint64_t>(S) * HD, AT_D, AT_S);
    if (rc) return rc;
    constexpr int SMEM = 7 * TILE_BYTES + 64;  // tiles + 3 mbarriers + the TMEM base word
    static DeviceOnce once;
    DCAD_CUDA_TRY(ensure_dynamic_smem(attn_bwd_kernel, SMEM, once, /*max_carveout=*/true));
    const float ks = p_drop > 0.f ? 1.0f / (1.0f - p_drop) : 1.0f;
    return (int)launch_pdl(attn_bwd_kernel, dim3(B * H), dim3(AT_THREADS), SMEM, reinterpret_cast<cudaStream_t>(stream),
                           tq, td, reinterpret_cast<const __nv_bfloat16*>(ctx),
                           reinterpret_cast<const __nv_bfloat16*>(dctx), lse2, reinterpret_cast<__nv_bfloat16*>(dqkv),
                           lens, S, H, 0.125f, drop_thresh_attn(p_drop), ks, rng, stream_id, dbias);
}
