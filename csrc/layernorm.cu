// Fused (dropout +) residual-add + LayerNorm, forward and backward, and the embedding variant
// (word gather + position add + LayerNorm + dropout, and its scatter-add backward).
//
// Replaces, per DistilBERT block, HF's `sa_layer_norm(sa_output + x)` / `output_layer_norm(ffn_output + sa_output)`
// (transformers modeling_distilbert.py:257,261), the FFN dropout (:227) and the Embeddings module (:96-122) that the
// reference reaches through client1.py:61; backward replaces their autograd (client1.py:109).
//
// One warp owns one row (cols = NV*256, 8 bf16 per 16-byte access), statistics by warp shuffles.  The backward needs
// x_hat: the forward can save it (bf16, `xhat_out`; its rounding error is 2^-9 |x_hat| whatever gamma and beta are),
// or the backward rebuilds it from the saved LN *output*, x_hat = (y - beta) / gamma, which costs no extra tensor but
// amplifies the bf16 rounding of y by |beta| / |gamma| -- fine for gamma ~ 1 (random init), poor for channels of a
// pretrained encoder whose gamma is small.  The training engine saves x_hat.
#include "common.cuh"
#include "../include/dcad_b200.h"

namespace dcad {

constexpr int LN_WARPS = 8;

__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}
__device__ __forceinline__ void load8f(const float* p, float (&v)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p));
    const float4 b = __ldg(reinterpret_cast<const float4*>(p + 4));
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
// sum of `splits` fp32 partial tiles left by a split-K GEMM (slice s at part + s * slice_elems), plus an optional bias
struct PartialIn {
    const float* part;   // null: the operand is the bf16 tensor instead
    const float* bias;   // [cols] or null
    int64_t slice_elems;
    int splits;
};
__device__ __forceinline__ void load8_partials(const PartialIn& in, size_t off, int col, float (&v)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(in.part + off);
    const float4 b = *reinterpret_cast<const float4*>(in.part + off + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    for (int s = 1; s < in.splits; ++s) {
        const float4 c = *reinterpret_cast<const float4*>(in.part + s * in.slice_elems + off);
        const float4 d = *reinterpret_cast<const float4*>(in.part + s * in.slice_elems + off + 4);
        v[0] += c.x; v[1] += c.y; v[2] += c.z; v[3] += c.w; v[4] += d.x; v[5] += d.y; v[6] += d.z; v[7] += d.w;
    }
    if (in.bias) {
        const float4 c = __ldg(reinterpret_cast<const float4*>(in.bias + col));
        const float4 d = __ldg(reinterpret_cast<const float4*>(in.bias + col + 4));
        v[0] += c.x; v[1] += c.y; v[2] += c.z; v[3] += c.w; v[4] += d.x; v[5] += d.y; v[6] += d.z; v[7] += d.w;
    }
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&v)[8]) {
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = o;
}
__device__ __forceinline__ uint64_t rng_seed(const uint64_t* rng) {
    return rng ? rng[0] + rng[1] * 0xD1B54A32D192ED03ULL : 0ULL;
}
__device__ __forceinline__ float safe_gamma(float g) {
    return fabsf(g) < 1e-6f ? (g < 0.f ? -1e-6f : 1e-6f) : g;
}

// ------------------------------------------------------------------------------------------------
// forward: out = LN(dropout(a) + res)
// ------------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32)
add_ln_fwd_kernel(const __nv_bfloat16* __restrict__ a, const PartialIn a_part, const __nv_bfloat16* __restrict__ res,
                  const float* __restrict__ gamma, const float* __restrict__ beta, __nv_bfloat16* __restrict__ out,
                  float* __restrict__ rstd_out, __nv_bfloat16* __restrict__ xhat_out, int rows, float eps,
                  uint32_t thresh, float keep_scale, const uint64_t* __restrict__ rng, uint32_t stream_id) {
    pdl_trigger();
    pdl_wait();
    constexpr int COLS = NV * 256;
    const int row = blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
    if (row >= rows) return;
    const uint32_t lane = lane_id();
    const uint64_t seed = rng_seed(rng);
    float x[NV][8];
    float sum = 0.f;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int col = (v * 32 + lane) * 8;
        if (a_part.part) load8_partials(a_part, static_cast<size_t>(row) * COLS + col, col, x[v]);
        else load8(a + static_cast<size_t>(row) * COLS + col, x[v]);
        if (thresh) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                x[v][j] = dropout_keep(seed, stream_id, row * COLS + col + j, thresh) ? x[v][j] * keep_scale : 0.f;
        }
        if (res) {
            float r[8];
            load8(res + static_cast<size_t>(row) * COLS + col, r);
#pragma unroll
            for (int j = 0; j < 8; ++j) x[v][j] += r[j];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) sum += x[v][j];
    }
    const float mean = warp_sum(sum) * (1.0f / COLS);
    float var = 0.f;
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float d = x[v][j] - mean; var += d * d; }
    const float rstd = rsqrtf(warp_sum(var) * (1.0f / COLS) + eps);
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int col = (v * 32 + lane) * 8;
        float g[8], b[8], y[8], xh[8];
        load8f(gamma + col, g);
        load8f(beta + col, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) { xh[j] = (x[v][j] - mean) * rstd; y[j] = xh[j] * g[j] + b[j]; }
        store8(out + static_cast<size_t>(row) * COLS + col, y);
        if (xhat_out) store8(xhat_out + static_cast<size_t>(row) * COLS + col, xh);
    }
    if (lane == 0 && rstd_out) rstd_out[row] = rstd;
}

// ------------------------------------------------------------------------------------------------
// backward.  dz = grad wrt (dropout(a)+res) (== grad wrt res); da = dz * mask * keep_scale.
// dgamma/dbeta/dbias are accumulated with atomics into zero-initialised fp32 buffers.
// ------------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32)
add_ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy1, const PartialIn dy1_part,
                  const __nv_bfloat16* __restrict__ dy2,
                  const __nv_bfloat16* __restrict__ y, const float* __restrict__ gamma,
                  const float* __restrict__ beta, const float* __restrict__ rstd, __nv_bfloat16* __restrict__ dz,
                  __nv_bfloat16* __restrict__ da, float* __restrict__ dgamma, float* __restrict__ dbeta,
                  float* __restrict__ dbias, int rows, uint32_t thresh, float keep_scale,
                  const uint64_t* __restrict__ rng, uint32_t stream_id) {
    pdl_trigger();
    pdl_wait();
    constexpr int COLS = NV * 256;
    // one slot per warp: [warp][3][COLS] partial column sums (dgamma, dbeta, dbias); one row per warp
    extern __shared__ float s_part[];
    const uint32_t warp = threadIdx.x >> 5, lane = lane_id();
    const uint64_t seed = rng_seed(rng);
    const int row = blockIdx.x * LN_WARPS + warp;
    float* slot = s_part + warp * 3 * COLS;
    if (row < rows) {
        float g[NV][8], xh[NV][8], d[NV][8];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int col = (v * 32 + lane) * 8;
            const size_t off = static_cast<size_t>(row) * COLS + col;
            float yy[8], gm[8], bt[8];
            if (dy1_part.part) load8_partials(dy1_part, off, col, d[v]);
            else load8(dy1 + off, d[v]);
            if (dy2) {
                float d2[8];
                load8(dy2 + off, d2);
#pragma unroll
                for (int j = 0; j < 8; ++j) d[v][j] += d2[j];
            }
            load8(y + off, yy);
            load8f(gamma + col, gm);
            if (beta) load8f(beta + col, bt);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                // beta == nullptr: `y` holds the x_hat the forward saved; else x_hat is rebuilt from the LN output
                xh[v][j] = beta ? (yy[j] - bt[j]) / safe_gamma(gm[j]) : yy[j];
                g[v][j] = d[v][j] * gm[j];
                c1 += g[v][j];
                c2 += g[v][j] * xh[v][j];
            }
        }
        c1 = warp_sum(c1) * (1.0f / COLS);
        c2 = warp_sum(c2) * (1.0f / COLS);
        const float rs = rstd[row];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int col = (v * 32 + lane) * 8;
            const size_t off = static_cast<size_t>(row) * COLS + col;
            float o[8], pg[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) { o[j] = rs * (g[v][j] - c1 - xh[v][j] * c2); pg[j] = d[v][j] * xh[v][j]; }
            store8(dz + off, o);
            if (da) {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    o[j] = dropout_keep(seed, stream_id, row * COLS + col + j, thresh) ? o[j] * keep_scale : 0.f;
                store8(da + off, o);
            }
            *reinterpret_cast<float4*>(slot + col) = make_float4(pg[0], pg[1], pg[2], pg[3]);
            *reinterpret_cast<float4*>(slot + col + 4) = make_float4(pg[4], pg[5], pg[6], pg[7]);
            *reinterpret_cast<float4*>(slot + COLS + col) = make_float4(d[v][0], d[v][1], d[v][2], d[v][3]);
            *reinterpret_cast<float4*>(slot + COLS + col + 4) = make_float4(d[v][4], d[v][5], d[v][6], d[v][7]);
            *reinterpret_cast<float4*>(slot + 2 * COLS + col) = make_float4(o[0], o[1], o[2], o[3]);
            *reinterpret_cast<float4*>(slot + 2 * COLS + col + 4) = make_float4(o[4], o[5], o[6], o[7]);
        }
    } else {
        for (int c = lane; c < 3 * COLS; c += 32) slot[c] = 0.f;
    }
    __syncthreads();
    for (int c = threadIdx.x * 4; c < 3 * COLS; c += blockDim.x * 4) {  // four columns per thread, one vector RED
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int w = 0; w < LN_WARPS; ++w) {
            const float4 a = *reinterpret_cast<const float4*>(s_part + w * 3 * COLS + c);
            t.x += a.x; t.y += a.y; t.z += a.z; t.w += a.w;
        }
        float* dst = c < COLS ? dgamma : (c < 2 * COLS ? dbeta : dbias);
        if (dst) red_add_v4(dst + (c % COLS), t.x, t.y, t.z, t.w);
    }
}

// ------------------------------------------------------------------------------------------------
// embeddings forward: out = dropout(LN(word[ids] + pos[row % seq]))
// ------------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32)
embed_ln_fwd_kernel(const int32_t* __restrict__ ids, const float* __restrict__ word, const float* __restrict__ pos,
                    const float* __restrict__ gamma, const float* __restrict__ beta, __nv_bfloat16* __restrict__ out,
                    float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int seq, float eps,
                    uint32_t thresh, float keep_scale, const uint64_t* __restrict__ rng, uint32_t stream_id) {
    pdl_trigger();
    pdl_wait();
    constexpr int COLS = NV * 256;
    const int row = blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
    if (row >= rows) return;
    const uint32_t lane = lane_id();
    const uint64_t seed = rng_seed(rng);
    const int id = ids[row];
    const int ps = row % seq;
    float x[NV][8];
    float sum = 0.f;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int col = (v * 32 + lane) * 8;
        float w[8], pp[8];
        load8f(word + static_cast<size_t>(id) * COLS + col, w);
        load8f(pos + static_cast<size_t>(ps) * COLS + col, pp);
#pragma unroll
        for (int j = 0; j < 8; ++j) { x[v][j] = w[j] + pp[j]; sum += x[v][j]; }
    }
    const float mean = warp_sum(sum) * (1.0f / COLS);
    float var = 0.f;
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float d = x[v][j] - mean; var += d * d; }
    const float rstd = rsqrtf(warp_sum(var) * (1.0f / COLS) + eps);
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int col = (v * 32 + lane) * 8;
        float g[8], b[8], yv[8];
        load8f(gamma + col, g);
        load8f(beta + col, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float t = (x[v][j] - mean) * rstd * g[j] + b[j];
            if (thresh) t = dropout_keep(seed, stream_id, row * COLS + col + j, thresh) ? t * keep_scale : 0.f;
            yv[j] = t;
        }
        store8(out + static_cast<size_t>(row) * COLS + col, yv);
    }
    if (lane == 0) {
        if (mean_out) mean_out[row] = mean;
        if (rstd_out) rstd_out[row] = rstd;
    }
}

// embeddings backward: recompute x from the tables, LN backward, scatter-add into word/pos gradients.
// Rows at or past the sample's valid length are padding: nothing downstream depends on them (keys are masked, only
// the CLS row feeds the head), so their incoming gradient is exactly zero and they are skipped -- this also removes
// the atomic pile-up on the [PAD] embedding row.
template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32)
embed_ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy1, const __nv_bfloat16* __restrict__ dy2,
                    const int32_t* __restrict__ ids, const float* __restrict__ word, const float* __restrict__ pos,
                    const float* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ rstd,
                    float* __restrict__ dword, float* __restrict__ dpos, float* __restrict__ dgamma,
                    float* __restrict__ dbeta, const int32_t* __restrict__ lens, int rows, int seq, uint32_t thresh,
                    float keep_scale, const uint64_t* __restrict__ rng, uint32_t stream_id,
                    uint8_t* __restrict__ touched) {
    pdl_trigger();
    pdl_wait();
    constexpr int COLS = NV * 256;
    extern __shared__ float s_part[];  // [warp][2][COLS]
    const uint32_t warp = threadIdx.x >> 5, lane = lane_id();
    const uint64_t seed = rng_seed(rng);
    const int row = blockIdx.x * LN_WARPS + warp;
    float* slot = s_part + warp * 2 * COLS;
    bool active = row < rows;
    if (active && lens) active = (row % seq) < lens[row / seq];
    if (active) {
        const int id = ids[row];
        const int ps = row % seq;
        if (touched && lane == 0) touched[id] = 1;  // this vocabulary row now has optimizer state (see adam.cu)
        const float mu = mean[row], rs = rstd[row];
        float g[NV][8], xh[NV][8], d[NV][8];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int col = (v * 32 + lane) * 8;
            const size_t off = static_cast<size_t>(row) * COLS + col;
            float w[8], pp[8], gm[8];
            load8(dy1 + off, d[v]);
            if (dy2) {
                float d2[8];
                load8(dy2 + off, d2);
#pragma unroll
                for (int j = 0; j < 8; ++j) d[v][j] += d2[j];
            }
            load8f(word + static_cast<size_t>(id) * COLS + col, w);
            load8f(pos + static_cast<size_t>(ps) * COLS + col, pp);
            load8f(gamma + col, gm);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (thresh)
                    d[v][j] = dropout_keep(seed, stream_id, row * COLS + col + j, thresh) ? d[v][j] * keep_scale : 0.f;
                xh[v][j] = (w[j] + pp[j] - mu) * rs;
                g[v][j] = d[v][j] * gm[j];
                c1 += g[v][j];
                c2 += g[v][j] * xh[v][j];
            }
        }
        c1 = warp_sum(c1) * (1.0f / COLS);
        c2 = warp_sum(c2) * (1.0f / COLS);
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int col = (v * 32 + lane) * 8;
            float pg[8], dx[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                dx[j] = rs * (g[v][j] - c1 - xh[v][j] * c2);
                pg[j] = d[v][j] * xh[v][j];
            }
            float* dw = dword + static_cast<size_t>(id) * COLS + col;
            float* dp = dpos + static_cast<size_t>(ps) * COLS + col;
            red_add_v4(dw, dx[0], dx[1], dx[2], dx[3]);
            red_add_v4(dw + 4, dx[4], dx[5], dx[6], dx[7]);
            red_add_v4(dp, dx[0], dx[1], dx[2], dx[3]);
            red_add_v4(dp + 4, dx[4], dx[5], dx[6], dx[7]);
            *reinterpret_cast<float4*>(slot + col) = make_float4(pg[0], pg[1], pg[2], pg[3]);
            *reinterpret_cast<float4*>(slot + col + 4) = make_float4(pg[4], pg[5], pg[6], pg[7]);
            *reinterpret_cast<float4*>(slot + COLS + col) = make_float4(d[v][0], d[v][1], d[v][2], d[v][3]);
            *reinterpret_cast<float4*>(slot + COLS + col + 4) = make_float4(d[v][4], d[v][5], d[v][6], d[v][7]);
        }
    } else {
        for (int c = lane; c < 2 * COLS; c += 32) slot[c] = 0.f;
    }
    __syncthreads();
    for (int c = threadIdx.x * 4; c < 2 * COLS; c += blockDim.x * 4) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int w = 0; w < LN_WARPS; ++w) {
            const float4 a = *reinterpret_cast<const float4*>(s_part + w * 2 * COLS + c);
            t.x += a.x; t.y += a.y; t.z += a.z; t.w += a.w;
        }
        if (t.x != 0.f || t.y != 0.f || t.z != 0.f || t.w != 0.f)
            red_add_v4((c < COLS ? dgamma : dbeta) + (c % COLS), t.x, t.y, t.z, t.w);
    }
}

// column sum of a bf16 [rows][ld] matrix into fp32 (atomic accumulate): bias gradients
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, int rows, int cols, int ld,
                   int rows_per_cta) {
    // block (32 x 8): x-dim covers 32*8 = 256 columns (8 per thread), y-dim strides rows
    pdl_trigger();
    pdl_wait();
    __shared__ float s[8][256];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int col = blockIdx.x * 256 + tx * 8;
    const int r0 = blockIdx.y * rows_per_cta;
    const int r1 = min(rows, r0 + rows_per_cta);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (col < cols) {
        for (int r = r0 + ty; r < r1; r += 8) {
            float v[8];
            load8(x + static_cast<size_t>(r) * ld + col, v);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += v[j];
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) s[ty][tx * 8 + j] = acc[j];
    __syncthreads();
    const int c = threadIdx.x;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += s[w][c];
    if (blockIdx.x * 256 + c < cols) atomicAdd(out + blockIdx.x * 256 + c, t);
}

static inline uint32_t drop_thresh(float p) {
    if (p <= 0.f) return 0u;
    double t = static_cast<double>(p) * 4294967296.0;
    if (t > 4294967295.0) t = 4294967295.0;
    return static_cast<uint32_t>(t);
}

}  // namespace dcad

using namespace dcad;

#define DCAD_NV_DISPATCH(cols, ...)                            \
    switch ((cols) / 256) {                                    \
        case 1: { constexpr int NV = 1; __VA_ARGS__; } break;  \
        case 2: { constexpr int NV = 2; __VA_ARGS__; } break;  \
        case 3: { constexpr int NV = 3; __VA_ARGS__; } break;  \
        case 4: { constexpr int NV = 4; __VA_ARGS__; } break;  \
        default: return DCAD_ERR_BAD_SHAPE;                    \
    }

static int add_ln_fwd_run(const void* a, PartialIn a_part, const void* res, const float* gamma, const float* beta,
                          void* out, float* rstd, int rows, int cols, float eps, float p_drop, const uint64_t* rng,
                          uint32_t stream_id, void* stream, void* xhat_out = nullptr) {
    if (rows <= 0) return 0;
    if (cols % 256 || cols > 1024) return DCAD_ERR_BAD_SHAPE;
    if (p_drop > 0.f && rng == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    const uint32_t th = drop_thresh(p_drop);
    const float ks = p_drop > 0.f ? 1.0f / (1.0f - p_drop) : 1.0f;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int grid = (rows + LN_WARPS - 1) / LN_WARPS;
    DCAD_NV_DISPATCH(cols, return (int)launch_pdl(
        add_ln_fwd_kernel<NV>, dim3(grid), dim3(LN_WARPS * 32), 0, s, reinterpret_cast<const __nv_bfloat16*>(a),
        a_part, reinterpret_cast<const __nv_bfloat16*>(res), gamma, beta, reinterpret_cast<__nv_bfloat16*>(out), rstd,
        reinterpret_cast<__nv_bfloat16*>(xhat_out), rows, eps, th, ks, rng, stream_id));
    return 0;
}

extern "C" int dcad_add_ln_fwd(const void* a, const void* res, const float* gamma, const float* beta, void* out,
                               float* rstd, int rows, int cols, float eps, float p_drop, const uint64_t* rng,
                               uint32_t stream_id, void* stream) {
    return add_ln_fwd_run(a, PartialIn{nullptr, nullptr, 0, 0}, res, gamma, beta, out, rstd, rows, cols, eps, p_drop,
                          rng, stream_id, stream);
}

// Same as dcad_add_ln_fwd, and also stores x_hat = (x - mean) * rstd (bf16, same shape as out) for a backward that does
// not want to rebuild it from the output (dcad_add_ln_bwd with beta == NULL and y = that x_hat).
extern "C" int dcad_add_ln_fwd_x(const void* a, const void* res, const float* gamma, const float* beta, void* out,
                                 float* rstd, void* xhat_out, int rows, int cols, float eps, float p_drop,
                                 const uint64_t* rng, uint32_t stream_id, void* stream) {
    return add_ln_fwd_run(a, PartialIn{nullptr, nullptr, 0, 0}, res, gamma, beta, out, rstd, rows, cols, eps, p_drop,
                          rng, stream_id, stream, xhat_out);
}

extern "C" int dcad_add_ln_fwd_partials_x(const float* a_partials, int splits, int64_t slice_elems,
                                          const float* a_bias, const void* res, const float* gamma, const float* beta,
                                          void* out, float* rstd, void* xhat_out, int rows, int cols, float eps,
                                          float p_drop, const uint64_t* rng, uint32_t stream_id, void* stream) {
    if (a_partials == nullptr || splits < 1 || (slice_elems & 3) ||
        (reinterpret_cast<uintptr_t>(a_partials) & 15u))
        return DCAD_ERR_BAD_ARGUMENT;
    return add_ln_fwd_run(nullptr, PartialIn{a_partials, a_bias, slice_elems, splits}, res, gamma, beta, out, rstd,
                          rows, cols, eps, p_drop, rng, stream_id, stream, xhat_out);
}

extern "C" int dcad_add_ln_fwd_partials(const float* a_partials, int splits, int64_t slice_elems, const float* a_bias,
                                        const void* res, const float* gamma, const float* beta, void* out,
                                        float* rstd, int rows, int cols, float eps, float p_drop, const uint64_t* rng,
                                        uint32_t stream_id, void* stream) {
    if (a_partials == nullptr || splits < 1 || (slice_elems & 3) ||
        (reinterpret_cast<uintptr_t>(a_partials) & 15u))
        return DCAD_ERR_BAD_ARGUMENT;
    return add_ln_fwd_run(nullptr, PartialIn{a_partials, a_bias, slice_elems, splits}, res, gamma, beta, out, rstd,
                          rows, cols, eps, p_drop, rng, stream_id, stream);
}

static int add_ln_bwd_run(const void* dy1, PartialIn dy1_part, const void* dy2, const void* y, const float* gamma,
                          const float* beta, const float* rstd, void* dz, void* da, float* dgamma, float* dbeta,
                          float* dbias, int rows, int cols, float p_drop, const uint64_t* rng, uint32_t stream_id,
                          void* stream) {
    if (rows <= 0) return 0;
    if (cols % 256 || cols > 1024) return DCAD_ERR_BAD_SHAPE;
    if (p_drop > 0.f && (rng == nullptr || da == nullptr)) return DCAD_ERR_BAD_ARGUMENT;
    const uint32_t th = drop_thresh(p_drop);
    const float ks = p_drop > 0.f ? 1.0f / (1.0f - p_drop) : 1.0f;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int grid = (rows + LN_WARPS - 1) / LN_WARPS;
    const int smem = LN_WARPS * 3 * cols * (int)sizeof(float);
    DCAD_NV_DISPATCH(cols, {
        static DeviceOnce once;
        DCAD_CUDA_TRY(ensure_dynamic_smem(add_ln_bwd_kernel<NV>, LN_WARPS * 3 * NV * 256 * (int)sizeof(float), once));
        return (int)launch_pdl(
            add_ln_bwd_kernel<NV>, dim3(grid), dim3(LN_WARPS * 32), smem, s,
            reinterpret_cast<const __nv_bfloat16*>(dy1), dy1_part, reinterpret_cast<const __nv_bfloat16*>(dy2),
            reinterpret_cast<const __nv_bfloat16*>(y), gamma, beta, rstd, reinterpret_cast<__nv_bfloat16*>(dz),
            reinterpret_cast<__nv_bfloat16*>(p_drop > 0.f ? da : nullptr), dgamma, dbeta, dbias, rows, th, ks, rng,
            stream_id);
    });
    return 0;
}

extern "C" int dcad_add_ln_bwd(const void* dy1, const void* dy2, const void* y, const float* gamma, const float* beta,
                               const float* rstd, void* dz, void* da, float* dgamma, float* dbeta, float* dbias,
                               int rows, int cols, float p_drop, const uint64_t* rng, uint32_t stream_id,
                               void* stream) {
    return add_ln_bwd_run(dy1, PartialIn{nullptr, nullptr, 0, 0}, dy2, y, gamma, beta, rstd, dz, da, dgamma, dbeta,
                          dbias, rows, cols, p_drop, rng, stream_id, stream);
}

extern "C" int dcad_add_ln_bwd_partials(const float* dy1_partials, int splits, int64_t slice_elems, const void* dy2,
                                        const void* y, const float* gamma, const float* beta, const float* rstd,
                                        void* dz, void* da, float* dgamma, float* dbeta, float* dbias, int rows,
                                        int cols, float p_drop, const uint64_t* rng, uint32_t stream_id,
                                        void* stream) {
    if (dy1_partials == nullptr || splits < 1 || (slice_elems & 3) ||
        (reinterpret_cast<uintptr_t>(dy1_partials) & 15u))
        return DCAD_ERR_BAD_ARGUMENT;
    return add_ln_bwd_run(nullptr, PartialIn{dy1_partials, nullptr, slice_elems, splits}, dy2, y, gamma, beta, rstd,
                          dz, da, dgamma, dbeta, dbias, rows, cols, p_drop, rng, stream_id, stream);
}

extern "C" int dcad_embed_ln_fwd(const int32_t* ids, const float* word, const float* pos, const float* gamma,
                                 const float* beta, void* out, float* mean, float* rstd, int rows, int seq, int cols,
                                 float eps, float p_drop, const uint64_t* rng, uint32_t stream_id, void* stream) {
    if (rows <= 0) return 0;
    if (cols % 256 || cols > 1024 || seq <= 0) return DCAD_ERR_BAD_SHAPE;
    if (p_drop > 0.f && rng == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    const uint32_t th = drop_thresh(p_drop);
    const float ks = p_drop > 0.f ? 1.0f / (1.0f - p_drop) : 1.0f;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int grid = (rows + LN_WARPS - 1) / LN_WARPS;
    DCAD_NV_DISPATCH(cols, return (int)launch_pdl(
        embed_ln_fwd_kernel<NV>, dim3(grid), dim3(LN_WARPS * 32), 0, s, ids, word, pos, gamma, beta,
        reinterpret_cast<__nv_bfloat16*>(out), mean, rstd, rows, seq, eps, th, ks, rng, stream_id));
    return 0;
}

extern "C" int dcad_embed_ln_bwd(const void* dy1, const void* dy2, const int32_t* ids, const float* word,
                                 const float* pos, const float* gamma, const float* mean, const float* rstd,
                                 float* dword, float* dpos, float* dgamma, float* dbeta, const int32_t* lens, int rows,
                                 int seq, int cols, float p_drop, const uint64_t* rng, uint32_t stream_id,
                                 uint8_t* touched_rows, void* stream) {
    if (rows <= 0) return 0;
    if (cols % 256 || cols > 1024 || seq <= 0) return DCAD_ERR_BAD_SHAPE;
    if (p_drop > 0.f && rng == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    const uint32_t th = drop_thresh(p_drop);
    const float ks = p_drop > 0.f ? 1.0f / (1.0f - p_drop) : 1.0f;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int grid = (rows + LN_WARPS - 1) / LN_WARPS;
    const int smem = LN_WARPS * 2 * cols * (int)sizeof(float);
    DCAD_NV_DISPATCH(cols, {
        static DeviceOnce once;
        DCAD_CUDA_TRY(ensure_dynamic_smem(embed_ln_bwd_kernel<NV>, LN_WARPS * 2 * NV * 256 * (int)sizeof(float), once));
        return (int)launch_pdl(
            embed_ln_bwd_kernel<NV>, dim3(grid), dim3(LN_WARPS * 32), smem, s,
            reinterpret_cast<const __nv_bfloat16*>(dy1), reinterpret_cast<const __nv_bfloat16*>(dy2), ids, word, pos,
            gamma, mean, rstd, dword, dpos, dgamma, dbeta, lens, rows, seq, th, ks, rng, stream_id, touched_rows);
    });
    return 0;
}

extern "C" int dcad_colsum_bf16(const void* x, float* out, int rows, int cols, int ld, void* stream) {
    if (rows <= 0 || cols <= 0) return 0;
    if (cols % 8 || ld % 8) return DCAD_ERR_BAD_ALIGNMENT;
    const int rows_per_cta = 64;
    dim3 grid((cols + 255) / 256, (rows + rows_per_cta - 1) / rows_per_cta);
    return (int)launch_pdl(colsum_bf16_kernel, grid, dim3(256), 0, reinterpret_cast<cudaStream_t>(stream),
                           reinterpret_cast<const __nv_bfloat16*>(x), out, rows, cols, ld, rows_per_cta);
}
