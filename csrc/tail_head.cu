// Fused end of the forward pass and start of the backward pass for the CLS rows of the last block:
//
//     y      = LayerNorm(dropout(ffn_out + bias) + residual)          output_layer_norm of the last block, B rows
//     logits = Linear(dropout_0.3(y));  loss = CrossEntropy(logits)   classifier head (reference client1.py:62-64,108)
//     ... and their backward: d(logits), dW_cls, db_cls, dy, then the LayerNorm backward (dz, d(ffn_out),
//     dgamma, dbeta, d(lin2 bias)).
//
// In the unfused step these are three dependent launches (add_ln_fwd, cls_head_train, add_ln_bwd) that together move
// 16 rows of 768 numbers; each costs 8 us of launch / dependency latency.  One CTA does all of it: warp per sample
// for the row-wise work (fp32 throughout: y and x_hat never round to bf16 in between), then the whole CTA for the
// column sums over the samples.  Everything the column pass needs except x_hat and a few per-row scalars is
// recomputed (dropout masks are a pure function of the index).
#include "common.cuh"
#include "../include/dcad_b200.h"

namespace dcad {

constexpr int TH_THREADS = 512;
constexpr int TH_WARPS = TH_THREADS / 32;

struct TailHeadArgs {
    // LayerNorm input: either fp32 split-K partial tiles (+ bias) or a bf16 tensor
    const float* a_part;
    const float* a_bias;
    int64_t slice_elems;
    int splits;
    const __nv_bfloat16* a;
    const __nv_bfloat16* res;   // residual [B, COLS]
    const float* gamma;
    const float* beta;
    float eps;
    uint32_t ln_thresh;         // dropout on the LayerNorm input
    float ln_keep_scale;
    uint32_t ln_stream;
    // classifier
    const float* W;             // [2, COLS]
    const float* cls_bias;      // [2]
    const int32_t* labels;
    uint32_t head_thresh;
    float head_keep_scale;
    uint32_t head_stream;
    const uint64_t* rng;
    int batch;
    // outputs
    float* loss_accum;          // {sum of batch-mean losses, batches}
    float* logits;              // [B, 2] or null
    float* dW;                  // [2, COLS]  (accumulated)
    float* db;                  // [2]
    __nv_bfloat16* dz;          // [B, COLS] grad wrt (dropout(a) + res)
    __nv_bfloat16* da;          // [B, COLS] grad wrt a (null when ln_thresh == 0: equals dz)
    float* dgamma;
    float* dbeta;
    float* dbias;               // grad of the bias folded into a (lin2 bias)
};

__device__ __forceinline__ uint64_t th_seed(const uint64_t* rng) {
    return rng ? rng[0] + rng[1] * 0xD1B54A32D192ED03ULL : 0ULL;
}

template <int NV>
__global__ void __launch_bounds__(TH_THREADS)
tail_head_kernel(const TailHeadArgs p) {
    pdl_trigger();
    pdl_wait();
    constexpr int COLS = NV * 256;
    extern __shared__ float s_xhat[];            // [batch][COLS]
    __shared__ float s_row[64][6];               // per sample: g0, g1, c1, c2, rstd, loss
    const int warp = threadIdx.x >> 5, lane = lane_id();
    const uint64_t seed = th_seed(p.rng);
    const float inv_b = 1.0f / static_cast<float>(p.batch);

    // ---------------- row-wise pass: one warp per sample ----------------
    for (int s = warp; s < p.batch; s += TH_WARPS) {
        float x[NV][8];
        float sum = 0.f;
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int col = (v * 32 + lane) * 8;
            const size_t off = static_cast<size_t>(s) * COLS + col;
            if (p.a_part) {
                const float4 a0 = *reinterpret_cast<const float4*>(p.a_part + off);
                const float4 a1 = *reinterpret_cast<const float4*>(p.a_part + off + 4);
                x[v][0] = a0.x; x[v][1] = a0.y; x[v][2] = a0.z; x[v][3] = a0.w;
                x[v][4] = a1.x; x[v][5] = a1.y; x[v][6] = a1.z; x[v][7] = a1.w;
                for (int k = 1; k < p.splits; ++k) {
                    const float4 b0 = *reinterpret_cast<const float4*>(p.a_part + k * p.slice_elems + off);
                    const float4 b1 = *reinterpret_cast<const float4*>(p.a_part + k * p.slice_elems + off + 4);
                    x[v][0] += b0.x; x[v][1] += b0.y; x[v][2] += b0.z; x[v][3] += b0.w;
                    x[v][4] += b1.x; x[v][5] += b1.y; x[v][6] += b1.z; x[v][7] += b1.w;
                }
                if (p.a_bias) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) x[v][j] += __ldg(p.a_bias + col + j);
                }
            } else {
                const uint4 u = *reinterpret_cast<const uint4*>(p.a + off);
                const float2 q0 = unpack_bf16x2(u.x), q1 = unpack_bf16x2(u.y), q2 = unpack_bf16x2(u.z), q3 = unpack_bf16x2(u.w);
                x[v][0] = q0.x; x[v][1] = q0.y; x[v][2] = q1.x; x[v][3] = q1.y;
                x[v][4] = q2.x; x[v][5] = q2.y; x[v][6] = q3.x; x[v][7] = q3.y;
            }
            const uint4 r = *reinterpret_cast<const uint4*>(p.res + off);
            const float2 r0 = unpack_bf16x2(r.x), r1 = unpack_bf16x2(r.y), r2 = unpack_bf16x2(r.z), r3 = unpack_bf16x2(r.w);
            const float rr[8] = {r0.x, r0.y, r1.x, r1.y, r2.x, r2.y, r3.x, r3.y};
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (p.ln_thresh)
                    x[v][j] = dropout_keep(seed, p.ln_stream, s * COLS + col + j, p.ln_thresh) ? x[v][j] * p.ln_keep_scale : 0.f;
                x[v][j] += rr[j];
                sum += x[v][j];
            }
        }
        const float mean = warp_sum(sum) * (1.0f / COLS);
        float var = 0.f;
#pragma unroll
        for (int v = 0; v < NV; ++v)
#pragma unroll
            for (int j = 0; j < 8; ++j) { const float d = x[v][j] - mean; var += d * d; }
        const float rstd = rsqrtf(warp_sum(var) * (1.0f / COLS) + p.eps);

        // x <- x_hat; classifier forward on dropout(y)
        float z0 = 0.f, z1 = 0.f;
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int col = (v * 32 + lane) * 8;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                x[v][j] = (x[v][j] - mean) * rstd;
                const float y = x[v][j] * __ldg(p.gamma + col + j) + __ldg(p.beta + col + j);
                const bool keep = p.head_thresh ? dropout_keep(seed, p.head_stream, s * COLS + col + j, p.head_thresh) : true;
                const float yd = keep ? y * p.head_keep_scale : 0.f;
                z0 += yd * __ldg(p.W + col + j);
                z1 += yd * __ldg(p.W + COLS + col + j);
            }
            *reinterpret_cast<float4*>(s_xhat + static_cast<size_t>(s) * COLS + col) = make_float4(x[v][0], x[v][1], x[v][2], x[v][3]);
            *reinterpret_cast<float4*>(s_xhat + static_cast<size_t>(s) * COLS + col + 4) = make_float4(x[v][4], x[v][5], x[v][6], x[v][7]);
        }
        z0 = warp_sum(z0) + p.cls_bias[0];
        z1 = warp_sum(z1) + p.cls_bias[1];
        const int lab = p.labels[s];
        const float m = fmaxf(z0, z1);
        const float e0 = expf(z0 - m), e1 = expf(z1 - m);
        const float lse = m + logf(e0 + e1);
        const float g0 = (e0 / (e0 + e1) - (lab == 0 ? 1.f : 0.f)) * inv_b;
        const float g1 = (e1 / (e0 + e1) - (lab == 1 ? 1.f : 0.f)) * inv_b;
        if (lane == 0 && p.logits) { p.logits[2 * s] = z0; p.logits[2 * s + 1] = z1; }

        // LayerNorm backward of this row: dy = grad wrt y (through the head dropout)
        float gsum = 0.f, gxsum = 0.f;
        float g[NV][8];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int col = (v * 32 + lane) * 8;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const bool keep = p.head_thresh ? dropout_keep(seed, p.head_stream, s * COLS + col + j, p.head_thresh) : true;
                const float dy = keep ? (g0 * __ldg(p.W + col + j) + g1 * __ldg(p.W + COLS + col + j)) * p.head_keep_scale : 0.f;
                g[v][j] = dy * __ldg(p.gamma + col + j);
                gsum += g[v][j];
                gxsum += g[v][j] * x[v][j];
            }
        }
        const float c1 = warp_sum(gsum) * (1.0f / COLS);
        const float c2 = warp_sum(gxsum) * (1.0f / COLS);
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int col = (v * 32 + lane) * 8;
            const size_t off = static_cast<size_t>(s) * COLS + col;
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = rstd * (g[v][j] - c1 - x[v][j] * c2);
            uint4 u;
            u.x = pack_bf16x2(o[0], o[1]); u.y = pack_bf16x2(o[2], o[3]);
            u.z = pack_bf16x2(o[4], o[5]); u.w = pack_bf16x2(o[6], o[7]);
            *reinterpret_cast<uint4*>(p.dz + off) = u;
            if (p.da) {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    o[j] = dropout_keep(seed, p.ln_stream, s * COLS + col + j, p.ln_thresh) ? o[j] * p.ln_keep_scale : 0.f;
                u.x = pack_bf16x2(o[0], o[1]); u.y = pack_bf16x2(o[2], o[3]);
                u.z = pack_bf16x2(o[4], o[5]); u.w = pack_bf16x2(o[6], o[7]);
                *reinterpret_cast<uint4*>(p.da + off) = u;
            }
        }
        if (lane == 0) {
            s_row[s][0] = g0; s_row[s][1] = g1; s_row[s][2] = c1; s_row[s][3] = c2; s_row[s][4] = rstd;
            s_row[s][5] = lse - (lab == 0 ? z0 : z1);
        }
    }
    __syncthreads();

    // ---------------- column pass: sums over the samples ----------------
    for (int c = threadIdx.x; c < COLS; c += TH_THREADS) {
        const float gm = __ldg(p.gamma + c), bt = __ldg(p.beta + c);
        const float w0 = __ldg(p.W + c), w1 = __ldg(p.W + COLS + c);
        float dw0 = 0.f, dw1 = 0.f, dg = 0.f, dbt = 0.f, dbs = 0.f;
        for (int s = 0; s < p.batch; ++s) {
            const float xh = s_xhat[static_cast<size_t>(s) * COLS + c];
            const float g0 = s_row[s][0], g1 = s_row[s][1];
            const bool keep = p.head_thresh ? dropout_keep(seed, p.head_stream, s * COLS + c, p.head_thresh) : true;
            const float yd = keep ? (xh * gm + bt) * p.head_keep_scale : 0.f;
            dw0 += g0 * yd;
            dw1 += g1 * yd;
            const float dy = keep ? (g0 * w0 + g1 * w1) * p.head_keep_scale : 0.f;
            dg += dy * xh;
            dbt += dy;
            float dzv = s_row[s][4] * (dy * gm - s_row[s][2] - xh * s_row[s][3]);
            if (p.ln_thresh) dzv = dropout_keep(seed, p.ln_stream, s * COLS + c, p.ln_thresh) ? dzv * p.ln_keep_scale : 0.f;
            dbs += dzv;
        }
        atomicAdd(p.dW + c, dw0);
        atomicAdd(p.dW + COLS + c, dw1);
        if (p.dgamma) atomicAdd(p.dgamma + c, dg);
        if (p.dbeta) atomicAdd(p.dbeta + c, dbt);
        if (p.dbias) atomicAdd(p.dbias + c, dbs);
    }
    if (threadIdx.x == 0) {
        float l = 0.f, a = 0.f, b = 0.f;
        for (int s = 0; s < p.batch; ++s) { l += s_row[s][5]; a += s_row[s][0]; b += s_row[s][1]; }
        atomicAdd(p.db, a);
        atomicAdd(p.db + 1, b);
        if (p.loss_accum) {
            atomicAdd(p.loss_accum, l * inv_b);
            atomicAdd(p.loss_accum + 1, 1.0f);
        }
    }
}

}  // namespace dcad

using namespace dcad;

static uint32_t th_thresh(float p) {
    if (p <= 0.f) return 0u;
    const double t = static_cast<double>(p) * 4294967296.0;
    return t >= 4294967295.0 ? 4294967295u : static_cast<uint32_t>(t);
}

// a_partials != NULL: the LayerNorm input is sum_k a_partials[k * slice_elems + row * cols + col] + a_bias[col]
// (split-K pass 1 output); otherwise `a` (bf16 [batch, cols]).  da may be NULL only when p_ln == 0.
extern "C" int dcad_tail_head(const float* a_partials, int splits, int64_t slice_elems, const float* a_bias,
                              const void* a, const void* res, const float* gamma, const float* beta, float eps,
                              float p_ln, uint32_t ln_stream, const float* W, const float* cls_bias,
                              const int32_t* labels, float p_head, uint32_t head_stream, const uint64_t* rng,
                              int batch, int cols, float* loss_accum, float* logits, float* dW, float* db, void* dz,
                              void* da, float* dgamma, float* dbeta, float* dbias, void* stream) {
    if (batch <= 0) return 0;
    if (cols % 256 || cols > 1024 || batch > 64) return DCAD_ERR_BAD_SHAPE;
    if ((a_partials == nullptr) == (a == nullptr)) return DCAD_ERR_BAD_ARGUMENT;
    if (a_partials && (splits < 1 || (slice_elems & 3))) return DCAD_ERR_BAD_ARGUMENT;
    if ((p_ln > 0.f || p_head > 0.f) && rng == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (p_ln > 0.f && da == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    const size_t smem = static_cast<size_t>(batch) * cols * sizeof(float);
    if (smem > 200 * 1024) return DCAD_ERR_TOO_LARGE;
    TailHeadArgs p;
    p.a_part = a_partials; p.a_bias = a_bias; p.slice_elems = slice_elems; p.splits = splits;
    p.a = reinterpret_cast<const __nv_bfloat16*>(a);
    p.res = reinterpret_cast<const __nv_bfloat16*>(res);
    p.gamma = gamma; p.beta = beta; p.eps = eps;
    p.ln_thresh = th_thresh(p_ln); p.ln_keep_scale = p_ln > 0.f ? 1.0f / (1.0f - p_ln) : 1.0f; p.ln_stream = ln_stream;
    p.W = W; p.cls_bias = cls_bias; p.labels = labels;
    p.head_thresh = th_thresh(p_head); p.head_keep_scale = p_head > 0.f ? 1.0f / (1.0f - p_head) : 1.0f;
    p.head_stream = head_stream; p.rng = rng; p.batch = batch;
    p.loss_accum = loss_accum; p.logits = logits; p.dW = dW; p.db = db;
    p.dz = reinterpret_cast<__nv_bfloat16*>(dz);
    p.da = reinterpret_cast<__nv_bfloat16*>(p_ln > 0.f ? da : nullptr);
    p.dgamma = dgamma; p.dbeta = dbeta; p.dbias = dbias;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
#define DCAD_TH_CASE(NV_)                                                                                          \
    case NV_: {                                                                                                    \
        static DeviceOnce once;                                                                                    \
        DCAD_CUDA_TRY(ensure_dynamic_smem(tail_head_kernel<NV_>, 200 * 1024, once));                               \
        return (int)launch_pdl(tail_head_kernel<NV_>, dim3(1), dim3(TH_THREADS), smem, s, p);                      \
    }
    switch (cols / 256) {
        DCAD_TH_CASE(1)
        DCAD_TH_CASE(2)
        DCAD_TH_CASE(3)
        DCAD_TH_CASE(4)
        default: return DCAD_ERR_BAD_SHAPE;
    }
#undef DCAD_TH_CASE
}
