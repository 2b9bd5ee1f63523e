// Classifier head fused with softmax cross-entropy (training: forward + backward in one launch) and the
// evaluation variant that keeps all metrics on the device.
//
// Training replaces reference client1.py:62-64 (CLS pooling, Dropout(0.3), Linear(768,2)) + :108
// (nn.CrossEntropyLoss, mean) + their backward (:109).  Evaluation replaces client1.py:134-142: loss, softmax[:,1],
// argmax and the per-batch host round trips; instead a 2x2 confusion matrix and the loss sum stay on the GPU and are
// read once per evaluation.
#include "common.cuh"
#include "../include/dcad_b200.h"

namespace dcad {

constexpr int CLS_THREADS = 256;
constexpr int CLS_TRAIN_THREADS = 512;
constexpr int CLS_MAX_COLS = 1024;

__device__ __forceinline__ uint64_t cls_seed(const uint64_t* rng) {
    return rng ? rng[0] + rng[1] * 0xD1B54A32D192ED03ULL : 0ULL;
}

__device__ __forceinline__ void cls_load8(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}

// One CTA of 16 warps; warp w handles samples w, w+16, ...  Lane l owns columns (v*32 + l)*8 .. +7 for v < NV
// (16-byte accesses).  Per-warp dW partials live in registers, are parked in a per-warp smem slot and summed by the
// whole CTA; no shared-memory atomics.
template <int NV>
__global__ void __launch_bounds__(CLS_TRAIN_THREADS)
cls_head_train_kernel(const __nv_bfloat16* __restrict__ h, int sample_stride, const float* __restrict__ W,
                      const float* __restrict__ bias, const int32_t* __restrict__ labels, int batch,
                      uint32_t thresh, float keep_scale, const uint64_t* __restrict__ rng, uint32_t stream_id,
                      float* __restrict__ loss_accum, float* __restrict__ logits, float* __restrict__ dW,
                      float* __restrict__ db, __nv_bfloat16* __restrict__ dcls) {
    pdl_trigger();
    pdl_wait();
    constexpr int COLS = NV * 256;
    constexpr int WARPS = CLS_TRAIN_THREADS / 32;
    extern __shared__ float s_slot[];  // [WARPS][2*COLS]
    __shared__ float s_red[WARPS][3];
    const int warp = threadIdx.x >> 5, lane = lane_id();
    const uint64_t seed = cls_seed(rng);
    float w0[NV][8], w1[NV][8], a0[NV][8], a1[NV][8];
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int col = (v * 32 + lane) * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            w0[v][j] = __ldg(W + col + j);
            w1[v][j] = __ldg(W + COLS + col + j);
            a0[v][j] = 0.f;
            a1[v][j] = 0.f;
        }
    }
    float loss_sum = 0.f, db0 = 0.f, db1 = 0.f;
    const float inv_b = 1.0f / static_cast<float>(batch);
    for (int s = warp; s < batch; s += WARPS) {
        const __nv_bfloat16* x = h + static_cast<size_t>(s) * sample_stride;
        float xv[NV][8];
        bool keep[NV][8];
        float z0 = 0.f, z1 = 0.f;
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int col = (v * 32 + lane) * 8;
            cls_load8(x + col, xv[v]);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                keep[v][j] = thresh ? dropout_keep(seed, stream_id, s * COLS + col + j, thresh) : true;
                xv[v][j] = keep[v][j] ? xv[v][j] * keep_scale : 0.f;
                z0 += xv[v][j] * w0[v][j];
                z1 += xv[v][j] * w1[v][j];
            }
        }
        z0 = warp_sum(z0) + bias[0];
        z1 = warp_sum(z1) + bias[1];
        const int y = labels[s];
        const float m = fmaxf(z0, z1);
        const float e0 = expf(z0 - m), e1 = expf(z1 - m);
        const float lse = m + logf(e0 + e1);
        const float g0 = (e0 / (e0 + e1) - (y == 0 ? 1.f : 0.f)) * inv_b;
        const float g1 = (e1 / (e0 + e1) - (y == 1 ? 1.f : 0.f)) * inv_b;
        if (lane == 0) {
            loss_sum += lse - (y == 0 ? z0 : z1);
            db0 += g0;
            db1 += g1;
            if (logits) { logits[2 * s] = z0; logits[2 * s + 1] = z1; }
        }
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int col = (v * 32 + lane) * 8;
            float dx[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                a0[v][j] += g0 * xv[v][j];
                a1[v][j] += g1 * xv[v][j];
                const float t = g0 * w0[v][j] + g1 * w1[v][j];
                dx[j] = keep[v][j] ? t * keep_scale : 0.f;
            }
            if (dcls) {
                uint4 o;
                o.x = pack_bf16x2(dx[0], dx[1]); o.y = pack_bf16x2(dx[2], dx[3]);
                o.z = pack_bf16x2(dx[4], dx[5]); o.w = pack_bf16x2(dx[6], dx[7]);
                *reinterpret_cast<uint4*>(dcls + static_cast<size_t>(s) * COLS + col) = o;
            }
        }
    }
    float* slot = s_slot + warp * 2 * COLS;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int col = (v * 32 + lane) * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) { slot[col + j] = a0[v][j]; slot[COLS + col + j] = a1[v][j]; }
    }
    if (lane == 0) { s_red[warp][0] = loss_sum; s_red[warp][1] = db0; s_red[warp][2] = db1; }
    __syncthreads();
    for (int c = threadIdx.x; c < 2 * COLS; c += blockDim.x) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) t += s_slot[w * 2 * COLS + c];
        atomicAdd(dW + c, t);
    }
    if (threadIdx.x == 0) {
        float l = 0.f, a = 0.f, b = 0.f;
        for (int w = 0; w < WARPS; ++w) { l += s_red[w][0]; a += s_red[w][1]; b += s_red[w][2]; }
        atomicAdd(db, a);
        atomicAdd(db + 1, b);
        if (loss_accum) {
            atomicAdd(loss_accum, l * inv_b);  // mean loss of this batch (reference: running_loss += loss.item())
            atomicAdd(loss_accum + 1, 1.0f);   // number of batches
        }
    }
}

__global__ void __launch_bounds__(CLS_THREADS)
cls_head_eval_kernel(const __nv_bfloat16* __restrict__ h, int sample_stride, const float* __restrict__ W,
                     const float* __restrict__ bias, const int32_t* __restrict__ labels, int batch, int cols,
                     float* __restrict__ loss_accum, int32_t* __restrict__ confusion, float* __restrict__ probs,
                     int32_t* __restrict__ preds) {
    __shared__ float s_loss[8];
    const int warp = threadIdx.x >> 5, lane = lane_id();
    float loss_sum = 0.f;
    for (int s = warp; s < batch; s += CLS_THREADS / 32) {
        const __nv_bfloat16* x = h + static_cast<size_t>(s) * sample_stride;
        float z0 = 0.f, z1 = 0.f;
        for (int c = lane; c < cols; c += 32) {
            const float xv = __bfloat162float(x[c]);
            z0 += xv * __ldg(W + c);
            z1 += xv * __ldg(W + cols + c);
        }
        z0 = warp_sum(z0) + bias[0];
        z1 = warp_sum(z1) + bias[1];
        if (lane == 0) {
            const int y = labels[s];
            const float m = fmaxf(z0, z1);
            const float e0 = expf(z0 - m), e1 = expf(z1 - m);
            loss_sum += m + logf(e0 + e1) - (y == 0 ? z0 : z1);
            const int pred = z1 > z0 ? 1 : 0;  // torch.max returns the first maximal index on ties
            if (confusion && (y == 0 || y == 1)) atomicAdd(confusion + y * 2 + pred, 1);
            if (probs) probs[s] = e1 / (e0 + e1);
            if (preds) preds[s] = pred;
        }
    }
    if (lane == 0) s_loss[warp] = loss_sum;
    __syncthreads();
    if (threadIdx.x == 0 && loss_accum) {
        float l = 0.f;
        for (int w = 0; w < CLS_THREADS / 32; ++w) l += s_loss[w];
        atomicAdd(loss_accum, l / static_cast<float>(batch));  // reference: total_loss += loss.item() (batch mean)
        atomicAdd(loss_accum + 1, 1.0f);
    }
}

}  // namespace dcad

using namespace dcad;

template <int NV>
static int launch_cls_train(const void* h, int sample_stride, const float* W, const float* bias,
                            const int32_t* labels, int batch, uint32_t th, float ks, const uint64_t* rng,
                            uint32_t stream_id, float* loss_accum, float* logits, float* dW, float* db, void* dcls,
                            cudaStream_t s) {
    constexpr int SMEM = (CLS_TRAIN_THREADS / 32) * 2 * NV * 256 * (int)sizeof(float);
    static DeviceOnce once;
    DCAD_CUDA_TRY(ensure_dynamic_smem(cls_head_train_kernel<NV>, SMEM, once));
    return (int)launch_pdl(cls_head_train_kernel<NV>, dim3(1), dim3(CLS_TRAIN_THREADS), SMEM, s,
                           reinterpret_cast<const __nv_bfloat16*>(h), sample_stride, W, bias, labels, batch, th, ks, rng,
                           stream_id, loss_accum, logits, dW, db, reinterpret_cast<__nv_bfloat16*>(dcls));
}

extern "C" int dcad_cls_head_train(const void* h, int sample_stride, const float* W, const float* bias,
                                   const int32_t* labels, int batch, float p_drop, const uint64_t* rng,
                                   uint32_t stream_id, float* loss_accum, float* logits, float* dW, float* db,
                                   void* dcls, int cols, void* stream) {
    if (batch <= 0) return 0;
    if (cols <= 0 || cols > CLS_MAX_COLS || cols % 256 || sample_stride % 8) return DCAD_ERR_BAD_SHAPE;
    if (p_drop > 0.f && rng == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    uint32_t th = 0;
    if (p_drop > 0.f) {
        double t = static_cast<double>(p_drop) * 4294967296.0;
        th = static_cast<uint32_t>(t > 4294967295.0 ? 4294967295.0 : t);
    }
    const float ks = p_drop > 0.f ? 1.0f / (1.0f - p_drop) : 1.0f;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    switch (cols / 256) {
        case 1: return launch_cls_train<1>(h, sample_stride, W, bias, labels, batch, th, ks, rng, stream_id, loss_accum, logits, dW, db, dcls, s);
        case 2: return launch_cls_train<2>(h, sample_stride, W, bias, labels, batch, th, ks, rng, stream_id, loss_accum, logits, dW, db, dcls, s);
        case 3: return launch_cls_train<3>(h, sample_stride, W, bias, labels, batch, th, ks, rng, stream_id, loss_accum, logits, dW, db, dcls, s);
        default: return launch_cls_train<4>(h, sample_stride, W, bias, labels, batch, th, ks, rng, stream_id, loss_accum, logits, dW, db, dcls, s);
    }
}

extern "C" int dcad_cls_head_eval(const void* h, int sample_stride, const float* W, const float* bias,
                                  const int32_t* labels, int batch, int cols, float* loss_accum, int32_t* confusion,
                                  float* probs, int32_t* preds, void* stream) {
    if (batch <= 0) return 0;
    if (cols <= 0 || cols > CLS_MAX_COLS) return DCAD_ERR_BAD_SHAPE;
    cls_head_eval_kernel<<<1, CLS_THREADS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(h), sample_stride, W, bias, labels, batch, cols, loss_accum, confusion,
        probs, preds);
    return (int)cudaGetLastError();
}
