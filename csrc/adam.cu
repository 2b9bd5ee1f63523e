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
                 const uint64_t* __restrict__ state) {
    __shared__ float s_c[2];
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
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x * 4;
    for (int64_t i = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
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
    }
}

__global__ void step_advance_kernel(uint64_t* state) {
    if (threadIdx.x == 0 && blockIdx.x == 0) state[1] += 1;
}

}  // namespace dcad

using namespace dcad;

// n, bf16_begin/end, zero_begin/end are element offsets into the arena and must be multiples of 4.
extern "C" int dcad_adam_step(float* p, float* g, float* m, float* v, void* p_bf16, int64_t n, int64_t bf16_begin,
                              int64_t bf16_end, int64_t zero_begin, int64_t zero_end, float lr, float beta1,
                              float beta2, float eps, const uint64_t* state, void* stream) {
    if (n <= 0) return 0;
    if ((n | bf16_begin | bf16_end | zero_begin | zero_end) & 3) return DCAD_ERR_BAD_ALIGNMENT;
    if ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
         reinterpret_cast<uintptr_t>(v)) & 15u)
        return DCAD_ERR_BAD_ALIGNMENT;
    int64_t blocks = (n / 4 + 255) / 256;
    const int64_t cap = 148 * 16;
    if (blocks > cap) blocks = cap;
    adam_flat_kernel<<<static_cast<int>(blocks), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        p, g, m, v, reinterpret_cast<__nv_bfloat16*>(p_bf16), n, bf16_begin, bf16_end, zero_begin, zero_end, lr, beta1,
        beta2, eps, state);
    return (int)cudaGetLastError();
}

extern "C" int dcad_step_advance(uint64_t* state, void* stream) {
    step_advance_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(state);
    return (int)cudaGetLastError();
}
