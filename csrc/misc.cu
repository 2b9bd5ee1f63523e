// Small utility kernels: version query, fp32 -> bf16 cast, row gather/scatter (last-layer CLS pruning).
#include "common.cuh"
#include "../include/dcad_b200.h"

namespace dcad {

int g_pdl_enabled = 1;

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n) {
    pdl_trigger();
    pdl_wait();
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x * 8;
    for (int64_t i = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 8; i < n; i += stride) {
        if (i + 8 <= n) {
            const float4 a = *reinterpret_cast<const float4*>(src + i);
            const float4 b = *reinterpret_cast<const float4*>(src + i + 4);
            uint4 o;
            o.x = pack_bf16x2(a.x, a.y); o.y = pack_bf16x2(a.z, a.w);
            o.z = pack_bf16x2(b.x, b.y); o.w = pack_bf16x2(b.z, b.w);
            *reinterpret_cast<uint4*>(dst + i) = o;
        } else {
            for (int64_t j = i; j < n; ++j) dst[j] = __float2bfloat16(src[j]);
        }
    }
}

// dst[r, :] = src[r * row_stride_rows, :]   (pick every `every`-th row: the CLS token of each sequence)
__global__ void gather_rows_kernel(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst, int rows,
                                   int cols, int every, int ld_src, int ld_dst) {
    pdl_trigger();
    pdl_wait();
    const int r = blockIdx.x;
    if (r >= rows) return;
    const uint4* s = reinterpret_cast<const uint4*>(src + static_cast<size_t>(r) * every * ld_src);
    uint4* d = reinterpret_cast<uint4*>(dst + static_cast<size_t>(r) * ld_dst);
    for (int c = threadIdx.x; c < cols / 8; c += blockDim.x) d[c] = s[c];
}
__global__ void scatter_rows_kernel(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst, int rows,
                                    int cols, int every, int ld_src, int ld_dst) {
    // dst is [rows*every, cols]: row r*every receives src[r], all other rows are zeroed
    pdl_trigger();
    pdl_wait();
    const int r = blockIdx.x;
    if (r >= rows * every) return;
    uint4* d = reinterpret_cast<uint4*>(dst + static_cast<size_t>(r) * ld_dst);
    if (r % every == 0) {
        const uint4* s = reinterpret_cast<const uint4*>(src + static_cast<size_t>(r / every) * ld_src);
        for (int c = threadIdx.x; c < cols / 8; c += blockDim.x) d[c] = s[c];
    } else {
        for (int c = threadIdx.x; c < cols / 8; c += blockDim.x) d[c] = make_uint4(0, 0, 0, 0);
    }
}

}  // namespace dcad

using namespace dcad;

extern "C" int dcad_version(void) { return 101; }

extern "C" int dcad_set_pdl(int enabled) {
    const int prev = g_pdl_enabled;
    g_pdl_enabled = enabled ? 1 : 0;
    return prev;
}

extern "C" int dcad_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream) {
    if (n <= 0) return 0;
    if ((reinterpret_cast<uintptr_t>(src) & 15u) || (reinterpret_cast<uintptr_t>(dst) & 15u)) return DCAD_ERR_BAD_ALIGNMENT;
    int64_t blocks = (n / 8 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    return (int)launch_pdl(cast_f32_bf16_kernel, dim3((unsigned)blocks), dim3(256), 0,
                           reinterpret_cast<cudaStream_t>(stream), src, reinterpret_cast<__nv_bfloat16*>(dst), n);
}

extern "C" int dcad_gather_rows_bf16(const void* src, void* dst, int rows, int cols, int every, int ld_src, void* stream) {
    if (cols % 8) return DCAD_ERR_BAD_ALIGNMENT;
    return (int)launch_pdl(gather_rows_kernel, dim3(rows), dim3(96), 0, reinterpret_cast<cudaStream_t>(stream),
                           reinterpret_cast<const __nv_bfloat16*>(src), reinterpret_cast<__nv_bfloat16*>(dst), rows,
                           cols, every, ld_src, cols);
}

extern "C" int dcad_scatter_rows_bf16(const void* src, void* dst, int rows, int cols, int every, int ld_dst, void* stream) {
    if (cols % 8) return DCAD_ERR_BAD_ALIGNMENT;
    return (int)launch_pdl(scatter_rows_kernel, dim3(rows * every), dim3(96), 0,
                           reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(src),
                           reinterpret_cast<__nv_bfloat16*>(dst), rows, cols, every, cols, ld_dst);
}
