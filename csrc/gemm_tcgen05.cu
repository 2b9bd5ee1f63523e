// Persistent, warp-specialised bf16 GEMM for sm_100a: TMA -> 128B-swizzled smem ring -> tcgen05.mma with fp32
// accumulators in TMEM (double-buffered) -> tcgen05.ld epilogue (bias / GELU / dGELU / fp32 store).
//
// Replaces the cuBLAS GEMMs that the reference reaches through HF DistilBERT's nn.Linear layers
// (reference client1.py:61 -> transformers modeling_distilbert.py:187-189,205,224-226) and their autograd
// backward (client1.py:109).  Three operand-layout combinations cover forward, dgrad and wgrad without ever
// materialising a transposed tensor:
//     forward  Y[T,N]  = X[T,K]  W[N,K]^T        A K-major , B K-major
//     dgrad    dX[T,K] = dY[T,N] W[N,K]          A K-major , B MN-major
//     wgrad    dW[N,K] = dY[T,N]^T X[T,K]        A MN-major, B MN-major
//
// Roles (192 threads): warp 0 = TMA producer, warp 1 = single-thread MMA issuer, warps 2..5 = epilogue
// (warp 2 also owns the TMEM allocation).  Three mbarrier pipelines: smem full/empty, TMEM full/empty.
#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "../include/dcad_b200.h"

namespace dcad {

constexpr int BM = 128;  // UMMA M (cta_group::1, full-rate shape)
constexpr int BK = 64;   // one 128-byte swizzle atom of bf16 along the reduction dim per stage
constexpr int UMMA_K = 16;
constexpr int SLAB_BYTES = BK * 128;  // one [BK][64] MN-major slab

struct GemmParams {
    int M, N, K;
    int ldd, ldaux;
    const float* bias;
    const __nv_bfloat16* aux;  // EPI_DGELU: pre-activation (same shape as D)
    __nv_bfloat16* aux_out;    // EPI_BIAS_GELU: where to keep the pre-activation for backward (may be null)
    void* D;
    int epi;
    int out_fp32;
    int accumulate;  // fp32 output only: D += acc
    int num_m_blocks, num_n_blocks, num_k_blocks;
};

__host__ __device__ constexpr int pow2_ceil(int v) {
    int p = 32;
    while (p < v) p <<= 1;
    return p;
}

template <int BN>
struct GemmCfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES_RAW = (200 * 1024) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
    static constexpr int ACC_STRIDE = pow2_ceil(BN);  // TMEM columns per accumulator stage
    static constexpr int TMEM_COLS = 2 * ACC_STRIDE;  // two accumulator stages
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024;
};

constexpr int GEMM_THREADS = 384;        // 12 warps: 0 = TMA, 1 = MMA, 2 = TMEM owner, 3 = idle, 4..11 = epilogue
constexpr int EPI_THREADS = 256;

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

template <int N>
struct TmemChunk;
template <>
struct TmemChunk<32> {
    static __device__ __forceinline__ void ld(uint32_t taddr, uint32_t (&r)[32]) { tmem_ld_32x32b_x32(taddr, r); }
};
template <>
struct TmemChunk<16> {
    static __device__ __forceinline__ void ld(uint32_t taddr, uint32_t (&r)[16]) { tmem_ld_32x32b_x16(taddr, r); }
};

template <int BN, int A_MAJOR, int B_MAJOR>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const GemmParams p) {
    using Cfg = GemmCfg<BN>;
    constexpr int STAGES = Cfg::STAGES;
    static_assert(BN % 32 == 0 && BN >= 32 && BN <= 256, "BLOCK_N must be a multiple of 32 in [32,256]");
    static_assert(B_MAJOR == kMajorK || BN % 64 == 0, "MN-major B needs 64-column slabs");

    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[STAGES];
    __shared__ __align__(8) uint64_t empty_bar[STAGES];
    __shared__ __align__(8) uint64_t tmem_full_bar[2];
    __shared__ __align__(8) uint64_t tmem_empty_bar[2];
    __shared__ __align__(16) float s_bias[2][BN];
    __shared__ uint32_t tmem_base_smem;

    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = lane_id();
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B wants 1024-B alignment

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
    }
    if (warp == 1 && lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(smem_u32(&full_bar[s]), 1);
            mbar_init(smem_u32(&empty_bar[s]), 1);
        }
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            mbar_init(smem_u32(&tmem_full_bar[s]), 1);
            mbar_init(smem_u32(&tmem_empty_bar[s]), EPI_THREADS);
        }
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc(smem_u32(&tmem_base_smem), Cfg::TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    const int num_tiles = p.num_m_blocks * p.num_n_blocks;

    if (warp == 0) {
        // ===================== TMA producer =====================
        uint32_t stage = 0, phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const int m_blk = tile % p.num_m_blocks;
            const int n_blk = tile / p.num_m_blocks;
            for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
                if (lane == 0) {
                    const uint32_t bar = smem_u32(&full_bar[stage]);
                    const uint32_t sA = smem_base + stage * Cfg::STAGE_BYTES;
                    const uint32_t sB = sA + Cfg::A_BYTES;
                    mbar_arrive_expect_tx(bar, Cfg::STAGE_BYTES);
                    if constexpr (A_MAJOR == kMajorK) {
                        tma_load_2d(sA, &tmA, bar, kb * BK, m_blk * BM);
                    } else {
#pragma unroll
                        for (int c = 0; c < BM / 64; ++c)
                            tma_load_2d(sA + c * SLAB_BYTES, &tmA, bar, m_blk * BM + c * 64, kb * BK);
                    }
                    if constexpr (B_MAJOR == kMajorK) {
                        tma_load_2d(sB, &tmB, bar, kb * BK, n_blk * BN);
                    } else {
#pragma unroll
                        for (int c = 0; c < BN / 64; ++c)
                            tma_load_2d(sB + c * SLAB_BYTES, &tmB, bar, n_blk * BN + c * 64, kb * BK);
                    }
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (one thread) =====================
        constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MAJOR, B_MAJOR);
        uint32_t stage = 0, phase = 0, acc_stage = 0, acc_phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            mbar_wait(smem_u32(&tmem_empty_bar[acc_stage]), acc_phase ^ 1u);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc_stage * Cfg::ACC_STRIDE;
            for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                mbar_wait(smem_u32(&full_bar[stage]), phase);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t sA = smem_base + stage * Cfg::STAGE_BYTES;
                    const uint32_t sB = sA + Cfg::A_BYTES;
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        const uint64_t adesc =
                            (A_MAJOR == kMajorK)
                                ? make_smem_desc_sw128(sA + k * (UMMA_K * 2), 16, 1024)
                                : make_smem_desc_sw128(sA + k * (UMMA_K * 128), SLAB_BYTES, 1024);
                        const uint64_t bdesc =
                            (B_MAJOR == kMajorK)
                                ? make_smem_desc_sw128(sB + k * (UMMA_K * 2), 16, 1024)
                                : make_smem_desc_sw128(sB + k * (UMMA_K * 128), SLAB_BYTES, 1024);
                        umma_bf16_ss(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(smem_u32(&empty_bar[stage]));  // frees the smem slot when the MMAs retire
                    if (kb == p.num_k_blocks - 1) umma_commit(smem_u32(&tmem_full_bar[acc_stage]));
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
            if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1u; }
        }
    } else if (warp >= 4) {
        // ===================== epilogue: 8 warps.  TMEM lane quarter = warp % 4, column half = (warp - 4) / 4 =====
        constexpr int HALF_N = BN / 2;
        constexpr int CH = (HALF_N % 32 == 0) ? 32 : 16;
        constexpr int NCH = HALF_N / CH;
        const uint32_t q = warp & 3u;
        const int half = static_cast<int>(warp - 4) >> 2;
        const int ep_tid = static_cast<int>(threadIdx.x) - 128;
        const bool has_bias = (p.epi == DCAD_EPI_BIAS || p.epi == DCAD_EPI_BIAS_GELU);
        uint32_t acc_stage = 0, acc_phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const int m_blk = tile % p.num_m_blocks;
            const int n_blk = tile / p.num_m_blocks;
            const int n_base = n_blk * BN;
            const int row = m_blk * BM + q * 32 + lane;
            const bool row_ok = row < p.M;
            const int col_base = n_base + half * HALF_N;
            // stage this tile's bias while the MMAs are still running
            if (has_bias) {
                if (ep_tid < BN) s_bias[acc_stage][ep_tid] = (n_base + ep_tid < p.N) ? __ldg(p.bias + n_base + ep_tid) : 0.f;
                epi_bar_sync();
            }
            // prefetch the first aux chunk (dGELU pre-activations)
            uint4 aux_cur[CH / 8];
            const __nv_bfloat16* aux_row =
                (p.epi == DCAD_EPI_DGELU && row_ok) ? p.aux + static_cast<size_t>(row) * p.ldaux + col_base : nullptr;
            if (aux_row) {
#pragma unroll
                for (int j = 0; j < CH / 8; ++j)
                    aux_cur[j] = (col_base + j * 8 < p.N) ? __ldg(reinterpret_cast<const uint4*>(aux_row) + j)
                                                          : make_uint4(0, 0, 0, 0);
            }
            mbar_wait(smem_u32(&tmem_full_bar[acc_stage]), acc_phase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + acc_stage * Cfg::ACC_STRIDE + ((q * 32u) << 16) + half * HALF_N;
#pragma unroll 1
            for (int c = 0; c < NCH; ++c) {
                uint32_t r[CH];
                TmemChunk<CH>::ld(t_row + c * CH, r);
                // prefetch the next aux chunk while the TMEM load is in flight
                uint4 aux_nxt[CH / 8];
                if (aux_row && c + 1 < NCH) {
#pragma unroll
                    for (int j = 0; j < CH / 8; ++j)
                        aux_nxt[j] = (col_base + (c + 1) * CH + j * 8 < p.N)
                                         ? __ldg(reinterpret_cast<const uint4*>(aux_row + (c + 1) * CH) + j)
                                         : make_uint4(0, 0, 0, 0);
                }
                tmem_ld_wait();
                const int n0 = col_base + c * CH;
                if (row_ok && n0 < p.N) {
                    float v[CH];
#pragma unroll
                    for (int j = 0; j < CH; ++j) v[j] = __uint_as_float(r[j]);
                    if (has_bias) {
                        const float4* bs = reinterpret_cast<const float4*>(&s_bias[acc_stage][half * HALF_N + c * CH]);
#pragma unroll
                        for (int j = 0; j < CH; j += 4) {
                            const float4 b = bs[j >> 2];
                            v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
                        }
                    }
                    if (p.epi == DCAD_EPI_BIAS_GELU) {
                        if (p.aux_out != nullptr) {
                            __nv_bfloat16* ao = p.aux_out + static_cast<size_t>(row) * p.ldaux + n0;
#pragma unroll
                            for (int j = 0; j < CH; j += 8) {
                                if (n0 + j < p.N) {
                                    uint4 o;
                                    o.x = pack_bf16x2(v[j], v[j + 1]); o.y = pack_bf16x2(v[j + 2], v[j + 3]);
                                    o.z = pack_bf16x2(v[j + 4], v[j + 5]); o.w = pack_bf16x2(v[j + 6], v[j + 7]);
                                    *reinterpret_cast<uint4*>(ao + j) = o;
                                }
                            }
                        }
#pragma unroll
                        for (int j = 0; j < CH; ++j) v[j] = gelu_erf(v[j]);
                    } else if (p.epi == DCAD_EPI_DGELU) {
#pragma unroll
                        for (int j = 0; j < CH; j += 8) {
                            const uint4 a = aux_cur[j >> 3];
                            const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y);
                            const float2 a2 = unpack_bf16x2(a.z), a3 = unpack_bf16x2(a.w);
                            v[j] *= dgelu_erf(a0.x); v[j + 1] *= dgelu_erf(a0.y);
                            v[j + 2] *= dgelu_erf(a1.x); v[j + 3] *= dgelu_erf(a1.y);
                            v[j + 4] *= dgelu_erf(a2.x); v[j + 5] *= dgelu_erf(a2.y);
                            v[j + 6] *= dgelu_erf(a3.x); v[j + 7] *= dgelu_erf(a3.y);
                        }
                    }
                    if (p.out_fp32) {
                        float* d = reinterpret_cast<float*>(p.D) + static_cast<size_t>(row) * p.ldd + n0;
#pragma unroll
                        for (int j = 0; j < CH; j += 4) {
                            if (n0 + j < p.N) {
                                float4 o = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                                if (p.accumulate) {
                                    const float4 old = *reinterpret_cast<const float4*>(d + j);
                                    o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
                                }
                                *reinterpret_cast<float4*>(d + j) = o;
                            }
                        }
                    } else {
                        __nv_bfloat16* d =
                            reinterpret_cast<__nv_bfloat16*>(p.D) + static_cast<size_t>(row) * p.ldd + n0;
#pragma unroll
                        for (int j = 0; j < CH; j += 8) {
                            if (n0 + j < p.N) {
                                uint4 o;
                                o.x = pack_bf16x2(v[j], v[j + 1]); o.y = pack_bf16x2(v[j + 2], v[j + 3]);
                                o.z = pack_bf16x2(v[j + 4], v[j + 5]); o.w = pack_bf16x2(v[j + 6], v[j + 7]);
                                *reinterpret_cast<uint4*>(d + j) = o;
                            }
                        }
                    }
                }
                if (aux_row) {
#pragma unroll
                    for (int j = 0; j < CH / 8; ++j) aux_cur[j] = aux_nxt[j];
                }
            }
            tc_fence_before();
            mbar_arrive(smem_u32(&tmem_empty_bar[acc_stage]));
            if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1u; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------------
// host side: tensor-map cache + dispatch
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

struct MapKey {
    const void* ptr;
    uint64_t rows, cols, ld;
    uint32_t box_rows, box_cols;
    bool operator==(const MapKey& o) const {
        return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows &&
               box_cols == o.box_cols;
    }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        size_t h = reinterpret_cast<size_t>(k.ptr);
        h = h * 1000003u ^ k.rows; h = h * 1000003u ^ k.cols; h = h * 1000003u ^ k.ld;
        h = h * 1000003u ^ k.box_rows; h = h * 1000003u ^ k.box_cols;
        return h;
    }
};

// 2D bf16 row-major [rows][ld] tensor, box = box_rows x box_cols (box_cols == 64 => 128-byte swizzle span)
int make_tensor_map_bf16_2d(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                            uint32_t box_rows, uint32_t box_cols) {
    static std::mutex mu;
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    MapKey key{ptr, rows, cols, ld, box_rows, box_cols};
    {
        std::lock_guard<std::mutex> g(mu);
        auto it = cache.find(key);
        if (it != cache.end()) { *out = it->second; return 0; }
    }
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return DCAD_ERR_NO_DRIVER_ENTRY;
    if ((reinterpret_cast<uintptr_t>(ptr) & 15u) || (ld * 2) % 16 != 0) return DCAD_ERR_BAD_ALIGNMENT;
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstride[1] = {ld * 2};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return DCAD_ERR_TENSOR_MAP;
    std::lock_guard<std::mutex> g(mu);
    if (cache.size() > 4096) cache.clear();
    cache.emplace(key, *out);
    return 0;
}

// 3D bf16 tensor [d2][d1][d0] (d0 contiguous), box = 1 x b1 x b0; used by the attention kernels so that rows past the
// end of a sequence are zero-filled instead of spilling into the next sample.
int make_tensor_map_bf16_3d(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t ld1,
                            uint64_t ld2, uint32_t b0, uint32_t b1) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return DCAD_ERR_NO_DRIVER_ENTRY;
    if ((reinterpret_cast<uintptr_t>(ptr) & 15u) || (ld1 * 2) % 16 != 0 || (ld2 * 2) % 16 != 0)
        return DCAD_ERR_BAD_ALIGNMENT;
    cuuint64_t gdim[3] = {d0, d1, d2};
    cuuint64_t gstride[2] = {ld1 * 2, ld2 * 2};
    cuuint32_t box[3] = {b0, b1, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : DCAD_ERR_TENSOR_MAP;
}

static int num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    }
    return n;
}

template <int BN, int A_MAJOR, int B_MAJOR>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int max_ctas,
                       cudaStream_t stream) {
    using Cfg = GemmCfg<BN>;
    auto kern = gemm_bf16_tcgen05_kernel<BN, A_MAJOR, B_MAJOR>;
    static bool attr_set = false;
    if (!attr_set) {
        DCAD_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
        attr_set = true;
    }
    int grid = p.num_m_blocks * p.num_n_blocks;
    int cap = max_ctas > 0 ? max_ctas : num_sms();
    if (grid > cap) grid = cap;
    kern<<<grid, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(tmA, tmB, p);
    return (int)cudaGetLastError();
}

}  // namespace dcad

using namespace dcad;

extern "C" int dcad_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int a_major, int b_major,
                              int lda, int ldb, int ldd, const float* bias, const void* aux, void* aux_out, int ldaux,
                              int epi, int out_fp32, int accumulate, int block_n, int max_ctas, void* stream) {
    if (M <= 0 || N <= 0 || K <= 0) return DCAD_ERR_BAD_SHAPE;
    if ((N % 8) || (ldd % 8) || (lda % 8) || (ldb % 8)) return DCAD_ERR_BAD_ALIGNMENT;
    if ((epi == DCAD_EPI_BIAS || epi == DCAD_EPI_BIAS_GELU) && bias == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (epi == DCAD_EPI_DGELU && aux == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (accumulate && !out_fp32) return DCAD_ERR_BAD_ARGUMENT;

    int bn = block_n;
    if (bn <= 0) bn = (N % 128 == 0) ? 128 : 64;

    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.ldd = ldd; p.ldaux = ldaux > 0 ? ldaux : ldd;
    p.bias = bias;
    p.aux = reinterpret_cast<const __nv_bfloat16*>(aux);
    p.aux_out = reinterpret_cast<__nv_bfloat16*>(aux_out);
    p.D = D; p.epi = epi; p.out_fp32 = out_fp32; p.accumulate = accumulate;
    p.num_m_blocks = (M + BM - 1) / BM;
    p.num_n_blocks = (N + bn - 1) / bn;
    p.num_k_blocks = (K + BK - 1) / BK;

    CUtensorMap tmA, tmB;
    int rc;
    if (a_major == kMajorK) rc = make_tensor_map_bf16_2d(&tmA, A, M, K, lda, BM, BK);   // A stored [M][lda]
    else                    rc = make_tensor_map_bf16_2d(&tmA, A, K, M, lda, BK, 64);   // A stored [K][lda]
    if (rc) return rc;
    if (b_major == kMajorK) rc = make_tensor_map_bf16_2d(&tmB, B, N, K, ldb, bn, BK);   // B stored [N][ldb]
    else                    rc = make_tensor_map_bf16_2d(&tmB, B, K, N, ldb, BK, 64);   // B stored [K][ldb]
    if (rc) return rc;

    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
#define DCAD_GEMM_CASE(BN_, AM_, BM_) \
    if (bn == BN_ && a_major == AM_ && b_major == BM_) return launch_gemm<BN_, AM_, BM_>(tmA, tmB, p, max_ctas, s);
    DCAD_GEMM_CASE(32, 0, 0)
    DCAD_GEMM_CASE(64, 0, 0)
    DCAD_GEMM_CASE(96, 0, 0)
    DCAD_GEMM_CASE(128, 0, 0)
    DCAD_GEMM_CASE(192, 0, 0)
    DCAD_GEMM_CASE(256, 0, 0)
    DCAD_GEMM_CASE(64, 0, 1)
    DCAD_GEMM_CASE(128, 0, 1)
    DCAD_GEMM_CASE(192, 0, 1)
    DCAD_GEMM_CASE(256, 0, 1)
    DCAD_GEMM_CASE(64, 1, 1)
    DCAD_GEMM_CASE(128, 1, 1)
    DCAD_GEMM_CASE(192, 1, 1)
    DCAD_GEMM_CASE(256, 1, 1)
#undef DCAD_GEMM_CASE
    return DCAD_ERR_UNSUPPORTED_TILE;
}
