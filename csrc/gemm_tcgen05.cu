// Persistent, warp-specialised bf16 GEMM for sm_100a: TMA -> 128B-swizzled smem ring -> tcgen05.mma with fp32
// accumulators in TMEM (double-buffered) -> tcgen05.ld epilogue (bias / GELU / dGELU) -> swizzled smem staging ->
// TMA store (or TMA reduce-add for split-K weight gradients).
//
// Replaces the cuBLAS GEMMs that the reference reaches through HF DistilBERT's nn.Linear layers
// (reference client1.py:61 -> transformers modeling_distilbert.py:187-189,205,224-226) and their autograd
// backward (client1.py:109).  Three operand-layout combinations cover forward, dgrad and wgrad without ever
// materialising a transposed tensor:
//     forward  Y[T,N]  = X[T,K]  W[N,K]^T        A K-major , B K-major
//     dgrad    dX[T,K] = dY[T,N] W[N,K]          A K-major , B MN-major
//     wgrad    dW[N,K] = dY[T,N]^T X[T,K]        A MN-major, B MN-major   (fp32 out, optional split-K)
//
// Roles (640 threads): warp 0 = TMA producer, warp 1 = single-thread MMA issuer, warp 2 = TMEM owner,
// warps 4..19 = epilogue (TMEM lane quarter = warp % 4).  Three mbarrier pipelines: smem full/empty, TMEM full/empty,
// plus bulk-group tracking of the TMA stores.
//
// Two epilogues are compiled (template parameter EPI, chosen per call by dcad_gemm_set_epilogue / DCAD_GEMM_EPI):
//   EPI 0 "block": the 16 warps fill one [128 rows][128 B] slab together (the four warps of a lane quarter split its
//                  64 columns), two block-wide barriers per slab, one thread issues the slab's TMA store, 2 slabs in
//                  flight.
//   EPI 1 "warp" : every epilogue warp owns a private [32 rows][128 B] staging box and issues its own TMA stores
//                  (32-row boxes), so the epilogue has no block-wide barrier and no store waits for another warp's.
//                  It also carries the 128 x 384 tile (BLOCK_N 384): two N = 192 accumulators fed by two MMAs per
//                  k-step, which turns the N = 3072 GEMMs from 1.3 waves of 148 SMs (192 tiles of 128 x 256) into one
//                  (128 tiles); its staging boxes alias the operand ring, idle once the tile's last MMA has retired.
//
// Measured on B200 (profiles/gemm_mainloop_anatomy.jsonl): a tcgen05.mma of N <= 128 costs ~80-110 cycles to issue regardless of
// N, only N = 256 runs at the tensor-pipe rate (128 cycles) -- hence wide tiles, and split-K (not narrower tiles)
// when a problem has too few tiles for 148 SMs.
#include <stdlib.h>

#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "../include/dcad_b200.h"

namespace dcad {

constexpr int BM = 128;  // UMMA M (cta_group::1, full-rate shape)
constexpr int BK = 64;   // one 128-byte swizzle atom of bf16 along the reduction dim per stage
constexpr int UMMA_K = 16;
constexpr int SLAB_BYTES = BK * 128;       // one [BK][64] MN-major operand slab
constexpr int STAGING_BYTES = BM * 128;    // one [128 rows][128 B] output slab (64 bf16 or 32 fp32 columns)
constexpr int NUM_STAGING = 4;             // D ping-pong + aux_out ping-pong
constexpr int BOX_ROWS = 32;               // EPI 1: output leaves in [32 rows][128 B] boxes (one TMEM lane quarter)
constexpr int BOX_BYTES = BOX_ROWS * 128;  // 4 KB staging box
constexpr int EPI_WARPS = 16;

struct GemmParams {
    int M, N, K;
    int ldaux;
    const float* bias;
    const int* row_lens;       // optional, rows come in sequences of BM: row r of tile m is padding when r >= row_lens[m]
    float* colsum;             // optional [N] fp32: += column sums of the bf16 output (bias gradient of the consumer)
    const __nv_bfloat16* aux;  // EPI_DGELU: pre-activation (same shape as D)
    int has_aux_out;           // EPI_BIAS_GELU: also store the pre-activation through tmAux
    int epi;
    int out_fp32;              // 0 = bf16, 1 = fp32 store, 2 = fp32 TMA reduce-add even without split-K (D may be a peer GPU's buffer)
    int split_k;               // > 1: fp32 output accumulated with TMA reduce-add (output must be zero-initialised)
    int ws_row_stride;         // > 0: workspace mode, split s stores its fp32 partial tile at row s*ws_row_stride + m
    int num_m_blocks, num_n_blocks, num_k_blocks, kb_per_split;
    unsigned long long* timing;  // optional: per-role globaltimer stamps of CTA 0 (profiling aid), else null
};

__host__ __device__ constexpr int pow2_ceil(int v) {
    int p = 32;
    while (p < v) p <<= 1;
    return p;
}

// CG = 1: one CTA per 128 x BN tile.  CG = 2: a CTA pair (cluster of 2, one TPC) per 256 x BN tile with
// tcgen05.mma.cta_group::2 -- each CTA stages its own 128 A rows and only BN/2 of the B rows, so the operand bytes
// read from smem per MMA (and fetched from L2 per tile) drop by a third for BN = 256.
// Up to G independent problems of one tile configuration run in ONE persistent launch (G = 4: the four
// weight-gradient GEMMs of a transformer block): one prologue and one pipeline fill instead of four, and their work
// units are spread over the SMs together.  Unit u belongs to the problem i with unit_begin[i] <= u < unit_begin[i+1].
template <int G>
struct GemmGroup {
    CUtensorMap tmA[G], tmB[G], tmD[G], tmAux[G];
    GemmParams p[G];
    int unit_begin[G + 1];
    int n;
};

// EPI 0: BN <= 256, one accumulator of N = BN per tile, two TMEM accumulator stages (the epilogue of a tile overlaps
// the main loop of the next), four 16 KB staging slabs beside the operand ring.
// EPI 1: per-warp 4 KB staging boxes.  BN <= 256 as above; BN = 384: two N = 192 accumulators per tile fill the TMEM
// (one accumulator stage), and the staging boxes (two per warp) alias the operand ring -- the producer does not refill
// it before the epilogue's stores have read it (stg_free_bar).
template <int BN, int CG = 1, int EPI = 0>
struct GemmCfg {
    static constexpr int NSUB = BN > 256 ? 2 : 1;   // MMAs (accumulators) per k-step
    static constexpr int SUBN = BN / NSUB;          // UMMA N
    static constexpr bool ALIAS = BN > 256 || EPI == 2;  // EPI 2: per-warp epilogue, staging always in the ring (deeper ring)
    static constexpr int ACC_STAGES = BN > 256 ? 1 : 2;
    static constexpr int NBUF = ALIAS ? 2 : 1;      // EPI 1: staging boxes per epilogue warp
    static constexpr int STAGING_TOTAL = EPI == 0 ? NUM_STAGING * STAGING_BYTES : EPI_WARPS * NBUF * BOX_BYTES;
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_ROWS = BN / CG;          // B rows (N) staged by one CTA
    static constexpr int B_BOX_ROWS = B_ROWS > 256 ? B_ROWS / 2 : B_ROWS;  // a TMA box has at most 256 rows
    static constexpr int B_BYTES = B_ROWS * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    // 227 KB per CTA minus the staging area, the alignment slack and the static part (bias copy + barriers)
    static constexpr int RING_BYTES = 232448 - (ALIAS ? 0 : STAGING_TOTAL) - 1024 - (2 * BN * 4 + 512);
    static constexpr int STAGES_RAW = RING_BYTES / STAGE_BYTES;
    static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
    static constexpr int ACC_STRIDE = pow2_ceil(BN);  // TMEM columns per accumulator stage
    static constexpr int TMEM_COLS = ACC_STAGES * ACC_STRIDE;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + (ALIAS ? 0 : STAGING_TOTAL) + 1024;
    static_assert(EPI >= 1 || BN <= 256, "the 384-wide tile needs the per-warp epilogue");
    static_assert(TMEM_COLS <= 512, "accumulators exceed the 512 TMEM columns");
    static_assert(!ALIAS || STAGES * STAGE_BYTES >= STAGING_TOTAL, "staging does not fit in the ring");
};

__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define DCAD_STAMP(i) do { if (g.p[0].timing && blockIdx.x == 0 && lane == 0) g.p[0].timing[i] = gtimer(); } while (0)

constexpr int GEMM_THREADS = 640;  // 4 control warps + 16 epilogue warps
constexpr int EPI_THREADS = 512;

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 512;" ::: "memory"); }

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// selects the problem a work unit belongs to and rebases the unit index
#define DCAD_SELECT_PROBLEM(unit)                                                             \
    int prob = 0;                                                                             \
    if constexpr (G > 1) {                                                                    \
        _Pragma("unroll") for (int i = 1; i < G; ++i)                                         \
            prob += (i < g.n && (unit) >= g.unit_begin[i]) ? 1 : 0;                           \
    }                                                                                         \
    const GemmParams& p = g.p[prob];                                                          \
    const CUtensorMap& tmA = g.tmA[prob]; const CUtensorMap& tmB = g.tmB[prob];               \
    const CUtensorMap& tmD = g.tmD[prob]; const CUtensorMap& tmAux = g.tmAux[prob];           \
    (void)tmA; (void)tmB; (void)tmD; (void)tmAux;                                             \
    const int lu = (unit) - g.unit_begin[prob];

template <int BN, int A_MAJOR, int B_MAJOR, int CG, int G, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ GemmGroup<G> g) {
    using Cfg = GemmCfg<BN, CG, EPI>;
    constexpr int STAGES = Cfg::STAGES;
    constexpr int ACC_STAGES = Cfg::ACC_STAGES;
    static_assert(BN % 64 == 0 && BN >= 64 && BN <= 512, "BLOCK_N must be a multiple of 64 in [64,512]");
    static_assert(Cfg::SUBN % 16 == 0 && Cfg::SUBN <= 256, "UMMA N is a multiple of 16 up to 256");
    static_assert(CG == 1 || CG == 2, "cta_group is 1 or 2");
    static_assert(CG == 1 || BN <= 256, "CTA-pair tiles carry one accumulator");
    static_assert(CG == 1 || B_MAJOR == kMajorK || Cfg::B_ROWS % 64 == 0, "MN-major B halves are whole 64-column slabs");
    static_assert(STAGES >= 2, "need at least a double-buffered operand ring");

    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[STAGES];
    __shared__ __align__(8) uint64_t empty_bar[STAGES];
    __shared__ __align__(8) uint64_t tmem_full_bar[2];
    __shared__ __align__(8) uint64_t tmem_empty_bar[2];
    __shared__ __align__(8) uint64_t stg_free_bar;   // ALIAS tiles only: the epilogue's stores have read the ring
    __shared__ __align__(16) float s_bias[2][BN];
    __shared__ uint32_t tmem_base_smem;

    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = lane_id();
    const uint32_t cta_rank = (CG == 2) ? cluster_ctarank() : 0u;  // 0 = leader (issues the MMAs), 1 = peer
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B wants 1024-B alignment
    const uint32_t staging_base = Cfg::ALIAS ? smem_base : smem_base + STAGES * Cfg::STAGE_BYTES;

    if (warp == 0 && lane < static_cast<uint32_t>(g.n)) {
        tma_prefetch_desc(&g.tmA[lane]);
        tma_prefetch_desc(&g.tmB[lane]);
        tma_prefetch_desc(&g.tmD[lane]);
        if (g.p[lane].has_aux_out) tma_prefetch_desc(&g.tmAux[lane]);
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
            // CG 1: every epilogue thread arrives; CG 2: one lane per epilogue warp of BOTH CTAs arrives on the leader
            mbar_init(smem_u32(&tmem_empty_bar[s]), CG == 1 ? EPI_THREADS : 2 * (EPI_THREADS / 32));
        }
        mbar_init(smem_u32(&stg_free_bar), EPI_WARPS);
        fence_barrier_init();
    }
    if (warp == 2) {
        if constexpr (CG == 1) {
            tmem_alloc(smem_u32(&tmem_base_smem), Cfg::TMEM_COLS);
            tmem_relinquish();
        } else {  // collective over the pair: the same columns in both SMs
            tmem_alloc_cg2(smem_u32(&tmem_base_smem), Cfg::TMEM_COLS);
            tmem_relinquish_cg2();
        }
    }
    pdl_trigger();  // the next kernel may start its own prologue as soon as every CTA of this grid is resident
    tc_fence_before();
    if constexpr (CG == 1) __syncthreads();
    else cluster_sync_all();  // the peer's barriers must be initialised before anything signals them
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;
    pdl_wait();  // everything above overlapped the previous kernel's tail; global memory is touched only below

    // work unit = (output tile of BM*CG x BN, k-split); units are strided over the persistent CTAs (CTA pairs)
    const int num_units = g.unit_begin[g.n];
    const int unit0 = static_cast<int>(blockIdx.x) / CG, unit_step = static_cast<int>(gridDim.x) / CG;
    if (warp == 0) DCAD_STAMP(0);

    if (warp == 0) {
        // ===================== TMA producer =====================
        uint32_t stage = 0, phase = 0, tiles_done = 0;
        for (int unit = unit0; unit < num_units; unit += unit_step) {
            DCAD_SELECT_PROBLEM(unit)
            const int tile = lu / p.split_k, split = lu % p.split_k;
            const int m_blk = (tile % p.num_m_blocks) * CG + static_cast<int>(cta_rank);
            const int n_blk = tile / p.num_m_blocks;
            const int n_row0 = n_blk * BN + static_cast<int>(cta_rank) * Cfg::B_ROWS;  // this CTA's share of B
            const int kb0 = split * p.kb_per_split;
            const int kb1 = min(p.num_k_blocks, kb0 + p.kb_per_split);
            if constexpr (Cfg::ALIAS) {  // the previous tile's output is still being read out of the ring
                if (tiles_done > 0) mbar_wait(smem_u32(&stg_free_bar), (tiles_done - 1u) & 1u);
                ++tiles_done;
            }
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
                if (lane == 0) {
                    const uint32_t bar = smem_u32(&full_bar[stage]);
                    const uint32_t sA = smem_base + stage * Cfg::STAGE_BYTES;
                    const uint32_t sB = sA + Cfg::A_BYTES;
                    // CG 2: both CTAs' loads are credited to the leader's barrier, which expects both stages
                    auto load = [&](uint32_t dst, const CUtensorMap* tm, int c0, int c1) {
                        if constexpr (CG == 1) tma_load_2d(dst, tm, bar, c0, c1);
                        else tma_load_2d_cg2(dst, tm, bar, c0, c1);
                    };
                    if (CG == 1 || cta_rank == 0) mbar_arrive_expect_tx(bar, CG * Cfg::STAGE_BYTES);
                    if constexpr (A_MAJOR == kMajorK) {
                        load(sA, &tmA, kb * BK, m_blk * BM);
                    } else {
#pragma unroll
                        for (int c = 0; c < BM / 64; ++c) load(sA + c * SLAB_BYTES, &tmA, m_blk * BM + c * 64, kb * BK);
                    }
                    if constexpr (B_MAJOR == kMajorK) {
#pragma unroll
                        for (int h = 0; h < Cfg::B_ROWS / Cfg::B_BOX_ROWS; ++h)
                            load(sB + h * Cfg::B_BOX_ROWS * 128, &tmB, kb * BK, n_row0 + h * Cfg::B_BOX_ROWS);
                    } else {
#pragma unroll
                        for (int c = 0; c < Cfg::B_ROWS / 64; ++c)
                            load(sB + c * SLAB_BYTES, &tmB, n_row0 + c * 64, kb * BK);
                    }
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
        }
        DCAD_STAMP(1);
    } else if (warp == 1 && (CG == 1 || cta_rank == 0)) {
        // ===================== MMA issuer (one thread; of the leader CTA when CG = 2) =====================
        // Descriptors are built once; per MMA only the 14-bit start-address field (16-byte units) advances.
        constexpr uint32_t idesc = make_idesc_bf16(BM * CG, Cfg::SUBN, A_MAJOR, B_MAJOR);
        // second accumulator of a wide tile: B rows SUBN.. of the stage (K-major: SUBN rows of 128 B; MN-major:
        // SUBN/64 slabs of 8 KB) -- the same byte offset either way
        constexpr uint32_t B_SUBSTEP = (Cfg::SUBN * 128) >> 4;
        constexpr uint32_t A_KSTEP = (A_MAJOR == kMajorK ? UMMA_K * 2 : UMMA_K * 128) >> 4;
        constexpr uint32_t B_KSTEP = (B_MAJOR == kMajorK ? UMMA_K * 2 : UMMA_K * 128) >> 4;
        const uint64_t a_desc0 = (A_MAJOR == kMajorK) ? make_smem_desc_sw128(smem_base, 16, 1024)
                                                      : make_smem_desc_sw128(smem_base, SLAB_BYTES, 1024);
        const uint64_t b_desc0 = (B_MAJOR == kMajorK) ? make_smem_desc_sw128(smem_base + Cfg::A_BYTES, 16, 1024)
                                                      : make_smem_desc_sw128(smem_base + Cfg::A_BYTES, SLAB_BYTES, 1024);
        const uint32_t a_hi = static_cast<uint32_t>(a_desc0 >> 32), b_hi = static_cast<uint32_t>(b_desc0 >> 32);
        const uint32_t a_lo0 = static_cast<uint32_t>(a_desc0), b_lo0 = static_cast<uint32_t>(b_desc0);
        uint32_t stage = 0, phase = 0, acc_stage = 0, acc_phase = 0;
        bool first = true;
        for (int unit = unit0; unit < num_units; unit += unit_step) {
            DCAD_SELECT_PROBLEM(unit)
            const int split = lu % p.split_k;
            const int kb0 = split * p.kb_per_split;
            const int kb1 = min(p.num_k_blocks, kb0 + p.kb_per_split);
            mbar_wait(smem_u32(&tmem_empty_bar[acc_stage]), acc_phase ^ 1u);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc_stage * Cfg::ACC_STRIDE;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(smem_u32(&full_bar[stage]), phase);
                tc_fence_after();
                if (first) { DCAD_STAMP(2); first = false; }
                if (lane == 0) {
                    const uint32_t a_lo = a_lo0 + stage * (Cfg::STAGE_BYTES >> 4);
                    const uint32_t b_lo = b_lo0 + stage * (Cfg::STAGE_BYTES >> 4);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        const uint64_t adesc = (static_cast<uint64_t>(a_hi) << 32) | (a_lo + k * A_KSTEP);
                        const uint32_t acc = (kb != kb0 || k != 0) ? 1u : 0u;
#pragma unroll
                        for (int h = 0; h < Cfg::NSUB; ++h) {
                            const uint64_t bdesc =
                                (static_cast<uint64_t>(b_hi) << 32) | (b_lo + k * B_KSTEP + h * B_SUBSTEP);
                            if constexpr (CG == 1) umma_bf16_ss(tmem_d + h * Cfg::SUBN, adesc, bdesc, idesc, acc);
                            else umma_bf16_ss_cg2(tmem_d, adesc, bdesc, idesc, acc);
                        }
                    }
                    if constexpr (CG == 1) {
                        umma_commit(smem_u32(&empty_bar[stage]));  // frees the smem slot when the MMAs retire
                        if (kb == kb1 - 1) umma_commit(smem_u32(&tmem_full_bar[acc_stage]));
                    } else {  // both CTAs' producers / epilogues are released
                        umma_commit_cg2(smem_u32(&empty_bar[stage]));
                        if (kb == kb1 - 1) umma_commit_cg2(smem_u32(&tmem_full_bar[acc_stage]));
                    }
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
            if (++acc_stage == ACC_STAGES) { acc_stage = 0; acc_phase ^= 1u; }
        }
        DCAD_STAMP(3);
    } else if (warp >= 4) {
      if constexpr (EPI == 0) {
            // ===================== epilogue (16 warps) =====================
            // Output leaves through [128 rows][128 B] slabs staged in 128B-swizzled smem and written by TMA: 64 bf16 or
            // 32 fp32 columns per slab.  Warp (q, part) owns rows 32q..32q+31 and quarter `part` of every slab, so a
            // thread converts 16 bf16 (or moves 8 fp32) columns per slab: the GELU / dGELU epilogues are ALU-bound and
            // need all the lanes they can get.
            const uint32_t q = warp & 3u;
            const int part = static_cast<int>(warp - 4) >> 2;  // 0..3
            const int ep_tid = static_cast<int>(threadIdx.x) - 128;
            const int r_in_tile = q * 32 + lane;
            uint32_t acc_stage = 0, acc_phase = 0, slab_ctr = 0;
            for (int unit = unit0; unit < num_units; unit += unit_step) {
                DCAD_SELECT_PROBLEM(unit)
                const bool has_bias = (p.epi == DCAD_EPI_BIAS || p.epi == DCAD_EPI_BIAS_GELU);
                const int tile = lu / p.split_k;
                const int m_blk = (tile % p.num_m_blocks) * CG + static_cast<int>(cta_rank);
                const int n_blk = tile / p.num_m_blocks;
                const int n_base = n_blk * BN;
                const bool tile_ok = m_blk * BM < p.M;  // the peer's half of a pair tile may lie wholly past M
                const int out_row0 = (lu % p.split_k) * p.ws_row_stride + m_blk * BM;  // ws_row_stride == 0 unless workspace mode
                const int row = m_blk * BM + r_in_tile;
                const bool row_ok = row < p.M;
                // Padded token rows (tile = one 128-token sequence): the GELU / dGELU arithmetic is skipped for them and
                // zeros are written -- nothing downstream reads them and their gradients are exactly zero.  The epilogue
                // is ALU-bound (two MUFU per element), and a third of the rows of a flow-sentence batch are padding.
                const bool row_live = row_ok && (p.row_lens == nullptr || r_in_tile < __ldg(p.row_lens + m_blk));
                if (has_bias) {  // stage this tile's bias while the MMAs are still running
                    if (ep_tid < BN)
                        s_bias[acc_stage][ep_tid] = (n_base + ep_tid < p.N) ? __ldg(p.bias + n_base + ep_tid) : 0.f;
                }
                const __nv_bfloat16* aux_row =
                    (p.epi == DCAD_EPI_DGELU && row_ok) ? p.aux + static_cast<size_t>(row) * p.ldaux + n_base : nullptr;
                mbar_wait(smem_u32(&tmem_full_bar[acc_stage]), acc_phase);
                tc_fence_after();
                if (warp == 4) DCAD_STAMP(4);
                const uint32_t t_row = tmem_base + acc_stage * Cfg::ACC_STRIDE + ((q * 32u) << 16);

                if (!p.out_fp32) {
                    // ---- bf16 output: 64-column slabs, this thread converts 16 columns per slab ----
    #pragma unroll 1
                    for (int s = 0; s < BN / 64; ++s) {
                        const uint32_t buf = slab_ctr & 1u;
                        const int c0 = s * 64 + part * 16;  // first accumulator column handled by this thread
                        uint32_t r[16];
                        tmem_ld_32x32b_x16(t_row + c0, r);
                        uint4 av[2];
    #pragma unroll
                        for (int j = 0; j < 2; ++j)
                            av[j] = (aux_row && n_base + c0 + j * 8 < p.N)
                                        ? __ldg(reinterpret_cast<const uint4*>(aux_row + c0) + j) : make_uint4(0, 0, 0, 0);
                        if (ep_tid == 0) bulk_wait_group_read<1>();  // the store that last read staging[buf] is done
                        epi_bar_sync();                               // (also publishes s_bias on the first slab)
                        tmem_ld_wait();
                        float v[16];
    #pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
                        if (has_bias) {
                            const float4* bs = reinterpret_cast<const float4*>(&s_bias[acc_stage][c0]);
    #pragma unroll
                            for (int j = 0; j < 16; j += 4) {
                                const float4 b = bs[j >> 2];
                                v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
                            }
                        }
                        const uint32_t sD = staging_base + buf * STAGING_BYTES;
                        if (p.epi == DCAD_EPI_BIAS_GELU) {
                            if (p.has_aux_out) {
                                const uint32_t sX = staging_base + (2 + buf) * STAGING_BYTES;
    #pragma unroll
                                for (int j = 0; j < 2; ++j)
                                    st_shared_v4(sX + sw128_offset(r_in_tile, part * 2 + j),
                                                 pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                                                 pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
                            }
                            if (row_live) {
    #pragma unroll
                                for (int j = 0; j < 16; ++j) v[j] = gelu_erf(v[j]);
                            } else {
    #pragma unroll
                                for (int j = 0; j < 16; ++j) v[j] = 0.f;
                            }
                        } else if (p.epi == DCAD_EPI_DGELU && !row_live) {
    #pragma unroll
                            for (int j = 0; j < 16; ++j) v[j] = 0.f;
                        } else if (p.epi == DCAD_EPI_DGELU) {
    #pragma unroll
                            for (int j = 0; j < 2; ++j) {
                                const float2 a0 = unpack_bf16x2(av[j].x), a1 = unpack_bf16x2(av[j].y);
                                const float2 a2 = unpack_bf16x2(av[j].z), a3 = unpack_bf16x2(av[j].w);
                                v[8 * j] *= dgelu_erf(a0.x); v[8 * j + 1] *= dgelu_erf(a0.y);
                                v[8 * j + 2] *= dgelu_erf(a1.x); v[8 * j + 3] *= dgelu_erf(a1.y);
                                v[8 * j + 4] *= dgelu_erf(a2.x); v[8 * j + 5] *= dgelu_erf(a2.y);
                                v[8 * j + 6] *= dgelu_erf(a3.x); v[8 * j + 7] *= dgelu_erf(a3.y);
                            }
                        }
    #pragma unroll
                        for (int j = 0; j < 2; ++j)
                            st_shared_v4(sD + sw128_offset(r_in_tile, part * 2 + j),
                                         pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                                         pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
                        if (p.colsum) {
                            // bias gradient of the layer that consumes this output: the warp's 32 rows x 16 columns are
                            // reduced with a shuffle butterfly, one fp32 atomic per column and warp
                            if (!row_ok) {
    #pragma unroll
                                for (int j = 0; j < 16; ++j) v[j] = 0.f;
                            }
                            warp_colsum<16>(v, lane);
                            const int col = n_base + c0 + static_cast<int>(lane);
                            if (lane < 16 && col < p.N) atomicAdd(p.colsum + col, v[0]);
                        }
                        fence_proxy_async_smem();
                        epi_bar_sync();
                        if (ep_tid == 0) {
                            if (tile_ok && n_base + s * 64 < p.N) {
                                tma_store_2d(&tmD, sD, n_base + s * 64, m_blk * BM);
                                if (p.epi == DCAD_EPI_BIAS_GELU && p.has_aux_out)
                                    tma_store_2d(&tmAux, staging_base + (2 + buf) * STAGING_BYTES, n_base + s * 64, m_blk * BM);
                            }
                            bulk_commit_group();
                        }
                        ++slab_ctr;
                    }
                } else {
                    // ---- fp32 output (weight gradients): 32-column slabs, this thread moves 8 columns per slab ----
    #pragma unroll 1
                    for (int s = 0; s < BN / 32; ++s) {
                        const uint32_t buf = slab_ctr & 1u;
                        const int c0 = s * 32 + part * 8;
                        uint32_t r[8];
                        tmem_ld_32x32b_x8(t_row + c0, r);
                        if (ep_tid == 0) bulk_wait_group_read<1>();
                        epi_bar_sync();
                        tmem_ld_wait();
                        const uint32_t sD = staging_base + buf * STAGING_BYTES;
    #pragma unroll
                        for (int j = 0; j < 2; ++j)
                            st_shared_v4(sD + sw128_offset(r_in_tile, part * 2 + j), r[4 * j], r[4 * j + 1], r[4 * j + 2],
                                         r[4 * j + 3]);
                        fence_proxy_async_smem();
                        epi_bar_sync();
                        if (ep_tid == 0) {
                            if (tile_ok && n_base + s * 32 < p.N) {
                                if ((p.split_k > 1 && p.ws_row_stride == 0) || p.out_fp32 == 2) tma_reduce_add_2d(&tmD, sD, n_base + s * 32, out_row0);
                                else tma_store_2d(&tmD, sD, n_base + s * 32, out_row0);
                            }
                            bulk_commit_group();
                        }
                        ++slab_ctr;
                    }
                }
                tc_fence_before();
                if constexpr (CG == 1) {
                    mbar_arrive(smem_u32(&tmem_empty_bar[acc_stage]));
                } else {  // the leader's MMA thread waits for the epilogues of both CTAs
                    __syncwarp();
                    if (lane == 0) mbar_arrive_remote(smem_u32(&tmem_empty_bar[acc_stage]), 0);
                }
                if (warp == 4) DCAD_STAMP(5);
                if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1u; }
            }
            if (ep_tid == 0) bulk_wait_group<0>();  // all output writes have landed before the CTA retires
    
      } else {
        // ===================== epilogue, EPI 1 (16 independent warps) =====================
        // Warp (q, part): TMEM lanes (= tile rows) 32q..32q+31, and the 64-column (bf16 out) or 32-column (fp32 out)
        // slabs part, part+4, ... of the tile.  A slab leaves as one [32 rows][128 B] box: tcgen05.ld -> registers
        // (bias / GELU / dGELU) -> the warp's private swizzled staging box -> one TMA store issued by lane 0.  Nothing
        // here synchronises with another warp, and a box is reused only after its own store has read it.
        const uint32_t q = warp & 3u;
        const int part = static_cast<int>(warp - 4) >> 2;  // 0..3
        const int ep_tid = static_cast<int>(threadIdx.x) - 128;
        const int r_in_tile = q * 32 + lane;
        const uint32_t stg0 = staging_base + (warp - 4) * (Cfg::NBUF * BOX_BYTES);
        uint32_t acc_stage = 0, acc_phase = 0, st_ctr = 0, tile_ctr = 0;
        auto box_acquire = [&]() -> uint32_t {  // a staging box that no TMA store is still reading
            if (lane == 0) bulk_wait_group_read<Cfg::NBUF - 1>();
            __syncwarp();
            return stg0 + (st_ctr % Cfg::NBUF) * BOX_BYTES;
        };
        for (int unit = unit0; unit < num_units; unit += unit_step, ++tile_ctr) {
            DCAD_SELECT_PROBLEM(unit)
            const bool has_bias = (p.epi == DCAD_EPI_BIAS || p.epi == DCAD_EPI_BIAS_GELU);
            const int tile = lu / p.split_k;
            const int m_blk = (tile % p.num_m_blocks) * CG + static_cast<int>(cta_rank);
            const int n_blk = tile / p.num_m_blocks;
            const int n_base = n_blk * BN;
            const bool tile_ok = m_blk * BM < p.M;  // the peer's half of a pair tile may lie wholly past M
            const int box_row0 = m_blk * BM + static_cast<int>(q) * BOX_ROWS;
            const int out_row0 = (lu % p.split_k) * p.ws_row_stride + box_row0;  // ws_row_stride == 0 unless workspace mode
            const int row = m_blk * BM + r_in_tile;
            const bool row_ok = row < p.M;
            const bool row_live = row_ok && (p.row_lens == nullptr || r_in_tile < __ldg(p.row_lens + m_blk));
            float* bias_s = s_bias[tile_ctr & 1u];
            if (has_bias) {  // stage this tile's bias while the MMAs are still running
                if (ep_tid < BN) bias_s[ep_tid] = (n_base + ep_tid < p.N) ? __ldg(p.bias + n_base + ep_tid) : 0.f;
                // publishes the bias; no warp can be two tiles ahead of another (it would have to pass the next
                // tile's barrier), so the two copies never collide
                epi_bar_sync();
            }
            const __nv_bfloat16* aux_row =
                (p.epi == DCAD_EPI_DGELU && row_ok) ? p.aux + static_cast<size_t>(row) * p.ldaux + n_base : nullptr;
            mbar_wait(smem_u32(&tmem_full_bar[acc_stage]), acc_phase);
            tc_fence_after();
            if (warp == 4) DCAD_STAMP(4);
            const uint32_t t_row = tmem_base + acc_stage * Cfg::ACC_STRIDE + ((q * 32u) << 16);

            if (!p.out_fp32) {
                // ---- bf16 output: 64-column slabs, converted 16 columns at a time ----
                const bool two_pass = (p.epi == DCAD_EPI_BIAS_GELU && p.has_aux_out);
#pragma unroll 1
                for (int s = part; s < BN / 64; s += 4) {
                    const int c0 = s * 64;
                    const bool do_store = tile_ok && n_base + c0 < p.N;
#pragma unroll 1
                    for (int pass = two_pass ? 0 : 1; pass < 2; ++pass) {
                        // pass 0 (lin1 forward in training): the pre-activation, kept for the backward; pass 1: D
                        const uint32_t sD = box_acquire();
#pragma unroll 1
                        for (int h = 0; h < 4; ++h) {
                            const int ch = c0 + h * 16;  // first accumulator column of this quarter-slab
                            uint32_t r[16];
                            tmem_ld_32x32b_x16(t_row + ch, r);
                            uint4 av[2];
                            if (p.epi == DCAD_EPI_DGELU) {
#pragma unroll
                                for (int j = 0; j < 2; ++j)
                                    av[j] = (aux_row && row_live && n_base + ch + j * 8 < p.N)
                                                ? __ldg(reinterpret_cast<const uint4*>(aux_row + ch) + j)
                                                : make_uint4(0, 0, 0, 0);
                            }
                            tmem_ld_wait();
                            float v[16];
#pragma unroll
                            for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
                            if (has_bias) {
                                const float4* bs = reinterpret_cast<const float4*>(bias_s + ch);
#pragma unroll
                                for (int j = 0; j < 16; j += 4) {
                                    const float4 b = bs[j >> 2];
                                    v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
                                }
                            }
                            if (p.epi == DCAD_EPI_BIAS_GELU && pass == 1) {
                                if (row_live) {
#pragma unroll
                                    for (int j = 0; j < 16; ++j) v[j] = gelu_erf(v[j]);
                                } else {
#pragma unroll
                                    for (int j = 0; j < 16; ++j) v[j] = 0.f;
                                }
                            } else if (p.epi == DCAD_EPI_DGELU) {
                                if (row_live) {
#pragma unroll
                                    for (int j = 0; j < 2; ++j) {
                                        const float2 a0 = unpack_bf16x2(av[j].x), a1 = unpack_bf16x2(av[j].y);
                                        const float2 a2 = unpack_bf16x2(av[j].z), a3 = unpack_bf16x2(av[j].w);
                                        v[8 * j] *= dgelu_erf(a0.x); v[8 * j + 1] *= dgelu_erf(a0.y);
                                        v[8 * j + 2] *= dgelu_erf(a1.x); v[8 * j + 3] *= dgelu_erf(a1.y);
                                        v[8 * j + 4] *= dgelu_erf(a2.x); v[8 * j + 5] *= dgelu_erf(a2.y);
                                        v[8 * j + 6] *= dgelu_erf(a3.x); v[8 * j + 7] *= dgelu_erf(a3.y);
                                    }
                                } else {
#pragma unroll
                                    for (int j = 0; j < 16; ++j) v[j] = 0.f;
                                }
                            }
#pragma unroll
                            for (int j = 0; j < 2; ++j)
                                st_shared_v4(sD + sw128_offset(lane, h * 2 + j),
                                             pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                                             pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
                            if (p.colsum && pass == 1) {
                                // bias gradient of the layer that consumes this output: the warp's 32 rows x 16 columns
                                // are reduced with a shuffle butterfly, one fp32 atomic per column and warp
                                if (!row_ok) {
#pragma unroll
                                    for (int j = 0; j < 16; ++j) v[j] = 0.f;
                                }
                                warp_colsum<16>(v, lane);
                                const int col = n_base + ch + static_cast<int>(lane);
                                if (lane < 16 && col < p.N) atomicAdd(p.colsum + col, v[0]);
                            }
                        }
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0) {
                            if (do_store) {
                                if (pass == 0) tma_store_2d(&tmAux, sD, n_base + c0, box_row0);
                                else tma_store_2d(&tmD, sD, n_base + c0, out_row0);
                            }
                            bulk_commit_group();  // (also when nothing was stored: box reuse counts groups)
                        }
                        ++st_ctr;
                    }
                }
            } else {
                // ---- fp32 output (weight gradients, split-K partials): 32-column slabs ----
#pragma unroll 1
                for (int s = part; s < BN / 32; s += 4) {
                    const int c0 = s * 32;
                    uint32_t r[32];
                    tmem_ld_32x32b_x32(t_row + c0, r);
                    const uint32_t sD = box_acquire();
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        st_shared_v4(sD + sw128_offset(lane, j), r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        if (tile_ok && n_base + c0 < p.N) {
                            if ((p.split_k > 1 && p.ws_row_stride == 0) || p.out_fp32 == 2) tma_reduce_add_2d(&tmD, sD, n_base + c0, out_row0);
                            else tma_store_2d(&tmD, sD, n_base + c0, out_row0);
                        }
                        bulk_commit_group();
                    }
                    ++st_ctr;
                }
            }
            tc_fence_before();
            if constexpr (CG == 1) {
                mbar_arrive(smem_u32(&tmem_empty_bar[acc_stage]));
            } else {  // the leader's MMA thread waits for the epilogues of both CTAs
                __syncwarp();
                if (lane == 0) mbar_arrive_remote(smem_u32(&tmem_empty_bar[acc_stage]), 0);
            }
            if constexpr (Cfg::ALIAS) {  // hand the ring back to the producer once this warp's stores have read it
                if (lane == 0) {
                    bulk_wait_group_read<0>();
                    mbar_arrive(smem_u32(&stg_free_bar));
                }
                __syncwarp();
            }
            if (warp == 4) DCAD_STAMP(5);
            if (++acc_stage == ACC_STAGES) { acc_stage = 0; acc_phase ^= 1u; }
        }
        if (lane == 0) bulk_wait_group<0>();  // this warp's output writes have landed before the CTA retires
      }
    }

    tc_fence_before();
    if constexpr (CG == 1) __syncthreads();
    else cluster_sync_all();  // neither CTA may retire while the other can still read its smem or signal its barriers
    if (warp == 0) DCAD_STAMP(6);
    if (warp == 2) {
        tc_fence_after();
        if constexpr (CG == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
        else tmem_dealloc_cg2(tmem_base, Cfg::TMEM_COLS);
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
    uint32_t box_rows, box_cols, esize;
    bool operator==(const MapKey& o) const {
        return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows &&
               box_cols == o.box_cols && esize == o.esize;
    }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        size_t h = reinterpret_cast<size_t>(k.ptr);
        h = h * 1000003u ^ k.rows; h = h * 1000003u ^ k.cols; h = h * 1000003u ^ k.ld;
        h = h * 1000003u ^ k.box_rows; h = h * 1000003u ^ k.box_cols; h = h * 1000003u ^ k.esize;
        return h;
    }
};

// 2D row-major [rows][ld] tensor of bf16 (esize 2) or fp32 (esize 4); box = box_rows x box_cols elements with
// box_cols * esize == 128 bytes (the 128-byte swizzle span)
static thread_local bool t_tensor_map_cache_enabled = true;  // off => maps are encoded on the stack, no host heap use

int make_tensor_map_2d(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                       uint32_t box_cols, uint32_t esize) {
    static std::mutex mu;
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    MapKey key{ptr, rows, cols, ld, box_rows, box_cols, esize};
    const bool use_cache = t_tensor_map_cache_enabled;
    if (use_cache) {
        std::lock_guard<std::mutex> g(mu);
        auto it = cache.find(key);
        if (it != cache.end()) { *out = it->second; return 0; }
    }
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return DCAD_ERR_NO_DRIVER_ENTRY;
    if ((reinterpret_cast<uintptr_t>(ptr) & 15u) || (ld * esize) % 16 != 0) return DCAD_ERR_BAD_ALIGNMENT;
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstride[1] = {ld * esize};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, esize == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                     const_cast<void*>(ptr), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return DCAD_ERR_TENSOR_MAP;
    if (!use_cache) return 0;
    std::lock_guard<std::mutex> g(mu);
    if (cache.size() > 4096) cache.clear();
    cache.emplace(key, *out);
    return 0;
}

int make_tensor_map_bf16_2d(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                            uint32_t box_rows, uint32_t box_cols) {
    return make_tensor_map_2d(out, ptr, rows, cols, ld, box_rows, box_cols, 2);
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

// SM count of the CURRENT device (cached per device ordinal: one process may drive several GPUs)
static int num_sms() {
    static std::atomic<int> cache[128] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    const bool cacheable = dev >= 0 && dev < 128;
    int n = cacheable ? cache[dev].load(std::memory_order_relaxed) : 0;
    if (n == 0) {
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (cacheable) cache[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

template <int BN, int A_MAJOR, int B_MAJOR, int CG, int G, int EPI>
static int launch_gemm(const GemmGroup<G>& g, int max_ctas, cudaStream_t stream) {
    using Cfg = GemmCfg<BN, CG, EPI>;
    auto kern = gemm_bf16_tcgen05_kernel<BN, A_MAJOR, B_MAJOR, CG, G, EPI>;
    static DeviceOnce once;
    DCAD_CUDA_TRY(ensure_dynamic_smem(kern, Cfg::SMEM_BYTES, once));
    int grid = g.unit_begin[g.n];  // work units = CTAs (CG 1) or CTA pairs (CG 2)
    int cap = (max_ctas > 0 ? max_ctas : num_sms()) / CG;
    if (cap < 1) cap = 1;
    if (grid > cap) grid = cap;
    return (int)launch_pdl_cluster(kern, dim3(grid * CG), dim3(GEMM_THREADS), Cfg::SMEM_BYTES, stream, CG, g);
}

}  // namespace dcad

using namespace dcad;

// which epilogue the dispatch picks: 0 = block-wide slabs, 1 = per-warp boxes.  DCAD_GEMM_EPI presets it.
static int g_gemm_epi = -1;
static int gemm_epi_mode() {
    if (g_gemm_epi < 0) {
        const char* e = getenv("DCAD_GEMM_EPI");
        g_gemm_epi = (e && e[0] == '1') ? 1 : ((e && e[0] == '2') ? 2 : 0);
    }
    return g_gemm_epi;
}
extern "C" int dcad_gemm_set_epilogue(int mode) {
    const int prev = gemm_epi_mode();
    if (mode >= 0 && mode <= 2) g_gemm_epi = mode;
    return prev;
}

static unsigned long long* g_gemm_timing = nullptr;
// profiling aid: when set, CTA 0 of every following GEMM writes 7 globaltimer stamps (ns) to this device buffer
extern "C" void dcad_gemm_set_timing_buffer(void* dev_ptr) { g_gemm_timing = reinterpret_cast<unsigned long long*>(dev_ptr); }

// ws_row_stride > 0: D is an fp32 workspace of split_k row-slices of ws_row_stride rows each (plain stores);
// ws_row_stride == 0 with split_k > 1: D is the fp32 result, accumulated with TMA reduce-add.
// Validates one problem and fills its parameters and tensor maps for tile width bn on cg CTAs per tile.
static int build_problem(GemmParams& p, CUtensorMap& tmA, CUtensorMap& tmB, CUtensorMap& tmD, CUtensorMap& tmAux,
                         const void* A, const void* B, void* D, int M, int N, int K, int a_major, int b_major, int lda,
                         int ldb, int ldd, const float* bias, const void* aux, void* aux_out, int ldaux, int epi,
                         int out_fp32, int split_k, int ws_row_stride, int bn, int cg, int epi_mode,
                         float* colsum = nullptr, const int* row_lens = nullptr) {
    if (M <= 0 || N <= 0 || K <= 0) return DCAD_ERR_BAD_SHAPE;
    if (colsum && out_fp32) return DCAD_ERR_BAD_ARGUMENT;
    if ((N % 8) || (ldd % 8) || (lda % 8) || (ldb % 8)) return DCAD_ERR_BAD_ALIGNMENT;
    if ((epi == DCAD_EPI_BIAS || epi == DCAD_EPI_BIAS_GELU) && bias == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (epi == DCAD_EPI_DGELU && aux == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (split_k < 1) split_k = 1;
    if (split_k > 1 && (!out_fp32 || epi != DCAD_EPI_NONE)) return DCAD_ERR_BAD_ARGUMENT;
    if (out_fp32 && epi != DCAD_EPI_NONE) return DCAD_ERR_BAD_ARGUMENT;

    p.M = M; p.N = N; p.K = K; p.ldaux = ldaux > 0 ? ldaux : ldd;
    p.bias = bias;
    p.colsum = colsum;
    p.row_lens = (cg == 1) ? row_lens : nullptr;  // (a CTA-pair tile spans two sequences: not supported there)
    p.aux = reinterpret_cast<const __nv_bfloat16*>(aux);
    p.has_aux_out = (epi == DCAD_EPI_BIAS_GELU && aux_out != nullptr) ? 1 : 0;
    p.epi = epi; p.out_fp32 = out_fp32;
    p.num_m_blocks = (M + BM * cg - 1) / (BM * cg);  // tile rows: 128 (one CTA) or 256 (a CTA pair)
    p.num_n_blocks = (N + bn - 1) / bn;
    p.num_k_blocks = (K + BK - 1) / BK;
    if (split_k > p.num_k_blocks) split_k = p.num_k_blocks;
    p.kb_per_split = (p.num_k_blocks + split_k - 1) / split_k;
    split_k = (p.num_k_blocks + p.kb_per_split - 1) / p.kb_per_split;  // no empty splits
    p.split_k = split_k;
    p.ws_row_stride = ws_row_stride;
    p.timing = g_gemm_timing;

    int rc;
    if (a_major == kMajorK) rc = make_tensor_map_2d(&tmA, A, M, K, lda, BM, BK, 2);   // A stored [M][lda]
    else                    rc = make_tensor_map_2d(&tmA, A, K, M, lda, BK, 64, 2);   // A stored [K][lda]
    if (rc) return rc;
    const int b_box_rows = (bn / cg > 256) ? bn / cg / 2 : bn / cg;  // GemmCfg::B_BOX_ROWS
    if (b_major == kMajorK) rc = make_tensor_map_2d(&tmB, B, N, K, ldb, b_box_rows, BK, 2);  // B stored [N][ldb]
    else                    rc = make_tensor_map_2d(&tmB, B, K, N, ldb, BK, 64, 2);   // B stored [K][ldb]
    if (rc) return rc;
    const uint64_t d_rows = ws_row_stride > 0 ? static_cast<uint64_t>(ws_row_stride) * split_k : static_cast<uint64_t>(M);
    const int d_box_rows = epi_mode >= 1 ? BOX_ROWS : BM;  // per-warp boxes or whole-tile slabs
    rc = out_fp32 ? make_tensor_map_2d(&tmD, D, d_rows, N, ldd, d_box_rows, 32, 4)
                  : make_tensor_map_2d(&tmD, D, d_rows, N, ldd, d_box_rows, 64, 2);
    if (rc) return rc;
    if (p.has_aux_out) {
        rc = make_tensor_map_2d(&tmAux, aux_out, M, N, p.ldaux, d_box_rows, 64, 2);
        if (rc) return rc;
    } else {
        tmAux = tmD;
    }
    return 0;
}

static void parse_tile(int block_n, int N, int& bn, int& cg) {
    cg = (block_n > 0 && (block_n & DCAD_TILE_PAIR)) ? 2 : 1;
    bn = block_n > 0 ? (block_n & ~DCAD_TILE_PAIR) : 0;
    if (bn <= 0) bn = (N >= 256) ? 256 : (N > 128 ? 192 : (N > 64 ? 128 : 64));
}

static int gemm_dispatch(const void* A, const void* B, void* D, int M, int N, int K, int a_major, int b_major, int lda,
                         int ldb, int ldd, const float* bias, const void* aux, void* aux_out, int ldaux, int epi,
                         int out_fp32, int split_k, int ws_row_stride, int block_n, int max_ctas, cudaStream_t s,
                         float* colsum = nullptr, const int* row_lens = nullptr) {
    int bn, cg;
    parse_tile(block_n, N, bn, cg);
    const int em = (bn > 256) ? 1 : gemm_epi_mode();  // the 384-wide tile exists with the per-warp epilogue only
    GemmGroup<1> g;
    int rc = build_problem(g.p[0], g.tmA[0], g.tmB[0], g.tmD[0], g.tmAux[0], A, B, D, M, N, K, a_major, b_major, lda,
                           ldb, ldd, bias, aux, aux_out, ldaux, epi, out_fp32, split_k, ws_row_stride, bn, cg, em, colsum,
                           row_lens);
    if (rc) return rc;
    g.n = 1;
    g.unit_begin[0] = 0;
    g.unit_begin[1] = g.p[0].num_m_blocks * g.p[0].num_n_blocks * g.p[0].split_k;

#define DCAD_GEMM_CASE_CG(BN_, AM_, BM_, CG_)                                             \
    if (cg == CG_ && bn == BN_ && a_major == AM_ && b_major == BM_) {                     \
        if (em == 1) return launch_gemm<BN_, AM_, BM_, CG_, 1, 1>(g, max_ctas, s);        \
        if (em == 2) return launch_gemm<BN_, AM_, BM_, CG_, 1, 2>(g, max_ctas, s);        \
        return launch_gemm<BN_, AM_, BM_, CG_, 1, 0>(g, max_ctas, s);                     \
    }
#define DCAD_GEMM_CASE(BN_, AM_, BM_) DCAD_GEMM_CASE_CG(BN_, AM_, BM_, 1)
#define DCAD_GEMM_CASE2(BN_, AM_, BM_) DCAD_GEMM_CASE_CG(BN_, AM_, BM_, 2)
    DCAD_GEMM_CASE2(128, 0, 0)
    DCAD_GEMM_CASE2(192, 0, 0)
    DCAD_GEMM_CASE2(256, 0, 0)
    DCAD_GEMM_CASE2(128, 0, 1)
    DCAD_GEMM_CASE2(256, 0, 1)
    DCAD_GEMM_CASE2(128, 1, 1)
    DCAD_GEMM_CASE2(256, 1, 1)
    DCAD_GEMM_CASE(64, 0, 0)
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
#undef DCAD_GEMM_CASE2
#undef DCAD_GEMM_CASE_CG
    if (cg == 1 && bn == 384 && a_major == 0 && b_major == 0) return launch_gemm<384, 0, 0, 1, 1, 1>(g, max_ctas, s);
    if (cg == 1 && bn == 384 && a_major == 0 && b_major == 1) return launch_gemm<384, 0, 1, 1, 1, 1>(g, max_ctas, s);
    return DCAD_ERR_UNSUPPORTED_TILE;
}

// Up to DCAD_GEMM_GROUP_MAX independent fp32-output GEMMs (no epilogue) with the same operand layouts and tile width in
// one persistent launch.  split_k[i] > 1 (or out_fp32 == 2) accumulates problem i with TMA reduce-add.
extern "C" int dcad_gemm_bf16_group(int n, const void* const* A, const void* const* B, void* const* D, const int* M,
                                    const int* N, const int* K, const int* lda, const int* ldb, const int* ldd,
                                    const int* split_k, int a_major, int b_major, int out_fp32, int block_n,
                                    int max_ctas, void* stream) {
    if (n < 1 || n > DCAD_GEMM_GROUP_MAX) return DCAD_ERR_BAD_ARGUMENT;
    if (out_fp32 != 1 && out_fp32 != 2) return DCAD_ERR_BAD_ARGUMENT;
    if (block_n <= 0 || (block_n & DCAD_TILE_PAIR)) return DCAD_ERR_UNSUPPORTED_TILE;
    const int bn = block_n;
    const int em = gemm_epi_mode();
    GemmGroup<DCAD_GEMM_GROUP_MAX> g;
    g.n = n;
    g.unit_begin[0] = 0;
    for (int i = 0; i < DCAD_GEMM_GROUP_MAX; ++i) {
        const int j = i < n ? i : 0;  // unused slots mirror problem 0 (never selected)
        int rc = build_problem(g.p[i], g.tmA[i], g.tmB[i], g.tmD[i], g.tmAux[i], A[j], B[j], D[j], M[j], N[j], K[j],
                               a_major, b_major, lda[j], ldb[j], ldd[j], nullptr, nullptr, nullptr, 0, DCAD_EPI_NONE,
                               out_fp32, split_k[j], 0, bn, 1, em);
        if (rc) return rc;
        g.unit_begin[i + 1] = g.unit_begin[i] +
                              (i < n ? g.p[i].num_m_blocks * g.p[i].num_n_blocks * g.p[i].split_k : 0);
    }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
#define DCAD_GROUP_CASE(BN_, AM_, BM_)                                                              \
    if (bn == BN_ && a_major == AM_ && b_major == BM_) {                                            \
        if (em >= 1) return launch_gemm<BN_, AM_, BM_, 1, DCAD_GEMM_GROUP_MAX, 1>(g, max_ctas, s);  \
        return launch_gemm<BN_, AM_, BM_, 1, DCAD_GEMM_GROUP_MAX, 0>(g, max_ctas, s);               \
    }
    DCAD_GROUP_CASE(128, 1, 1)
    DCAD_GROUP_CASE(256, 1, 1)
    DCAD_GROUP_CASE(256, 0, 1)
    DCAD_GROUP_CASE(256, 0, 0)
#undef DCAD_GROUP_CASE
    return DCAD_ERR_UNSUPPORTED_TILE;
}

extern "C" int dcad_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int a_major, int b_major,
                              int lda, int ldb, int ldd, const float* bias, const void* aux, void* aux_out, int ldaux,
                              int epi, int out_fp32, int split_k, int block_n, int max_ctas, void* stream) {
    return gemm_dispatch(A, B, D, M, N, K, a_major, b_major, lda, ldb, ldd, bias, aux, aux_out, ldaux, epi, out_fp32,
                         split_k, 0, block_n, max_ctas, reinterpret_cast<cudaStream_t>(stream));
}

// Same, plus colsum (fp32 [N], may be NULL): the epilogue also accumulates the column sums of the bf16 output D into
// it with atomics -- the bias gradient of the Linear layer whose output gradient D is (no separate pass over D).
extern "C" int dcad_gemm_bf16_colsum(const void* A, const void* B, void* D, int M, int N, int K, int a_major,
                                     int b_major, int lda, int ldb, int ldd, const float* bias, const void* aux,
                                     void* aux_out, int ldaux, int epi, int block_n, int max_ctas, float* colsum,
                                     const int32_t* row_lens, void* stream) {
    return gemm_dispatch(A, B, D, M, N, K, a_major, b_major, lda, ldb, ldd, bias, aux, aux_out, ldaux, epi, 0, 1, 0,
                         block_n, max_ctas, reinterpret_cast<cudaStream_t>(stream), colsum, row_lens);
}

// ------------------------------------------------------------------------------------------------
// Split-K with an fp32 workspace, for bf16-output GEMMs whose output has too few tiles for 148 SMs but whose
// reduction is long (FFN lin2 forward and lin1 dgrad: 2048 x 768 output, K = 3072).  Pass 1: the tcgen05 GEMM with
// the reduction cut into split_k slices, each slice storing its fp32 partial tile into its own row-slice of the
// workspace (no atomics, no ordering between slices).  Pass 2: one elementwise kernel sums the slices in slice order
// (deterministic), applies the epilogue and writes bf16.
// ------------------------------------------------------------------------------------------------
namespace dcad {

__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ ws, int splits, int64_t slice_elems, int M, int N,
                     const float* __restrict__ bias, const __nv_bfloat16* __restrict__ aux, int ldaux,
                     __nv_bfloat16* __restrict__ aux_out, __nv_bfloat16* __restrict__ D, int ldd, int epi) {
    pdl_trigger();
    pdl_wait();
    const int vec_per_row = N >> 3;
    const int64_t total = static_cast<int64_t>(M) * vec_per_row;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int row = static_cast<int>(i / vec_per_row);
        const int col = static_cast<int>(i % vec_per_row) << 3;
        const float* src = ws + static_cast<int64_t>(row) * N + col;
        float v[8];
        {
            const float4 a = *reinterpret_cast<const float4*>(src);
            const float4 b = *reinterpret_cast<const float4*>(src + 4);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        }
        for (int s = 1; s < splits; ++s) {
            const float4 a = *reinterpret_cast<const float4*>(src + s * slice_elems);
            const float4 b = *reinterpret_cast<const float4*>(src + s * slice_elems + 4);
            v[0] += a.x; v[1] += a.y; v[2] += a.z; v[3] += a.w; v[4] += b.x; v[5] += b.y; v[6] += b.z; v[7] += b.w;
        }
        if (epi == DCAD_EPI_BIAS || epi == DCAD_EPI_BIAS_GELU) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(bias + col));
            const float4 b = __ldg(reinterpret_cast<const float4*>(bias + col + 4));
            v[0] += a.x; v[1] += a.y; v[2] += a.z; v[3] += a.w; v[4] += b.x; v[5] += b.y; v[6] += b.z; v[7] += b.w;
        }
        if (epi == DCAD_EPI_BIAS_GELU) {
            if (aux_out) {
                uint4 o;
                o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
                o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
                *reinterpret_cast<uint4*>(aux_out + static_cast<int64_t>(row) * ldaux + col) = o;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = gelu_erf(v[j]);
        } else if (epi == DCAD_EPI_DGELU) {
            const uint4 a = __ldg(reinterpret_cast<const uint4*>(aux + static_cast<int64_t>(row) * ldaux + col));
            const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), a2 = unpack_bf16x2(a.z), a3 = unpack_bf16x2(a.w);
            v[0] *= dgelu_erf(a0.x); v[1] *= dgelu_erf(a0.y); v[2] *= dgelu_erf(a1.x); v[3] *= dgelu_erf(a1.y);
            v[4] *= dgelu_erf(a2.x); v[5] *= dgelu_erf(a2.y); v[6] *= dgelu_erf(a3.x); v[7] *= dgelu_erf(a3.y);
        }
        uint4 o;
        o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
        o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
        *reinterpret_cast<uint4*>(D + static_cast<int64_t>(row) * ldd + col) = o;
    }
}

// shared by the workspace entry points: pass 1 + pass 2 on `ws` (device memory, >= dcad_gemm_splitk_workspace_bytes)
static int gemm_splitk_run(const void* A, const void* B, void* D, int M, int N, int K, int a_major, int b_major,
                           int lda, int ldb, int ldd, const float* bias, const void* aux, void* aux_out, int ldaux,
                           int epi, int split_k, int block_n, int max_ctas, float* ws, cudaStream_t s) {
    const int m_pad = (M + BM - 1) / BM * BM;
    int rc = gemm_dispatch(A, B, ws, M, N, K, a_major, b_major, lda, ldb, N, nullptr, nullptr, nullptr, 0, DCAD_EPI_NONE, 1,
                           split_k, m_pad, block_n, max_ctas, s);
    if (rc) return rc;
    int kb = (K + BK - 1) / BK;
    int sk = split_k > kb ? kb : split_k;
    const int per = (kb + sk - 1) / sk;
    sk = (kb + per - 1) / per;  // the split count the GEMM actually used
    const int64_t total = static_cast<int64_t>(M) * (N >> 3);
    int blocks = static_cast<int>((total + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    return (int)launch_pdl(splitk_reduce_kernel, dim3(blocks), dim3(256), 0, s, ws, sk,
                           static_cast<int64_t>(m_pad) * N, M, N, bias, reinterpret_cast<const __nv_bfloat16*>(aux),
                           ldaux > 0 ? ldaux : ldd, reinterpret_cast<__nv_bfloat16*>(aux_out),
                           reinterpret_cast<__nv_bfloat16*>(D), ldd, epi);
}

}  // namespace dcad

extern "C" size_t dcad_gemm_splitk_workspace_bytes(int M, int N, int split_k) {
    const size_t m_pad = static_cast<size_t>((M + BM - 1) / BM * BM);
    return m_pad * static_cast<size_t>(N) * static_cast<size_t>(split_k < 1 ? 1 : split_k) * sizeof(float);
}

extern "C" int dcad_gemm_bf16_splitk(const void* A, const void* B, void* D, int M, int N, int K, int a_major,
                                     int b_major, int lda, int ldb, int ldd, const float* bias, const void* aux,
                                     void* aux_out, int ldaux, int epi, int split_k, int block_n, int max_ctas,
                                     void* workspace, size_t workspace_bytes, void* stream) {
    if (split_k < 2) return DCAD_ERR_BAD_ARGUMENT;
    if (workspace == nullptr || (reinterpret_cast<uintptr_t>(workspace) & 255u)) return DCAD_ERR_BAD_ALIGNMENT;
    if ((epi == DCAD_EPI_BIAS || epi == DCAD_EPI_BIAS_GELU) && bias == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (epi == DCAD_EPI_DGELU && aux == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (workspace_bytes < dcad_gemm_splitk_workspace_bytes(M, N, split_k)) return DCAD_ERR_TOO_LARGE;
    return gemm_splitk_run(A, B, D, M, N, K, a_major, b_major, lda, ldb, ldd, bias, aux, aux_out, ldaux, epi, split_k,
                           block_n, max_ctas, reinterpret_cast<float*>(workspace), reinterpret_cast<cudaStream_t>(stream));
}

// Pass 1 only: leaves the fp32 partial tiles in the workspace for a consumer that sums them itself (the fused
// residual + LayerNorm kernels read them directly, which removes the reduce launch and a bf16 round trip).
// Slice s of the result is workspace[s * slice_elems + m * N + n]; *splits_out slices were written.
extern "C" int dcad_gemm_bf16_splitk_partials(const void* A, const void* B, int M, int N, int K, int a_major,
                                              int b_major, int lda, int ldb, int split_k, int block_n, int max_ctas,
                                              void* workspace, size_t workspace_bytes, int* splits_out,
                                              int64_t* slice_elems_out, void* stream) {
    if (split_k < 1) return DCAD_ERR_BAD_ARGUMENT;
    if (workspace == nullptr || (reinterpret_cast<uintptr_t>(workspace) & 255u)) return DCAD_ERR_BAD_ALIGNMENT;
    if (workspace_bytes < dcad_gemm_splitk_workspace_bytes(M, N, split_k)) return DCAD_ERR_TOO_LARGE;
    const int m_pad = (M + BM - 1) / BM * BM;
    const int kb = (K + BK - 1) / BK;
    int sk = split_k > kb ? kb : split_k;
    const int per = (kb + sk - 1) / sk;
    sk = (kb + per - 1) / per;  // the split count the GEMM actually uses
    if (splits_out) *splits_out = sk;
    if (slice_elems_out) *slice_elems_out = static_cast<int64_t>(m_pad) * N;
    return gemm_dispatch(A, B, workspace, M, N, K, a_major, b_major, lda, ldb, N, nullptr, nullptr, nullptr, 0,
                         DCAD_EPI_NONE, 1, split_k, m_pad, block_n, max_ctas, reinterpret_cast<cudaStream_t>(stream));
}

// ------------------------------------------------------------------------------------------------
// Split-K GEMM for hosts that cannot hand over a workspace.
//
// dcad_gemm_bf16_splitk needs fp32 scratch that somebody must have allocated.  Not every host can: a runtime that
// issues this GEMM while its stream is being captured into a CUDA graph, an inference server whose device memory pool
// is carved up once at start-up and then frozen, a plug-in running inside an application that forbids cudaMalloc
// after initialisation because an allocation stalls every stream on the device.  For them the scratch is part of the
// library image: one fixed region per GPU that has loaded the library, kept until the process exits.  The call below
// allocates and frees nothing on the device or the host (tensor maps are encoded on the stack, the map cache is
// bypassed) and calls no allocating library routine.
//
// g_splitk_static_ws is a DEVICE object.  Host code may not pass its name to a kernel: the usable address is
// obtained with cudaGetSymbolAddress, once per GPU, while that GPU is current, and cached.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_splitk_static_ws[DCAD_SPLITK_STATIC_BYTES];

namespace {
constexpr int kMaxDevicesForStaticWs = 64;
float* g_static_ws_ptr[kMaxDevicesForStaticWs] = {};

// device address of the CURRENT GPU's built-in workspace; the first call on a GPU resolves and caches it
int static_workspace(float** out) {
    int dev = -1;
    DCAD_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevicesForStaticWs) return DCAD_ERR_BAD_ARGUMENT;
    if (g_static_ws_ptr[dev] == nullptr) {
        void* p = nullptr;
        DCAD_CUDA_TRY(cudaGetSymbolAddress(&p, g_splitk_static_ws));
        g_static_ws_ptr[dev] = reinterpret_cast<float*>(p);
    }
    *out = g_static_ws_ptr[dev];
    return 0;
}
}  // namespace

extern "C" size_t dcad_gemm_splitk_static_capacity(void) { return DCAD_SPLITK_STATIC_BYTES; }

extern "C" int dcad_gemm_bf16_splitk_static(const void* A, const void* B, void* D, int M, int N, int K, int a_major,
                                            int b_major, int lda, int ldb, int ldd, const float* bias,
                                            const void* aux, void* aux_out, int ldaux, int epi, int split_k,
                                            int block_n, int max_ctas, void* stream) {
    if (split_k < 2) return DCAD_ERR_BAD_ARGUMENT;
    if ((epi == DCAD_EPI_BIAS || epi == DCAD_EPI_BIAS_GELU) && bias == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (epi == DCAD_EPI_DGELU && aux == nullptr) return DCAD_ERR_BAD_ARGUMENT;
    if (M <= 0 || N <= 0 || K <= 0) return DCAD_ERR_BAD_SHAPE;
    // the built-in scratch has a fixed size: refuse anything that does not fit rather than spill past it
    if (dcad_gemm_splitk_workspace_bytes(M, N, split_k) > static_cast<size_t>(DCAD_SPLITK_STATIC_BYTES))
        return DCAD_ERR_TOO_LARGE;
    float* ws = nullptr;
    int rc = static_workspace(&ws);
    if (rc) return rc;
    // Nothing in the scratch has to be cleared: pass 1 stores every partial element that pass 2 reads, so whatever
    // an earlier call left behind is overwritten or never looked at.
    t_tensor_map_cache_enabled = false;
    rc = gemm_splitk_run(A, B, D, M, N, K, a_major, b_major, lda, ldb, ldd, bias, aux, aux_out, ldaux, epi, split_k,
                         block_n, max_ctas, ws, reinterpret_cast<cudaStream_t>(stream));
    t_tensor_map_cache_enabled = true;
    return rc;
}
