/* C ABI of libdcad_b200.so — the hand-written sm_100a kernels behind the B200-native federated DDoS-detection
 * framework.  Every entry point takes raw device pointers plus a CUDA stream handle (void* == cudaStream_t),
 * launches asynchronously on that stream and returns 0 on success, a positive cudaError_t, or a negative
 * DCAD_ERR_* code.  Unless an entry point says otherwise it allocates nothing: callers own all buffers.
 *
 * "reference" in the comments below is /root/reference (client1.py, client2.py, server.py); HF is the
 * transformers DistilBERT implementation the reference delegates its model body to.
 */
#ifndef DCAD_B200_H_
#define DCAD_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error codes (negative; positive values are cudaError_t) ---- */
#define DCAD_ERR_BAD_SHAPE (-1)
#define DCAD_ERR_BAD_ALIGNMENT (-2)
#define DCAD_ERR_BAD_ARGUMENT (-3)
#define DCAD_ERR_UNSUPPORTED_TILE (-4)
#define DCAD_ERR_NO_DRIVER_ENTRY (-5)
#define DCAD_ERR_TENSOR_MAP (-6)
#define DCAD_ERR_TOO_LARGE (-7)
#define DCAD_ERR_TIMEOUT (-8)

/* ---- GEMM epilogues ---- */
#define DCAD_EPI_NONE 0      /* D = acc                                   */
#define DCAD_EPI_BIAS 1      /* D = acc + bias[n]                         */
#define DCAD_EPI_BIAS_GELU 2 /* pre = acc + bias; aux_out = pre; D = gelu(pre)   (HF FFN lin1 + GELU(erf)) */
#define DCAD_EPI_DGELU 3     /* D = acc * gelu'(aux[m,n])                 (backward of the above)          */

/* operand majors */
#define DCAD_MAJOR_K 0  /* operand stored [MN][K]: reduction dimension contiguous */
#define DCAD_MAJOR_MN 1 /* operand stored [K][MN]: M (or N) dimension contiguous  */

/* OR into block_n: run the tile on a CTA pair (cluster of two SMs, tcgen05.mma.cta_group::2, 256 x block_n tile);
 * each SM then stages only half of the B operand.  block_n must be 128, 192 (K-major B only) or 256. */
#define DCAD_TILE_PAIR 0x1000

int dcad_version(void);

/* Programmatic dependent launch for the training-step kernels (default on): each kernel releases its successor at
 * once and waits for its predecessor right before its first global-memory access, so launch latency and prologues
 * overlap the previous kernel's tail.  Returns the previous setting.  Takes effect for launches (and graph captures)
 * made after the call. */
int dcad_set_pdl(int enabled);

/* Epilogue of the tcgen05 GEMMs below (replaces what cuBLAS's epilogue does for the reference's nn.Linear calls,
 * client1.py:61 -> transformers modeling_distilbert.py:187-226).  0 = block-wide: the 16 epilogue warps fill one
 * [128 rows][128 B] slab together and one thread stores it; 1 = per-warp: every epilogue warp stages and TMA-stores its
 * own [32 rows][128 B] boxes, no block-wide barrier.  block_n = 384 (128 x 384 tiles, two N = 192 accumulators) always
 * uses 1.  The environment variable DCAD_GEMM_EPI presets it (default 0).  Returns the previous mode; takes effect
 * for launches and graph captures made after the call. */
int dcad_gemm_set_epilogue(int mode);

/* bf16 x bf16 -> fp32-accumulated GEMM on tcgen05 tensor cores (TMA-fed, TMEM accumulators, persistent, TMA-store
 * epilogue).
 *   D[M,N] = epilogue( sum_k A(m,k) * B(n,k) )
 * A is [M][lda] when a_major == DCAD_MAJOR_K, [K][lda] when DCAD_MAJOR_MN; likewise B with N.
 * D is bf16 (out_fp32 == 0) or fp32 (out_fp32 == 1 store, 2 = TMA reduce-add into D even without split-K, so D may
 * be a buffer on a peer GPU that several GPUs accumulate into); fp32 output requires DCAD_EPI_NONE.
 * split_k > 1 (fp32 D only): the reduction is split over several CTAs that accumulate into D with TMA reduce-add;
 * D must be zero on entry.  block_n: 0 = auto, else 64/128/192/256.  max_ctas: 0 = #SMs.
 * Replaces the cuBLAS calls behind nn.Linear forward/backward in the reference (client1.py:61,109).
 */
int dcad_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int a_major, int b_major, int lda,
                   int ldb, int ldd, const float* bias, const void* aux, void* aux_out, int ldaux, int epi,
                   int out_fp32, int split_k, int block_n, int max_ctas, void* stream);

/* dcad_gemm_bf16 with a bf16 output whose epilogue ALSO accumulates the column sums of D into colsum [N] (fp32,
 * atomics): the bias gradient of the Linear layer whose output gradient D is, without a second pass over D.
 * row_lens (may be NULL; rows must come in sequences of 128 = one tile): rows at or past row_lens[sequence] are
 * padding -- the GELU / dGELU epilogues skip their arithmetic and store zeros. */
int dcad_gemm_bf16_colsum(const void* A, const void* B, void* D, int M, int N, int K, int a_major, int b_major, int lda,
                          int ldb, int ldd, const float* bias, const void* aux, void* aux_out, int ldaux, int epi,
                          int block_n, int max_ctas, float* colsum, const int32_t* row_lens, void* stream);

/* Up to DCAD_GEMM_GROUP_MAX independent fp32-output GEMMs (no epilogue) that share operand layouts and tile width, in
 * ONE persistent launch: one prologue and pipeline fill, work units of all problems spread over the SMs together.
 * The four weight-gradient GEMMs of a transformer block use it.  split_k[i] > 1 accumulates problem i into D[i] with
 * TMA reduce-add (D[i] zero beforehand); out_fp32 = 2 forces reduce-add for every problem (peer-GPU targets). */
#define DCAD_GEMM_GROUP_MAX 4
int dcad_gemm_bf16_group(int n, const void* const* A, const void* const* B, void* const* D, const int* M, const int* N,
                         const int* K, const int* lda, const int* ldb, const int* ldd, const int* split_k, int a_major,
                         int b_major, int out_fp32, int block_n, int max_ctas, void* stream);
/* Split-K for bf16-output GEMMs with a long reduction and few output tiles (FFN lin2 forward / lin1 dgrad).  The
 * reduction is cut into split_k slices; every slice writes its fp32 partial tile to `workspace` (scratch, contents
 * undefined on return), then a second kernel sums the slices in order, applies the epilogue and writes bf16 D.
 * workspace: device memory, 256-byte aligned, at least dcad_gemm_splitk_workspace_bytes(M, N, split_k) bytes
 * (ceil(M/128)*128 * N * split_k * 4); too small -> DCAD_ERR_TOO_LARGE.  split_k >= 2. */
size_t dcad_gemm_splitk_workspace_bytes(int M, int N, int split_k);
int dcad_gemm_bf16_splitk(const void* A, const void* B, void* D, int M, int N, int K, int a_major, int b_major, int lda,
                          int ldb, int ldd, const float* bias, const void* aux, void* aux_out, int ldaux, int epi,
                          int split_k, int block_n, int max_ctas, void* workspace, size_t workspace_bytes,
                          void* stream);
/* Pass 1 of the split-K GEMM only: the fp32 partial tiles stay in the workspace (slice s at
 * workspace + s * *slice_elems_out floats, row stride N, *splits_out slices) for a consumer that sums them itself. */
int dcad_gemm_bf16_splitk_partials(const void* A, const void* B, int M, int N, int K, int a_major, int b_major,
                                   int lda, int ldb, int split_k, int block_n, int max_ctas, void* workspace,
                                   size_t workspace_bytes, int* splits_out, int64_t* slice_elems_out, void* stream);
/* Split-K GEMM without a caller-supplied workspace (no counterpart in the reference, which never touches device
 * memory itself).  Same arguments, kernels and results as dcad_gemm_bf16_splitk, except that the fp32 scratch is a
 * fixed DCAD_SPLITK_STATIC_BYTES (24 MiB) region built into the library, one per GPU that has loaded it, held until
 * the process exits.  It exists for hosts that are not allowed to allocate once they are running: code issuing this
 * GEMM during stream capture, servers with a device pool frozen at start-up, plug-ins in applications that ban
 * run-time cudaMalloc because it synchronises the device.  The call performs no cudaMalloc / cudaFree, no host heap
 * allocation and no allocating library call; A, B, bias, aux (inputs) and D, aux_out (outputs) remain the caller's.
 *   limit       ceil(M/128)*128 * N * split_k * 4 bytes must fit in dcad_gemm_splitk_static_capacity(); a larger
 *               problem is refused with DCAD_ERR_TOO_LARGE (2048 x 768 with split_k = 3 needs 18.9 MB).
 *   first call  the first call made on a GPU looks up that GPU's scratch address (cudaGetSymbolAddress on the current
 *               device).  Make that first call outside of, and before, any stream capture.
 *   sharing     every call on a GPU uses the same region and finds the previous call's partial sums in it; none of it
 *               needs clearing because the first pass rewrites everything the second pass reads.  Calls whose
 *               execution overlaps on one GPU (different streams or threads) overwrite each other's partials: run
 *               them one after another. */
#define DCAD_SPLITK_STATIC_BYTES (24u << 20)
size_t dcad_gemm_splitk_static_capacity(void);
int dcad_gemm_bf16_splitk_static(const void* A, const void* B, void* D, int M, int N, int K, int a_major, int b_major,
                                 int lda, int ldb, int ldd, const float* bias, const void* aux, void* aux_out,
                                 int ldaux, int epi, int split_k, int block_n, int max_ctas, void* stream);
/* profiling aid: CTA 0 of every following GEMM writes 7 globaltimer stamps (ns) into this device buffer (NULL = off) */
void dcad_gemm_set_timing_buffer(void* dev_ptr);

/* ---- LayerNorm family (warp per row; cols must be 256/512/768/1024) ----
 * rng points at two uint64 in device memory: {seed, step}; dropout masks are a pure function of
 * (seed, step, stream_id, element index), so backward regenerates them and CUDA-graph replays get fresh masks.
 */
/* out = LayerNorm(dropout_p(a) + res) * gamma + beta; rstd[rows] saved for backward.  res may be NULL.
 * Replaces HF sa_layer_norm / output_layer_norm + FFN dropout (modeling_distilbert.py:227,257,261). */
int dcad_add_ln_fwd(const void* a, const void* res, const float* gamma, const float* beta, void* out, float* rstd,
                    int rows, int cols, float eps, float p_drop, const uint64_t* rng, uint32_t stream_id,
                    void* stream);
/* dcad_add_ln_fwd that also stores x_hat = (x - mean) * rstd (bf16 [rows, cols]) for the backward: pass it to
 * dcad_add_ln_bwd as `y` with beta == NULL.  HF keeps the LayerNorm input for autograd (client1.py:109 through
 * modeling_distilbert.py:257,261); rebuilding x_hat from the bf16 output instead amplifies its rounding by
 * |beta| / |gamma|, which pretrained encoders with small gamma channels do not tolerate. */
int dcad_add_ln_fwd_x(const void* a, const void* res, const float* gamma, const float* beta, void* out, float* rstd,
                      void* xhat_out, int rows, int cols, float eps, float p_drop, const uint64_t* rng,
                      uint32_t stream_id, void* stream);
int dcad_add_ln_fwd_partials_x(const float* a_partials, int splits, int64_t slice_elems, const float* a_bias,
                               const void* res, const float* gamma, const float* beta, void* out, float* rstd,
                               void* xhat_out, int rows, int cols, float eps, float p_drop, const uint64_t* rng,
                               uint32_t stream_id, void* stream);

/* backward of dcad_add_ln_fwd.  dy = dy1 (+ dy2 when non-NULL).  dz = grad wrt res (and wrt the pre-dropout sum);
 * da = grad wrt a (only written when p_drop > 0, otherwise it equals dz).  dgamma/dbeta/dbias (dbias may be NULL;
 * it is colsum(da), the bias gradient of the Linear that produced a) are ACCUMULATED into zero-initialised buffers. */
int dcad_add_ln_bwd(const void* dy1, const void* dy2, const void* y, const float* gamma, const float* beta,
                    const float* rstd, void* dz, void* da, float* dgamma, float* dbeta, float* dbias, int rows,
                    int cols, float p_drop, const uint64_t* rng, uint32_t stream_id, void* stream);
/* Variants whose first operand is the sum of the fp32 partial tiles a split-K GEMM left in its workspace
 * (dcad_gemm_bf16_splitk_partials): slice s at partials + s * slice_elems, row stride = cols, plus an optional bias.
 * They replace "reduce the slices to bf16, then LayerNorm" with one launch and no rounding of the GEMM result. */
int dcad_add_ln_fwd_partials(const float* a_partials, int splits, int64_t slice_elems, const float* a_bias,
                             const void* res, const float* gamma, const float* beta, void* out, float* rstd, int rows,
                             int cols, float eps, float p_drop, const uint64_t* rng, uint32_t stream_id, void* stream);
int dcad_add_ln_bwd_partials(const float* dy1_partials, int splits, int64_t slice_elems, const void* dy2,
                             const void* y, const float* gamma, const float* beta, const float* rstd, void* dz,
                             void* da, float* dgamma, float* dbeta, float* dbias, int rows, int cols, float p_drop,
                             const uint64_t* rng, uint32_t stream_id, void* stream);
/* out = dropout_p(LayerNorm(word[ids[r]] + pos[r % seq])).  Replaces HF Embeddings (modeling_distilbert.py:96-122). */
int dcad_embed_ln_fwd(const int32_t* ids, const float* word, const float* pos, const float* gamma, const float* beta,
                      void* out, float* mean, float* rstd, int rows, int seq, int cols, float eps, float p_drop,
                      const uint64_t* rng, uint32_t stream_id, void* stream);
/* backward: scatter-adds into dword [vocab, cols] / dpos [max_pos, cols] and accumulates dgamma / dbeta.
 * lens [rows / seq] (may be NULL): rows at or past a sample's valid length are padding with an exactly-zero
 * gradient and are skipped.  touched_rows [vocab] (may be NULL): byte flags, set to 1 for every vocabulary row that
 * receives a gradient -- dcad_adam_step_rows() uses them to skip rows that never had one. */
int dcad_embed_ln_bwd(const void* dy1, const void* dy2, const int32_t* ids, const float* word, const float* pos,
                      const float* gamma, const float* mean, const float* rstd, float* dword, float* dpos,
                      float* dgamma, float* dbeta, const int32_t* lens, int rows, int seq, int cols, float p_drop,
                      const uint64_t* rng, uint32_t stream_id, uint8_t* touched_rows, void* stream);
/* out[c] += sum_r x[r, c]  (bf16 in, fp32 accumulate): bias gradients (autograd of nn.Linear bias). */
int dcad_colsum_bf16(const void* x, float* out, int rows, int cols, int ld, void* stream);

/* ---- classifier head + cross-entropy (reference client1.py:62-64,108 and, for eval, :134-142) ----
 * h: bf16, CLS vector of sample s starts at h + s * sample_stride.  loss_accum[0] += batch-mean loss,
 * loss_accum[1] += 1.  dW [2, cols] / db [2] are accumulated; dcls [batch, cols] bf16 is written. */
int dcad_cls_head_train(const void* h, int sample_stride, const float* W, const float* bias, const int32_t* labels,
                        int batch, float p_drop, const uint64_t* rng, uint32_t stream_id, float* loss_accum,
                        float* logits, float* dW, float* db, void* dcls, int cols, void* stream);
/* confusion[true * 2 + pred] += 1; probs[s] = softmax(logits)[1]; preds[s] = argmax (any output may be NULL). */
int dcad_cls_head_eval(const void* h, int sample_stride, const float* W, const float* bias, const int32_t* labels,
                       int batch, int cols, float* loss_accum, int32_t* confusion, float* probs, int32_t* preds,
                       void* stream);

/* Fused output LayerNorm + classifier head + cross-entropy, forward AND backward, for the CLS rows of the last block
 * (one launch instead of add_ln_fwd -> cls_head_train -> add_ln_bwd; fp32 between the stages).
 * LayerNorm input: split-K partial tiles (a_partials != NULL, plus a_bias) or the bf16 tensor a, with dropout p_ln, plus
 * the residual res.  Outputs: loss_accum {sum of batch-mean losses, batches}, logits [batch,2] (may be NULL), dW / db
 * of the classifier, dz = grad wrt the LayerNorm input sum, da = grad wrt a (only when p_ln > 0), dgamma / dbeta /
 * dbias accumulated with atomics.  batch <= 64, batch * cols * 4 bytes <= 200 KB. */
int dcad_tail_head(const float* a_partials, int splits, int64_t slice_elems, const float* a_bias, const void* a,
                   const void* res, const float* gamma, const float* beta, float eps, float p_ln, uint32_t ln_stream,
                   const float* W, const float* cls_bias, const int32_t* labels, float p_head, uint32_t head_stream,
                   const uint64_t* rng, int batch, int cols, float* loss_accum, float* logits, float* dW, float* db,
                   void* dz, void* da, float* dgamma, float* dbeta, float* dbias, void* stream);

/* Grid size of the flat Adam kernels: 148 * blocks 256-thread blocks (default 16).  Returns the previous value. */
int dcad_set_adam_blocks_per_sm(int blocks);

/* ---- optimizer (reference client1.py:106,110,380: Adam(lr=2e-5) + zero_grad) ----
 * One launch over the flat arena.  Elements in [bf16_begin, bf16_end) are also written as bf16 to
 * p_bf16[i - bf16_begin]; gradients in [zero_begin, zero_end) are cleared after use.  The step number is
 * state[1] + 1 (device memory), advanced by dcad_step_advance at the end of each training step. */
int dcad_adam_step(float* p, float* g, float* m, float* v, void* p_bf16, int64_t n, int64_t bf16_begin,
                   int64_t bf16_end, int64_t zero_begin, int64_t zero_end, float lr, float beta1, float beta2,
                   float eps, const uint64_t* state, void* stream);
/* Same update, but elements [rows_begin, rows_end) form a table of rows of row_len elements (the word embeddings) and
 * a row is skipped while row_flags[row] == 0.  Exact, not an approximation: a row that never received a gradient has
 * g = m = v = 0, so Adam leaves it unchanged (no weight decay in the reference, client1.py:380).
 * advance_ticket (may be NULL): a zero-initialised device word; when given, the block that finishes last also does
 * state[1] += 1 (what dcad_step_advance does), for the last Adam launch of a step. */
int dcad_adam_step_rows(float* p, float* g, float* m, float* v, void* p_bf16, int64_t n, int64_t bf16_begin,
                        int64_t bf16_end, int64_t zero_begin, int64_t zero_end, float lr, float beta1, float beta2,
                        float eps, const uint64_t* state, const uint8_t* row_flags, int64_t rows_begin,
                        int64_t rows_end, int row_len, unsigned int* advance_ticket, void* stream);
int dcad_step_advance(uint64_t* state, void* stream);
/* Intra-client data parallelism (no counterpart in the reference, which has no DP): sharded Adam over the arena
 * elements [begin, end) this rank owns, gradients scaled by grad_scale (1/G after a sum over the group), with the
 * parameter all-gather fused in: updated bf16 GEMM weights are stored into every peer's mirror and updated fp32
 * values in [push32_begin, push32_end) into every peer's arena.  peer_p / peer_bf16: HOST arrays of n_peers pointers. */
int dcad_adam_step_dp(float* p, float* g, float* m, float* v, int64_t begin, int64_t end, int64_t bf16_begin,
                      int64_t bf16_end, int64_t push32_begin, int64_t push32_end, float lr, float beta1, float beta2,
                      float eps, float grad_scale, const uint64_t* state, const void* const* peer_p,
                      const void* const* peer_bf16, int n_peers, int self, void* stream);
/* Row-sparse form of the above for the word-embedding table (no counterpart in the reference): rows [row_lo, row_hi)
 * belong to this rank; a row is processed only if some GPU's flags[k][row] is set (it has ever received a gradient:
 * otherwise g = m = v = 0 everywhere and Adam is the identity).  For such a row the n locally accumulated gradient
 * rows are summed through the peer mappings, Adam is applied with this rank's moments, and the new row is stored into
 * all n parameter arenas.  p_ptrs / g_ptrs: HOST arrays of n pointers to the word table inside each GPU's parameter /
 * gradient arena (p_ptrs[0] = this rank's own); flag_ptrs: each GPU's touched-row bytes. */
int dcad_adam_step_dp_rows(float* m, float* v, float* g_own, void* const* p_ptrs, const void* const* g_ptrs,
                           const void* const* flag_ptrs, int n, int row_lo, int row_hi, int row_len, float lr,
                           float beta1, float beta2, float eps, float grad_scale, const uint64_t* state, void* stream);
/* One-block barrier across the GPUs of a group through their signal pads (words [48,64)); the epoch is
 * state[1] * 8 + phase + 1, read on the device, so the launch can be replayed from a CUDA graph.  n <= 8. */
int dcad_xgpu_barrier(const void* const* pads, int rank, int n, const uint64_t* state, uint32_t phase, void* stream);

/* ---- fused attention, head_dim 64 (HF modeling_distilbert.py:126-151 + attention dropout) ----
 * qkv [B*S, 3*H*64] bf16 (Q | K | V), ctx [B*S, H*64] bf16, lse2 [B*H*max(S,128)] fp32, lens [B] valid key counts or
 * NULL.  S <= 128 (the reference's max_len, client1.py:27): tcgen05 whole-sequence kernels; 128 < S <= 512: CUDA-core
 * fallback kernels (csrc/attention_generic.cu). */
int dcad_attn_fwd(const void* qkv, void* ctx, float* lse2, const int32_t* lens, int B, int S, int H, float p_drop,
                  const uint64_t* rng, uint32_t stream_id, void* stream);
/* dbias (may be NULL): [3*H*64] fp32, += column sums of dqkv = gradient of the fused QKV bias (atomics). */
int dcad_attn_bwd(const void* qkv, const void* ctx, const void* dctx, const float* lse2, void* dqkv,
                  const int32_t* lens, int B, int S, int H, float p_drop, const uint64_t* rng, uint32_t stream_id,
                  float* dbias, void* stream);

/* ---- FedAvg (reference server.py:67-79) ---- */
#define DCAD_FED_MAX_CLIENTS 16
#define DCAD_FED_TWO_SHOT 0
#define DCAD_FED_ONE_SHOT 1
#define DCAD_FED_PAD_WORDS 64 /* uint32 words per signal pad */
/* dst = mean over clients of srcs[k] (fp32 accumulate, client order, then / N; or sum_k weights[k]*srcs[k]).
 * srcs / weights are HOST arrays.  Optional bf16 mirror of elements [bf16_begin, bf16_end). */
int dcad_fedavg_local(const void* const* srcs, void* dst, void* dst_bf16, int n_clients, int64_t count, int in_bf16,
                      const float* weights, int64_t bf16_begin, int64_t bf16_end, int max_ctas, void* stream);
/* One launch per rank; the N kernels (one per GPU) handshake through `pads` (peer-mapped, DCAD_FED_PAD_WORDS uint32
 * each, zero-initialised, epoch strictly increasing per call).  Two-shot works in place on `arenas`; one-shot writes
 * outs[rank].  mirrors[k] (may be NULL) receive the bf16 copy.  All pointer arrays are HOST arrays of N entries. */
int dcad_fedavg_p2p(const void* const* arenas, const void* const* outs, const void* const* mirrors,
                    const void* const* pads, const float* weights, int rank, int n_clients, int64_t count,
                    int in_bf16, int mode, int64_t bf16_begin, int64_t bf16_end, uint32_t epoch, int max_ctas,
                    void* stream);

/* Two-shot FedAvg with separate source and destination sets: the n_src arenas in `srcs` (one per client) are
 * averaged, all n_ranks participating GPUs (dsts / mirrors / pads, one entry per rank) receive the result.  For
 * clients that span several GPUs. */
int dcad_fedavg_p2p_grouped(const void* const* srcs, int n_src, const void* const* dsts, const void* const* mirrors,
                            const void* const* pads, int n_ranks, const float* weights, int rank, int64_t count,
                            int in_bf16, int64_t bf16_begin, int64_t bf16_end, uint32_t epoch, int max_ctas,
                            void* stream);
/* NVLS variant: mc_arena / mc_mirror are MULTICAST addresses mapping every client's fp32 arena / bf16 mirror
 * (cuMulticast* objects, e.g. from torch symmetric memory); the NVSwitch sums the N copies (multimem.ld_reduce) and
 * replicates the averaged shard (multimem.st).  fp32 arenas only, unweighted; in-switch summation order is
 * unspecified (fp32 tolerance vs the reference loop).  pads: unicast signal pads as for dcad_fedavg_p2p. */
int dcad_fedavg_nvls(void* mc_arena, void* mc_mirror, const void* const* pads, int rank, int n_clients, int64_t count,
                     int64_t bf16_begin, int64_t bf16_end, uint32_t epoch, int max_ctas, void* stream);

/* ---- peer-memory runtime (set-up calls; these allocate / map) ---- */
int dcad_p2p_alloc(void** ptr, size_t bytes);
int dcad_p2p_free(void* ptr);
int dcad_p2p_ipc_get(void* ptr, void* handle64);
int dcad_p2p_ipc_open(const void* handle64, void** ptr);
int dcad_p2p_ipc_close(void* ptr);
int dcad_p2p_enable_peer_access(int peer_device);
int dcad_p2p_can_access(int device, int peer, int* out);

/* ---- misc ---- */
int dcad_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream);
int dcad_gather_rows_bf16(const void* src, void* dst, int rows, int cols, int every, int ld_src, void* stream);
int dcad_scatter_rows_bf16(const void* src, void* dst, int rows, int cols, int every, int ld_dst, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DCAD_B200_H_ */
