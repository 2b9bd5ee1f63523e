"""B200 training/evaluation engine for ``DDoSClassifier``: a fixed sequence of hand-written sm_100a kernels over
pre-allocated buffers, captured once into a CUDA graph and replayed per step.

What it replaces in the reference: the body of ``train_model``'s inner loop (client1.py:102-112: H2D of the batch,
``zero_grad``, forward through HF DistilBERT, ``CrossEntropyLoss``, ``backward``, ``Adam.step``, two ``loss.item()``
syncs) and ``evaluate_model``'s per-batch work (client1.py:128-142).  There is no autograd here: forward and backward
are explicit kernel sequences:

    embed_ln_fwd -> 6 x [QKV GEMM(+bias) -> attn_fwd -> out GEMM(+bias) -> add_ln -> lin1 GEMM(+bias+GELU)
                         -> lin2 GEMM(+bias) -> (dropout)+add_ln] -> cls_head(+CE, fwd+bwd)
    backward mirrors it with dgrad (K-major x MN-major) and wgrad (MN-major x MN-major) GEMMs, fused dGELU epilogue,
    LayerNorm backward producing the bias gradients, then ONE flat Adam launch that also refreshes the bf16 weights
    and clears the atomically-accumulated gradients.

Losses and evaluation metrics stay on the device (loss sum, batch count, 2x2 confusion matrix) and are read once
per epoch / evaluation, not twice per step.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Dict, Optional, Tuple

import torch
from types import SimpleNamespace

from ..ops import native as nv
from .arena import ModelDims, ParamArena

# dropout stream ids: one independent mask stream per dropout site
_STREAM_EMBED = 1
_STREAM_HEAD = 2


_WORD_EMB = "distilbert.embeddings.word_embeddings.weight"


def _layer_stream(layer: int, site: int) -> int:
    return 16 + layer * 4 + site  # site 0 = attention probs, 1 = FFN output


@dataclass
class GemmPlan:
    """(BLOCK_N, split_k) per GEMM of the step (M = tokens).  A tcgen05.mma of N <= 128 costs as much issue time as
    one of N = 256 (profiles/gemm_mainloop_anatomy.jsonl), so tiles are as wide as the problem allows and weight-gradient GEMMs,
    whose output has few tiles, split the token (reduction) dimension instead.  Measured with tools/tune_gemm.py."""
    qkv_fwd: int = 256
    out_fwd: int = 128
    # N = 3072: 128 x 256 tiles are 192 tiles = 1.3 waves of 148 SMs, so 192-wide tiles on 128 CTAs x 2 tiles were the
    # measured best.  DCAD_GEMM_WIDE=1 selects 128 x 384 tiles (two N = 192 accumulators per k-step): 128 tiles = one
    # wave, one tile per CTA (per-warp epilogue; see csrc/gemm_tcgen05.cu).
    lin1_fwd: int = field(default_factory=lambda: int(os.environ.get(
        "DCAD_LIN1_FWD_BN", "384" if os.environ.get("DCAD_GEMM_WIDE", "0") != "0" else "192")))
    lin2_fwd: tuple = (256, 3)     # (BLOCK_N, split_k): K = 3072, only 48 output tiles -> workspace split-K
    lin2_dgrad: int = field(default_factory=lambda: int(os.environ.get(
        "DCAD_LIN2_DGRAD_BN", "384" if os.environ.get("DCAD_GEMM_WIDE", "0") != "0" else "192")))   # -> [T, 3072]
    # 256 tiles: 128 CTAs x 2 tiles costs the same as 148 CTAs with a ragged second wave (tools/tune_wide.py) and
    # leaves 20 SMs to the weight-gradient branch
    lin2_dgrad_ctas: int = field(default_factory=lambda: int(os.environ.get("DCAD_LIN2_DGRAD_CTAS", "128")))
    lin1_dgrad: tuple = field(default_factory=lambda: (256, int(os.environ.get("DCAD_LIN1_DGRAD_SPLIT", "3"))))  # -> [T, 768], K = 3072
    out_dgrad: int = 128
    # -> [T, 768], K = 2304: only 48 tiles.  Cutting the reduction (split-K 2/3/4 -> 96/144/192 units) makes this
    # GEMM faster in isolation but the STEP slower (1.588 / 1.622 / 1.635 / 1.659 ms, same box): with 48 CTAs it leaves
    # 100 SMs to the weight-gradient / Adam branch running next to it.
    qkv_dgrad: tuple = field(default_factory=lambda: (256, int(os.environ.get("DCAD_QKV_DGRAD_SPLIT", "1"))))
    # Weight gradients (reduction over the T tokens).  Launched one by one they need split-K (36-54 tiles each); in
    # the grouped launch the four problems of a block have 144 tiles together = one wave of 148 SMs, so they run
    # unsplit: plain TMA stores instead of reduce-add (no read-modify-write), and Adam need not clear those gradients.
    wgrad_lin1: tuple = field(default_factory=lambda: (256, int(os.environ.get("DCAD_WGRAD_SPLIT", "1"))))   # [3072, 768]
    wgrad_lin2: tuple = field(default_factory=lambda: (256, int(os.environ.get("DCAD_WGRAD_SPLIT", "1"))))   # [768, 3072]
    wgrad_qkv: tuple = field(default_factory=lambda: (256, int(os.environ.get("DCAD_WGRAD_SPLIT", "1"))))    # [2304, 768]
    wgrad_out: tuple = field(default_factory=lambda: (256, int(os.environ.get("DCAD_WGRAD_SPLIT", "1"))))    # [768, 768]
    # CLS-row tail of the last block (M = batch rows): latency-bound, so narrow tiles spread the weight matrix over
    # many SMs and the K = 3072 reductions are cut eight ways (TOKENS=16 tools/tune_gemm.py)
    tail_bn: int = 64
    tail_long_k: tuple = (64, 8)
    tail_wgrad_bn: int = 128
    # the four weight-gradient GEMMs of a block go out as ONE grouped persistent launch (BLOCK_N 256, the split_k
    # values above) at the end of the block's backward instead of four launches spread through it
    group_wgrad: bool = field(default_factory=lambda: os.environ.get("DCAD_GROUP_WGRAD", "1") != "0")
    # block 0 (the end of the backward pass): weight gradients + Adam in two halves instead of one group at the very end
    split_last_wgrad: bool = field(default_factory=lambda: os.environ.get("DCAD_SPLIT_LAST_WGRAD", "1") != "0")
    # last LayerNorm + classifier + cross-entropy (forward and backward) of the pruned tail as one single-CTA launch
    # (dcad_tail_head) instead of three.  Correct and tested, but measured neutral (1.584 vs 1.582 ms per step, same
    # box): one CTA doing three kernels' work serially costs what the two saved launches cost.  Off by default.
    fuse_tail_head: bool = field(default_factory=lambda: os.environ.get("DCAD_FUSE_TAIL_HEAD", "0") != "0")
    # bias gradients of lin1 and of the fused QKV projection are accumulated by the kernels that PRODUCE dpre / dqkv
    # (dGELU GEMM epilogue, attention backward epilogue) instead of two column-sum passes over those tensors per block
    fuse_colsum: bool = field(default_factory=lambda: os.environ.get("DCAD_FUSE_COLSUM", "1") != "0")
    # persistent-grid cap of the grouped weight-gradient launch (0 = one CTA per SM)
    wgrad_max_ctas: int = field(default_factory=lambda: int(os.environ.get("DCAD_WGRAD_CTAS", "0")))
    # which split-K GEMMs hand their fp32 partial tiles straight to the LayerNorm kernels (see _gemm_then):
    # 0 none, 1 the CLS-row tail only, 2 all
    fuse_ln_partials: int = field(default_factory=lambda: int(os.environ.get("DCAD_FUSE_LN", "2")))


def plan_for(tokens: int) -> GemmPlan:
    """Tile plan for ``tokens`` = batch * executed sequence length rows.  The defaults of ``GemmPlan`` were tuned at
    2048 rows (16 x 128); at 1536 rows (16 x 96: flow sentences with the all-padding columns dropped) there are 12
    row-blocks, so 256-wide tiles make the N = 3072 GEMMs exactly one wave (144 tiles) and a 4-way split of the
    K-long N = 768 GEMMs gives 144 units (`profiles/gemm_round2/tune_gemm_T1536.json`).  ``DCAD_PLAN`` overrides single
    fields, e.g. ``DCAD_PLAN="qkv_fwd=192,lin2_fwd=256/4"``."""
    pl = GemmPlan()
    if tokens <= 1536:
        pl.qkv_fwd = 192      # 12 x 12 = 144 tiles (in-step A/B at 1536 rows: 1.2747 vs 1.2828 ms with 256-wide tiles)
        pl.lin1_fwd = int(os.environ.get("DCAD_LIN1_FWD_BN", "256"))
        pl.lin2_dgrad = int(os.environ.get("DCAD_LIN2_DGRAD_BN", "256"))
        pl.lin2_dgrad_ctas = int(os.environ.get("DCAD_LIN2_DGRAD_CTAS", "0"))
    for item in filter(None, os.environ.get("DCAD_PLAN", "").split(",")):
        k, _, v = item.partition("=")
        k = k.strip()
        if not hasattr(pl, k):
            raise ValueError(f"DCAD_PLAN: unknown field {k!r}")
        setattr(pl, k, tuple(int(x) for x in v.split("/")) if "/" in v else int(v))
    return pl


class B200Engine:
    def __init__(self, dims: ModelDims = ModelDims(), batch: int = 16, seq: int = 128, device: int | str = 0,
                 lr: float = 2e-5, betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8, seed: int = 0,
                 flat_params: Optional[torch.Tensor] = None, bf16_params: Optional[torch.Tensor] = None,
                 plan: Optional[GemmPlan] = None, init_seed: Optional[int] = 0, dp=None,
                 prune_last_layer: bool = True, overlap_wgrad: bool = True):
        if not torch.cuda.is_available():
            raise nv.NativeError("B200Engine needs a CUDA device; use models.distilbert.DDoSClassifier on CPU")
        nv.lib()  # fail loudly if the native library is missing
        if seq > dims.max_position or seq > 512:
            raise ValueError(f"seq {seq} exceeds the position table ({dims.max_position}) / the attention kernels (512)")
        # seq <= 128 (the reference's max_len, client1.py:27): the tcgen05 whole-sequence attention kernels;
        # 128 < seq <= 512: the same entry points fall back to the CUDA-core kernels of csrc/attention_generic.cu
        if dims.head_dim != 64 or dims.dim % 256 != 0:
            raise ValueError("kernels are specialised for head_dim 64 and dim in {256,512,768,1024}")
        self.dims, self.B, self.S = dims, batch, seq
        self.device = torch.device("cuda", device) if not isinstance(device, torch.device) else device
        torch.cuda.set_device(self.device)
        self.lr, self.betas, self.eps = lr, betas, eps
        self.plan = plan or plan_for(batch * seq)
        self.dp = dp  # parallel.dp.DPGroup when this client spans several GPUs
        # Only the CLS row of the LAST block reaches the classifier (reference client1.py:62), so after the last
        # attention (whose keys/values still need every token) out_lin, both LayerNorms and the FFN are evaluated for
        # the B CLS rows instead of all B*S tokens, forward and backward.  Same math, ~12 % fewer FLOPs per step.
        self.prune_last_layer = prune_last_layer
        dev = self.device
        # Weight-gradient GEMMs and bias column sums are off the critical path of the backward pass (nothing in the
        # same step reads them before Adam), so they are enqueued on a second, lower-priority stream: in the captured
        # graph they become a parallel branch whose CTAs fill the SMs that the data-gradient chain leaves idle
        # (partial waves, prologue/epilogue tails of its small dependent kernels).
        self.overlap_wgrad = overlap_wgrad = overlap_wgrad and os.environ.get("DCAD_OVERLAP_WGRAD", "1") != "0"
        self._side = (torch.cuda.Stream(device=dev, priority=int(os.environ.get("DCAD_SIDE_PRIO", "0")))
                      if overlap_wgrad else None)
        # Programmatic dependent launch pays on a single chain (1.918 -> 1.878 ms when it was introduced) but not
        # next to the second branch: early-launched CTAs of the chain park in griddepcontrol.wait on SMs the branch
        # could be using (same box: PDL everywhere 1.533 ms, main chain only 1.523 ms, nowhere 1.517 ms per step).
        self.use_pdl = os.environ.get("DCAD_PDL", "0" if overlap_wgrad else "1") != "0"
        # ... but the branch is idle during the forward pass and the CLS-row tail, where the chain is alone again:
        # there PDL is on ("auto"); DCAD_PDL_FWD=0/1 forces it
        self.use_pdl_fwd = os.environ.get("DCAD_PDL_FWD", os.environ.get("DCAD_PDL", "1")) != "0"
        if dp is not None:
            flat_params, bf16_params = dp.param_tensor(), dp.mirror_tensor()
        self.arena = ParamArena(dims, dev, flat=flat_params, with_optimizer_state=True, with_bf16=bf16_params is None)
        if bf16_params is not None:
            assert bf16_params.numel() == self.arena.layout.gemm_end and bf16_params.dtype == torch.bfloat16
            self.arena.bf16 = bf16_params
        if dp is not None:
            self.arena.grad = dp.grad_tensor()  # peers reduce-add into it and read it: must be peer-mappable
            self.arena.grad.zero_()
        if init_seed is not None:
            self.arena.init_random_(init_seed)
        self.rng = torch.tensor([seed, 0], dtype=torch.int64, device=dev)  # {seed, step}
        T, D, Fh, H, L = batch * seq, dims.dim, dims.hidden_dim, dims.n_heads, dims.n_layers
        self.T = T
        bf = dict(device=dev, dtype=torch.bfloat16)
        f32 = dict(device=dev, dtype=torch.float32)
        # static inputs
        # pinned staging for the end-to-end path: ids | lens | labels in one buffer -> one H2D copy per step
        self.h_in = torch.zeros(T + 2 * batch, dtype=torch.int32).pin_memory()
        self._h_in_np = self.h_in.numpy()
        self.d_in = torch.zeros(T + 2 * batch, device=dev, dtype=torch.int32)
        # the kernels read ids | lens | labels straight out of the staging buffer: the one H2D copy of the end-to-end
        # path is all the input handling there is (no unpack kernels)
        self.ids, self.lens, self.labels = self.d_in[:T], self.d_in[T: T + batch], self.d_in[T + batch:]
        self.lens.fill_(seq)
        self.h_loss = torch.zeros(2, dtype=torch.float32).pin_memory()
        self._h_loss_np = self.h_loss.numpy()
        # saved activations
        self.x = [torch.empty(T, D, **bf) for _ in range(L + 1)]
        self.qkv = [torch.empty(T, 3 * D, **bf) for _ in range(L)]
        self.ctx = [torch.empty(T, D, **bf) for _ in range(L)]
        self.sa = [torch.empty(T, D, **bf) for _ in range(L)]
        self.pre = [torch.empty(T, Fh, **bf) for _ in range(L)]
        self.hmid = [torch.empty(T, Fh, **bf) for _ in range(L)]
        # LayerNorm backward needs x_hat.  Saved by the forward in bf16 (3 MB per LayerNorm and step of extra HBM writes)
        # rather than rebuilt from the bf16 output as (y - beta) / gamma: that amplifies y's rounding by |beta| / |gamma|
        # and a pretrained DistilBERT -- what the reference fine-tunes, client1.py:55 -- has channels with small gamma
        # (DCAD_SAVE_XHAT=0 restores the rebuild)
        self.save_xhat = os.environ.get("DCAD_SAVE_XHAT", "1") != "0"
        self.xh1 = [torch.empty(T, D, **bf) if self.save_xhat else None for _ in range(L)]
        self.xh2 = [torch.empty(T, D, **bf) if self.save_xhat else None for _ in range(L)]
        self.rstd1 = [torch.empty(T, **f32) for _ in range(L)]
        self.rstd2 = [torch.empty(T, **f32) for _ in range(L)]
        self.lse2 = [torch.empty(batch * H * max(seq, 128), **f32) for _ in range(L)]
        self.emb_mean = torch.empty(T, **f32)
        self.emb_rstd = torch.empty(T, **f32)
        # transient forward / backward buffers
        self.tmp_d = torch.empty(T, D, **bf)        # attn_out / ffn_out
        # Two sets of backward scratch, alternated by layer parity: the weight-gradient branch of layer l may still be
        # reading set l%2 while layer l-1 (other set) runs; layer l-2 waits for it before reusing the set.
        #   a: dz1 (grad wrt attn_out / residual x)   b: dsa, then dx via QKV   c: dz2   d: da2 (dropped dz2)
        #   e: dctx   wide: dpre   qkv: dqkv
        self.gset = [SimpleNamespace(a=torch.empty(T, D, **bf), b=torch.empty(T, D, **bf), c=torch.empty(T, D, **bf),
                                     d=torch.empty(T, D, **bf), e=torch.empty(T, D, **bf),
                                     wide=torch.empty(T, Fh, **bf), qkv=torch.empty(T, 3 * D, **bf))
                     for _ in range(2)]
        self.dcls = torch.empty(batch, D, **bf)
        # compact (one row per sample) buffers for the pruned tail of the last block
        self.c_x, self.c_ctx, self.c_tmp, self.c_sa, self.c_out = (torch.empty(batch, D, **bf) for _ in range(5))
        self.c_pre, self.c_hmid, self.c_gwide = (torch.empty(batch, Fh, **bf) for _ in range(3))
        self.c_ga, self.c_gb, self.c_gc, self.c_gd, self.c_ge = (torch.empty(batch, D, **bf) for _ in range(5))
        self.c_rstd1, self.c_rstd2 = torch.empty(batch, **f32), torch.empty(batch, **f32)
        self.c_xh1, self.c_xh2 = ((torch.empty(batch, D, **bf), torch.empty(batch, D, **bf)) if self.save_xhat
                                  else (None, None))
        self.logits = torch.empty(batch, dims.n_classes, **f32)
        # scratch for the split-K GEMMs (fp32 partial tiles), sized for the larger of the two users
        sk = max(self.plan.lin2_fwd[1], self.plan.lin1_dgrad[1], self.plan.qkv_dgrad[1], 2)
        ws_bytes = max(nv.gemm_splitk_workspace_bytes(T, D, sk),
                       nv.gemm_splitk_workspace_bytes(batch, D, self.plan.tail_long_k[1]))
        self.splitk_ws = torch.empty(ws_bytes // 4, **f32)
        # device-side accumulators
        # vocabulary rows that have ever received a gradient since the optimizer state was last reset: Adam skips the
        # others (their g, m and v are exactly zero, so the update is the identity) -- flow sentences use a small
        # part of the 30522-row table, which is more than a third of all parameters
        self.touched_rows = torch.zeros(dims.vocab_size, device=dev, dtype=torch.uint8)
        # data-parallel client: the flags live in peer-mapped memory (the owner of a row reads every GPU's flag) and the
        # word-embedding table is reduced / updated / re-distributed row-sparsely (DCAD_DP_SPARSE_EMB=0: dense)
        self.dp_sparse_emb = dp is not None and os.environ.get("DCAD_DP_SPARSE_EMB", "1") != "0"
        if self.dp_sparse_emb:
            self.touched_rows = dp.flags_tensor()
            self.touched_rows.zero_()
        # ... and each block's GEMM weights are updated under the backward pass, behind a per-block barrier
        self.dp_early_adam = (dp is not None and os.environ.get("DCAD_DP_EARLY_ADAM", "1") != "0"
                              and dims.n_layers <= 6)
        # ... on a stream of its own: the barrier of block l waits for the SLOWEST GPU's weight gradients and the owner's
        # Adam follows it; on the weight-gradient stream both would sit in front of the next block's wgrad GEMM of
        # every GPU (measured on 4 GPUs: 30-125 us of barrier wait per block, 1.82 ms per step)
        # Single-GPU option (DCAD_ADAM_STREAM=1): the per-block Adam launches on that third stream as well, so that the
        # next block's weight-gradient GEMMs do not queue behind a bandwidth-bound Adam on the weight-gradient stream.
        adam_stream = dp is None and os.environ.get("DCAD_ADAM_STREAM", "0") != "0"
        self._opt_stream = (torch.cuda.Stream(device=dev, priority=int(os.environ.get("DCAD_SIDE_PRIO", "0")))
                            if ((self.dp_early_adam or adam_stream) and overlap_wgrad) else None)
        self._step_ticket = torch.zeros(1, device=dev, dtype=torch.int32)
        self.loss_accum = torch.zeros(2, **f32)                       # {sum of batch-mean losses, #batches}
        self.eval_loss_accum = torch.zeros(2, **f32)
        self.confusion = torch.zeros(4, device=dev, dtype=torch.int32)  # [true*2 + pred]
        self.probs = torch.empty(batch, **f32)
        self.preds = torch.empty(batch, device=dev, dtype=torch.int32)
        self._graph_train: Optional[torch.cuda.CUDAGraph] = None
        self._graph_eval: Optional[torch.cuda.CUDAGraph] = None
        self._graph_train_e2e: Optional[torch.cuda.CUDAGraph] = None
        self.launches_per_train_step = 0
        self._count_launches = False
        self._n_launch = 0
        self._bind_params()

    # ------------------------------------------------------------------------------------------------ parameters
    def _bind_params(self) -> None:
        a, L = self.arena, self.dims.n_layers
        pre = "distilbert.transformer.layer.{}."
        self.p: Dict[str, list] = {k: [] for k in (
            "wqkv", "bqkv", "wo", "bo", "g1", "b1", "w1", "bb1", "w2", "bb2", "g2", "b2",
            "d_wqkv", "d_bqkv", "d_wo", "d_bo", "d_g1", "d_b1", "d_w1", "d_bb1", "d_w2", "d_bb2", "d_g2", "d_b2")}
        lay = a.layout
        for l in range(L):
            p = pre.format(l)
            off = lay.qkv_weight_offset(l)
            D = self.dims.dim
            self.p["wqkv"].append(a.bf16[off: off + 3 * D * D].view(3 * D, D))
            self.p["d_wqkv"].append(self._wgrad_target(off, (3 * D, D), a.qkv_weight(l, a.grad)))
            self.p["bqkv"].append(a.qkv_bias(l))
            self.p["d_bqkv"].append(a.qkv_bias(l, a.grad))
            for short, key, is_w in (("wo", "attention.out_lin.weight", True), ("bo", "attention.out_lin.bias", False),
                                     ("g1", "sa_layer_norm.weight", False), ("b1", "sa_layer_norm.bias", False),
                                     ("w1", "ffn.lin1.weight", True), ("bb1", "ffn.lin1.bias", False),
                                     ("w2", "ffn.lin2.weight", True), ("bb2", "ffn.lin2.bias", False),
                                     ("g2", "output_layer_norm.weight", False), ("b2", "output_layer_norm.bias", False)):
                self.p[short].append(a.bf16_view(p + key) if is_w else a.view(p + key))
                local = a.view(p + key, a.grad)
                self.p["d_" + short].append(self._wgrad_target(lay.offsets[p + key], tuple(local.shape), local) if is_w else local)
        e = "distilbert.embeddings."
        self.word, self.pos = a.view(e + "word_embeddings.weight"), a.view(e + "position_embeddings.weight")
        self.eg, self.eb = a.view(e + "LayerNorm.weight"), a.view(e + "LayerNorm.bias")
        self.d_word, self.d_pos = a.view(e + "word_embeddings.weight", a.grad), a.view(e + "position_embeddings.weight", a.grad)
        self.d_eg, self.d_eb = a.view(e + "LayerNorm.weight", a.grad), a.view(e + "LayerNorm.bias", a.grad)
        self.cw, self.cb = a.view("classifier.weight"), a.view("classifier.bias")
        self.d_cw, self.d_cb = a.view("classifier.weight", a.grad), a.view("classifier.bias", a.grad)

    def _wgrad_target(self, offset: int, shape, local: torch.Tensor) -> torch.Tensor:
        """Where a weight-gradient GEMM writes: the local gradient arena, or (data-parallel client) the arena of the
        GPU that owns this weight, reached through its peer mapping."""
        return local if self.dp is None else self.dp.owner_grad_view(offset, tuple(shape))

    def state_dict(self, clone: bool = True):
        return self.arena.state_dict(clone=clone)

    def load_state_dict(self, sd) -> None:
        self.arena.load_state_dict(sd)

    def owned_optimizer_ranges(self):
        """Arena element ranges [lo, hi) whose Adam moments THIS GPU maintains: everything on a single GPU; on a
        data-parallel client the GEMM weights it owns, its share of the vocabulary rows and (first / last rank) the
        dense pieces around the word table -- the same split ``_optimizer`` walks."""
        lay = self.arena.layout
        if self.dp is None:
            return [(0, lay.total)]
        dp = self.dp
        out = list(dp.gemm_ranges(dp.rank))
        if self.dp_sparse_emb:
            w0 = lay.offsets[_WORD_EMB]
            w1 = w0 + lay.numel(_WORD_EMB)
            if dp.rank == 0:
                out.append((lay.gemm_end, w0))
            if dp.rank == dp.world - 1:
                out.append((w1, lay.total))
            row_lo, row_hi = dp.word_rows(dp.rank)
            out.append((w0 + row_lo * self.dims.dim, w0 + row_hi * self.dims.dim))
        else:
            out.append(dp.r2_range(dp.rank))
        return [(lo, hi) for lo, hi in out if hi > lo]

    def gathered_optimizer_state(self):
        """``(exp_avg, exp_avg_sq, touched_rows)`` of the whole model.  Single GPU: the arenas themselves.  Data-parallel
        client: a COLLECTIVE over the client's group -- every rank contributes the ranges it owns (zeros elsewhere),
        one sum all-reduce per moment and a max all-reduce of the touched-row flags; the result is complete on every
        rank and has the single-GPU layout, so a snapshot taken on G GPUs resumes on one and vice versa."""
        a = self.arena
        if self.dp is None:
            return a.exp_avg, a.exp_avg_sq, self.touched_rows
        import torch.distributed as dist

        torch.cuda.synchronize(self.device)
        out = []
        for src in (a.exp_avg, a.exp_avg_sq):
            full = torch.zeros_like(src)
            for lo, hi in self.owned_optimizer_ranges():
                full[lo:hi].copy_(src[lo:hi])
            dist.all_reduce(full, op=dist.ReduceOp.SUM, group=self.dp._dist_group)
            out.append(full)
        flags = self.touched_rows.clone()
        dist.all_reduce(flags, op=dist.ReduceOp.MAX, group=self.dp._dist_group)
        torch.cuda.synchronize(self.device)
        return out[0], out[1], flags

    def load_optimizer_state(self, exp_avg, exp_avg_sq, touched_rows, seed: int, step: int) -> None:
        """Inverse of ``gathered_optimizer_state`` (+ the dropout / Adam step counter).  Every rank of a data-parallel
        client loads the same full snapshot and keeps the moments of the ranges it owns; the flags become the union
        over the group (a superset of what this GPU touched: rows are then reduced whose local gradient is zero,
        which is exact).  The step counter moves, so the cross-GPU barrier words are cleared -- collective."""
        a = self.arena
        if exp_avg.numel() != a.exp_avg.numel():
            raise ValueError("optimizer arena size does not match this model")
        a.exp_avg.zero_()
        a.exp_avg_sq.zero_()
        for lo, hi in self.owned_optimizer_ranges():
            a.exp_avg[lo:hi].copy_(exp_avg[lo:hi])
            a.exp_avg_sq[lo:hi].copy_(exp_avg_sq[lo:hi])
        self.touched_rows.copy_(touched_rows)
        a.grad.zero_()
        self.rng.copy_(torch.tensor([seed, step], dtype=torch.int64))
        torch.cuda.synchronize(self.device)
        if self.dp is not None:
            self.dp.reset_barrier_flags()

    def refresh_bf16(self) -> None:
        nv.cast_f32_bf16(self.arena.flat[: self.arena.layout.gemm_end], self.arena.bf16)

    def reset_optimizer(self) -> None:
        self.arena.exp_avg.zero_()
        self.arena.exp_avg_sq.zero_()
        self.arena.grad.zero_()
        self.touched_rows.zero_()
        self.rng[1] = 0

    # ------------------------------------------------------------------------------------------------ helpers
    # Every launch goes through _count(): it keeps the per-step launch count, and -- profiling aid used by
    # tools/prefix_profile.py -- can log launch names or cut the step short after ``_launch_limit`` launches.
    _launch_limit: Optional[int] = None
    _launch_log: Optional[list] = None

    def _count(self, n: int, name: str) -> bool:
        if self._launch_limit is not None and self._n_launch + n > self._launch_limit:
            return False
        self._n_launch += n
        if self._launch_log is not None:
            side = self._side is not None and torch.cuda.current_stream() == self._side
            self._launch_log.extend([name + (" [side]" if side else "")] * n)
        return True

    def _gemm(self, a, b, out, **kw):
        if self._count(1, f"gemm {tuple(out.shape)} k={a.shape[1] if kw.get('a_major', 0) == 0 else a.shape[0]}"):
            nv.gemm_bf16(a, b, out, **kw)

    def _gemm_splitk(self, a, b, out, cfg, **kw):
        """(BLOCK_N, split_k) GEMM: plain persistent GEMM when split_k == 1, else two launches through the workspace."""
        bn, sk = cfg
        if sk <= 1:
            return self._gemm(a, b, out, block_n=bn, **kw)
        if self._count(2, f"gemm_splitk{sk} {tuple(out.shape)} k={a.shape[1]}"):
            nv.gemm_bf16_splitk(a, b, out, self.splitk_ws, split_k=sk, block_n=bn, **kw)

    def _gemm_then(self, a, b, M, N, cfg, scratch, fuse: bool, **kw):
        """GEMM whose consumer is a residual + LayerNorm kernel.  With ``fuse`` and split_k > 1 only pass 1 runs and
        the consumer sums the fp32 partial tiles itself (returns ops.native.Partials); otherwise the result lands in
        the bf16 ``scratch`` (returned).  ``bias`` is applied by whoever produces the final sum.

        Measured back to back on one B200 (box-to-box spread is ~1 %): fusing everywhere 1.660 ms per step, no fusion
        1.675 ms: one launch and one bf16 round trip less per LayerNorm, against more loads per LayerNorm warp."""
        bn, sk = cfg
        epi = nv.EPI_BIAS if kw.get("bias") is not None else nv.EPI_NONE
        if sk <= 1:
            self._gemm(a, b, scratch, block_n=bn, epi=epi, **kw)
            return scratch
        if not fuse:
            self._gemm_splitk(a, b, scratch, cfg, epi=epi, **kw)
            return scratch
        if not self._count(1, f"gemm_partials{sk} ({M}, {N}) k={a.shape[1]}"):
            return scratch
        return nv.gemm_bf16_splitk_partials(a, b, M, N, self.splitk_ws, split_k=sk, block_n=bn, **kw)

    def _row_lens(self, rows_b: int):
        """Per-sequence valid lengths for the GELU / dGELU GEMM epilogues (padded rows skip the activation math);
        only when a 128-row tile is exactly one sequence.  Measured neutral (1.444 vs 1.445 ms per step: those
        epilogues are bound by their TMA stores, not by the arithmetic), so it is off unless DCAD_SKIP_PAD_GELU=1."""
        if self.S != 128 or os.environ.get("DCAD_SKIP_PAD_GELU", "0") == "0":
            return None
        return self.lens[:rows_b]

    @staticmethod
    def _xh(buf, rows: int, train: bool):
        """where a LayerNorm forward keeps x_hat for the backward (training steps only)"""
        return buf[:rows] if (train and buf is not None) else None

    def _ln_saved(self, y, xhat, beta):
        """(tensor, beta) operands of add_ln_bwd: the saved x_hat (beta = None) or the LayerNorm output to rebuild from"""
        return (xhat, None) if (self.save_xhat and xhat is not None) else (y, beta)

    def _k(self, fn, *args, **kw):
        if self._count(1, fn.__name__):
            return fn(*args, **kw)

    # ------------------------------------------------------------------------------------------------ forward
    def _forward(self, rows_b: int, train: bool, fuse_head: bool = False):
        """Runs the encoder on the first ``rows_b`` samples of the static input buffers; returns the tensor holding the
        CLS vectors of the last block and the element stride between consecutive samples' CLS rows."""
        d, S, pl = self.dims, self.S, self.plan
        T = rows_b * S
        H = d.n_heads
        p_h = d.dropout if train else 0.0
        p_a = d.attention_dropout if train else 0.0
        rng = self.rng if train else None
        x0 = self.x[0][:T]
        self._k(nv.embed_ln_fwd, self.ids[:T], self.word, self.pos, self.eg, self.eb, x0, self.emb_mean[:T],
                self.emb_rstd[:T], S, eps=d.ln_eps, p_drop=p_h, rng=rng, stream_id=_STREAM_EMBED)
        for l in range(d.n_layers):
            x, qkv, ctx, sa = self.x[l][:T], self.qkv[l][:T], self.ctx[l][:T], self.sa[l][:T]
            pre, hmid, tmp = self.pre[l][:T], self.hmid[l][:T], self.tmp_d[:T]
            self._gemm(x, self.p["wqkv"][l], qkv, bias=self.p["bqkv"][l], epi=nv.EPI_BIAS, block_n=pl.qkv_fwd)
            self._k(nv.attn_fwd, qkv, ctx, self.lse2[l], self.lens, rows_b, S, H, p_drop=p_a, rng=rng,
                    stream_id=_layer_stream(l, 0))
            if self.prune_last_layer and l == d.n_layers - 1:
                return self._forward_tail_pruned(l, rows_b, train, p_h, rng, fuse_head), d.dim
            self._gemm(ctx, self.p["wo"][l], tmp, bias=self.p["bo"][l], epi=nv.EPI_BIAS, block_n=pl.out_fwd)
            self._k(nv.add_ln_fwd, tmp, x, self.p["g1"][l], self.p["b1"][l], sa, self.rstd1[l][:T], eps=d.ln_eps,
                    xhat_out=self._xh(self.xh1[l], T, train))
            self._gemm(sa, self.p["w1"][l], hmid, bias=self.p["bb1"][l], aux_out=pre if train else None,
                       epi=nv.EPI_BIAS_GELU, block_n=pl.lin1_fwd, row_lens=self._row_lens(rows_b))
            ffn = self._gemm_then(hmid, self.p["w2"][l], T, d.dim, pl.lin2_fwd, tmp, pl.fuse_ln_partials >= 2,
                                  bias=self.p["bb2"][l])
            self._k(nv.add_ln_fwd, ffn, sa, self.p["g2"][l], self.p["b2"][l], self.x[l + 1][:T], self.rstd2[l][:T],
                    eps=d.ln_eps, p_drop=p_h, rng=rng, stream_id=_layer_stream(l, 1),
                    xhat_out=self._xh(self.xh2[l], T, train))
        return self.x[d.n_layers][:T], S * d.dim

    def _forward_tail_pruned(self, l: int, rows_b: int, train: bool, p_h: float, rng, fuse_head: bool = False):
        """out_lin -> LN -> FFN -> LN of the last block on the CLS rows only; returns ``[rows_b, D]``.  With
        ``fuse_head`` the last LayerNorm is left to ``_tail_head`` (fused with the classifier and their backward) and
        the not-yet-normalised FFN output is returned instead."""
        d, S, pl, P = self.dims, self.S, self.plan, self.p
        Bc = rows_b
        xc, ctxc, tmpc, sac, outc = (t[:Bc] for t in (self.c_x, self.c_ctx, self.c_tmp, self.c_sa, self.c_out))
        prec, hmidc = self.c_pre[:Bc], self.c_hmid[:Bc]
        self._k(nv.gather_rows_bf16, self.ctx[l], ctxc, Bc, d.dim, S)
        self._k(nv.gather_rows_bf16, self.x[l], xc, Bc, d.dim, S)
        self._gemm(ctxc, P["wo"][l], tmpc, bias=P["bo"][l], epi=nv.EPI_BIAS, block_n=pl.tail_bn)
        self._k(nv.add_ln_fwd, tmpc, xc, P["g1"][l], P["b1"][l], sac, self.c_rstd1[:Bc], eps=d.ln_eps,
                xhat_out=self._xh(self.c_xh1, Bc, train))
        self._gemm(sac, P["w1"][l], hmidc, bias=P["bb1"][l], aux_out=prec if train else None, epi=nv.EPI_BIAS_GELU,
                   block_n=pl.tail_bn)
        ffn = self._gemm_then(hmidc, P["w2"][l], Bc, d.dim, pl.tail_long_k, tmpc, pl.fuse_ln_partials >= 1, bias=P["bb2"][l])
        # dropout indices differ from the unpruned layout (row = sample instead of token); masks stay consistent
        # between forward and backward because both use the compact indexing
        if fuse_head:
            return ffn
        self._k(nv.add_ln_fwd, ffn, sac, P["g2"][l], P["b2"][l], outc, self.c_rstd2[:Bc], eps=d.ln_eps, p_drop=p_h,
                rng=rng, stream_id=_layer_stream(l, 1), xhat_out=self._xh(self.c_xh2, Bc, train))
        return outc

    def _tail_head(self, ffn) -> None:
        """Output LayerNorm of the last block + classifier head + cross-entropy, forward and backward, in one launch
        on the CLS rows (replaces add_ln_fwd -> cls_head_train -> add_ln_bwd)."""
        d, P, l = self.dims, self.p, self.dims.n_layers - 1
        self._k(nv.tail_head, ffn, self.c_sa, P["g2"][l], P["b2"][l], self.cw, self.cb, self.labels, self.B,
                self.loss_accum, self.logits, self.d_cw, self.d_cb, self.c_gc, self.c_gd if d.dropout > 0 else None,
                P["d_g2"][l], P["d_b2"][l], P["d_bb2"][l], cols=d.dim, eps=d.ln_eps, p_ln=d.dropout,
                ln_stream=_layer_stream(l, 1), p_head=d.head_dropout, head_stream=_STREAM_HEAD, rng=self.rng)

    # ------------------------------------------------------------------------------------------------ backward
    def _fork(self, fn) -> None:
        """Enqueue ``fn()`` (weight-gradient work that nothing on the main chain waits for) on the side stream, ordered
        after everything issued on the current stream so far."""
        if self._side is None:
            fn()
            return
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self._side.wait_event(ev)
        # Side-branch kernels are always launched WITHOUT programmatic dependent launch: an early-launched
        # persistent GEMM would sit in griddepcontrol.wait holding a whole SM's shared memory that the chain needs.
        side_pdl = os.environ.get("DCAD_SIDE_PDL", "0") != "0"
        prev = None if side_pdl else nv.set_pdl(False)
        try:
            with torch.cuda.stream(self._side):
                fn()
        finally:
            if prev is not None:
                nv.set_pdl(prev)

    def _fork_opt(self, fn) -> None:
        """Enqueue ``fn()`` (data-parallel client: per-block barrier + sharded Adam; single GPU with DCAD_ADAM_STREAM:
        the block's Adam) on the optimizer stream, ordered after everything issued so far on the weight-gradient
        stream and on the current one.  Without an optimizer stream this is ``_fork``."""
        if self._opt_stream is None or self._side is None:
            self._fork(fn)
            return
        ev = torch.cuda.Event()
        ev.record(self._side)
        self._opt_stream.wait_event(ev)
        # ... and after everything issued on the chain so far: the block's last data-gradient GEMM reads the bf16
        # weights this update rewrites (it usually ends before the weight gradients do; this makes it an ordering)
        ev_main = torch.cuda.Event()
        ev_main.record(torch.cuda.current_stream())
        self._opt_stream.wait_event(ev_main)
        prev = nv.set_pdl(False)
        try:
            with torch.cuda.stream(self._opt_stream):
                fn()
        finally:
            nv.set_pdl(prev)
        self._opt_pending = True

    def _side_mark(self):
        if self._side is None:
            return None
        ev = torch.cuda.Event()
        ev.record(self._side)
        return ev

    @staticmethod
    def _join(ev) -> None:
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)

    def _wgrad(self, dy, act, dw, cfg, dbias_src=None, dbias=None) -> None:
        """dW += dy^T act (fp32, TMA reduce-add split-K) and optionally dbias = colsum(dbias_src)."""
        self._gemm(dy, act, dw, a_major=nv.MAJOR_MN, b_major=nv.MAJOR_MN, block_n=cfg[0], split_k=cfg[1],
                   reduce_add=self.dp is not None)
        if dbias is not None:
            self._k(nv.colsum_bf16, dbias_src, dbias)

    def _wgrad_block(self, l: int, da2, G) -> None:
        """All weight gradients (and the two bias column sums) of block l in one grouped launch; side stream."""
        P, pl = self.p, self.plan
        if self._count(1, "gemm_group wgrad x4"):
            nv.gemm_bf16_group([(da2, self.hmid[l], P["d_w2"][l], pl.wgrad_lin2[1]),
                                (G.wide, self.sa[l], P["d_w1"][l], pl.wgrad_lin1[1]),
                                (G.qkv, self.x[l], P["d_wqkv"][l], pl.wgrad_qkv[1]),
                                (G.a, self.ctx[l], P["d_wo"][l], pl.wgrad_out[1])],
                               block_n=256, reduce_add=self.dp is not None, max_ctas=pl.wgrad_max_ctas)
        if not pl.fuse_colsum:
            self._k(nv.colsum_bf16, G.wide, P["d_bb1"][l])
            self._k(nv.colsum_bf16, G.qkv, P["d_bqkv"][l])

    def _wgrad_ffn(self, l: int, da2, G) -> None:
        """First half of a split block: the two FFN weight gradients (and db1) in one grouped launch; side stream."""
        P, pl = self.p, self.plan
        if self._count(1, "gemm_group wgrad ffn x2"):
            nv.gemm_bf16_group([(da2, self.hmid[l], P["d_w2"][l], pl.wgrad_lin2[1]),
                                (G.wide, self.sa[l], P["d_w1"][l], pl.wgrad_lin1[1])],
                               block_n=256, reduce_add=self.dp is not None, max_ctas=pl.wgrad_max_ctas)
        if not pl.fuse_colsum:
            self._k(nv.colsum_bf16, G.wide, P["d_bb1"][l])

    def _wgrad_attn(self, l: int, G) -> None:
        """Second half of a split block: the fused-QKV and out_lin weight gradients (and dbqkv); side stream."""
        P, pl = self.p, self.plan
        if self._count(1, "gemm_group wgrad attn x2"):
            nv.gemm_bf16_group([(G.qkv, self.x[l], P["d_wqkv"][l], pl.wgrad_qkv[1]),
                                (G.a, self.ctx[l], P["d_wo"][l], pl.wgrad_out[1])],
                               block_n=256, reduce_add=self.dp is not None, max_ctas=pl.wgrad_max_ctas)
        if not pl.fuse_colsum:
            self._k(nv.colsum_bf16, G.qkv, P["d_bqkv"][l])

    def _backward(self, early_adam: bool = False, head_fused: bool = False) -> None:
        """Backward pass.  With ``early_adam`` (training step, single-GPU client, side stream on) the Adam update of a
        block's GEMM weights is enqueued on the side stream as soon as that block's last data-gradient GEMM has been
        issued: the HBM-bound optimizer pass then runs underneath the tensor-bound backward of the earlier blocks, and
        ``_optimizer`` only has the embeddings / biases / LayerNorm / classifier region left."""
        d, S, B, pl = self.dims, self.S, self.B, self.plan
        in_step = early_adam  # called from the training step (as opposed to a bare backward in tests / tools)
        early_adam = early_adam and self._side is not None and (self.dp is None or self.dp_early_adam)
        self._adam_early_done = early_adam
        H, D = d.n_heads, d.dim
        MN = nv.MAJOR_MN
        top = d.n_layers - 1
        P = self.p
        side_done = {}  # layer -> event marking the end of its side-stream work
        if self.prune_last_layer:
            dy1, dy2 = self._backward_tail_pruned(top, head_fused)
            if in_step:  # training step: from here on the second branch has work (see use_pdl)
                nv.set_pdl(self.use_pdl)
        else:
            # gradient wrt the last hidden states: only the CLS rows are non-zero
            dy1, dy2 = self.gset[(top + 1) % 2].a, None
            self._k(nv.scatter_rows_bf16, self.dcls, dy1, B, D, S)
        for l in reversed(range(d.n_layers)):
            x, qkv, ctx, sa, pre, hmid = self.x[l], self.qkv[l], self.ctx[l], self.sa[l], self.pre[l], self.hmid[l]
            G = self.gset[l % 2]
            self._join(side_done.pop(l + 2, None))  # the side work of layer l+2 read this scratch set
            # Block 0 is the last one of the backward pass: its weight gradients and its Adam update are what the step
            # ends with, so they are issued in two halves -- (lin2, lin1) as soon as their inputs exist, with their
            # Adam behind them, and only (qkv, out_lin) after the attention backward -- instead of one grouped launch
            # plus one Adam over the whole block after it.
            split = (l == 0 and top > 0 and pl.group_wgrad and pl.split_last_wgrad and early_adam and self.dp is None)
            lin1_off = self.arena.layout.offsets[f"distilbert.transformer.layer.{l}.ffn.lin1.weight"]
            if self.prune_last_layer and l == top:
                # the tail of this block already ran on the CLS rows: dy1 = d(ctx) scattered to all tokens, dy2 = the
                # residual gradient wrt x (non-zero on the CLS rows only)
                dctx, dz1 = dy1, dy2
            else:
                # output_layer_norm(dropout(ffn_out) + sa)
                dz2 = G.c
                da2 = G.d if d.dropout > 0 else dz2  # without dropout the two gradients coincide
                y2, beta2 = self._ln_saved(self.x[l + 1], self.xh2[l], P["b2"][l])
                self._k(nv.add_ln_bwd, dy1, dy2, y2, P["g2"][l], beta2, self.rstd2[l], dz2,
                        da2 if d.dropout > 0 else None, P["d_g2"][l], P["d_b2"][l], P["d_bb2"][l], p_drop=d.dropout,
                        rng=self.rng, stream_id=_layer_stream(l, 1))
                # lin2: dW2 = da2^T hmid (side) ; dpre = (da2 W2) * gelu'(pre)
                if not pl.group_wgrad:
                    self._fork(lambda: self._wgrad(da2, hmid, P["d_w2"][l], pl.wgrad_lin2))
                self._gemm(da2, P["w2"][l], G.wide, b_major=MN, aux=pre, epi=nv.EPI_DGELU, block_n=pl.lin2_dgrad,
                           max_ctas=pl.lin2_dgrad_ctas, colsum=P["d_bb1"][l] if pl.fuse_colsum else None,
                           row_lens=self._row_lens(B))
                # lin1: dW1 = dpre^T sa, db1 = colsum(dpre) (side) ; dsa = dpre W1
                if not pl.group_wgrad:
                    self._fork(lambda: self._wgrad(G.wide, sa, P["d_w1"][l], pl.wgrad_lin1, G.wide,
                                                   None if pl.fuse_colsum else P["d_bb1"][l]))
                if split:
                    self._fork(lambda: self._wgrad_ffn(l, da2, G))
                dsa = self._gemm_then(G.wide, P["w1"][l], self.T, D, pl.lin1_dgrad, G.b, pl.fuse_ln_partials >= 2,
                                      b_major=MN)
                if split:  # issued after the lin1 data-gradient GEMM so that the chain's kernel is ahead in the queue
                    self._fork_opt(lambda: self._adam_range(lin1_off, self.arena.layout.qkv_weight_offset(l + 1)))
                # sa_layer_norm(attn_out + x): dz1 is the gradient wrt both attn_out and the residual x
                dz1 = G.a
                y1, beta1 = self._ln_saved(sa, self.xh1[l], P["b1"][l])
                self._k(nv.add_ln_bwd, dsa, dz2, y1, P["g1"][l], beta1, self.rstd1[l], dz1, None, P["d_g1"][l],
                        P["d_b1"][l], P["d_bo"][l])
                # out_lin: dWo = dz1^T ctx (side) ; dctx = dz1 Wo
                if not pl.group_wgrad:
                    self._fork(lambda: self._wgrad(dz1, ctx, P["d_wo"][l], pl.wgrad_out))
                dctx = G.e
                self._gemm(dz1, P["wo"][l], dctx, b_major=MN, block_n=pl.out_dgrad)
            # attention
            self._k(nv.attn_bwd, qkv, ctx, dctx, self.lse2[l], G.qkv, self.lens, B, S, H,
                    p_drop=d.attention_dropout, rng=self.rng, stream_id=_layer_stream(l, 0),
                    dbias=P["d_bqkv"][l] if pl.fuse_colsum else None)
            # QKV: dWqkv = dqkv^T x, dbqkv = colsum(dqkv) (side) ; dx = dqkv Wqkv
            if split:
                self._fork(lambda: self._wgrad_attn(l, G))
            elif pl.group_wgrad and not (self.prune_last_layer and l == top):
                self._fork(lambda: self._wgrad_block(l, da2, G))  # dz1 of this block is G.a
            else:
                self._fork(lambda: self._wgrad(G.qkv, x, P["d_wqkv"][l], pl.wgrad_qkv, G.qkv,
                                               None if pl.fuse_colsum else P["d_bqkv"][l]))
            # block 0 feeds the embedding backward, which wants a materialised bf16 tensor
            dxq = self._gemm_then(G.qkv, P["wqkv"][l], self.T, D, pl.qkv_dgrad, G.b,
                                  pl.fuse_ln_partials >= 2 and l > 0, b_major=MN)
            if early_adam:  # nothing reads this block's weights any more, and its weight gradients precede on the side stream
                lay = self.arena.layout
                lo = lay.qkv_weight_offset(l)
                hi = lay.qkv_weight_offset(l + 1) if l + 1 < d.n_layers else lay.gemm_end
                if self.dp is None:
                    self._fork_opt(lambda: self._adam_range(lo, lin1_off if split else hi))
                else:
                    self._fork_opt(lambda: self._adam_block_dp(l, lo, hi))
            side_done[l] = self._side_mark()
            dy1, dy2 = dxq, dz1  # gradient wrt x_l = via QKV projection + via the residual
        self._k(nv.embed_ln_bwd, dy1, dy2, self.ids, self.word, self.pos, self.eg, self.emb_mean, self.emb_rstd,
                self.d_word, self.d_pos, self.d_eg, self.d_eb, S, lens=self.lens, p_drop=d.dropout, rng=self.rng,
                stream_id=_STREAM_EMBED, touched=self.touched_rows)
        for ev in side_done.values():  # every weight gradient is complete before the optimizer reads it
            self._join(ev)
        if getattr(self, "_opt_pending", False):  # ... and so are the per-block updates of a data-parallel client
            ev = torch.cuda.Event()
            ev.record(self._opt_stream)
            self._join(ev)
            self._opt_pending = False

    def _backward_tail_pruned(self, l: int, head_fused: bool = False):
        """Backward of the pruned tail (LN2, FFN, LN1, out_lin of the last block) on the B CLS rows.  Returns
        (d ctx scattered to [T, D], residual gradient wrt x scattered to [T, D])."""
        d, S, B, pl, P = self.dims, self.S, self.B, self.plan, self.p
        D = d.dim
        MN = nv.MAJOR_MN
        G = self.gset[l % 2]
        dz2, da2 = self.c_gc, (self.c_gd if d.dropout > 0 else self.c_gc)
        if not head_fused:  # otherwise _tail_head already produced dz2 / da2 and the LayerNorm parameter gradients
            y2, beta2 = self._ln_saved(self.c_out, self.c_xh2, P["b2"][l])
            self._k(nv.add_ln_bwd, self.dcls, None, y2, P["g2"][l], beta2, self.c_rstd2, dz2,
                    da2 if d.dropout > 0 else None, P["d_g2"][l], P["d_b2"][l], P["d_bb2"][l], p_drop=d.dropout,
                    rng=self.rng, stream_id=_layer_stream(l, 1))
        self._gemm(da2, P["w2"][l], self.c_gwide, b_major=MN, aux=self.c_pre, epi=nv.EPI_DGELU, block_n=pl.tail_bn,
                   colsum=P["d_bb1"][l] if pl.fuse_colsum else None)
        dsa = self._gemm_then(self.c_gwide, P["w1"][l], B, D, pl.tail_long_k, self.c_gb, pl.fuse_ln_partials >= 1, b_major=MN)
        dz1 = self.c_ga
        y1, beta1 = self._ln_saved(self.c_sa, self.c_xh1, P["b1"][l])
        self._k(nv.add_ln_bwd, dsa, dz2, y1, P["g1"][l], beta1, self.c_rstd1, dz1, None, P["d_g1"][l],
                P["d_b1"][l], P["d_bo"][l])
        dctx_c = self.c_ge
        self._gemm(dz1, P["wo"][l], dctx_c, b_major=MN, block_n=pl.tail_bn)
        self._k(nv.scatter_rows_bf16, dctx_c, G.e, B, D, S)
        self._k(nv.scatter_rows_bf16, dz1, G.a, B, D, S)

        # The weight gradients of this block come from the CLS rows only (plain stores; reduce-add into the owner
        # GPU's arena for a data-parallel client).  Nothing on the chain above needs them: one grouped launch on the
        # side stream, which is idle at this point of the step.
        def tail_wgrads():
            if self._count(1, "gemm_group tail wgrad x3"):
                nv.gemm_bf16_group([(da2, self.c_hmid, P["d_w2"][l], 1), (self.c_gwide, self.c_sa, P["d_w1"][l], 1),
                                    (dz1, self.c_ctx, P["d_wo"][l], 1)],
                                   block_n=pl.tail_wgrad_bn, reduce_add=self.dp is not None)
            if not pl.fuse_colsum:
                self._k(nv.colsum_bf16, self.c_gwide, P["d_bb1"][l])

        self._fork(tail_wgrads)
        return G.e, G.a

    def _adam_range(self, lo: int, hi: int, advance: bool = False) -> None:
        """Adam + zero_grad (+ bf16 mirror refresh for the GEMM-weight part) on arena elements [lo, hi)."""
        a = self.arena
        bf_hi = min(hi, a.layout.gemm_end)
        w0 = a.layout.offsets[_WORD_EMB]
        w1 = w0 + a.layout.numel(_WORD_EMB)
        # GEMM-weight gradients that are written with plain stores every step need no clearing (4 of 34 bytes per
        # parameter); everything accumulated with atomics / reduce-add does
        pl = self.plan
        stored = (self.dp is None and hi <= a.layout.gemm_end and os.environ.get("DCAD_ADAM_SKIP_ZERO", "1") != "0"
                  and max(pl.wgrad_lin1[1], pl.wgrad_lin2[1], pl.wgrad_qkv[1], pl.wgrad_out[1]) == 1)
        args = (a.flat[lo:hi], a.grad[lo:hi], a.exp_avg[lo:hi], a.exp_avg_sq[lo:hi],
                a.bf16[lo:bf_hi] if bf_hi > lo else None, hi - lo, 0, max(bf_hi - lo, 0), 0, 0 if stored else hi - lo,
                self.lr, self.betas[0], self.betas[1], self.eps, self.rng)
        ticket = self._step_ticket if advance else None  # last launch of the step also advances the step counter
        if (lo <= w0 and w1 <= hi and lo >= a.layout.gemm_end and self.dp is None
                and os.environ.get("DCAD_ADAM_ROW_KERNEL", "0") != "0"):
            # Option (off): the word-embedding table through the row-indexed kernel of the data-parallel path (one warp
            # per row) and the two small dense pieces around it through the flat kernel, instead of ONE flat launch.
            # Measured on one B200: 1.2825 vs 1.2704 ms per step (two more launches at the end of the step, and one
            # warp per row is latency-bound); the flat kernel has since got its own row-group walk (csrc/adam.cu).
            def dense(d_lo: int, d_hi: int, tk) -> None:
                if d_hi <= d_lo:
                    return
                d_args = (a.flat[d_lo:d_hi], a.grad[d_lo:d_hi], a.exp_avg[d_lo:d_hi], a.exp_avg_sq[d_lo:d_hi], None,
                          d_hi - d_lo, 0, 0, 0, d_hi - d_lo, self.lr, self.betas[0], self.betas[1], self.eps, self.rng)
                if tk is not None:
                    self._k(nv.adam_step_rows, *d_args, None, 0, 0, 4, tk)
                else:
                    self._k(nv.adam_step, *d_args)

            self._k(nv.adam_step_dp_rows, a.exp_avg[w0:w1], a.exp_avg_sq[w0:w1], a.grad[w0:w1],
                    [a.flat.data_ptr() + w0 * 4], [a.grad.data_ptr() + w0 * 4], [self.touched_rows.data_ptr()],
                    0, self.dims.vocab_size, self.dims.dim, self.lr, self.betas[0], self.betas[1], self.eps, 1.0,
                    self.rng)
            last_is_tail = hi > w1
            dense(lo, w0, None if last_is_tail else ticket)
            dense(w1, hi, ticket)
        elif lo <= w0 and w1 <= hi:  # the range holds the word-embedding table: skip rows without optimizer state
            self._k(nv.adam_step_rows, *args, self.touched_rows, w0 - lo, w1 - lo, self.dims.dim, ticket)
        elif advance:
            self._k(nv.adam_step_rows, *args, None, 0, 0, 4, ticket)
        else:
            self._k(nv.adam_step, *args)

    def _optimizer(self) -> None:
        a = self.arena
        lay = a.layout
        if self.dp is None:
            # GEMM weights were already updated block by block under the backward pass when early_adam was on
            self._adam_range(lay.gemm_end if getattr(self, "_adam_early_done", False) else 0, lay.total, advance=True)
            self._adam_early_done = False
            return
        # ---- data-parallel client: reduce-scatter (already done tile by tile for the GEMM weights by the wgrad
        # epilogues) -> sharded Adam with the parameter all-gather fused in -> barrier ----
        dp = self.dp
        G, r = dp.world, dp.rank
        scale = 1.0 / G
        adam = (self.lr, self.betas[0], self.betas[1], self.eps, scale, self.rng)
        # barrier phases 0..5 belong to the per-block updates of _adam_block_dp; the two of the step's end are 6 and 7
        self._k(nv.xgpu_barrier, dp.pad_ptrs, r, self.rng, 6)          # every peer's backward (and its pushes) is done
        if not getattr(self, "_adam_early_done", False):
            for glo, ghi in dp.gemm_ranges(r):
                self._k(nv.adam_step_dp, a.flat, a.grad, a.exp_avg, a.exp_avg_sq, glo, ghi, 0, lay.gemm_end, 0, 0, *adam,
                        None, dp.mirror_ptrs, r)
        self._adam_early_done = False

        def dense_range(lo: int, hi: int) -> None:
            """biases / LayerNorm / classifier / position table: sum the G locally accumulated copies of [lo, hi),
            Adam, push the fp32 values into every peer's arena"""
            if hi <= lo:
                return
            self._k(nv.fedavg_local, [p + lo * 4 for p in dp.grad_ptrs], dp.grad_ptrs[r] + lo * 4, hi - lo,
                    weights=[1.0] * G)
            self._k(nv.adam_step_dp, a.flat, a.grad, a.exp_avg, a.exp_avg_sq, lo, hi, 0, 0, lo, hi, *adam,
                    dp.param_ptrs, None, r)

        if self.dp_sparse_emb:
            w0 = lay.offsets[_WORD_EMB]
            w1 = w0 + lay.numel(_WORD_EMB)
            # the small dense pieces around the word table: first piece on rank 0, second on the last rank
            if r == 0:
                dense_range(lay.gemm_end, w0)
            if r == G - 1:
                dense_range(w1, lay.total)
            # the word table, row-sparse: only rows some GPU has ever touched are reduced, updated and re-distributed
            row_lo, row_hi = dp.word_rows(r)
            first = dp.self_first
            self._k(nv.adam_step_dp_rows, a.exp_avg[w0:w1], a.exp_avg_sq[w0:w1], a.grad[w0:w1],
                    first([p + w0 * 4 for p in dp.param_ptrs]), first([p + w0 * 4 for p in dp.grad_ptrs]),
                    first(dp.flag_ptrs), row_lo, row_hi, self.dims.dim, *adam)
        else:
            lo, hi = dp.r2_range(r)
            dense_range(lo, hi)
        self._k(nv.xgpu_barrier, dp.pad_ptrs, r, self.rng, 7)          # all pushed parameters have landed everywhere
        # the locally accumulated gradients outside the GEMM region have been consumed by their owners: clear them
        a.grad[lay.gemm_end: lay.total].zero_()
        self._k(nv.step_advance, self.rng)

    def _adam_block_dp(self, l: int, lo: int, hi: int) -> None:
        """Data-parallel client, side stream: once EVERY GPU of the group has reduce-added its weight gradients of
        block ``l`` into the owners' arenas (one cross-GPU barrier; nothing reads this block's weights again in this
        step), the owner updates its share of the block and stores the new bf16 weights into every peer's mirror --
        the optimizer and the parameter all-gather of block l run under the backward pass of blocks l-1 .. 0."""
        dp, a = self.dp, self.arena
        self._k(nv.xgpu_barrier, dp.pad_ptrs, dp.rank, self.rng, self.dims.n_layers - 1 - l)
        for glo, ghi in dp.gemm_ranges(dp.rank):
            b0, b1 = max(lo, glo), min(hi, ghi)
            if b1 > b0:
                self._k(nv.adam_step_dp, a.flat, a.grad, a.exp_avg, a.exp_avg_sq, b0, b1, 0, a.layout.gemm_end, 0, 0,
                        self.lr, self.betas[0], self.betas[1], self.eps, 1.0 / dp.world, self.rng, None, dp.mirror_ptrs,
                        dp.rank)

    # ------------------------------------------------------------------------------------------------ public API
    def _train_step_kernels(self) -> None:
        prev_pdl = nv.set_pdl(self.use_pdl_fwd)
        try:
            self._train_step_kernels_inner()
        finally:
            nv.set_pdl(prev_pdl)

    def _train_step_kernels_inner(self) -> None:
        self._n_launch = 0
        fuse_head = self.prune_last_layer and self.plan.fuse_tail_head and self.B <= 64
        h, stride = self._forward(self.B, train=True, fuse_head=fuse_head)
        if fuse_head:
            self._tail_head(h)
        else:
            self._k(nv.cls_head_train, h, stride, self.cw, self.cb, self.labels, self.B, self.loss_accum,
                    self.logits, self.d_cw, self.d_cb, self.dcls, cols=self.dims.dim, p_drop=self.dims.head_dropout,
                    rng=self.rng, stream_id=_STREAM_HEAD)
        self._backward(early_adam=True, head_fused=fuse_head)
        self._optimizer()
        self.launches_per_train_step = self._n_launch

    def set_batch(self, input_ids: torch.Tensor, lens: torch.Tensor, labels: torch.Tensor) -> None:
        """Device-side batch set-up (inputs already on the GPU, e.g. from the device-resident dataset)."""
        self.ids.copy_(input_ids.reshape(-1))
        self.lens.copy_(lens)
        self.labels.copy_(labels)

    def capture(self) -> None:
        """Warm up eagerly, then capture the training step (and its end-to-end variant) into CUDA graphs."""
        # the critical chain outranks the weight-gradient branch
        s = torch.cuda.Stream(device=self.device, priority=int(os.environ.get("DCAD_MAIN_PRIO", "-1")))
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            saved = (self.arena.flat.clone(), self.rng.clone())
            for _ in range(2):
                self._train_step_kernels()
            s.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                self._train_step_kernels()
            self._graph_train = g
            g2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g2, stream=s):
                self.d_in.copy_(self.h_in, non_blocking=True)
                self._train_step_kernels()
                self.h_loss.copy_(self.loss_accum, non_blocking=True)
            self._graph_train_e2e = g2
            # evaluation forward (dropout off) on a full batch; partial last batches run eagerly
            self._eval_kernels(self.B, True)
            s.synchronize()
            g3 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g3, stream=s):
                self._eval_kernels(self.B, True)
            self._graph_eval = g3
            self.eval_loss_accum.zero_()
            self.confusion.zero_()
            # undo the warm-up steps
            self.arena.flat.copy_(saved[0])
            self.rng.copy_(saved[1])
            self.reset_optimizer()
            self.refresh_bf16()
            self.loss_accum.zero_()
            if self.dp is not None:  # the step counter was rewound: the barrier epochs restart too
                s.synchronize()
                self.dp.reset_barrier_flags()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize(self.device)

    def train_step(self) -> None:
        """One optimisation step on the batch currently in the static input buffers (no host sync)."""
        if self._graph_train is not None:
            self._graph_train.replay()
        else:
            self._train_step_kernels()

    def train_step_from_host(self, ids: torch.Tensor, lens: torch.Tensor, labels: torch.Tensor) -> float:
        """End-to-end step as a user calls it: host tensors in, running mean loss out.  Copies the batch through
        pinned memory (one H2D), runs the step, reads back the loss accumulators (one D2H, one sync)."""
        T, B = self.T, self.B
        hn = self._h_in_np  # numpy views of the pinned staging buffers: three memcpys, no tensor-op dispatch
        hn[:T] = ids.numpy().reshape(-1)
        hn[T: T + B] = lens.numpy()
        hn[T + B:] = labels.numpy()
        if self._graph_train_e2e is not None:
            self._graph_train_e2e.replay()
        else:
            self.d_in.copy_(self.h_in, non_blocking=True)
            self._train_step_kernels()
            self.h_loss.copy_(self.loss_accum, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        hl = self._h_loss_np
        return float(hl[0]) / max(float(hl[1]), 1.0)

    @property
    def h2d_bytes_per_step(self) -> int:
        return self.h_in.numel() * 4

    @property
    def d2h_bytes_per_step(self) -> int:
        return self.h_loss.numel() * 4

    def read_loss(self, reset: bool = True) -> Tuple[float, int]:
        """(mean of batch-mean losses since the last reset, number of batches) — one D2H sync."""
        v = self.loss_accum.tolist()
        if reset:
            self.loss_accum.zero_()
        n = int(v[1])
        return (v[0] / n if n else 0.0), n

    # ---- evaluation ----
    def eval_reset(self) -> None:
        self.eval_loss_accum.zero_()
        self.confusion.zero_()

    def eval_batch(self, rows_b: Optional[int] = None, want_probs: bool = False) -> None:
        """Forward (dropout off) + classifier metrics on the first ``rows_b`` samples of the static inputs."""
        rows_b = self.B if rows_b is None else rows_b
        g = getattr(self, "_graph_eval", None)
        if g is not None and rows_b == self.B:   # full batches replay a captured graph (probs / preds always written)
            g.replay()
            return
        self._eval_kernels(rows_b, want_probs)

    def _eval_kernels(self, rows_b: int, want_probs: bool) -> None:
        h, stride = self._forward(rows_b, train=False)
        self._k(nv.cls_head_eval, h, stride, self.cw, self.cb, self.labels, rows_b,
                self.eval_loss_accum, self.confusion, self.probs if want_probs else None,
                self.preds if want_probs else None, cols=self.dims.dim)

    def eval_read(self) -> Tuple[float, int, torch.Tensor]:
        """(sum of batch-mean losses, number of batches, 2x2 confusion matrix [true, pred]) — one sync."""
        la = self.eval_loss_accum.tolist()
        cm = self.confusion.cpu().view(2, 2).to(torch.int64)
        return la[0], int(la[1]), cm

    def logits_eval(self, rows_b: Optional[int] = None) -> torch.Tensor:
        """Logits in eval mode for the current static batch (used by parity tests)."""
        rows_b = self.B if rows_b is None else rows_b
        h, stride = self._forward(rows_b, train=False)
        cls = (h if stride == self.dims.dim else h.view(rows_b, self.S, self.dims.dim)[:, 0, :]).float()
        return cls @ self.cw.t() + self.cb
