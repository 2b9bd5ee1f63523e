"""Per-role timeline (globaltimer stamps of CTA 0) of selected GEMMs: where does a tile's time go?"""
import ctypes
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.ops import native as nv  # noqa: E402
from tools.tune_gemm import GEMMS  # noqa: E402

L = nv.lib()
L.dcad_gemm_set_timing_buffer.argtypes = [ctypes.c_void_p]
L.dcad_gemm_set_timing_buffer.restype = None
tb = torch.zeros(8, dtype=torch.int64, device="cuda")
names = ["start", "producer_done", "mma_first_full", "mma_issued_all", "epi_acc_ready(last)", "epi_done(last)", "exit"]
# (gemm name, block_n, epilogue override or None, keep aux_out?)
import os  # noqa: E402

if os.environ.get("TIMELINE_MODES"):   # compare the epilogue variants on a few shapes
    MODES = [int(x) for x in os.environ["TIMELINE_MODES"].split(",")]
    CASES = [("qkv_fwd", 256, None, False), ("out_fwd", 128, None, False), ("lin1_fwd", 256, 2, True),
             ("lin1_fwd", 192, 2, True), ("lin1_fwd", 384, 2, True), ("lin1_wgrad", 256, None, False)]
else:
    MODES = [None]
    CASES = None
CASES = CASES or [("lin1_fwd", 256, 0, False), ("lin1_fwd", 256, 1, False), ("lin1_fwd", 256, 2, False), ("lin1_fwd", 256, 2, True),
         ("lin2_dgrad", 256, 0, False), ("lin2_dgrad", 256, 3, False), ("lin2_fwd", 256, None, False),
         ("out_fwd", 128, None, False), ("lin1_wgrad", 256, None, False), ("qkv_fwd", 256, None, False)]
for mode in MODES:
  if mode is not None:
    nv.set_gemm_epilogue(mode)
    print(f"---- epilogue mode {mode}")
  for (gname, bn, epi_o, keep_aux) in CASES:
      name, M, N, K, am, bm, epi = [g for g in GEMMS if g[0] == gname][0]
      if epi_o is not None:
          epi = epi_o
      A = torch.randn((M, K) if am == 0 else (K, M), device="cuda").bfloat16()
      B = (torch.randn((N, K) if bm == 0 else (K, N), device="cuda") * 0.05).bfloat16()
      out = torch.empty(M, N, device="cuda", dtype=torch.float32 if am == 1 else torch.bfloat16)
      bias = torch.randn(N, device="cuda") if epi in (1, 2) else None
      aux = torch.randn(M, N, device="cuda").bfloat16() if epi == 3 else None
      aux_out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16) if (epi == 2 and keep_aux) else None
      for it in range(3):
          L.dcad_gemm_set_timing_buffer(tb.data_ptr() if it == 2 else None)
          e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
          e0.record()
          nv.gemm_bf16(A, B, out, a_major=am, b_major=bm, bias=bias, aux=aux, aux_out=aux_out, epi=epi, block_n=bn)
          e1.record()
          torch.cuda.synchronize()
      L.dcad_gemm_set_timing_buffer(None)
      t = tb.tolist()
      base = t[0]
      print(f"{gname} bn={bn} epi={epi} aux_out={keep_aux} ({M}x{N}x{K}) :: " +
            " ".join(f"{n}={(v-base)/1e3:.2f}us" for n, v in zip(names[1:], t[1:7])) +
          f" | launch-to-end {e0.elapsed_time(e1)*1e3:.1f}us", flush=True)
