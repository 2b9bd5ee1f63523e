"""Launch the non-GEMM kernels of the training step once each at the bench shapes (batch 16, seq 128, DistilBERT-base)
-- the target of `ncu --set full` captures of attention / LayerNorm / head / embedding kernels."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
import importlib  # noqa: E402

eng_mod = importlib.import_module(f"{PKG}.models.engine")
arena_mod = importlib.import_module(f"{PKG}.models.arena")
synth = importlib.import_module(f"{PKG}.data.synth")

dims = arena_mod.ModelDims()
eng = eng_mod.B200Engine(dims, batch=16, seq=128, init_seed=0, overlap_wgrad=False)
ids, lens, labels = synth.synth_token_batch(16, 128, dims.vocab_size, seed=1)
eng.set_batch(ids.cuda(), lens.cuda(), labels.cuda())
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    eng._train_step_kernels()
torch.cuda.synchronize()
print("ok", eng.launches_per_train_step)
