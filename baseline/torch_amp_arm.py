"""Secondary, builder-run baseline: the reference's model and training loop at MATCHED precision.

The driver's reference arm (``bench.py --impl reference``, ``baseline/reference_arm.py``) runs the reference exactly as
published: strict fp32, no TF32, no autocast (client1.py:379-380).  This repository computes in bf16 on the tensor
cores with fp32 master weights and Adam moments, so next to the stock ratio the fair question is how fast the
reference's OWN code is when PyTorch is allowed the same arithmetic.  This arm answers it: the unmodified
``DDoSClassifier`` and ``train_model`` from ``baseline/_ref/client1.py``, same synthetic data, tokenizer, batch size and
optimizer, with the forward pass wrapped in ``torch.autocast(device_type='cuda', dtype=torch.bfloat16)`` and TF32
enabled.  Nothing else changes (per-sample tokenisation, H2D copies and the two ``loss.item()`` syncs per step stay).

    python bench.py --impl reference-amp --steps 100 --warmup 10

prints the same JSON line with ``"impl": "reference-bf16-autocast"``; ``bench.py`` (ours, one GPU) also runs it in
process after its own measurement and reports ``vs_reference_bf16_autocast``.
"""
from __future__ import annotations

import reference_arm


def run(args, rank: int, world: int, local_rank: int):
    return reference_arm.run(args, rank, world, local_rank, amp=True)
