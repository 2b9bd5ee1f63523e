"""`bench.py --impl reference`: the UNMODIFIED reference (baseline/_ref/client1.py, installed from /root/reference —
see DESIGN.md "Reference install") measured through its own public API and stock code path:

    model  = client1.DDoSClassifier('./distilbert-base-uncased')        (client1.py:53-58, 374)
    loader = DataLoader(client1.CICIDS2017Dataset(texts, labels, DistilBertTokenizer), batch_size=16, shuffle=True)
    client1.train_model(model, loader, nn.CrossEntropyLoss(), Adam(lr=2e-5), num_epochs=1, device)   (client1.py:96-115)

None of this repository's models, kernels or engine are on that path.  The harness only supplies what the reference
expects to find on disk (it has no network either): a ``./distilbert-base-uncased`` directory holding a random-init
DistilBERT-base (HF ``DistilBertModel(DistilBertConfig()).save_pretrained``) and a synthetic 30,522-entry
``vocab.txt``, synthetic CICIDS-shaped flow rows rendered by the reference's own ``features_to_text``, and stub
``matplotlib`` / ``seaborn`` modules (not installed in this image; the reference imports them at module scope but
training never calls them).

The timed region brackets the reference's ``train_model`` call (W warm-up batches in a first call, K timed batches
in a second) with CUDA events + synchronize; it therefore includes what the reference's loop includes every step: the
per-sample tokenisation inside ``Dataset.__getitem__``, three H2D copies and two ``loss.item()`` syncs.
"""
from __future__ import annotations

import contextlib
import importlib.util
import io
import json
import os
import sys
import tempfile
import types

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
REF_DIRS = [os.path.join(HERE, "_ref"), "/root/reference"]


def find_reference():
    for d in REF_DIRS:
        if os.path.exists(os.path.join(d, "client1.py")):
            return d
    return None


def _stub_plotting():
    for name in ("matplotlib", "matplotlib.pyplot", "seaborn"):
        if name in sys.modules:
            continue
        try:
            __import__(name)
        except Exception:  # noqa: BLE001
            m = types.ModuleType(name)
            m.use = lambda *a, **k: None
            sys.modules[name] = m
    if not hasattr(sys.modules["matplotlib"], "pyplot"):
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]


def load_reference(name: str = "client1"):
    d = find_reference()
    if d is None:
        return None
    _stub_plotting()
    spec = importlib.util.spec_from_file_location(f"_reference_{name}", os.path.join(d, f"{name}.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def make_model_dir(path: str) -> str:
    """A local 'distilbert-base-uncased' the reference can ``from_pretrained``: random-init weights + synthetic vocab."""
    from transformers import DistilBertConfig, DistilBertModel

    sys.path.insert(0, ROOT)
    tok_mod = importlib.import_module(f"{PKG}.data.tokenizer")
    os.makedirs(path, exist_ok=True)
    if not os.path.exists(os.path.join(path, "config.json")):
        import torch

        torch.manual_seed(0)
        DistilBertModel(DistilBertConfig()).save_pretrained(path)
    tok_mod.write_vocab(os.path.join(path, "vocab.txt"))
    with open(os.path.join(path, "tokenizer_config.json"), "w") as f:
        json.dump({"do_lower_case": True, "model_max_length": 512, "tokenizer_class": "DistilBertTokenizer"}, f)
    return path


def run(args, rank: int, world: int, local_rank: int, amp: bool = False):
    """``amp=False`` (the driver's reference arm): the stock reference, nothing changed.  ``amp=True`` is the secondary,
    precision-matched arm of ``baseline/torch_amp_arm.py``: same model, data and ``train_model``, with the forward pass
    run under ``torch.autocast(bfloat16)`` and TF32 allowed for what stays in fp32."""
    impl = "reference-bf16-autocast" if amp else "reference"

    def unavailable(why: str):
        return {"impl": impl, "unavailable": why}  # bench.py prints it on rank 0 and exits 0

    if find_reference() is None:
        return unavailable("baseline/_ref/client1.py not found (run baseline/install_reference.sh)")
    try:
        import torch
        import torch.nn as nn
        from torch.utils.data import DataLoader
        from transformers import DistilBertTokenizer
    except Exception as e:  # noqa: BLE001
        return unavailable(f"import failed: {e!r}")

    sys.path.insert(0, ROOT)
    synth = importlib.import_module(f"{PKG}.data.synth")
    cic = importlib.import_module(f"{PKG}.data.cicids")
    clocks = importlib.import_module(f"{PKG}.utils.clocks")
    try:
        ref = load_reference("client1")
    except Exception as e:  # noqa: BLE001
        return unavailable(f"reference import failed: {e!r}")

    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank) if use_cuda else torch.device("cpu")
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl" if use_cuda else "gloo",
                                **({"device_id": device} if use_cuda else {}))

    work = tempfile.mkdtemp(prefix=f"ref_arm_{rank}_")
    cwd = os.getcwd()
    os.chdir(work)
    try:
        model_dir = make_model_dir(os.path.join(work, "distilbert-base-uncased"))
        tokenizer = DistilBertTokenizer.from_pretrained(model_dir)            # client1.py:364
        B = args.batch
        n_rows = (args.warmup + args.steps) * B
        df = cic.clean_frame(synth.synth_flows(n_rows, seed=100 + rank))
        texts = df.apply(ref.features_to_text, axis=1).tolist()               # the reference's own text rendering
        labels = [1 if x == "DDoS" else 0 for x in df["Label"]]
        nw = args.warmup * B
        warm = (DataLoader(ref.CICIDS2017Dataset(texts[:nw], labels[:nw], tokenizer), batch_size=B, shuffle=True)
                if nw > 0 else None)
        timed = DataLoader(ref.CICIDS2017Dataset(texts[nw:], labels[nw:], tokenizer), batch_size=B, shuffle=True)
        sample = [int(tokenizer(t, add_special_tokens=True, truncation=True, max_length=128,
                                return_tensors="pt")["input_ids"].shape[1]) for t in texts[:256]]
        token_lens = {"mean": sum(sample) / len(sample), "min": min(sample), "max": max(sample), "samples": len(sample)}
        model = ref.DDoSClassifier(model_dir).to(device)                      # client1.py:374
        if amp:
            # bf16 tensor-core matmuls with fp32 master weights / Adam (the mixed precision this repository's engine
            # uses): the reference's own forward under autocast; its train_model (loss, backward, Adam) is unchanged
            torch.backends.cuda.matmul.allow_tf32 = True
            torch.backends.cudnn.allow_tf32 = True
            stock_forward = model.forward

            def autocast_forward(*a, **k):
                with torch.autocast(device_type=device.type, dtype=torch.bfloat16):
                    return stock_forward(*a, **k).float()

            model.forward = autocast_forward
        opt = torch.optim.Adam(model.parameters(), lr=2e-5)                   # client1.py:380
        crit = nn.CrossEntropyLoss()

        def sync_all():
            if world > 1:
                import torch.distributed as dist
                dist.barrier()
            if use_cuda:
                torch.cuda.synchronize(device)

        quiet = io.StringIO()
        with contextlib.redirect_stdout(quiet), contextlib.redirect_stderr(quiet):
            if args.warmup > 0:
                ref.train_model(model, warm, crit, opt, 1, device)            # untimed warm-up
        sync_all()
        sampler = clocks.ClockSampler(gpu_index=local_rank).start() if (rank == 0 and use_cuda) else None
        import time as _t
        if use_cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        t0 = _t.perf_counter()
        with contextlib.redirect_stdout(quiet), contextlib.redirect_stderr(quiet):
            ref.train_model(model, timed, crit, opt, 1, device)               # the reference's stock training loop
        if use_cuda:
            e1.record()
            torch.cuda.synchronize(device)
            ms = e0.elapsed_time(e1)
        else:
            ms = (_t.perf_counter() - t0) * 1e3
        clk = sampler.stop() if sampler is not None else None
        if world > 1:
            import torch.distributed as dist
            t = torch.tensor([ms], device=device, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        value = world * B * args.steps / (ms / 1e3)
        out = {
            "metric": "train_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": value / (40.6 * world), "dtype": "bf16-autocast" if amp else "fp32",
            "data": "synthetic CICIDS-shaped flow sentences (reference features_to_text + HF tokenizer, synthetic "
                    "vocab), random-init weights",
            "impl": impl,
            "config": {"model": "reference DDoSClassifier (HF DistilBertModel + Linear(768,2))", "global_batch": B * world,
                       "per_client_batch": B, "seq_len": 128, "parallelism": f"fed{world} (independent clients)",
                       "token_lens": token_lens,
                       "path": ("client1.train_model, stock DataLoader/tokenizer/Adam, forward under "
                                "torch.autocast(bfloat16), TF32 allowed" if amp else
                                "client1.train_model, stock DataLoader/tokenizer/Adam, fp32 (no TF32/AMP)")},
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": B * 128 * 8 * 2 + B * 8,
                    "d2h_bytes_per_step": 8,
                    "note": "the reference's loop is end to end by construction (tokenise + H2D + 2x loss.item())"},
            "gpu_launches": None,
            "clocks": clk,
        }
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
            dist.destroy_process_group()
        return out
    except Exception as e:  # noqa: BLE001
        return unavailable(f"reference run failed: {e!r}"[:300])
    finally:
        os.chdir(cwd)
