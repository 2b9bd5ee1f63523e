"""Measures the UNMODIFIED reference's socket FedAvg round (server.py main + the clients' send_model /
receive_aggregated_model from client1.py / client2.py) on THIS machine, with random-init DistilBERT-base
DDoSClassifier state_dicts (265 MB each).  CPU-only: the reference's round never touches the GPU.

Prints one JSON line: per-client round latency (start of send -> aggregated weights received), server wall time.
The same measurement for this repository's socket fallback (gzip level 1, concurrent downloads) follows.
"""
import contextlib
import io
import json
import os
import sys
import tempfile
import threading
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import reference_arm  # noqa: E402

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"


def main():
    import importlib

    import torch

    arena_mod = importlib.import_module(f"{PKG}.models.arena")
    st = importlib.import_module(f"{PKG}.parallel.socket_transport")
    logmod = importlib.import_module(f"{PKG}.report.log")
    sds = [arena_mod.ParamArena().init_random_(s).state_dict(clone=True) for s in (1, 2)]

    class FakeModel:
        def __init__(self, sd):
            self.sd = sd

        def state_dict(self):
            return self.sd

    work = tempfile.mkdtemp(prefix="ref_socket_")
    os.chdir(work)
    out = {}
    # ---------------- the reference ----------------
    if reference_arm.find_reference() is not None:
        server = reference_arm.load_reference("server")
        c1 = reference_arm.load_reference("client1")
        c2 = reference_arm.load_reference("client2")
        quiet = io.StringIO()
        res = {}

        def run_server():
            t0 = time.perf_counter()
            server.main()
            res["server_s"] = time.perf_counter() - t0

        def run_client(mod, sd, key):
            time.sleep(1.0)  # let the server bind
            t0 = time.perf_counter()
            ok = mod.send_model(FakeModel(sd))
            agg = mod.receive_aggregated_model() if ok else None
            res[key] = time.perf_counter() - t0
            res[key + "_ok"] = agg is not None

        ts = [threading.Thread(target=run_server), threading.Thread(target=run_client, args=(c1, sds[0], "client1_s")),
              threading.Thread(target=run_client, args=(c2, sds[1], "client2_s"))]
        real_out, real_err = sys.stdout, sys.stderr
        sys.stdout = sys.stderr = quiet          # one global redirect: redirect_stdout is not thread-safe
        try:
            [t.start() for t in ts]
            [t.join() for t in ts]
        finally:
            sys.stdout, sys.stderr = real_out, real_err
        out["reference"] = res
    else:
        out["reference"] = {"unavailable": "baseline/_ref not installed"}
    # ---------------- this repository's socket fallback (same wire format) ----------------
    srv = st.FedServer(2, port_recv=22345, port_send=22346, save_path=None, log=logmod.EventLog("SERVER", quiet=True),
                       aggregate=importlib.import_module(f"{PKG}.parallel.fedavg").fedavg_state_dicts_cpu)
    srv.open()
    res2 = {}
    t = threading.Thread(target=srv.run_round)
    t.start()

    def ours(i):
        t0 = time.perf_counter()
        ok = st.send_model(sds[i], "localhost", 22345)
        agg = st.receive_aggregated_model("localhost", 22346) if ok else None
        res2[f"client{i + 1}_s"] = time.perf_counter() - t0
        res2[f"client{i + 1}_ok"] = agg is not None

    cs = [threading.Thread(target=ours, args=(i,)) for i in (0, 1)]
    t0 = time.perf_counter()
    [c.start() for c in cs]
    [c.join() for c in cs]
    t.join()
    res2["round_s"] = time.perf_counter() - t0
    srv.close()
    out["ours_socket_fallback"] = res2
    print(json.dumps(out))


if __name__ == "__main__":
    main()
