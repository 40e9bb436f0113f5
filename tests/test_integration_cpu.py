"""End-to-end on CPU: BASELINE config 1 (2 clients, 1 round, 128x128 synthetic, real gRPC, real local fits on the
oracle back-end), CLIs, trainer, measurement, Gloo federation runtime."""
import json
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

from crackfl.config import fast_test_config
from crackfl.models.spec import REFERENCE_SPEC as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tiny_cfg(tmp_path, size=32, n_train=8, n_val=4, epochs=1):
    cfg = fast_test_config()
    cfg.train.backend = "torch_ref"
    cfg.train.image_size = (size, size)
    cfg.train.batch_size = 4
    cfg.train.local_epochs = epochs
    cfg.train.tensorboard = False
    cfg.data.synthetic_train, cfg.data.synthetic_val = n_train, n_val
    cfg.client.log_root = str(tmp_path / "send_logs" / "logs")
    cfg.client.weights_path = str(tmp_path / "saved_weight" / "weights.pickle")
    cfg.server.weights_dir = str(tmp_path / "server_weights")
    return cfg


def test_local_trainer_fit_reduces_loss(tmp_path):
    from crackfl.trainer import LocalTrainer
    cfg = _tiny_cfg(tmp_path, epochs=3)
    tr = LocalTrainer(cfg, log=lambda s: None)
    w0 = S.init_weights(0)
    w1, m = tr.fit(w0, 1)
    assert len(w1) == 112 and m["num_samples"] == 8 and np.isfinite(m["loss"]) and m["val_loss"] is not None
    assert any(np.abs(a - b).max() > 0 for a, b in zip(w0, w1))
    # manage_train: bytes in (reference-compatible)
    from crackfl.io import weights as wio
    cfg.transport.allow_pickle = True
    out = tr.manage_train(wio.dumps_pickle(w0), cr=2)
    assert len(out) == 112


def test_config1_two_clients_one_round_128px_real_grpc(tmp_path):
    """2-client FedAvg, 1 round, crack U-Net 128x128 synthetic images on CPU (BASELINE.json config 1)."""
    from crackfl.protocol import client as C
    from crackfl.protocol.server import FLServer, make_grpc_server
    from crackfl.trainer import LocalTrainer
    scfg = _tiny_cfg(tmp_path, size=128, n_train=4, n_val=0)
    scfg.server.max_rounds = 1
    srv = FLServer(scfg, S.init_weights(0), S, log=lambda s: None)
    gs, port = make_grpc_server(srv, "127.0.0.1:0")
    gs.start()
    res = {}

    def run(i):
        cfg = _tiny_cfg(tmp_path / f"c{i}", size=128, n_train=4, n_val=0)
        cfg.train.validate = False
        cfg.client.name = f"client{i}"
        call, ch = C.make_stub(cfg, f"127.0.0.1:{port}")
        tr = LocalTrainer(cfg, client_index=i, log=lambda s: None)
        cl = C.FLClient(cfg, tr, call, log=lambda s: None)
        res[i] = (cl.run(), cl, tr)
        ch.close()
    ts = [threading.Thread(target=run, args=(i,)) for i in range(2)]
    [t.start() for t in ts]; [t.join(300) for t in ts]
    gs.stop(0)
    assert srv.finished and srv.history[0]["num_clients"] == 2
    w0 = res[0][2].backend.get_weights(); w1 = res[1][2].backend.get_weights()
    for g, a, b in zip(srv.global_weights, w0, w1):
        np.testing.assert_allclose(g, (a + b) / 2, rtol=1e-5, atol=1e-6)  # aggregated == mean of client weights
    assert res[0][0]["state"] == "FIN"


def test_measurement_defined_outputs(tmp_path):
    from crackfl.postproc.measure import contour, measure, predict_and_measure, prob_to_bgr
    p = np.zeros((64, 64, 1), np.float32)
    p[20:40, 10:50] = 0.9
    bgr = prob_to_bgr(p)
    area, round2, length, mount = contour(bgr, out_dir=str(tmp_path / "contour"))
    assert mount == 1 and abs(area - 19 * 39) < 1e-6 and abs(length - 2 * (19 + 39)) < 1e-6
    assert abs(round2 - 4 * np.pi * area / length ** 2) < 1e-9
    assert (tmp_path / "contour" / "img1.jpg").exists() and (tmp_path / "contour" / "img3.jpg").exists()
    assert measure(np.zeros((8, 8, 3), np.uint8)).mount == 0  # no contour: zeros, not IndexError

    class DS:
        def __len__(self): return 1
        def images_only(self, i): return np.zeros((2, 64, 64, 3), np.uint8)
    rows, amount = predict_and_measure(lambda x: np.repeat(p[None], 2, 0), DS(), None)
    assert len(rows) == 2 and amount == 2 and rows[0][0] == area


def test_generator_matches_reference_preprocessing(tmp_path):
    import cv2
    from crackfl.data.generator import Generator, gen_train_val_data
    (tmp_path / "img").mkdir(); (tmp_path / "msk").mkdir()
    rng = np.random.default_rng(0)
    for i in range(6):
        cv2.imwrite(str(tmp_path / "img" / f"a{i}.jpg"), rng.integers(0, 255, (40, 50, 3), dtype=np.uint8))
        m = np.zeros((40, 50), np.uint8); m[10:20, 5 * i:5 * i + 10] = 255
        cv2.imwrite(str(tmp_path / "msk" / f"a{i}.jpg"), m)
    tr, va = gen_train_val_data(str(tmp_path / "img"), str(tmp_path / "msk"), batch_size=2, img_size=(32, 32), count=4)
    assert len(tr) == 2 and len(va) == 1
    x, y = tr[0]
    assert x.shape == (2, 32, 32, 3) and x.dtype == np.uint8 and y.shape == (2, 32, 32, 1) and set(np.unique(y)) <= {0, 1}
    g = Generator(tr.x, tr.y, 2, (32, 32), as_float=True)
    xf, yf = g[0]
    fn = tr.x[0]
    ref = cv2.resize(cv2.cvtColor(cv2.imread(fn, -1), cv2.COLOR_BGR2RGB), (32, 32)) / 255  # client_fit_model.py:35-43
    np.testing.assert_allclose(xf[0], ref)
    assert xf.dtype == np.float64 and yf.max() == 1.0


def test_train_and_infer_clis(tmp_path):
    env = dict(os.environ, PYTHONPATH=ROOT)
    common = ["--train.backend=torch_ref", "--train.image_size=32x32", "--train.batch_size=4", "--train.tensorboard=false",
              "--data.synthetic_train=8", "--data.synthetic_val=4"]
    h5 = str(tmp_path / "crack_segmentation.h5")
    r = subprocess.run([sys.executable, "-m", "crackfl.cli.train", "--epochs", "2", "--out", h5, "--save-dir",
                        str(tmp_path / "my_model")] + common, cwd=str(tmp_path), env=env, capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert os.path.exists(h5) and os.path.exists(tmp_path / "my_model" / "manifest.json")
    r = subprocess.run([sys.executable, "-m", "crackfl.cli.infer", "--weights", h5, "--out-dir", str(tmp_path / "out")]
                       + common, cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    meas = json.loads(r.stdout.strip().splitlines()[-1])["measurements"]
    assert len(meas) == 4 and os.path.exists(tmp_path / "out" / "pred2.png")


def test_gloo_two_process_onbox_federation(tmp_path):
    """Host-side logic of the on-box runtime with 2 Gloo ranks (the in-kernel path needs GPUs)."""
    script = tmp_path / "run.py"
    script.write_text(f"""
import sys, json
sys.path.insert(0, {ROOT!r})
import numpy as np, torch, torch.distributed as dist
from crackfl.config import fast_test_config
from crackfl.parallel.runtime import init_distributed, run_federated_onbox
init_distributed("gloo")
cfg = fast_test_config()
cfg.train.backend = "torch_ref"; cfg.train.image_size = (32, 32); cfg.train.batch_size = 4; cfg.train.local_epochs = 1
cfg.train.tensorboard = False; cfg.train.validate = False
cfg.data.synthetic_train = 4; cfg.data.synthetic_val = 0
cfg.client.log_root = {str(tmp_path)!r} + "/logs%d" % dist.get_rank()
out = run_federated_onbox(cfg, rounds=2, clients_per_gpu=2, strategy="fedavg_weighted", log=lambda s: None)
if dist.get_rank() == 0:
    w = out["weights"]
    print("RESULT", json.dumps({{"rounds": len(out["rounds"]), "n": len(w), "finite": bool(all(np.isfinite(a).all() for a in w))}}))
dist.barrier(); dist.destroy_process_group()
""")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29541", str(script)], capture_output=True,
                       text=True, timeout=600, env=dict(os.environ, PYTHONPATH=ROOT))
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT")][0]
    res = json.loads(line.split(" ", 1)[1])
    assert res == {"rounds": 2, "n": 112, "finite": True}


def test_gloo_mixed_round_onbox_plus_remote_update(tmp_path):
    """SURVEY §5.8 item 6 on the host path: 2 Gloo ranks with one on-box client each + ONE off-box update staged on
    rank 0 -> every rank's global model is the sample-weighted mean of the three (fl_server.py:92-102 arithmetic)."""
    script = tmp_path / "mixed.py"
    script.write_text(f"""
import sys, json
sys.path.insert(0, {ROOT!r})
import numpy as np, torch, torch.distributed as dist
from crackfl.parallel.runtime import OnBoxFederation, init_distributed
init_distributed("gloo")
r = dist.get_rank()
n = 4096
fed = OnBoxFederation(clients_per_gpu=1, strategy="fedavg_weighted", numel=n, remote_slots=2)
g = torch.linspace(-1, 1, n)
fed.set_global(g)
ups = [g + 0.1 * (i + 1) * torch.cos(torch.arange(n) * 0.01 * (i + 1)) for i in range(3)]
ns = [10.0, 30.0, 60.0]
for rnd in range(2):   # second round: no remote update -> plain 2-client mean
    fed.publish(0, ups[r])
    if r == 0 and rnd == 0:
        fed.stage_remote(ups[2].numpy(), ns[2])
    fed.aggregate([ns[r]], weighted=True)
    if rnd == 0:
        want = (ns[0] * ups[0] + ns[1] * ups[1] + ns[2] * ups[2]) / 100.0
    else:
        want = (ns[0] * ups[0] + ns[1] * ups[1]) / 40.0
    err = float((fed.global_arena - want).abs().max())
    open({str(tmp_path)!r} + "/res_%d_%d.txt" % (r, rnd), "w").write(repr(err))
dist.barrier(); dist.destroy_process_group()
""")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29543", str(script)], capture_output=True,
                       text=True, timeout=600, env=dict(os.environ, PYTHONPATH=ROOT))
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    errs = [float(open(tmp_path / f"res_{rk}_{rnd}.txt").read()) for rk in (0, 1) for rnd in (0, 1)]
    assert all(e < 1e-6 for e in errs), errs


def test_bridge_server_plus_onbox_round_cpu(tmp_path, monkeypatch):
    """One process (world 1, oracle back-end): an FLServer whose aggregator is the OnBoxBridge serves ONE gRPC loopback
    client while run_federated_onbox hosts 2 local clients; each round's global model is the uniform mean of the three
    updates (fl_server.py:92-102) and the remote client receives it back."""
    import threading
    import numpy as np
    import torch
    from crackfl.config import fast_test_config
    from crackfl.io import weights as wio
    from crackfl.models.spec import REFERENCE_SPEC as S
    from crackfl.parallel.bridge import OnBoxBridge
    from crackfl.parallel.runtime import run_federated_onbox
    from crackfl.protocol import client as C
    from crackfl.protocol.server import FLServer, make_grpc_server
    monkeypatch.chdir(tmp_path)
    cfg = fast_test_config()
    cfg.train.backend = "torch_ref"; cfg.train.image_size = (32, 32); cfg.train.batch_size = 4; cfg.train.local_epochs = 1
    cfg.train.tensorboard = False; cfg.train.validate = False
    cfg.data.synthetic_train = 4; cfg.data.synthetic_val = 0
    cfg.client.log_root = str(tmp_path / "logs"); cfg.client.write_weights_file = False
    cfg.server.max_rounds = 2; cfg.server.weights_dir = str(tmp_path / "sw")
    bridge = OnBoxBridge(S, remote_slots=1, timeout_s=120)
    srv = FLServer(cfg, S.init_weights(cfg.server.seed), S, aggregator=bridge, log=lambda s: None)
    gs, port = make_grpc_server(srv, "127.0.0.1:0")
    gs.start()
    seen = {}

    class Remote:
        def fit(self, weights, cr):
            seen[cr] = [w.copy() for w in weights]
            return [w + 0.5 for w in weights], {"loss": 0.1, "accuracy": 0.9, "num_samples": 8}

    def run_remote():
        call, ch = C.make_stub(cfg, f"127.0.0.1:{port}")
        cl = C.FLClient(cfg, Remote(), call, log=lambda s: None)
        seen["result"] = cl.run()
        seen["final"] = cl.final_weights
        ch.close()
    th = threading.Thread(target=run_remote)
    th.start()
    try:
        out = run_federated_onbox(cfg, rounds=2, clients_per_gpu=2, strategy="fedavg", log=lambda s: None, bridge=bridge,
                                  remote_slots=1)
        th.join(120)
    finally:
        gs.stop(0)
    assert not th.is_alive() and seen["result"]["rounds"] == 2 and bridge.rounds_bridged == 2
    fed = out["federation"]
    # last round: updates[0..1] are the two local fits, updates[2] the remote update; uniform mean of the three
    want = fed.updates[:3].mean(0)
    assert torch.allclose(fed.global_arena, want, rtol=1e-5, atol=1e-6)
    # the remote client trained round 2 on the round-1 global model and got the final one back
    np.testing.assert_allclose(wio.flatten(seen["final"], S), fed.global_arena.numpy(), rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(wio.flatten(srv.global_weights, S), fed.global_arena.numpy(), rtol=1e-6, atol=1e-6)
    assert not np.allclose(wio.flatten(seen[2], S), wio.flatten(seen[1], S))
