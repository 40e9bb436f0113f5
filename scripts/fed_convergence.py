#!/usr/bin/env python
"""Multi-round FedAvg on ONE GPU, engine vs fp32 oracle on identical data ('round 2 loss jumps').

K clients train sequentially from the round's global model (fresh Adam each round, client_fit_model.py:155-157),
the global model is the plain mean (fl_server.py:92-102).  After every aggregation the global model is evaluated
(inference-mode BN, moving statistics) AND scored in training-mode BN on a held-out batch set.
Prints one JSON line per round and arm."""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from crackfl.data.generator import synth_crack_batch
from crackfl.io import weights as wio
from crackfl.models.spec import REFERENCE_SPEC as S
from crackfl.models import unet_ref as R


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--clients", type=int, default=2)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--size", type=int, default=64)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--nbatch", type=int, default=8)
    ap.add_argument("--arms", default="engine,oracle")
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--bce-clip", type=int, default=0)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    B, H = a.batch, a.size
    data = {k: [synth_crack_batch(B, H, H, 1000 * k + i) for i in range(a.nbatch)] for k in range(a.clients)}
    val = [synth_crack_batch(B, H, H, 777000 + i) for i in range(4)]
    w0 = S.init_weights(0)

    def prep(x, y):
        xt = torch.from_numpy(x).to(dev).permute(0, 3, 1, 2).float() / 255.0
        yt = torch.from_numpy(y).to(dev).permute(0, 3, 1, 2).float()
        return xt, yt

    for arm in a.arms.split(","):
        glob = [w.copy() for w in w0]
        if arm == "engine":
            from crackfl.models.unet import UNetEngine
            eng = UNetEngine(S, B, H, H, device=0, lr=a.lr, bce_clip=bool(a.bce_clip))
        for rnd in range(1, a.rounds + 1):
            ups, last = [], []
            for k in range(a.clients):
                if arm == "engine":
                    eng.load_weights(glob); eng.reset_optimizer(); eng.reset_metrics()
                    tail = []
                    for s in range(a.steps):
                        x, y = data[k][s % a.nbatch]
                        eng.reset_metrics()
                        eng.train_step(x, y[..., 0])
                        if s >= a.steps - 4:
                            tail.append(eng.read_metrics()["loss"])
                    ups.append(eng.get_weights()); last.append(float(np.mean(tail)))
                else:
                    tr = R.RefTrainer(glob, S, lr=a.lr, device=dev, bce_clip=bool(a.bce_clip))
                    tail = []
                    for s in range(a.steps):
                        x, y = data[k][s % a.nbatch]
                        l, _ = tr.train_step(*prep(x, y))
                        if s >= a.steps - 4:
                            tail.append(l)
                    ups.append(tr.get_weights()); last.append(float(np.mean(tail)))
            glob = [np.mean([u[i] for u in ups], axis=0) for i in range(len(glob))]
            # evaluate the aggregated model with the oracle (fp32): inference BN and training-mode BN
            wt = R.to_torch_weights(glob, device=dev)
            ev, tv = [], []
            with torch.no_grad():
                for x, y in val:
                    xt, yt = prep(x, y)
                    ev.append(float(R.bce_logits_loss(R.unet_forward(wt, xt, training=False, return_logits=True)[0], yt)))
                    tv.append(float(R.bce_logits_loss(R.unet_forward(wt, xt, training=True, return_logits=True)[0], yt)))
            print(json.dumps({"arm": arm, "round": rnd, "client_tail_loss": last, "global_eval_loss": float(np.mean(ev)),
                              "global_trainbn_loss": float(np.mean(tv)),
                              "drift": float(np.sqrt(sum(((ups[0][i] - ups[-1][i]) ** 2).sum() for i in range(len(glob)))))}),
                  flush=True)


if __name__ == "__main__":
    main()
