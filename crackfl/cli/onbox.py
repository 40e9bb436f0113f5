"""``torchrun --nproc-per-node G -m crackfl.cli.onbox`` - a whole federation on one box: every GPU hosts K clients,
rounds are aggregated and re-broadcast by the fused in-kernel NVLink path (no gRPC, no host hop)."""
from __future__ import annotations

import argparse
import json
import sys

from ..parallel.runtime import init_distributed, run_federated_onbox, shutdown_distributed
from .common import parse_with_config


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--rounds", type=int, default=5)
    ap.add_argument("--clients-per-gpu", type=int, default=1)
    ap.add_argument("--strategy", default="fedavg", choices=["fedavg", "fedavg_weighted", "fedadam"])
    ap.add_argument("--payload", default="fp32", choices=["fp32", "bf16_delta"])
    ap.add_argument("--eval-global", action="store_true", help="score every new global model on client 0's validation set")
    ap.add_argument("--save", default=None, help="write the final global model (Keras .h5) on rank 0")
    args, cfg = parse_with_config(ap, argv)
    init_distributed()
    out = run_federated_onbox(cfg, args.rounds, args.clients_per_gpu, args.strategy, args.payload,
                              eval_global=args.eval_global)
    if out["weights"] is not None:
        if args.save:
            from ..io.keras_h5 import save_keras_h5
            from ..trainer import spec_from_config
            save_keras_h5(args.save, spec_from_config(cfg), out["weights"])
        print(json.dumps({"rounds": out["rounds"]}))
    shutdown_distributed()
    return 0


if __name__ == "__main__":
    sys.exit(main())
