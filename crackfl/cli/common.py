"""Shared CLI plumbing: ``--config file`` plus ``--section.key=value`` overrides (crackfl/config.py)."""
from __future__ import annotations

import argparse
from typing import List, Tuple

from ..config import Config


def parse_with_config(ap: argparse.ArgumentParser, argv=None) -> Tuple[argparse.Namespace, Config]:
    ap.add_argument("--config", default=None, help="YAML/JSON config file")
    ap.add_argument("--fast", action="store_true", help="drop the reference's 5 s / 10 s / 20 s protocol sleeps")
    args, rest = ap.parse_known_args(argv)
    overrides: List[str] = []
    for r in rest:
        if r.startswith("--") and "=" in r and "." in r.split("=", 1)[0]:
            overrides.append(r)
        else:
            ap.error(f"unrecognised argument {r!r} (overrides look like --server.max_rounds=3)")
    cfg = Config.load(args.config, overrides)
    if args.fast:
        cfg.server.ready_delay_s = 0.0
        cfg.server.admission_window_s = min(cfg.server.admission_window_s, 2.0)
        cfg.client.poll_interval_s = 0.05
    return args, cfg
