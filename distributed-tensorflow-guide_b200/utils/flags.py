"""Command-line flags.  The reference has exactly two, parsed with
``parse_known_args`` (Hogwild/Hogwild.py:59-75): ``--job_name`` and
``--task_index``; everything else is a literal in its sources (cluster dict,
lr, window, last_step, log dir).  Both flags are kept verbatim and the literals
become optional flags."""
from __future__ import annotations

import argparse
import json
from typing import Optional, Sequence

from ..cluster import ClusterSpec

__all__ = ["add_reference_flags", "add_scheme_flags", "parse_cluster", "parse_flags"]


def add_reference_flags(parser: argparse.ArgumentParser) -> argparse.ArgumentParser:
    parser.add_argument("--job_name", type=str, default="", help="One of 'ps', 'worker'")
    parser.add_argument("--task_index", type=int, default=0, help="Index of task within the job")
    return parser


def _float_list(s: str):
    return [float(x) for x in s.split(",") if x]


def add_scheme_flags(parser: argparse.ArgumentParser) -> argparse.ArgumentParser:
    from ..parallel.engines import SCHEMES
    parser.add_argument("--scheme", choices=SCHEMES, default="hogwild")
    parser.add_argument("--cluster", type=str, default="ps=1,worker=2",
                        help="JSON dict or 'ps=N,worker=M' (ports from --base_port)")
    parser.add_argument("--base_port", type=int, default=2222)
    parser.add_argument("--model", type=str, default="toy",
                        choices=["toy", "toy_random", "mnist_mlp", "wide_mlp", "gpt2_medium",
                                 "bert_large", "resnet50"])
    parser.add_argument("--batch_size", type=int, default=32)
    parser.add_argument("--lr", type=float, default=1e-4)
    parser.add_argument("--local_lr", type=_float_list, default=None,
                        help="one value, or one per worker, comma separated")
    parser.add_argument("--worker_lrs", type=_float_list, default=None)
    parser.add_argument("--update_window", type=int, default=1)
    parser.add_argument("--elastic_alpha", type=float, default=None,
                        help="AEASGD/AEAMSGD moving rate (default 0.9 / n_workers)")
    parser.add_argument("--momentum", type=float, default=0.9, help="AEAMSGD Nesterov momentum")
    parser.add_argument("--replicas_to_aggregate", type=int, default=None)
    parser.add_argument("--last_step", type=int, default=None)
    parser.add_argument("--max_iters", type=int, default=None)
    parser.add_argument("--checkpoint_dir", type=str, default=None)
    parser.add_argument("--save_checkpoint_secs", type=float, default=None)
    parser.add_argument("--ps_sharding", choices=["round_robin", "balanced"], default="round_robin")
    parser.add_argument("--pace_secs", type=float, default=0.0,
                        help="sleep per step (the scripts sleep 0.1-1 s 'to observe training')")
    parser.add_argument("--log_every", type=int, default=10)
    parser.add_argument("--linger", action="store_true",
                        help="ps never exits, as in the reference (kill it yourself)")
    parser.add_argument("--device", type=str, default="auto", choices=["auto", "cpu", "cuda"])
    parser.add_argument("--profile_dir", type=str, default=None,
                        help="write timeline-<n>.json (chrome trace) + phases.jsonl here")
    parser.add_argument("--profile_steps", type=int, default=100,
                        help="with --profile_dir: one timeline every this many runs")
    parser.add_argument("--inject_fault", type=str, default=None,
                        help="'worker:K@STEP' -- that worker exits abruptly at that round")
    return parser


def parse_cluster(s: str, base_port: int = 2222) -> ClusterSpec:
    s = s.strip()
    if s.startswith("{"):
        return ClusterSpec(json.loads(s))
    return ClusterSpec(s, base_port=base_port)


def parse_flags(argv: Optional[Sequence[str]] = None, scheme_flags: bool = True):
    p = argparse.ArgumentParser()
    add_reference_flags(p)
    if scheme_flags:
        add_scheme_flags(p)
    return p.parse_known_args(argv)
