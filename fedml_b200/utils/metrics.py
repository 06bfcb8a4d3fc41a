"""Metric sinks: stdout log, in-memory history, optional wandb, optional MLOps publisher.

Key names follow the reference's wandb usage (``Train/Acc, Train/Loss, Test/Acc, Test/Loss`` with
``round``; fedavg_api.py:302-313, FedAVGAggregator.py:165-195).  wandb is only touched when
``enable_wandb`` is set *and* the module imports, so offline boxes never block on it.
"""
import json
import logging
import os
import time


class MetricsSink:
    def __init__(self, args=None):
        self.args = args
        self.records = []
        self._wandb = None
        if args is not None and getattr(args, "enable_wandb", False):
            try:
                import wandb  # noqa
                os.environ.setdefault("WANDB_MODE", "offline")
                self._wandb = wandb
            except Exception:  # pragma: no cover
                self._wandb = None
        self._jsonl = None
        out = getattr(args, "local_log_output_path", None) if args is not None else None
        if out:
            os.makedirs(out, exist_ok=True)
            self._jsonl = os.path.join(out, f"metrics-run-{getattr(args, 'run_id', 0)}.jsonl")

    def log(self, record: dict):
        rec = dict(record); rec.setdefault("timestamp", time.time())
        self.records.append(rec)
        logging.info("metrics %s", {k: (round(v, 5) if isinstance(v, float) else v) for k, v in record.items()})
        if self._wandb is not None and getattr(self._wandb, "run", None) is not None:
            self._wandb.log(record)
        if self._jsonl:
            with open(self._jsonl, "a") as fh:
                fh.write(json.dumps(rec) + "\n")

    def last(self, key, default=None):
        for rec in reversed(self.records):
            if key in rec:
                return rec[key]
        return default
