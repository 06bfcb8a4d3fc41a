"""Communication timing helpers.

The four ``log_*`` functions keep the reference's names and ``--Benchmark ...`` log lines (communication/utils.py:5-34) so log
scrapers written for it keep working; here a tick / tock pair also returns the measured latency, ``comm_timer`` wraps a block,
and ``latency_summary`` aggregates what was measured (count, mean, p50, p95, max per (sender, receiver))."""
import contextlib
import logging
import threading
import time
from collections import defaultdict
from typing import Dict, List, Optional, Tuple

_lock = threading.Lock()
_open: Dict[Tuple[object, object], float] = {}
_latencies: Dict[Tuple[object, object], List[float]] = defaultdict(list)


def _now(timestamp: Optional[float]) -> float:
    return time.time() if timestamp is None else float(timestamp)


def log_communication_tick(sender, receiver, timestamp=None):
    t = _now(timestamp)
    with _lock:
        _open[(sender, receiver)] = t
    logging.info("--Benchmark tick from %s to %s at %f", sender, receiver, t)


def log_communication_tock(sender, receiver, timestamp=None):
    t = _now(timestamp)
    with _lock:
        t0 = _open.pop((sender, receiver), None)
        if t0 is not None:
            _latencies[(sender, receiver)].append(t - t0)
    suffix = "" if t0 is None else f" (latency {t - t0:.6f}s)"
    logging.info("--Benchmark tock from %s to %s at %f%s", sender, receiver, t, suffix)
    return None if t0 is None else t - t0


def log_round_start(client_idx, round_number, timestamp=None):
    logging.info("--Benchmark start round %s for %s at %f", round_number, client_idx, _now(timestamp))


def log_round_end(client_idx, round_number, timestamp=None):
    logging.info("--Benchmark end round %s for %s at %f", round_number, client_idx, _now(timestamp))


@contextlib.contextmanager
def comm_timer(sender, receiver):
    """``with comm_timer(0, 3): cm.send_message(m)`` = tick before, tock after."""
    log_communication_tick(sender, receiver)
    try:
        yield
    finally:
        log_communication_tock(sender, receiver)


def latency_summary(reset: bool = False) -> Dict[Tuple[object, object], dict]:
    with _lock:
        items = {k: sorted(v) for k, v in _latencies.items() if v}
        if reset:
            _latencies.clear()
    pick = lambda v, q: v[min(len(v) - 1, int(q * len(v)))]      # noqa: E731
    return {k: {"count": len(v), "mean": sum(v) / len(v), "p50": pick(v, 0.5), "p95": pick(v, 0.95), "max": v[-1]} for k, v in items.items()}
