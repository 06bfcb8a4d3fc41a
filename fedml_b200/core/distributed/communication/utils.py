"""Communication timing helpers (reference: communication/utils.py:5-34, log lines ``--Benchmark ...``)."""
import logging
import time

_ticks = {}


def log_communication_tick(sender, receiver, timestamp=None):
    _ticks[(sender, receiver)] = time.time() if timestamp is None else timestamp
    logging.info("--Benchmark tick from %s to %s at %f", sender, receiver, _ticks[(sender, receiver)])


def log_communication_tock(sender, receiver, timestamp=None):
    now = time.time() if timestamp is None else timestamp
    t0 = _ticks.pop((sender, receiver), None)
    logging.info("--Benchmark tock from %s to %s at %f%s", sender, receiver, now,
                 "" if t0 is None else f" (latency {now - t0:.6f}s)")
    return None if t0 is None else now - t0


def log_round_start(client_idx, round_number, timestamp=None):
    logging.info("--Benchmark start round %s for %s at %f", round_number, client_idx, timestamp or time.time())


def log_round_end(client_idx, round_number, timestamp=None):
    logging.info("--Benchmark end round %s for %s at %f", round_number, client_idx, timestamp or time.time())
