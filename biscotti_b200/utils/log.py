"""Peer logging with the reference's observable markers.

The reference logs to stderr through two log.Logger instances with "[peer]" / "[err]" prefixes and
UTC microsecond timestamps (DistSys/main.go:136-137); its evaluation pipeline scrapes node 0's log
for substrings such as "Train Error" and "Attack Rate" (honest.go:147-154,
eval/eval_performance/parseLogs.py:25-194).  The same markers are emitted here so the scrapers'
concepts carry over; `biscotti_b200.eval.parse_logs` reads them back.
"""
from __future__ import annotations

import sys
from datetime import datetime, timezone


class PeerLogger:
    def __init__(self, rank: int = 0, stream=None, enabled: bool | None = None):
        self.rank = rank
        self.stream = stream
        self.enabled = (stream is not None) if enabled is None else enabled
        self.records: list[tuple[str, int, float]] = []

    def _emit(self, prefix: str, msg: str):
        if not self.enabled:
            return
        ts = datetime.now(timezone.utc).strftime("%Y/%m/%d %H:%M:%S.%f")
        (self.stream or sys.stderr).write(f"{prefix}{ts} engine.py: {msg}\n")

    def info(self, msg: str):
        self._emit("[peer] ", msg)

    def error(self, msg: str):
        self._emit("[err] ", msg)

    def train_error(self, iteration: int, err: float):
        self.records.append(("train_error", iteration, err))
        self._emit("[peer] ", f"Train Error: {err:.6f} Iteration: {iteration}")

    def attack_rate(self, iteration: int, rate: float):
        self.records.append(("attack_rate", iteration, rate))
        self._emit("[peer] ", f"Attack Rate: {rate:.6f} Iteration: {iteration}")

    def phase(self, name: str, iteration: int):
        self._emit("[peer] ", f"{name} Iteration: {iteration}")
