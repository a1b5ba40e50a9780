"""Per-block phase timeline as a Chrome / Perfetto trace (chrome://tracing, ui.perfetto.dev).

The reference has no tracer: its "profiling" is differences between timestamps of well-known log lines, scraped
after the run (eval/eval_performance/parseLogs.py:79-194, second resolution), and GoVector/ShiViz is
installed but never imported.  Here an instrumented engine (`eng._timing = True`, or `localtest --trace out.json`)
measures every phase of every block with CUDA events; this module lays those durations out on one track per rank
so that a run can be inspected visually.  Phases appear back to back in protocol order (the instrumented schedule
is in line by construction); the host time of the whole iteration is a second track.
"""
from __future__ import annotations

import json

PHASE_ORDER = ["local_step", "commit", "noise", "verify", "secagg", "witness", "miner_aggregate", "block"]


def chrome_trace(reports, rank: int = 0) -> dict:
    """reports: BlockReport list (phase_ms filled in) -> {"traceEvents": [...]} (durations in microseconds)."""
    ev = [{"name": "process_name", "ph": "M", "pid": rank, "args": {"name": f"rank {rank}"}},
          {"name": "thread_name", "ph": "M", "pid": rank, "tid": 0, "args": {"name": "device phases"}},
          {"name": "thread_name", "ph": "M", "pid": rank, "tid": 1, "args": {"name": "host iteration"}}]
    t = 0.0
    for r in reports:
        start = t
        for name in PHASE_ORDER:
            ms = r.phase_ms.get(name)
            if not ms:
                continue
            ev.append({"name": name, "cat": "phase", "ph": "X", "pid": rank, "tid": 0, "ts": t, "dur": ms * 1e3,
                       "args": {"iteration": r.iteration}})
            t += ms * 1e3
        host = r.phase_ms.get("total_host", 0.0) * 1e3
        ev.append({"name": f"block {r.iteration}", "cat": "block", "ph": "X", "pid": rank, "tid": 1, "ts": start,
                   "dur": max(host, t - start),
                   "args": {"contributors": r.n_contributors, "accepted": r.n_accepted, "krum_set": r.n_krum,
                            "empty": r.empty, "test_error": r.test_error, "attack_rate": r.attack_rate}})
        t = start + max(host, t - start)
    return {"traceEvents": ev, "displayTimeUnit": "ms"}


def write_chrome_trace(reports, path: str, rank: int = 0) -> None:
    with open(path, "w") as f:
        json.dump(chrome_trace(reports, rank), f)
