"""Log scraper with the reference's conventions (eval/eval_performance/parseLogs.py:25-210): substring
matching on "[peer] ... Train Error: <e> Iteration: <i>" / "Attack Rate" lines, completion time =
timestamp of the last "Train Error" minus the first, per-iteration seconds = that / blocks."""
from __future__ import annotations

from datetime import datetime


def _ts(line: str):
    # "[peer] 2026/11/06 21:00:00.123456 engine.py: ..."
    try:
        return datetime.strptime(line[7:33], "%Y/%m/%d %H:%M:%S.%f")
    except ValueError:
        return None


def parse(lines):
    train, attack = [], []
    for line in lines:
        if "Train Error" in line:
            v = float(line.split("Train Error:")[1].split()[0])
            it = int(line.rsplit("Iteration:", 1)[1])
            train.append((it, v, _ts(line)))
        elif "Attack Rate" in line:
            v = float(line.split("Attack Rate:")[1].split()[0])
            it = int(line.rsplit("Iteration:", 1)[1])
            attack.append((it, v, _ts(line)))
    out = {"train_error": [(i, v) for i, v, _ in train], "attack_rate": [(i, v) for i, v, _ in attack]}
    stamps = [t for _, _, t in train if t is not None]
    if len(stamps) >= 2:
        total = (stamps[-1] - stamps[0]).total_seconds()
        out["completion_s"] = total
        out["s_per_iteration"] = total / (len(stamps) - 1)
    return out


def parse_file(path: str):
    with open(path) as f:
        return parse(f.readlines())
