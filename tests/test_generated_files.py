"""The generated artefacts (C header, command-list dispatcher, opcode table) must match the sources they are generated from:
an entry point added to a .cu file without re-running tools/gen_header.py + tools/gen_dispatch.py would give the native
executor a stale opcode table."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FILES = ["include/biscotti_b200.h", "biscotti_b200/csrc/dispatch_gen.inc", "biscotti_b200/ops/_ops_table.py"]


def test_generated_header_and_dispatcher_are_up_to_date():
    before = {f: open(os.path.join(ROOT, f)).read() for f in FILES}
    try:
        for tool in ("tools/gen_header.py", "tools/gen_dispatch.py"):
            subprocess.run([sys.executable, os.path.join(ROOT, tool)], check=True, capture_output=True, cwd=ROOT)
        after = {f: open(os.path.join(ROOT, f)).read() for f in FILES}
    finally:
        for f, text in before.items():            # leave the tree as it was, whatever the outcome
            if open(os.path.join(ROOT, f)).read() != text:
                with open(os.path.join(ROOT, f), "w") as fh:
                    fh.write(text)
    stale = [f for f in FILES if before[f] != after[f]]
    assert not stale, f"regenerate with tools/gen_header.py and tools/gen_dispatch.py: {stale}"


def test_every_opcode_has_a_prototype_with_a_trailing_stream():
    from biscotti_b200.ops._ops_table import OPS

    header = open(os.path.join(ROOT, "include/biscotti_b200.h")).read()
    assert len(OPS) >= 50
    codes = sorted(op for op, _ in OPS.values())
    assert codes == list(range(len(codes)))                       # dense, no duplicates
    for name, (_, kinds) in OPS.items():
        assert f"int bsc_{name}(" in header
        assert set(kinds) <= set("pifdq")
