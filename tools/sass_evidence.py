#!/usr/bin/env python
"""cuobjdump -sass of the in-tree objects -> profiles/sass_evidence.md (tcgen05 / TMA / TMEM mnemonic counts)."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "biscotti_b200/lib/obj")
WANT = ["UTCHMMA", "UTMALDG", "LDTM", "UTCBAR", "SYNCS", "IMAD.WIDE.U32", "SHFL"]
KEEP = ("msm_table_kernel<4>", "ed25519_scalarmult_quad_kernel", "secagg_fused_kernel", "g1_comb_mul_kernel")
rows = []
for o in sorted(f for f in os.listdir(OBJ) if f.endswith(".o")):
    out = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, o)], capture_output=True, text=True).stdout
    cur, cnt = None, collections.defaultdict(collections.Counter)
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            continue
        if cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
            cnt[cur]["total"] += 1
            for w in WANT:
                if re.search(r"\b" + re.escape(w), line):
                    cnt[cur][w] += 1
    for k, c in cnt.items():
        if any(c[w] for w in ("UTCHMMA", "UTMALDG", "LDTM")) or k.endswith(KEEP):
            rows.append((o, k, c))
with open(os.path.join(ROOT, "profiles/sass_evidence.md"), "w") as f:
    f.write("# SASS evidence (cuobjdump -sass of the in-tree objects, sm_100a)\n\n")
    f.write("Mnemonics per the profiling recipe: `UTCHMMA` = tcgen05.mma, `UTMALDG` = TMA tensor load, `LDTM` = tcgen05.ld "
            "(TMEM -> registers), `UTCBAR` = tcgen05.commit, `SYNCS` = mbarrier operations.  The MSM / VRF kernels are integer "
            "code (IMAD.WIDE) and are listed for contrast.  Regenerate with `python tools/sass_evidence.py`.\n\n")
    f.write("| object | kernel | SASS instrs | UTCHMMA | UTMALDG | LDTM | UTCBAR | SYNCS | IMAD.WIDE.U32 | SHFL |\n|---|---|---|---|---|---|---|---|---|---|\n")
    for o, k, c in rows:
        f.write(f"| {o} | `{k}` | {c['total']} | {c['UTCHMMA']} | {c['UTMALDG']} | {c['LDTM']} | {c['UTCBAR']} | {c['SYNCS']} | "
                f"{c['IMAD.WIDE.U32']} | {c['SHFL']} |\n")
print(open(os.path.join(ROOT, "profiles/sass_evidence.md")).read())
