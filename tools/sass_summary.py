#!/usr/bin/env python
"""SASS evidence: mnemonic counts per kernel from `cuobjdump -sass` of the sm_100a objects built by `python -m fdt.build`
→ profiles/sass_summary.md (+ gzipped full listings of the tensor-core / collective sources)."""
import collections
import gzip
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "faster-distributed-training_b200", "build")
INTEREST = re.compile(r"^(UTCHMMA|UTCQMMA|UTCIMMA|UTCOMMA|UTC[A-Z]*MMA|UTMALDG|UTMASTG|UTMAPF|UTMACCTL|LDTM|STTM|UTCBAR|UTCATOMSWS|SYNCS|REDG|RED|ATOMG|ATOMS|"
                      r"LDG\.E\.128|STG\.E\.128|LDS\.128|STS\.128|MEMBAR|REDUX|ACQBULK|PREEXIT|MULTIMEM|LDGMC|HMMA|FFMA)")
KEEP_FULL = None      # full gzipped listing for every source file


def main():
    out = ["# SASS evidence (cuobjdump -sass of the sm_100a objects built by `python -m fdt.build`)\n",
           "Mnemonic counts per kernel.  `UTC*MMA` = tcgen05.mma (`UTCHMMA` bf16 / tf32 kinds), `UTMALDG` / `UTMASTG` = TMA tensor load / store,",
           "`LDTM` = tcgen05.ld, `UTCBAR` = tcgen05.commit, `SYNCS.*` = mbarrier ops, `REDG` / `ATOMG` = global reductions / atomics",
           "(`.SYS` + `MULTIMEM`/`MMEM` forms: NVLink / NVLS), `ACQBULK` / `PREEXIT` = griddepcontrol.wait / launch_dependents (PDL).\n"]
    for obj in sorted(f for f in os.listdir(OBJ) if f.endswith(".o")):
        name = obj[:-2]
        txt = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, obj)], capture_output=True, text=True).stdout
        if KEEP_FULL is None or name in KEEP_FULL:
            with gzip.open(os.path.join(ROOT, "profiles", f"sass_{name}.txt.gz"), "wt") as f:
                f.write(txt)
        out.append(f"\n## {name}.cu\n")
        func, counts, n = None, None, 0
        funcs = []
        for line in txt.splitlines():
            m = re.match(r"\s*Function : (\S+)", line)
            if m:
                if func:
                    funcs.append((func, counts, n))
                func, counts, n = m.group(1), collections.Counter(), 0
                continue
            m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
            if m and func:
                n += 1
                mn = m.group(1)
                if INTEREST.match(mn):
                    key = mn if not mn.startswith(("FFMA", "HMMA")) else mn.split(".")[0]
                    counts[key] += 1
        if func:
            funcs.append((func, counts, n))
        for fn, c, n in funcs:
            dem = subprocess.run(["cu++filt", fn], capture_output=True, text=True).stdout.strip() or fn
            c.pop("FFMA", None)
            items = ", ".join(f"{k}×{v}" for k, v in sorted(c.items()))
            out.append(f"* `{dem[:110]}` — {n} instr: {items}")
    with open(os.path.join(ROOT, "profiles", "sass_summary.md"), "w") as f:
        f.write("\n".join(out) + "\n")
    print("wrote profiles/sass_summary.md", sum(1 for l in out if l.startswith("* ")), "kernels")


if __name__ == "__main__":
    main()
