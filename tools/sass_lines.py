"""Attribute the SASS of one kernel to source lines (needs -lineinfo): a static instruction budget per source line,
the offline proxy for issue-bound kernels (no GPU needed).

    python tools/sass_lines.py sentinel2-landcover-classification_b200/build/dwconv2_bwd.o 'dw2_bwdx_kernelILi3ELi1E' [lo hi]

lo/hi (hex) restrict the count to an address window (e.g. the step loop found from the BRA targets)."""
from __future__ import annotations

import collections
import os
import re
import subprocess
import sys
import tempfile


def main() -> None:
    obj, pat = sys.argv[1], sys.argv[2]
    lo = int(sys.argv[3], 16) if len(sys.argv) > 3 else 0
    hi = int(sys.argv[4], 16) if len(sys.argv) > 4 else 1 << 30
    with tempfile.TemporaryDirectory() as d:
        subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, stdout=subprocess.DEVNULL)
        cubin = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(d, cubin)], capture_output=True,
                             text=True).stdout.splitlines()
    inside = False
    cur = "?"
    per_line = collections.Counter()
    per_op = collections.Counter()
    ops_by_line = collections.defaultdict(collections.Counter)
    total = 0
    for l in txt:
        if l.startswith("//--------------------- .text."):
            inside = pat in l
            continue
        if l.startswith("//--------------------- ") and inside:
            inside = False
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = f"{os.path.basename(m.group(1))}:{m.group(2)}"
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if not m:
            continue
        a = int(m.group(1), 16)
        if not lo <= a <= hi:
            continue
        ins = re.sub(r"^@!?U?P\w+\s+", "", m.group(2).strip())
        op = ins.split()[0].split(".")[0]
        per_line[cur] += 1
        per_op[op] += 1
        ops_by_line[cur][op] += 1
        total += 1
    print("total", total)
    print("by opcode:", ", ".join(f"{k} {v}" for k, v in per_op.most_common(25)))
    for k, v in sorted(per_line.items(), key=lambda kv: (kv[0].split(":")[0], int(kv[0].split(":")[1]))):
        top = " ".join(f"{o}:{n}" for o, n in ops_by_line[k].most_common(5))
        print(f"{k:28s} {v:5d}  {top}")


if __name__ == "__main__":
    main()
