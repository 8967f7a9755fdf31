"""Join an ncu SASS source page with nvdisasm line info: executed warp instructions and stall samples per source line.

    ncu -i rep.ncu-rep --page source --csv --print-source sass > sass.csv
    cuobjdump -xelf all libs2lc.so; nvdisasm -g -c file.cubin > file.dis
    python tools/ncu_lines.py sass.csv file.dis <mangled-kernel-substring> [top]
"""
import csv
import re
import sys
from collections import Counter


def main() -> None:
    sass, dis, sub = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    line_of = {}
    cur, inside = None, False
    for ln in open(dis):
        if ln.startswith(".text."):
            inside = sub in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
        if m:
            line_of[int(m.group(1), 16)] = cur
    rows = list(csv.reader(open(sass)))
    hdr = rows[1]
    ia, ie, ismp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    base = None
    inst, smp = Counter(), Counter()
    for r in rows[2:]:
        try:
            a = int(r[ia], 16) if not r[ia].isdigit() else int(r[ia])
        except ValueError:
            continue
        if base is None:
            base = a
        key = line_of.get(a - base)
        inst[key] += int(r[ie] or 0)
        smp[key] += int(r[ismp] or 0)
    tot, tots = sum(inst.values()), sum(smp.values())
    print(f"total warp instructions {tot}, samples {tots}")
    for key, n in inst.most_common(top):
        print(f"{str(key):40s} inst {n:11d} {n / tot:6.3f}   samples {smp[key]:7d} {smp[key] / max(tots, 1):6.3f}")


if __name__ == "__main__":
    main()
