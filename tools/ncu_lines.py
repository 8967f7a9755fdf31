"""Executed warp instructions and stall samples per CUDA source line of one captured kernel.

    python tools/ncu_lines.py gpurun_out/capture.ncu-rep [top]

Reads `ncu -i … --page source --csv --print-source cuda,sass` (needs -lineinfo and --import-source on)."""
import csv
import subprocess
import sys


def main() -> None:
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    cur, hdr, rows = None, None, []
    for r in csv.reader(out.splitlines()):
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif hdr and r[0].isdigit():
            ie, ismp = hdr.index("Instructions Executed"), hdr.index("# Samples")
            num = lambda v: int(v) if v.isdigit() else 0
            rows.append((cur, int(r[0]), r[1].strip(), num(r[ie]), num(r[ismp])))
    tot, tots = sum(x[3] for x in rows), sum(x[4] for x in rows)
    print(f"total warp instructions {tot}, samples {tots}")
    for f, ln, src, n, s in sorted(rows, key=lambda x: -x[3])[:top]:
        print(f"{f}:{ln:<5d} {n / tot:6.3f} {s / max(tots, 1):6.3f}  {src[:110]}")


if __name__ == "__main__":
    main()
