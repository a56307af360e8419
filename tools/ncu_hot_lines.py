"""Summarise an ncu report's source page: stall samples per source line of one kernel.  Usage: ncu_hot_lines.py REP KERNEL_REGEX [TOP]"""
import csv
import subprocess
import sys


def main(rep, kernel, top=40):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source=cuda,sass", "-k", f"regex:{kernel}"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    cur, h, idx, lines = None, None, None, []
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            h = r
            idx = {n: j for j, n in enumerate(h)}
        elif r[0].isdigit() and h:
            try:
                s = int(r[idx["# Samples"]])
            except (ValueError, KeyError):
                continue
            lines.append((s, cur, r[0], r[1].strip()[:100], r))
    tot = sum(x[0] for x in lines) or 1
    stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
    print("total samples", tot)
    for s, f, ln, src, r in sorted(lines, key=lambda x: -x[0])[:top]:
        st = []
        for n in stalls:
            try:
                st.append((int(r[idx[n]]), n[6:]))
            except ValueError:
                pass
        st = [f"{b}:{a}" for a, b in sorted(st, reverse=True)[:2] if a]
        print(f"{s:6d} {100 * s / tot:5.1f}% {f}:{ln} {src}  [{' '.join(st)}]")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40)
