"""Per-source-line hot spots of an .ncu-rep captured with --import-source on.
usage: python profiles/src_hot.py rep.ncu-rep [kernel-substring] [top-N]"""
import csv
import subprocess
import sys


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return 0.0


def main():
    rep = sys.argv[1]
    want = sys.argv[2] if len(sys.argv) > 2 else ""
    topn = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    secs, cur, fn, fp = [], None, "", ""
    for r in rows:
        if not r:
            continue
        if r[0] == "Function Name":
            fn = r[1]
        elif r[0] == "File Path":
            fp = r[1]
        elif r[0] == "Line No":
            cur = dict(fn=fn, fp=fp, hdr=r, rows=[])
            secs.append(cur)
        elif cur is not None and r[0].isdigit():
            cur["rows"].append(r)
    kernels = []
    for s in secs:
        if s["fn"] not in kernels:
            kernels.append(s["fn"])
    for k in kernels:
        if want not in k:
            continue
        mine = [s for s in secs if s["fn"] == k]
        h = mine[0]["hdr"]
        iS, iI, iT = h.index("# Samples"), h.index("Instructions Executed"), h.index("Thread Instructions Executed")
        iL = h.index("stall_long_sb")
        allr = [(s["fp"].split("/")[-1], r) for s in mine for r in s["rows"]]
        tot_s = sum(num(r[iS]) for _, r in allr) or 1.0
        tot_i = sum(num(r[iI]) for _, r in allr) or 1.0
        print(f"==== {k}\n     samples {tot_s:.0f}  warp-instructions {tot_i:.0f}")
        top = sorted(allr, key=lambda fr: -(num(fr[1][iS]) / tot_s + num(fr[1][iI]) / tot_i))[:topn]
        for f, r in sorted(top, key=lambda fr: (fr[0], int(fr[1][0]))):
            ins = num(r[iI])
            print(f"{f[:14]:14s}:{r[0]:>4s} {100 * num(r[iS]) / tot_s:5.1f}%smp {100 * num(r[iL]) / tot_s:5.1f}%longsb "
                  f"{100 * ins / tot_i:5.1f}%inst thr/inst {num(r[iT]) / max(ins, 1):4.1f} | {r[1].strip()[:100]}")


if __name__ == "__main__":
    main()
