"""Group an `ncu --page source --csv --print-source sass` export into runs of SASS instructions with the same execution
count (= loop nests), with warp-instructions per unit of work and the share of warp-stall samples.

    ncu -i rep.ncu-rep --page source --csv --print-source sass > sass.csv
    python tools/ncu_runs.py sass.csv <units per launch, e.g. frames>   [min instructions per unit to print]

Used for the per-frame instruction budgets of the Viterbi kernels in DESIGN.md."""
import collections
import csv
import sys


def main(path, per=1.0, floor=3.0):
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    h = rows[hi]
    col = {n: i for i, n in enumerate(h)}
    stall_cols = [i for i, n in enumerate(h) if n.startswith("stall_")]
    runs, W, T = [], 0, 0.0
    for r in rows[hi + 1:]:
        if len(r) < len(h):
            continue
        n = int(r[col["Instructions Executed"]] or 0) / per
        w = int(r[col["Warp Stall Sampling (All Samples)"]] or 0)
        W += w
        T += n
        s = r[col["Source"]].strip()
        op = (s.split()[1] if s.startswith("@") else s.split()[0]).split(".")[0]
        st = [int(r[i] or 0) for i in stall_cols]
        if runs and abs(runs[-1]["k"] - n) < 0.02 * max(n, 1.0) + 0.05:
            u = runs[-1]
            u["c"] += 1; u["n"] += n; u["w"] += w; u["ops"][op] += 1
            u["st"] = [a + b for a, b in zip(u["st"], st)]
        else:
            runs.append(dict(k=n, c=1, n=n, w=w, ops=collections.Counter([op]), a=r[0][-5:], st=st))
    print(f"warp-instructions per unit: {T:.1f}; stall samples {W}")
    for u in runs:
        if u["n"] >= floor or u["w"] > 0.01 * W:
            top = sorted(zip(u["st"], [h[i][6:] for i in stall_cols]), reverse=True)[:3]
            print(f"{u['a']} x{u['k']:6.2f}  {u['c']:3d} sass {u['n']:7.1f} inst  {100 * u['w'] / max(W, 1):5.1f}% samples  "
                  f"{' '.join(f'{k}:{v}' for k, v in u['ops'].most_common(6))} | "
                  f"{' '.join(f'{nm}:{100 * c / max(u['w'], 1):.0f}%' for c, nm in top if c)}")


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]) if len(sys.argv) > 2 else 1.0, float(sys.argv[3]) if len(sys.argv) > 3 else 3.0)
