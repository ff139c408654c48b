"""Aggregate an `ncu --page source --csv --print-source sass` export per barrier-delimited phase of a kernel.

    ncu -i rep.ncu-rep --page source --csv --print-source sass > src.csv
    python tools/ncu_phases.py src.csv

Phases are the SASS ranges between BAR.SYNC instructions, in address order (loops that span a barrier show up as
the phases they contain).  For each: warp-stall samples (share of the kernel), issued warp instructions, and the three
largest stall reasons.  Used for the phase tables in DESIGN.md section 4."""
import csv
import sys


def main(path):
    rows = list(csv.reader(open(path)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    col = {n: i for i, n in enumerate(hdr)}
    stall_cols = [i for i, n in enumerate(hdr) if n.startswith("stall_")]
    phases, cur = [], dict(first=None, samples=0, inst=0, n=0, stalls=[0] * len(stall_cols), ops={})
    for r in rows[hdr_i + 1:]:
        if len(r) < len(hdr):
            continue
        src = r[col["Source"]].strip()
        if cur["first"] is None:
            cur["first"] = src
        cur["samples"] += int(r[col["Warp Stall Sampling (All Samples)"]] or 0)
        cur["inst"] += int(r[col["Instructions Executed"]] or 0)
        cur["n"] += 1
        op = src.split()[1] if src.startswith("@") else src.split()[0]
        op = op.split(".")[0]
        cur["ops"][op] = cur["ops"].get(op, 0) + int(r[col["Instructions Executed"]] or 0)
        for k, i in enumerate(stall_cols):
            cur["stalls"][k] += int(r[i] or 0)
        if "BAR.SYNC" in src:
            phases.append(cur)
            cur = dict(first=None, samples=0, inst=0, n=0, stalls=[0] * len(stall_cols), ops={})
    phases.append(cur)
    tot_s = sum(p["samples"] for p in phases) or 1
    tot_i = sum(p["inst"] for p in phases) or 1
    print(f"{'phase':>5} {'sass':>5} {'samples%':>9} {'inst%':>7} {'inst':>12}  top stalls | top opcodes")
    for k, p in enumerate(phases):
        top = sorted(zip(p["stalls"], [hdr[i][6:] for i in stall_cols]), reverse=True)[:3]
        ops = sorted(p["ops"].items(), key=lambda kv: -kv[1])[:4]
        print(f"{k:5d} {p['n']:5d} {100 * p['samples'] / tot_s:9.1f} {100 * p['inst'] / tot_i:7.1f} {p['inst']:12d}  "
              + ", ".join(f"{n} {100 * s / max(p['samples'], 1):.0f}%" for s, n in top) + " | "
              + ", ".join(f"{o} {100 * c / max(p['inst'], 1):.0f}%" for o, c in ops))


if __name__ == "__main__":
    main(sys.argv[1])
