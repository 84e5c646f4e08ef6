"""Top stall sites of one kernel from `ncu -i X.ncu-rep --page source --csv` (SASS view).

    ncu -i gpurun_out/x.ncu-rep --page source --csv > /tmp/x.csv; python tests/tools/ncu_hot.py /tmp/x.csv [N]
Prints the N SASS instructions with the most stall samples, their dominant stall reason, and the
stall-reason totals of the kernel.
"""
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1], errors="replace")))
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    h = rows[hi]
    col = {n: i for i, n in enumerate(h)}
    stall_cols = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
    body = [r for r in rows[hi + 1:] if len(r) == len(h)]
    tot = {n: 0 for n in stall_cols}
    items = []
    for k, r in enumerate(body):
        s = int(r[col["# Samples"]] or 0)
        st = {n: int(r[col[n]] or 0) for n in stall_cols}
        for n in stall_cols:
            tot[n] += st[n]
        items.append((s, k, r[col["Source"]].strip(), st, int(r[col["Instructions Executed"]] or 0)))
    all_s = sum(i[0] for i in items)
    print(f"samples {all_s}  instructions {len(items)}")
    print("stall totals: " + ", ".join(f"{n[6:]} {v * 100 // max(1, all_s)}%" for n, v in
                                       sorted(tot.items(), key=lambda kv: -kv[1])[:8]))
    for s, k, src, st, ex in sorted(items, reverse=True)[:top]:
        dom = max(st.items(), key=lambda kv: kv[1])
        print(f"{s:7d} {s * 100.0 / max(1, all_s):5.1f}%  #{k:5d} x{ex:<8d} {dom[0][6:]:<12s} {src[:90]}")


if __name__ == "__main__":
    main()
