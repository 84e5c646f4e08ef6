"""Text summary of one `ncu --set full --import-source on` report (run here, on the CPU box, on a
.ncu-rep brought back in gpurun_out/): the roofline inputs (duration, DRAM bytes, tensor pipe),
occupancy limiters, the warp-stall breakdown and the hottest source lines by stall samples.

    python tests/tools/ncu_full_summary.py gpurun_out/X.ncu-rep > profiles/X_summary.txt
"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__grid_size", "launch__block_size",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_uniform.sum",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]


def page(rep, name):
    r = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True)
    return r.stdout


def main():
    rep = sys.argv[1]
    rows = list(csv.reader(io.StringIO(page(rep, "raw"))))
    head, units, vals = rows[0], rows[1], rows[2]
    col = {}
    for i, h in enumerate(head):
        col[h.split(".", 2)[-1] if h.count(".") >= 2 and h.split(".")[0].isupper() else h] = i
        col[h] = i
    print("#", rep)
    print("kernel:", vals[head.index("Kernel Name")][:120])
    print("grid", vals[head.index("Grid Size")], "block", vals[head.index("Block Size")])

    def find(key):
        for h, i in col.items():
            if h.endswith(key):
                return vals[i], units[i]
        return None, None
    for k in KEYS:
        v, u = find(k)
        if v not in (None, ""):
            print(f"{k:75s} {v} {u}")
    print("\n## warp stall samples (smsp__pcsamp_warps_issue_stalled_*), share of all samples")
    stalls = []
    for h, i in col.items():
        if "smsp__pcsamp_warps_issue_stalled_" in h and not h.endswith("_not_issued") and vals[i]:
            try:
                stalls.append((float(vals[i]), h.split("smsp__pcsamp_warps_issue_stalled_")[-1]))
            except ValueError:
                pass
    seen, uniq = set(), []
    for v, n in sorted(stalls, reverse=True):
        if n not in seen:
            seen.add(n)
            uniq.append((v, n))
    tot = sum(v for v, _ in uniq) or 1.0
    for v, n in uniq[:12]:
        print(f"  {n:40s} {v:12.0f}  {100 * v / tot:5.1f} %")
    for view in ("cuda", "sass"):
        r = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", view],
                           capture_output=True, text=True)
        rows = list(csv.reader(io.StringIO(r.stdout)))
        hi = next((i for i, row in enumerate(rows) if "Source" in row and "# Samples" in row), None)
        if hi is None:
            continue
        h = rows[hi]
        si, ci = h.index("Source"), h.index("# Samples")
        top = []
        for row in rows[hi + 1:]:
            try:
                top.append((float(row[ci] or 0), row[si].strip()[:150]))
            except (ValueError, IndexError):
                pass
        tot = sum(v for v, _ in top) or 1.0
        print(f"\n## hottest {view} lines by warp-stall samples (# Samples, share of {tot:.0f})")
        for v, src in sorted(top, reverse=True)[:22]:
            if v > 0:
                print(f"  {100 * v / tot:5.1f} %  {src}")


if __name__ == "__main__":
    main()
