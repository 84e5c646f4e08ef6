"""Collect the per-kernel ncu metric files of one tests/tools/gpu_r02_ncu.sh run (CSV, one row per
metric) into profiles/r02_ncu_counters.json -- the file bench.py reads `traffic` and the
tensor-pipe share from.

    python tests/tools/ncu_counters.py gpurun_out r02ah > profiles/r02_ncu_counters.json
"""
import csv
import json
import os
import sys

CASES = {  # file stem -> key in the json (bench.py looks up attn_bwd3_kernel / attn_fwd2_kernel / gemm_kernel / ln_*)
    "t_attn_bwd": "attn_bwd3_kernel", "t_attn_fwd": "attn_fwd2_kernel", "t_fc1_fwd": "gemm_kernel",
    "t_qkv_fwd": "gemm_kernel_qkv_fwd", "t_fc2_dw": "gemm_kernel_fc2_dw",
    "t_fc2_fwd_ln": "gemm_kernel_fc2_fwd_ln", "t_out_fwd_ln": "gemm_kernel_out_fwd_ln",
    "t_fc2_dx_cs": "gemm_kernel_fc2_dx_colsum", "ln_bwd_pieces": "ln_bwd_pieces_kernel",
    "ln_fwd_pieces": "ln_fwd_pieces_kernel"}
NAMES = {"gpu__time_duration.sum": ("duration_us", 1e-3), "dram__bytes_read.sum": ("dram_read", 1.0),
         "dram__bytes_write.sum": ("dram_write", 1.0),
         "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed": ("tensor_pipe_pct", 1.0),
         "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32.sum": ("utchmma_flops", 1.0),
         "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": ("smem_multi_wavefront_accesses", 1.0),
         "sm__warps_active.avg.pct_of_peak_sustained_active": ("warps_active_pct", 1.0),
         "launch__registers_per_thread": ("registers", 1.0),
         "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": ("dram_throughput_pct", 1.0)}


def main():
    d, tag = sys.argv[1], sys.argv[2]
    out = {"_source": f"ncu --metrics ... --clock-control none -c 1 (after 3 warm-up launches), tests/tools/gpu_r02_ncu.sh, run {tag}"}
    for stem, key in CASES.items():
        path = os.path.join(d, f"{tag}_ncu_{stem}.csv")
        if not os.path.exists(path):
            continue
        lines = [l for l in open(path) if l.startswith('"')]
        rows = list(csv.DictReader(lines))
        if not rows:
            continue
        e = {"case": stem, "kernel": rows[0]["Kernel Name"].split("(")[0]}
        for r in rows:
            if r["Metric Name"] in NAMES:
                name, scale = NAMES[r["Metric Name"]]
                v = float(r["Metric Value"].replace(",", "")) * scale
                if r["Metric Name"] == "gpu__time_duration.sum":
                    v = float(r["Metric Value"].replace(",", "")) * (1e-3 if r["Metric Unit"] in ("ns", "nsecond") else 1.0)
                e[name] = v
        if "dram_read" in e and "dram_write" in e:
            e["dram_bytes"] = e["dram_read"] + e["dram_write"]
        out[key] = e
    json.dump(out, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
