"""Summarise an `ncu --set full` report (read here, in the container, with `ncu -i REP --page raw --csv`) as a markdown
table + the DRAM traffic table bench.py reads for `roofline.traffic`.

    python tools/ncu_summary.py gpurun_out/r2/full.ncu-rep profiles/r2_full_ncu.md profiles/r2_traffic.json
"""
import csv
import io
import json
import subprocess
import sys

METRICS = [
    ("duration (us)", "gpu__time_duration.sum", 1e-3),
    ("DRAM read (MB)", "dram__bytes_read.sum", 1e-6),
    ("DRAM write (MB)", "dram__bytes_write.sum", 1e-6),
    ("tcgen05 bf16 MMA ops % of peak (elapsed)", "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed", 1),
    ("tcgen05 bf16 MMA ops executed (GFLOP)", "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.sum", 1e-9),
    ("smem->tensor operand cycles % (elapsed)", "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 1),
    ("SM throughput %", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("L2 throughput %", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("L2 hit rate %", "lts__t_sector_hit_rate.pct", 1),
    ("L2->SM read (MB)", "lts__t_sectors_srcunit_tex_op_read.sum", 32e-6),
    ("DRAM throughput %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("regs/thread", "launch__registers_per_thread", 1),
    ("threads/block", "launch__block_size", 1),
    ("warps active %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1),
    ("issue active %", "sm__inst_issued.avg.pct_of_peak_sustained_active", 1),
]
UNIT_SCALE = {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9, "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6, "second": 1e9, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rep, out_md, out_json = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else None)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, data = rows[0], rows[1], rows[2:]
    col = {n: i for i, n in enumerate(head)}
    kernels = []
    for r in data:
        name = r[col["Kernel Name"]]
        vals = {}
        for label, metric, scale in METRICS:
            if metric not in col or r[col[metric]] in ("", "n/a"):
                vals[label] = None
                continue
            v = float(r[col[metric]].replace(",", ""))
            v *= UNIT_SCALE.get(units[col[metric]], 1.0)
            vals[label] = v * scale
        kernels.append((name, vals))
    names = []
    conv = 0
    for name, _ in kernels:
        if "conv3d_bn_relu" in name:
            conv += 1
            names.append("conv-%d" % conv)
        elif "svfe" in name:
            names.append("k_vox_svfe_packed")
        else:
            names.append(name.split("(")[0][-40:])
    with open(out_md, "a") as f:
        f.write("| metric | " + " | ".join(names) + " |\n|---|" + "---|" * len(names) + "\n")
        for label, _, _ in METRICS:
            f.write("| %s | " % label + " | ".join("n/a" if v[label] is None else "%.3f" % v[label] for _, v in kernels) + " |\n")
    if out_json:
        tab = {}
        for n, (_, v) in zip(names, kernels):
            key = {"k_vox_svfe_packed": "svfe", "conv-1": "conv1", "conv-2": "conv2", "conv-3": "conv3"}.get(n)
            if key and v["DRAM read (MB)"] is not None:
                tab[key] = (v["DRAM read (MB)"] + v["DRAM write (MB)"]) * 1e6
        json.dump({"source": "%s (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch, config car16)" % out_md,
                   "car16": tab}, open(out_json, "w"), indent=1)
    print(open(out_md).read())


if __name__ == "__main__":
    main()
