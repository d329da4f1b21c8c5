"""Turn `ncu --set full` reports (.ncu-rep) into the small per-kernel JSON kept under profiles/.

  python benchmarks/summarize_ncu.py --what "..." --out profiles/rN_ncu_kernels.json a.ncu-rep [b.ncu-rep ...]

Per kernel (first launch found): duration, DRAM bytes read+written, warp instructions, issue-active %, threads per
instruction, achieved occupancy, registers, shared-memory wavefronts.  bench.py reads `dram_bytes` from this file
for roofline.traffic (ncu cannot run inside a timed region).
"""
import argparse
import csv
import io
import json
import re
import subprocess

METRICS = {
    "gpu__time_duration.sum": ("duration_us", 1e-3),  # reported in ns
    "dram__bytes_read.sum": ("dram_read", None),
    "dram__bytes_write.sum": ("dram_write", None),
    "smsp__inst_executed.sum": ("warp_instructions", 1.0),
    "smsp__issue_active.avg.pct_of_peak_sustained_active": ("issue_active_pct", 1.0),
    "smsp__thread_inst_executed_per_inst_executed.ratio": ("threads_per_instruction", 1.0),
    "sm__warps_active.avg.pct_of_peak_sustained_active": ("warps_active_pct", 1.0),
    "launch__registers_per_thread": ("registers", 1.0),
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum": ("smem_wavefronts", 1.0),
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed": ("l1tex_throughput_pct", 1.0),
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": ("dram_throughput_pct", 1.0),
    "LTS.TriageCompute.lts__throughput.avg.pct_of_peak_sustained_elapsed": ("l2_throughput_pct", 1.0),
    "l1tex__t_requests_pipe_lsu_mem_global_op_red.sum": ("global_red_requests", 1.0),
    "l1tex__m_l1tex2xbar_write_sectors_mem_global_op_red.sum": ("global_red_sectors_to_l2", 1.0),
    "l1tex__t_requests_pipe_lsu_mem_global_op_atom.sum": ("global_atom_requests", 1.0),
}
UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "ms": 1e6, "usecond": 1e3,
              "nsecond": 1.0, "msecond": 1e6}


def short_name(full):
    m = re.search(r"(\w+)(<[^>]*>)?\s*\(", full)
    if not m:
        return full[:60]
    args = (m.group(2) or "").replace("(int)", "").replace("(bool)", "").replace(" ", "")
    return m.group(1) + args


def summarize(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units = rows[0], rows[1]
    out = {}
    for r in rows[2:]:
        d = dict(zip(head, r))
        name = short_name(d["Kernel Name"])
        if name in out:
            continue
        k = {}
        for col, (key, scale) in METRICS.items():
            if col not in d or d[col] == "":
                continue
            try:
                v = float(d[col].replace(",", ""))
            except ValueError:      # "no data" / "n/a" for a metric the kernel did not produce
                continue
            unit = units[head.index(col)]
            if key in ("dram_read", "dram_write"):
                v *= UNIT_SCALE.get(unit, 1.0)
            elif key == "duration_us":
                v = v * UNIT_SCALE.get(unit, 1.0) * 1e-3
            k[key] = v
        k["dram_bytes"] = k.pop("dram_read", 0.0) + k.pop("dram_write", 0.0)
        out[name] = k
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", required=True)
    ap.add_argument("--out", required=True)
    ap.add_argument("reports", nargs="+")
    a = ap.parse_args()
    kernels = {}
    for p in a.reports:
        kernels.update(summarize(p))
    with open(a.out, "w") as f:
        json.dump({"what": a.what, "kernels": kernels}, f, indent=1)
    for k, v in kernels.items():
        print(k, v)


if __name__ == "__main__":
    main()
