#!/usr/bin/env python
"""Distils `ncu --set full` captures of k_render into the small JSON bench.py reads for its roofline object.

    python tools/ncu_metrics_json.py <build-tag> cfg2=gpurun_out/prof_X.ncu-rep [rollout=...] [cfg5=...] > profiles/r02_ncu_metrics.json

Per workload: DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum), L1 global-load wavefronts per launch, executed warp
instructions, kernel duration under ncu, tensor-pipe and issue-slot activity.  bench.py reports `roofline.traffic` and the L1-wavefront
floor from this file and only for the build tag it was made from."""
import csv
import json
import subprocess
import sys

WANT = {
    "dram_read_bytes": "dram__bytes_read.sum", "dram_write_bytes": "dram__bytes_write.sum",
    "l1_global_load_wavefronts": "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum",
    "warp_instructions": "smsp__inst_executed.sum", "duration_ms_under_ncu": "gpu__time_duration.sum",
    "tensor_pipe_active_pct": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "issue_active_pct": "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm_cycles": "sm__cycles_elapsed.avg",
    "l1_lsu_data_pipe_pct": "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1_hit_rate_pct": "l1tex__t_sector_hit_rate.pct", "l2_hit_rate_pct": "lts__t_sector_hit_rate.pct",
}
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ms": 1, "us": 1e-3, "s": 1e3, "ns": 1e-6}


def main():
    tag = sys.argv[1]
    out = {"build": tag, "source": "ncu --set full --clock-control none -k regex:k_render (tools/gpu_ncu.sh), one launch each"}
    for spec in sys.argv[2:]:
        wl, rep = spec.split("=", 1)
        rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
        hdr, units, vals = rows[0], rows[1], rows[2]
        d = {}
        for key, metric in WANT.items():
            i = hdr.index(metric)
            d[key] = float(vals[i]) * SCALE.get(units[i], 1)
        d["dram_bytes"] = d["dram_read_bytes"] + d["dram_write_bytes"]
        d["report"] = rep.split("/")[-1]
        out[wl] = d
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
