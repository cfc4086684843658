"""Summary of an `ncu --set full` capture: one metric,unit,value CSV per captured launch of a kernel and (with --traffic)
the DRAM / L2 bytes per step of all episodes that bench.py reports as roofline.traffic / roofline.l2.

    ncu -i gpurun_out/r2_final_c2.ncu-rep --page raw --csv > /tmp/raw.csv
    python profiles/ncu_summary.py /tmp/raw.csv episode_c2_kernel 30 profiles/r2_ncu_episode_c2_kernel_summary.csv \
        --traffic profiles/r2_ncu_traffic.json --commit <id>
(30 = env-steps of every episode in the captured launch; profiles/ep_c2_workload.py makes two 30-step launches)
"""
import argparse
import csv
import json

KEEP = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "lts__t_sectors.sum",
    "lts__t_sectors_srcunit_tex.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
    "lts__t_sectors_srcunit_tex_op_red.sum", "lts__t_sectors_srcunit_ltcfabric.sum", "lts__t_sector_hit_rate.pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "TPC.TriageCompute.sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size", "launch__cluster_size",
    "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]

ap = argparse.ArgumentParser()
ap.add_argument("raw"); ap.add_argument("kernel"); ap.add_argument("steps", type=int); ap.add_argument("out")
ap.add_argument("--launch", type=int, default=-1, help="which captured launch of the kernel (default: the last)")
ap.add_argument("--traffic"); ap.add_argument("--commit", default="")
a = ap.parse_args()
rows = list(csv.reader(open(a.raw)))
hdr, units = rows[0], rows[1]
name_col = hdr.index("Kernel Name")
mine = [r for r in rows[2:] if a.kernel in r[name_col]]
r = mine[a.launch]
col = {h: i for i, h in enumerate(hdr)}
with open(a.out, "w") as f:
    f.write("metric,unit,value\n")
    f.write(f"kernel,,\"{r[name_col]}\"\n")
    for k in KEEP:
        if k in col:
            f.write(f"{k},{units[col[k]]},{r[col[k]]}\n")
    for h in hdr:
        if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
            f.write(f"{h},{units[col[h]]},{r[col[h]]}\n")


def val(k):
    v, u = float(r[col[k]]), units[col[k]]
    return v * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0, "sector": 1.0, "": 1.0}.get(u, 1.0)


if a.traffic:
    try:
        tr = json.load(open(a.traffic))
    except FileNotFoundError:
        tr = {}
    dram = (val("dram__bytes_read.sum") + val("dram__bytes_write.sum")) / a.steps
    l2_all = val("lts__t_sectors.sum") * 32 / a.steps
    l2_sm = val("lts__t_sectors_srcunit_tex.sum") * 32 / a.steps
    tr[a.kernel] = dict(
        dram_bytes_per_env_step=dram, l2_bytes_per_env_step=l2_all, l2_bytes_from_sms_per_env_step=l2_sm,
        l2_bytes_between_partitions_per_env_step=val("lts__t_sectors_srcunit_ltcfabric.sum") * 32 / a.steps,
        launch_us=val("gpu__time_duration.sum"), steps_in_launch=a.steps,
        tensor_pipe_active_pct=float(r[col["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]]),
        source=f"ncu --set full --clock-control none of profiles/ep_c2_workload.py (4096 episodes, one {a.steps}-step launch), "
               f"build {a.commit}: (dram__bytes_read.sum + dram__bytes_write.sum) / {a.steps} and lts__t_sectors.sum x 32 B / {a.steps}, "
               f"i.e. bytes per step of ALL 4096 episodes; summary in {a.out}")
    json.dump(tr, open(a.traffic, "w"), indent=1)
    print(json.dumps(tr[a.kernel], indent=1))
