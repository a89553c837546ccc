#!/usr/bin/env python3
"""Summaries for profiles/: tools/ncu_summary.py full <raw.csv>  |  tools/ncu_summary.py launches <launches.csv>

full     : per-launch table of an `ncu --set full` capture exported with `ncu -i X.ncu-rep --page raw --csv`
launches : per-kernel totals / shares of an `ncu --metrics gpu__time_duration.sum` launch list"""
import csv
import sys
from collections import defaultdict

FULL = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"]


def full(path):
    rows = [r for r in csv.reader(open(path)) if r]
    while rows and "Kernel Name" not in rows[0]:
        rows.pop(0)
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print(f"== {r[hdr.index('Kernel Name')]}  (launch id {r[hdr.index('ID')]})")
        for m in FULL:
            if m in hdr:
                i = hdr.index(m)
                print(f"   {m:80s} {r[i]:>16s} {units[i]}")


def launches(path):
    rows = [r for r in csv.reader(open(path)) if r]
    while rows and "Kernel Name" not in rows[0]:
        rows.pop(0)
    hdr = rows[0]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows[1:]:
        if len(r) <= mv:
            continue
        v = float(r[mv].replace(",", ""))
        u = r[mu]
        ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u.replace("usecond", "us").replace("nsecond", "ns").replace("msecond", "ms").replace("second", "s"), 1e-6)
        name = r[kn].split("(")[0]
        tot[name] += ms
        cnt[name] += 1
    s = sum(tot.values())
    print(f"{'kernel':60s} {'launches':>8s} {'total ms':>10s} {'avg us':>9s} {'share':>7s}")
    for k in sorted(tot, key=lambda k: -tot[k]):
        print(f"{k:60s} {cnt[k]:8d} {tot[k]:10.3f} {1e3 * tot[k] / cnt[k]:9.1f} {100 * tot[k] / s:6.1f}%")
    print(f"{'total':60s} {sum(cnt.values()):8d} {s:10.3f}")


def traffic(path, source="profiles/"):
    """per-kernel-family mean DRAM bytes per launch of a `--set full` capture, as JSON for bench.py's roofline.traffic"""
    import json
    rows = [r for r in csv.reader(open(path)) if r]
    while rows and "Kernel Name" not in rows[0]:
        rows.pop(0)
    hdr, units = rows[0], rows[1]
    kn, rd, wr = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    fam = {"k_wave": "intra_lf_wavefront", "k_inter": "inter_pred", "k_itx": "itxfm_add"}
    acc = defaultdict(list)
    for r in rows[2:]:
        for k, name in fam.items():
            if k in r[kn]:
                acc[name].append(float(r[rd].replace(",", "")) * scale[units[rd]] + float(r[wr].replace(",", "")) * scale[units[wr]])
    out = {n: {"dram_bytes_per_launch": sum(v) / len(v), "launches_captured": len(v),
               "source": f"{source} (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, mean over the captured "
                         "launches of tools/ncu_frames.py c3 3: one key frame + two inter frames)"} for n, v in acc.items()}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    {"full": full, "launches": launches, "traffic": traffic}[sys.argv[1]](*sys.argv[2:])
