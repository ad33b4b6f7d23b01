#!/usr/bin/env python
"""Digest of an .ncu-rep for profiles/: per captured launch the metrics the roofline discussion uses (duration, DRAM bytes,
L1 / L2 hit rates, lanes per instruction, issue-slot use, L1 wavefront use, occupancy, stall reasons) and, per kernel, the
hottest straight-line instruction runs of the SASS with their execution counts and active lanes.

    python tools/ncu_digest.py gpurun_out/x.ncu-rep > profiles/x_digest.txt
"""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__thread_inst_executed_pred_on_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"]


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep] + list(args), capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    rows = list(csv.reader(ncu(rep, "--page", "raw", "--csv").splitlines()))
    hdr, units = rows[0], rows[1]
    print(f"# digest of {rep} (ncu --set full --clock-control none; cold-cache, serialised launches)")
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print(f"\n== launch {d.get('ID')}: {d['Kernel Name'][:110]}")
        for k in KEYS:
            if k in d:
                print(f"   {k} = {d[k]} {units[hdr.index(k)]}")
        try:
            dram = float(d["dram__bytes_read.sum"]) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[units[hdr.index("dram__bytes_read.sum")]] + \
                float(d["dram__bytes_write.sum"]) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[units[hdr.index("dram__bytes_write.sum")]]
            ms = float(d["gpu__time_duration.sum"]) * {"ms": 1, "us": 1e-3, "s": 1e3, "ns": 1e-6}[units[hdr.index("gpu__time_duration.sum")]]
            print(f"   -> DRAM traffic {dram * 1e-9:.3f} GB in {ms:.3f} ms = {dram / ms * 1e-6:.0f} GB/s")
        except Exception:
            pass
    src = list(csv.reader(ncu(rep, "--page", "source", "--csv").splitlines()))
    blocks, hdr2, body, name = [], None, [], None
    for r in src:
        if r and r[0] == "Kernel Name":
            if body:
                blocks.append((name, hdr2, body)); body = []
            name = r[1] if len(r) > 1 else ""
        elif r and r[0] == "Address":
            hdr2 = r
        elif r and r[0].startswith("0x"):
            body.append(r)
    if body:
        blocks.append((name, hdr2, body))
    done = set()
    for name, h, body in blocks:
        if name in done or h is None:
            continue
        done.add(name)
        ix = {k: i for i, k in enumerate(h)}
        num = lambda r, k: int(float(r[ix[k]] or 0))
        seen, b2 = set(), []
        for r in body:
            if r[0] not in seen:
                seen.add(r[0]); b2.append(r)
        tot = sum(num(r, "Instructions Executed") for r in b2) or 1
        tott = sum(num(r, "Thread Instructions Executed") for r in b2)
        samples = sum(num(r, "# Samples") for r in b2) or 1
        print(f"\n== SASS runs of {name[:100]}: {tot} warp instructions, {tott / tot:.2f} lanes per instruction, {len(b2)} static instructions")
        runs = []
        for r in b2:
            c = num(r, "Instructions Executed")
            if runs and abs(runs[-1]["c"] - c) <= 0.001 * max(c, 1):
                R = runs[-1]
            else:
                R = {"c": c, "n": 0, "t": 0, "s": 0, "first": r[ix["Source"]].strip()[:44], "ops": {}}; runs.append(R)
            R["n"] += 1; R["t"] += num(r, "Thread Instructions Executed"); R["s"] += num(r, "# Samples")
            o = r[ix["Source"]].strip().split(); o = (o[1] if o[0].startswith("@") and len(o) > 1 else o[0]).split(".")[0]
            R["ops"][o] = R["ops"].get(o, 0) + 1
        for R in runs:
            share = R["c"] * R["n"] / tot * 100
            if share < 1.0:
                continue
            ops = " ".join(f"{k}x{v}" for k, v in sorted(R["ops"].items(), key=lambda kv: -kv[1])[:6])
            print(f"   executed {R['c']:>9} x {R['n']:>3} instr = {share:5.1f} % of issue, {R['t'] / max(R['c'] * R['n'], 1):5.1f} lanes, {R['s'] / samples * 100:5.1f} % of stall samples | {R['first']} | {ops}")


if __name__ == "__main__":
    main()
