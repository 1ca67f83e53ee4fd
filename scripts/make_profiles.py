#!/usr/bin/env python
"""Regenerates the evidence files under profiles/ from the built library and from whatever ncu reports
and bench outputs gpurun brought back into gpurun_out/ (scratch).  Run here, on the CPU box."""
import collections
import csv
import glob
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "deepfm-tensorflow-distributed-training-on-amazon-sagemaker_b200", "_native", "libdfm_cuda.so")
OUT = os.path.join(ROOT, "profiles")
os.makedirs(OUT, exist_ok=True)

MNEMONICS = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMAPF", "LDTM", "UTCATOMSWS", "SYNCS", "LDGMC", "STG.E.128.STRONG.SYS",
             "STG.E.STRONG.SYS", "STG.E.ENL2.256", "REDG", "RED.E",
             "ATOMG", "LDG", "STG", "MEMBAR.SC.SYS", "MEMBAR.ALL.SYS", "HMMA", "FFMA"]


def sass_summary():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    per = collections.OrderedDict()
    cur = None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            per[cur] = collections.Counter()
            continue
        if cur and "/*" in line:
            m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
            if m:
                op = m.group(1)
                per[cur]["_total"] += 1
                for k in MNEMONICS:
                    if op.startswith(k):
                        per[cur][k] += 1
    with open(os.path.join(OUT, "sass_summary.txt"), "w") as f:
        f.write("SASS mnemonic counts per kernel of libdfm_cuda.so (cuobjdump -sass, sm_100a).\n"
                "UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA tensor load, UTCBAR = tcgen05.commit,\n"
                "UTCATOMSWS = TMEM alloc, SYNCS = mbarrier ops, STG.E.ENL2.256 = 256-bit store, REDG = red.global.\n"
                "comm.cu: LDGMC.E.ADD.F32x4 = multimem.ld_reduce (the sum is formed inside the NVSwitch), multimem.st is a\n"
                "STG.E.128.STRONG.SYS on the multicast address; peer loads / stores and the flag protocol are ordinary\n"
                "LDG / STG plus ld.acquire / st.release at system scope (MEMBAR.*.SYS).\n\n")
        for k, c in per.items():
            if c["_total"] < 8:
                continue
            parts = ", ".join(f"{m}={c[m]}" for m in MNEMONICS if c[m])
            f.write(f"{k}\n    instructions={c['_total']}  {parts}\n")
    return per


def ptxas_summary():
    rows = []
    for log in sorted(glob.glob(os.path.join(ROOT, "build", "obj", "*.ptxas.log"))):
        txt = open(log).read()
        for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'.*?Used (\d+) registers(.*?)\n", txt, re.S):
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            spill = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", m.group(0))
            rows.append((os.path.basename(log).replace(".ptxas.log", ".cu"), name, m.group(2),
                         spill.group(1) if spill else "?", spill.group(2) if spill else "?"))
    with open(os.path.join(OUT, "ptxas_registers.txt"), "w") as f:
        f.write("nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -Xptxas -v\nfile, kernel, registers, spill stores B, spill loads B\n")
        for r in rows:
            f.write(", ".join(r) + "\n")


def ncu_summaries():
    want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__registers_per_thread",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "lts__t_sector_hit_rate.pct", "lts__t_sectors_srcunit_tex_op_red.sum", "launch__waves_per_multiprocessor",
            "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
            "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"]
    for rep in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "*.ncu-rep"))):
        p = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True)
        rows = list(csv.reader(io.StringIO(p.stdout)))
        if len(rows) < 3:
            continue
        H = rows[0]
        idx = [H.index(w) if w in H else None for w in want]
        name = os.path.basename(rep).replace(".ncu-rep", "")
        with open(os.path.join(OUT, f"ncu_{name}.csv"), "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(want)
            w.writerow([rows[1][i] if i is not None else "" for i in idx])
            for r in rows[2:]:
                w.writerow([r[i] if i is not None else "" for i in idx])


def sass_listings():
    """Full SASS of the kernels BASELINE.json asks to see: the fused collective kernel and one tcgen05 GEMM."""
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    blocks = re.split(r"(?=\t\tFunction : )", txt)
    for key, fname in (("exchange_apply_kernelILb0", "sass_exchange_apply_kernel.txt"),
                       ("dense_sync_kernelILb0", "sass_dense_sync_kernel.txt"),
                       ("allreduce_kernel", "sass_allreduce_kernel.txt"),
                       ("mlp_layer_kernelILb0", "sass_mlp_layer_fwd_kernel.txt"),
                       ("embed_bwd_kernel", "sass_embed_bwd_kernel.txt")):
        for b in blocks:
            if key in b.split("\n", 1)[0]:
                lines = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", ln) for ln in b.splitlines()]
                lines = [ln for ln in lines if ln.strip() and not re.fullmatch(r"\s*/\* 0x[0-9a-f]+ \*/", ln)]
                lines = [re.sub(r"^\s*(/\*[0-9a-f]+\*/)\s+", r"\1 ", ln) for ln in lines]
                with open(os.path.join(OUT, fname), "w") as f:
                    f.write("cuobjdump -sass libdfm_cuda.so (sm_100a), encodings stripped\n" + "\n".join(lines) + "\n")
                break


if __name__ == "__main__":
    sass_listings()
    sass_summary()
    ptxas_summary()
    ncu_summaries()
    print("profiles/ updated:", sorted(os.listdir(OUT)))
