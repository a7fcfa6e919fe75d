"""Reads an `ncu --set full` report (here, on the CPU box) and writes what is committed under profiles/:

    python tools/ncu_summary.py gpurun_out/prof_targets_r02.ncu-rep profiles/ncu_full_r02_summary.txt

  * the per-launch summary of selected raw-page metrics for every kvdp_ kernel in the report, and
  * profiles/ncu_traffic.json: per kernel the measured DRAM bytes per launch TOGETHER WITH a hash of the source files
    the kernel is compiled from. bench.py prints `roofline.traffic` only while that hash still matches the tree it runs
    from, so a kernel change cannot keep quoting an old capture.
"""
import csv
import hashlib
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "dynamo_exp_b200", "csrc")
# kernel-name key -> the source files it is built from
KERNEL_SOURCES = {
    "kvdp_copy_tma_tma_kernel": ["strided_copy.cu", "strided_copy.cuh", "common.cuh"],
    "kvdp_copy_tma_in_kernel": ["strided_copy.cu", "strided_copy.cuh", "common.cuh"],
    "kvdp_copy_small_kernel": ["strided_copy.cu", "strided_copy.cuh", "common.cuh"],
    "kvdp_copy_vec_kernel": ["strided_copy.cu", "strided_copy.cuh", "common.cuh"],
    "kvdp_hash_blocks_tma": ["token_hash.cu", "common.cuh"],
    "kvdp_hash_chain": ["token_hash.cu", "common.cuh"],
}
METRICS = ["launch__grid_size", "launch__block_size", "launch__registers_per_thread", "gpu__time_duration.sum",
           "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
           "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
           "lts__t_sector_hit_rate.pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-9, "us": 1e-6, "ms": 1e-3,
         "s": 1.0, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}


def source_hash(kernel_key: str) -> str:
    h = hashlib.sha256()
    for f in KERNEL_SOURCES[kernel_key]:
        h.update(f.encode())
        h.update(open(os.path.join(CSRC, f), "rb").read())
    return h.hexdigest()[:16]


def kernel_key(name: str):
    for k in KERNEL_SOURCES:
        if k in name:
            return k
    return None


def main(rep: str, out_txt: str, note: str = ""):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr, units = rows[hdr_i], rows[hdr_i + 1]
    col = {name: i for i, name in enumerate(hdr)}
    out = ["# `ncu --set full --clock-control none --import-source on` summary (raw page, selected metrics) of %s" %
           os.path.basename(rep)]
    if note:
        out += ["# " + ln for ln in note.splitlines()]
    out += ["# Times under ncu are cold-cache and serialised; bench.py's CUDA-event numbers are the ones to quote.", ""]
    traffic = {}
    for r in rows[hdr_i + 2:]:
        if len(r) < len(hdr):
            continue
        name = r[col["Kernel Name"]]
        out.append("  Kernel Name = %s" % name)
        vals = {}
        for m in METRICS:
            if m in col:
                v, u = r[col[m]], units[col[m]]
                out.append("  %s = %s %s" % (m, v, u))
                try:
                    vals[m] = float(v.replace(",", "")) * SCALE.get(u, 1.0)
                except ValueError:
                    pass
        if "dram__bytes_read.sum" in vals and "gpu__time_duration.sum" in vals:
            tot = vals["dram__bytes_read.sum"] + vals["dram__bytes_write.sum"]
            out.append("  -> dram traffic (read+write) = %.4f GB per launch; %.0f GB/s under ncu (cold caches, serialised)" %
                       (tot / 1e9, tot / vals["gpu__time_duration.sum"] / 1e9))
            k = kernel_key(name)
            if k:
                traffic.setdefault(k, []).append(tot)
        out.append("")
    with open(out_txt, "w") as f:
        f.write("\n".join(out))
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    table = {}
    if os.path.exists(tpath):
        try:
            table = {k: v for k, v in json.load(open(tpath)).items() if isinstance(v, dict)}
        except Exception:
            table = {}
    for k, v in traffic.items():
        table[k] = {"dram_bytes_per_launch": sum(v) / len(v), "launches": len(v), "source_hash": source_hash(k),
                    "report": os.path.basename(rep), "summary": os.path.relpath(out_txt, ROOT)}
    with open(tpath, "w") as f:
        json.dump(table, f, indent=1)
    print("wrote", out_txt, "and", tpath, "for", sorted(traffic))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
