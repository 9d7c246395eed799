"""Turn gpurun_out ncu artefacts into the small tracked summaries under profiles/.
usage: python profiles/summarize.py launches <launches.csv> | full <report.ncu-rep> | fullcsv <raw.csv>... |
       traffic <out.json> key=<family>:<raw.csv>:<kernel substring>[:grid] ..."""
import collections
import csv
import re
import subprocess
import sys

KEEP = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
        "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.max", "launch__shared_mem_per_block_dynamic"]


def launches(path):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    h, data = rows[hi], rows[hi + 1:]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in data:
        if len(r) <= vi:
            continue
        name = re.sub(r"\(.*", "", r[ki])
        agg[name][0] += 1
        agg[name][1] += float(r[vi].replace(",", ""))
    tot = sum(v[1] for v in agg.values())
    print("kernel,launches,total_us,avg_us,share_pct")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k},{v[0]},{v[1] / 1e3:.1f},{v[1] / v[0] / 1e3:.2f},{100 * v[1] / tot:.1f}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = rows[0]
    keep = [k for k in KEEP if k in h]
    w = csv.writer(sys.stdout)
    w.writerow(keep)
    w.writerow([rows[1][h.index(k)] for k in keep])
    for r in rows[2:]:
        w.writerow([r[h.index(k)][:70] for k in keep])


EXTRA = ["sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_alu.sum",
         "sm__inst_executed_pipe_xu.sum", "sm__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


_SCALE = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6,
          "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}


def _rows(path):
    """Header and data rows of an `ncu --page raw --csv` export, durations in us and byte counts in MB
    (ncu picks a unit per file)."""
    rows = list(csv.reader(open(path)))
    h, units = rows[0], rows[1]
    data = rows[2:]
    for i, (name, u) in enumerate(zip(h, units)):
        if name in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum") and u in _SCALE and _SCALE[u] != 1.0:
            for r in data:
                r[i] = f"{float(r[i].replace(',', '')) * _SCALE[u]:.6f}"
    return h, data


def fullcsv(*paths):
    """`ncu -i rep --page raw --csv` outputs (exported on the GPU box) -> one row per distinct
    (kernel, grid): the LAST instance of each, i.e. warmed up.  Tag = the csv's section name."""
    keep = KEEP + EXTRA
    w = csv.writer(sys.stdout)
    first = True
    for path in paths:
        h, data = _rows(path)
        cols = [k for k in keep if k in h]
        if first:
            w.writerow(["section"] + [c + {"gpu__time_duration.sum": " [us]", "dram__bytes_read.sum": " [MB]",
                                                  "dram__bytes_write.sum": " [MB]"}.get(c, "") for c in cols])
            first = False
        seen = {}
        for r in data:
            key = (r[h.index("Kernel Name")], r[h.index("launch__grid_size")])
            seen[key] = r
        tag = re.sub(r".*ncu_|\.csv$", "", path)
        for r in seen.values():
            w.writerow([tag] + [re.sub(r"\(.*", "", r[h.index(k)])[:80] if k == "Kernel Name" else r[h.index(k)] for k in cols])


def _src_hash(root, names):
    import hashlib
    import os
    hsh = hashlib.sha256()
    for name in names:
        with open(os.path.join(root, name), "rb") as f:
            hsh.update(f.read())
    return hsh.hexdigest()[:16]


KERNEL_SOURCES = {"mlp": ["splitlearning_b200/csrc/mlp_kernels.cuh", "splitlearning_b200/csrc/common.cuh"],
                  "alice": ["splitlearning_b200/csrc/alice_kernels.cuh", "splitlearning_b200/csrc/common.cuh"],
                  "tc": ["splitlearning_b200/csrc/gemm_tc.cuh", "splitlearning_b200/csrc/common.cuh"]}


def traffic(out_json, *specs):
    """specs: key=family:csvpath:kernel-substring[:grid] ... (family = mlp | alice | tc: which source
    files the kernel lives in).  Writes the per-launch DRAM bytes of each named kernel together with a
    hash of those sources, so that bench.py can tell a stale figure (and then reports null)."""
    import json
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    table = {"kernels": {}}
    for spec in specs:
        key, rest = spec.split("=", 1)
        parts = rest.split(":")
        fam, path, sub, grid = parts[0], parts[1], parts[2], (parts[3] if len(parts) > 3 else None)
        h, data = _rows(path)
        hit = [r for r in data if sub in r[h.index("Kernel Name")] and (grid is None or r[h.index("launch__grid_size")] == grid)]
        if not hit:
            raise SystemExit(f"no kernel matching {sub!r} in {path}")
        r = hit[-1]
        rd = float(r[h.index("dram__bytes_read.sum")]) * 1e6
        wr = float(r[h.index("dram__bytes_write.sum")]) * 1e6
        table["kernels"][key] = {"kernel": re.sub(r"\(.*", "", r[h.index("Kernel Name")]), "grid": r[h.index("launch__grid_size")],
                                 "dram_bytes_read": rd, "dram_bytes_write": wr, "dram_bytes_per_launch": rd + wr,
                                 "us": float(r[h.index("gpu__time_duration.sum")]), "capture": os.path.basename(path),
                                 "sources": KERNEL_SOURCES[fam], "source_sha16": _src_hash(root, KERNEL_SOURCES[fam])}
    with open(out_json, "w") as f:
        json.dump(table, f, indent=1)
    print("wrote", out_json)


if __name__ == "__main__":
    {"launches": launches, "full": full, "fullcsv": fullcsv, "traffic": traffic}[sys.argv[1]](*sys.argv[2:])
