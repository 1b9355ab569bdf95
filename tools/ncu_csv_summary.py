"""One line per kernel launch out of an `ncu --page raw --csv` export (the export is made on the GPU box when the
.ncu-rep itself is too large to bring back):
   python tools/ncu_csv_summary.py raw.csv [figure-A bytes file] > profiles/rNN_..._ncu.txt
Columns: duration, DRAM bytes read + written (the `traffic` of bench.py's roofline), DRAM throughput % of peak,
achieved GB/s on the DRAM bytes, warps active %, issue active %, shared-memory bank conflicts % of shared wavefronts,
local-memory sectors (spills), registers, grid x block."""
import csv
import re
import sys

path = sys.argv[1]
with open(path, newline="") as f:
    rows = list(csv.reader(ln for ln in f if ln.startswith('"')))
head, units = rows[0], rows[1]
ix = {n: i for i, n in enumerate(head)}


def num(row, name, default=0.0):
    if name not in ix:
        return default
    v = row[ix[name]].replace(",", "")
    try:
        return float(v)
    except ValueError:
        return default


def scaled(row, name):
    """value in base units (bytes, ns): ncu prints Kbyte / Mbyte / usecond ... per column"""
    v = num(row, name)
    u = units[ix[name]] if name in ix else ""
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "nsecond": 1, "usecond": 1e3, "msecond": 1e6,
            "second": 1e9, "ns": 1, "us": 1e3, "ms": 1e6}.get(u, 1)
    return v * mult


print(f"# {path}: {len(rows) - 2} launches (ncu --set full, --clock-control none; cold caches, serialised)")
print(f"{'kernel':58s} {'us':>8s} {'dram rd MB':>10s} {'dram wr MB':>10s} {'dram %':>6s} {'GB/s':>7s} {'warps%':>6s} "
      f"{'issue%':>6s} {'bank%':>6s} {'local sect':>10s} {'regs':>4s} {'grid x block':>14s}")
for r in rows[2:]:
    name = re.sub(r"^void ", "", r[ix["Kernel Name"]])
    name = re.sub(r"^ps::", "", name)
    name = re.sub(r"\((?:[^()]|\([^()]*\))*\)$", "", name)
    if len(name) > 58:
        name = name[:55] + "..."
    dur = scaled(r, "gpu__time_duration.sum")
    rd, wr = scaled(r, "dram__bytes_read.sum"), scaled(r, "dram__bytes_write.sum")
    wave = num(r, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")
    conf = num(r, "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum")
    local = num(r, "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum") + num(r, "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum")
    grid = r[ix["Grid Size"]] if "Grid Size" in ix else ""
    block = r[ix["Block Size"]] if "Block Size" in ix else ""
    print(f"{name:58s} {dur / 1e3:8.1f} {rd / 1e6:10.1f} {wr / 1e6:10.1f} "
          f"{num(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):6.1f} {(rd + wr) / dur if dur else 0:7.0f} "
          f"{num(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):6.1f} "
          f"{num(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):6.1f} "
          f"{100 * conf / wave if wave else 0:6.1f} {local:10.0f} {int(num(r, 'launch__registers_per_thread')):4d} "
          f"{grid + ' x ' + block:>14s}")
