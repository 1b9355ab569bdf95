"""Summarise one bench step out of an ncu launch list (gpu__time_duration.sum, --csv):
   python tools/launch_list.py launches.csv [first-kernel regex] [must-contain regex] [must-not-contain regex]
The last complete step (from one launch matching the first regex to the next) that contains a kernel matching the
second regex (default: the hybrid coalesce's ls2_sort_dedup_kernel, i.e. the uniform configs[1] step rather than the
skewed variant that follows it) is printed launch by launch with its share; generator / torch kernels are left out."""
import collections
import csv
import re
import sys

path = sys.argv[1]
first = re.compile(sys.argv[2] if len(sys.argv) > 2 else r"^ls_bucket_hist_kernel")
skip = re.compile(r"^(synth_|void at::|at::|void \(anonymous|elementwise|vectorized)")
rows = []
with open(path, newline="") as f:
    lines = [ln for ln in f if ln.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        name = r["Kernel Name"]
        short = re.sub(r"^void ", "", re.sub(r"\(.*$", "", name))
        rows.append((short, float(r["Metric Value"].replace(",", "")) / 1e3))
starts = [i for i, (n, _) in enumerate(rows) if first.search(n)]
if len(starts) < 2:
    sys.exit("fewer than two steps in the list")
must = re.compile(sys.argv[3] if len(sys.argv) > 3 else r"^ls2?_sort_dedup_kernel")
# the skewed variant of the bench refines its oversized buckets (rf_* kernels): left out by default
must_not = re.compile(sys.argv[4] if len(sys.argv) > 4 else r"^rf_")
step = None
for a, b in reversed(list(zip(starts[:-1], starts[1:]))):  # steps that are certainly complete, latest first
    cand = [(n, us) for n, us in rows[a:b] if not skip.search(n)]
    if any(must.search(n) for n, _ in cand) and not (must_not.pattern and any(must_not.search(n) for n, _ in cand)):
        step = cand
        break
if step is None:
    sys.exit("no complete step contains a kernel matching " + must.pattern)
total = sum(us for _, us in step)
print("#        us   share  kernel")
for n, us in step:
    print(f"{us:11.1f}  {100 * us / total:5.1f}%  {n}")
print(f"# total {total:.1f} us in {len(step)} launches")
fam = collections.Counter()
for n, us in step:
    fam[re.sub(r"<.*$", "", n)] += us
print("# by family:")
for n, us in fam.most_common():
    print(f"#   {n:30s} {us:8.1f} us  {100 * us / total:5.1f}%")
