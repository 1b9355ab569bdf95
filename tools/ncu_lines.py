"""Per-source-line warp-stall summary of one kernel in an .ncu-rep (run here, no GPU needed):
   python tools/ncu_lines.py gpurun_out/prof.ncu-rep <kernel regex> [min_pct] [launches of that kernel to skip]
Reads `ncu --page source --print-source cuda,sass --csv` and prints the CUDA lines that hold the most
stall samples, with their top stall reasons and executed-instruction counts."""
import csv
import io
import subprocess
import sys


def main():
    rep, pat = sys.argv[1], sys.argv[2]
    min_pct = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
    skip = sys.argv[4] if len(sys.argv) > 4 else "0"
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                          "--kernel-name", "regex:" + pat, "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    files, cur, head = [], None, None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1]
            continue
        if r[0] == "Line No":
            head = r
            continue
        if head is None or len(r) != len(head) or r[0] == "":
            continue  # sass rows have an empty line number
        files.append((cur, r))
    ix = {n: i for i, n in enumerate(head)}
    S = ix["Warp Stall Sampling (All Samples)"]
    I = ix["Instructions Executed"]
    stall_cols = [(n, i) for i, n in enumerate(head) if n.startswith("stall_") and "Not Issued" not in n]
    tot = sum(int(r[S] or 0) for _, r in files)
    toti = sum(int(r[I] or 0) for _, r in files)
    print(f"total samples {tot}, warp instructions {toti}")
    for f, r in files:
        s = int(r[S] or 0)
        if s < tot * min_pct / 100:
            continue
        st = sorted(((int(r[i] or 0), n) for n, i in stall_cols), reverse=True)[:3]
        print(f"{f.split('/')[-1]}:{r[0]:>4} {100 * s / tot:5.1f}% inst {100 * int(r[I] or 0) / toti:4.1f}%  "
              f"{r[1].strip()[:70]:70s} {[(n[6:], c) for c, n in st if c]}")


if __name__ == "__main__":
    main()
