"""Opcode histogram (executed warp instructions, stall samples) of one kernel in an .ncu-rep:
   python tools/ncu_ops.py rep.ncu-rep <kernel regex>"""
import collections, csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                      "regex:" + sys.argv[2], "--launch-count", "1"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(out)))
hi = [i for i, x in enumerate(r) if x and x[0] == "Address"][0]
h = r[hi]
ix = {n: i for i, n in enumerate(h)}
cnt, st = collections.Counter(), collections.Counter()
for x in r[hi + 1:]:
    if len(x) != len(h) or not x[ix["Instructions Executed"]].isdigit():
        continue
    op = [o for o in x[ix["Source"]].split() if not o.startswith("@")][0].split(".")[0]
    cnt[op] += int(x[ix["Instructions Executed"]])
    st[op] += int(x[ix["Warp Stall Sampling (All Samples)"]] or 0)
tot, ts = sum(cnt.values()), sum(st.values())
print("total warp instructions", tot)
for op, c in cnt.most_common(28):
    print(f"{op:10s} {c:>12d} {100 * c / tot:5.1f}%  stalls {100 * st[op] / ts:5.1f}%")
