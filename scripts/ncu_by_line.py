"""Join an ncu SASS source page (ncu -i X.ncu-rep --page source --csv --kernel-name regex:K) with nvdisasm -g line
info of the same kernel: warp instructions executed and stall samples per CUDA source line.
    python scripts/ncu_by_line.py sass.csv disasm.txt kernel_substring [top]"""
import csv
import re
import sys

sass_csv, dis, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
rows = list(csv.reader(open(sass_csv)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
inst = [r for r in rows[2:] if len(r) > 5]
lines, cur, on = [], None, False
for ln in open(dis):
    if ln.startswith(".text.") or ln.lstrip().startswith(".section\t.text."):
        on = kname in ln
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.search(r"/\*[0-9a-f]{4,}\*/", ln) and not ln.strip().startswith("/*") is False or re.match(r"\s+/\*[0-9a-f]{4}\*/", ln):
        lines.append(cur)
print("sass rows %d, disasm instructions %d" % (len(inst), len(lines)))
n = min(len(inst), len(lines))
agg = {}
for r, l in zip(inst[:n], lines[:n]):
    a = agg.setdefault(l, [0, 0, 0, 0, 0.0])
    a[0] += int(r[ix["Instructions Executed"]])
    a[1] += int(r[ix["# Samples"]])
    a[2] += int(r[ix["stall_no_inst"]]) if "stall_no_inst" in ix else 0
    a[3] += 1
    a[4] += int(r[ix["Thread Instructions Executed"]])
tot = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print("total warp instructions %d, samples %d" % (tot, ts))
for l, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%-22s sass %4d  inst %6.2f%%  samples %6.2f%%  no_inst %5.1f%% of its samples  lanes %.1f" % (
        "%s:%d" % l if l else "?", a[3], 100.0 * a[0] / tot, 100.0 * a[1] / max(ts, 1), 100.0 * a[2] / max(a[1], 1), a[4] / max(a[0], 1)))
