"""Aggregate an `ncu --page source --csv` SASS export by CUDA source line, using `nvdisasm -g -c <cubin>` line markers.
   python tools/ncu_lines.py <source.csv> <nvdisasm.txt> <function-substring> [top]"""
import csv, re, sys, collections

src_csv, dis, fn = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
# --- line markers per instruction of the function
lines, cur, infn = [], None, False
for ln in open(dis):
    if ln.lstrip().startswith(".text."):
        infn = fn in ln
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
        lines.append(cur)
rows = list(csv.reader(open(src_csv)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ci = {h: i for i, h in enumerate(hdr)}
data = rows[hi + 1:]
print(f"# {len(data)} SASS rows, {len(lines)} disassembled instructions")
agg = collections.defaultdict(lambda: collections.Counter())
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for k, r in enumerate(data):
    key = lines[k] if k < len(lines) else None
    a = agg[key]
    a["samples"] += int(r[ci["# Samples"]] or 0)
    a["inst"] += int(r[ci["Instructions Executed"]] or 0)
    for s in stalls:
        a[s] += int(r[ci[s]] or 0)
tot = sum(a["samples"] for a in agg.values())
print(f"# total samples {tot}")
for key, a in sorted(agg.items(), key=lambda kv: -kv[1]["samples"])[:top]:
    st = ", ".join(f"{s[6:]}={a[s]}" for s in sorted(stalls, key=lambda s: -a[s])[:4] if a[s])
    print(f"{str(key):28s} samples {a['samples']:7d} ({100*a['samples']/tot:5.1f}%) inst {a['inst']:10d}  {st}")
