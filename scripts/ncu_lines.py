"""Join an ncu SASS source page (csv) with nvdisasm -g line info: per source line instruction counts and stall samples.
usage: ncu_lines.py <src.csv from `ncu -i rep --page source --csv`> <nvdisasm -g -c output> <kernel mangled substring>"""
import csv, re, sys, collections
src_csv, sass, kern = sys.argv[1:4]
# parse nvdisasm: track current function, file/line markers, instruction offsets
cur_fn, cur_loc, off2loc = None, None, {}
for line in open(sass, errors="ignore"):
    m = re.match(r"\s*\.text\.(\S+):", line)
    if m: cur_fn = m.group(1); continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m: cur_loc = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m and cur_fn and kern in cur_fn:
        off2loc[int(m.group(1), 16)] = cur_loc
rows = list(csv.reader(open(src_csv)))
hdr, data = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
base = int(data[0][0], 16)
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for r in data:
    loc = off2loc.get(int(r[0], 16) - base, ("?", 0))
    a = agg[loc]
    a[0] += int(r[ix["Instructions Executed"]] or 0)
    a[1] += int(r[ix["# Samples"]] or 0)
    for s in stalls:
        v = int(r[ix[s]] or 0)
        if v: a[2][s] += v
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print(f"total instructions {ti}, samples {ts}")
for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[4]) if len(sys.argv) > 4 else 40]:
    top = ", ".join(f"{k[6:]} {v}" for k, v in a[2].most_common(3))
    print(f"{loc[0]}:{loc[1]:<5d} inst {100*a[0]/ti:5.1f}%  samples {100*a[1]/ts:5.1f}%  [{top}]")
