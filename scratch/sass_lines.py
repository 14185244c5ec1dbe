"""Attribute the per-instruction counters of an ncu source-page CSV (SASS view) to source lines, using nvdisasm -g line markers
of the same cubin: python scratch/sass_lines.py <cubin.disasm> <kernel substring> <source.csv> <n_units> [top]"""
import collections, csv, re, sys
dis, kname, src_csv, units = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
top = int(sys.argv[5]) if len(sys.argv) > 5 else 50
lines = open(dis).read().splitlines()
start = [i for i, l in enumerate(lines) if l.startswith('.text.') and kname in l][0]
cur, seq = None, []
for l in lines[start + 1:]:
    if l.startswith('.section') or l.startswith('.text.'):
        break
    m = re.search(r'//## File "(.*?)", line (\d+)', l)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        seq.append((cur, m.group(2).strip()))
rows = list(csv.reader(open(src_csv)))
hdr, data = rows[1], rows[2:]
ix, isamp = hdr.index("Instructions Executed"), hdr.index("# Samples")
assert len(data) == len(seq), (len(data), len(seq))
by, bys, tot, tots = collections.Counter(), collections.Counter(), 0, 0
for (line, ins), r in zip(seq, data):
    by[line] += int(r[ix]); bys[line] += int(r[isamp]); tot += int(r[ix]); tots += int(r[isamp])
print("warp instructions", tot, "per unit", tot / units, "samples", tots)
cache = {}
for line, n in by.most_common(top):
    f, ln = line
    txt = ""
    try:
        if f not in cache:
            cache[f] = open("fit-slam_b200/csrc/" + f).read().splitlines()
        txt = cache[f][ln - 1].strip()[:100]
    except Exception:
        pass
    print(f"{f}:{ln:<5d} {n / units:9.0f} {100 * n / tot:5.1f}%  samp {100 * bys[line] / tots:5.1f}%  {txt}")
