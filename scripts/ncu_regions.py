"""Dynamic instruction / stall-sample attribution of one kernel to 10-line source regions.
usage: ncu_regions.py src.csv sass.txt kernel_substring [top]"""
import re, csv, collections, sys
src, sassf, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
sass = open(sassf).read().split('\n')
start = [i for i, l in enumerate(sass) if l.strip().startswith('.section') and '.text.' in l and kern in l][0]
r = re.compile(r'//## File "([^"]+)", line (\d+)')
cur, seq = None, []
for l in sass[start + 1:]:
    if l.strip().startswith('.section'): break
    m = r.search(l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', l): seq.append(cur)
rows = list(csv.reader(open(src)))
hdr = rows[1]; ie = hdr.index("Instructions Executed"); isamp = hdr.index("# Samples")
data = rows[2:]
assert len(data) >= len(seq) - 2, (len(data), len(seq))
inst, samp = collections.Counter(), collections.Counter()
for k, row in enumerate(data):
    if k >= len(seq) or len(row) <= isamp: break
    key = seq[k]
    key = (key[0], key[1] // 10 * 10) if key else None
    inst[key] += int(row[ie]); samp[key] += int(row[isamp])
ti, ts = sum(inst.values()), sum(samp.values())
print("total warp-inst", ti, "samples", ts)
for k, v in samp.most_common(top):
    print("%-40s samples %5.1f%%  inst %5.1f%%" % (k, 100 * v / ts, 100 * inst[k] / ti))
