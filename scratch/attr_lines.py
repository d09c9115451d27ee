"""Per source line (innermost frame) instruction / stall-sample shares for one file."""
import csv, re, sys, collections
sass_csv, dis_txt, want = sys.argv[1], sys.argv[2], sys.argv[3]
lo, hi = (int(sys.argv[4]), int(sys.argv[5])) if len(sys.argv) > 5 else (0, 10**9)
rows = list(csv.reader(open(sass_csv))); hdr = rows[1]; data = rows[2:]
ix = {n: i for i, n in enumerate(hdr)}
chain = []; last = [('?', 0)]; addr2 = {}; fresh = True
for l in open(dis_txt):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        if fresh: chain = []; fresh = False
        chain.append((m.group(1), int(m.group(2)))); continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S.*);', l)
    if m:
        if chain: last = chain
        addr2[int(m.group(1), 16)] = last; fresh = True
base = int(data[0][ix['Address']], 16)
agg = collections.defaultdict(lambda: [0, 0, 0]); tot = [0, 0]
for r in data:
    ch = addr2.get(int(r[ix['Address']], 16) - base, [('?', 0)])
    ie = int(r[ix['Instructions Executed']]); ss = int(r[ix['Warp Stall Sampling (All Samples)']])
    tot[0] += ie; tot[1] += ss
    for f, line in ch:
        if f.endswith(want):
            agg[line][0] += ie; agg[line][1] += ss; agg[line][2] += 1
            break
src = open([f for f in ['prisma_b200/csrc/' + want] ][0]).read().split('\n')
for line in sorted(agg):
    if lo <= line <= hi:
        v = agg[line]
        if v[1] * 1000 > tot[1] * 2 or v[0] * 1000 > tot[0] * 2:
            print("%4d inst %4.1f%% smp %4.1f%% n%4d | %s" % (line, 100.0 * v[0] / tot[0], 100.0 * v[1] / tot[1], v[2], src[line - 1][:110]))
