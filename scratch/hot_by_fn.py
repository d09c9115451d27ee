"""SASS of one kernel folded by source function: SASS count, hot instructions (> 2 % of the most executed one), executed
instructions, share of warp instructions and of stall samples.  usage: hot_by_fn.py <ncu sass csv> <nvdisasm text>"""
import csv, re, sys, collections, bisect
sass_csv, dis_txt = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(sass_csv))); hdr = rows[1]; data = rows[2:]
ix = {n: i for i, n in enumerate(hdr)}
def func_table(path):
    tab = []
    pat = re.compile(r'^\s{0,4}([A-Za-z_][A-Za-z0-9_]*)\s*\(')
    for i, l in enumerate(open(path).read().split('\n')):
        m = pat.match(l)
        if m and not l.strip().startswith(('if', 'for', 'while', 'switch', 'return', 'else', '#')) and not l.rstrip().endswith(';'):
            tab.append((i + 1, m.group(1)))
    return tab
tabs = {}
def func_of(f, line):
    if f not in tabs:
        try: tabs[f] = func_table(f)
        except Exception: tabs[f] = []
    t = tabs[f]
    k = bisect.bisect_right([x[0] for x in t], line) - 1
    return t[k][1] if k >= 0 else '?'
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
ie = [int(r[ix['Instructions Executed']]) for r in data]
head = max(ie); tot = sum(ie); tots = sum(int(r[ix['Warp Stall Sampling (All Samples)']]) for r in data)
agg = collections.defaultdict(lambda: [0, 0, 0, 0, 0])
for r in data:
    ch = addr2.get(int(r[ix['Address']], 16) - base, [('?', 0)])
    ch = [c for c in ch if 'csrc' in c[0]] or ch
    f, line = ch[0]
    fn = func_of(f, line) if f != '?' else '?'
    e = int(r[ix['Instructions Executed']])
    a = agg[fn]; a[0] += 1; a[1] += e > 0.02 * head; a[2] += e > 0; a[3] += e; a[4] += int(r[ix['Warp Stall Sampling (All Samples)']])
print("sass %d executed %d hot %d; fn, sass, hot, executed, inst %%, samples %%" % (len(ie), sum(x > 0 for x in ie), sum(x > 0.02 * head for x in ie)))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][4])[:28]:
    print("%-22s %5d %5d %5d %6.1f %6.1f" % (k, v[0], v[1], v[2], 100.0 * v[3] / tot, 100.0 * v[4] / tots))
