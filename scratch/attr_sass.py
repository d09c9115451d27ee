"""Attribute the per-instruction ncu counters of k_run_par (ncu --page source --print-source sass --csv) to
source functions through the line info of the cubin (nvdisasm --print-line-info-inline)."""
import csv, re, sys, collections, bisect
sass_csv, dis_txt = sys.argv[1], sys.argv[2]
MODE = sys.argv[3] if len(sys.argv) > 3 else 'inner'
rows = list(csv.reader(open(sass_csv)))
hdr = rows[1]; data = rows[2:]
ix = {n: i for i, n in enumerate(hdr)}
# function table from sources
def func_table(path):
    tab = []
    pat = re.compile(r'^\s{0,4}([A-Za-z_][A-Za-z0-9_]*)\s*\(')
    lines = open(path).read().split('\n')
    for i, l in enumerate(lines):
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
    return (t[k][1] if k >= 0 else '?')
chain = []; last = [('?', 0)]; addr2 = {}; fresh = True
for l in open(dis_txt):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        if fresh:
            chain = []; fresh = False
        chain.append((m.group(1), int(m.group(2)))); continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S.*);', l)
    if m:
        if chain: last = chain
        addr2[int(m.group(1), 16)] = last
        fresh = True
base = int(data[0][ix['Address']], 16)
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
tot = [0, 0]
for r in data:
    a = int(r[ix['Address']], 16) - base
    ch = addr2.get(a, [('?', 0)])
    ch = [c for c in ch if 'csrc' in c[0]] or ch
    which = 0 if MODE == 'inner' else max(len(ch) - 2, 0)
    f, line = ch[which]
    fn = func_of(f, line) if f != '?' else '?'
    ie = int(r[ix['Instructions Executed']]); ss = int(r[ix['Warp Stall Sampling (All Samples)']])
    key = fn
    agg[key][0] += ie; agg[key][1] += ss; agg[key][2] += 1
    tot[0] += ie; tot[1] += ss
print("total inst %d samples %d sass %d" % (tot[0], tot[1], len(data)))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    print("%-22s inst %5.1f%%  samples %5.1f%%  sass %4d" % (k, 100.0 * v[0] / tot[0], 100.0 * v[1] / tot[1], v[2]))
