"""Join an `ncu --page source --csv` SASS export with `nvdisasm -gi` line info of the same cubin (instruction order), so that
stall samples and executed instructions can be attributed to source lines (innermost and outermost = kernel-level line).
usage: python tools/ncu_join_lines.py <source.csv.gz> <disasm.txt> <kernel substring>"""
import collections, csv, gzip, re, sys
src, dis, kname = sys.argv[1], sys.argv[2], sys.argv[3]
ins, cur, infn = [], None, False
for line in open(dis):
    if line.startswith('.text.'):
        infn = kname in line
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', line)
    if m:
        inl = re.findall(r'inlined at "([^"]+)", line (\d+)', m.group(3))
        cur = (m.group(1).split('/')[-1], int(m.group(2)), [(a.split('/')[-1], int(b)) for a, b in inl]); continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*?);', line)
    if m:
        ins.append((int(m.group(1), 16), cur, m.group(2)))
rows = list(csv.reader(gzip.open(src, 'rt')))
tabs, t = [], None
for r in rows:
    if r and r[0] == "Kernel Name": t = {'name': r[1], 'rows': []}; tabs.append(t); continue
    if r and r[0] == "Address": t['hdr'] = r; continue
    if t is not None and r: t['rows'].append(r)
t = [x for x in tabs if kname in x['name']][0]
h = t['hdr']; ie = h.index("Instructions Executed"); iss = h.index("Warp Stall Sampling (All Samples)")
print("instructions: disasm", len(ins), "ncu", len(t['rows']))
inner, inner_i, outer, outer_i = (collections.Counter() for _ in range(4))
for (off, cur, txt), r in zip(ins, t['rows']):
    st, ex = int(r[iss] or 0), int(r[ie] or 0)
    if cur is None: continue
    inner[(cur[0], cur[1])] += st; inner_i[(cur[0], cur[1])] += ex
    top = cur[2][-1] if cur[2] else (cur[0], cur[1])
    outer[top] += st; outer_i[top] += ex
ts, ti = sum(inner.values()), sum(inner_i.values())
print("by kernel-level line:")
for k, v in outer.most_common(16): print(f"  {v / ts * 100:5.1f}% stall {outer_i[k] / ti * 100:5.1f}% inst  {k[0]}:{k[1]}")
print("by innermost line:")
for k, v in inner.most_common(24): print(f"  {v / ts * 100:5.1f}% stall {inner_i[k] / ti * 100:5.1f}% inst  {k[0]}:{k[1]}")
