import sys, re, collections
# summarise attn_trace output: per role, average delta from previous event of the same role, by tag name
d=collections.defaultdict(list)
for line in open(sys.argv[1]):
    m=re.match(r"\s*(\d+)\s+\(\+\s*(\d+)\)\s+(\S+)\s+(.*)", line)
    if not m: continue
    t,dt,role,name=int(m.group(1)),int(m.group(2)),m.group(3),m.group(4).strip()
    d[(role,name)].append(dt)
for (role,name),v in sorted(d.items()):
    v2=v[len(v)//4:]  # skip warm-up quarter
    print("%-10s %-22s n=%4d  mean %7.0f  med %7.0f" % (role,name,len(v),sum(v2)/max(len(v2),1),sorted(v2)[len(v2)//2]))
