import re,sys
rows=[]
for line in open(sys.argv[1]):
    m=re.match(r"\[timeline\] (.+?)\s+entry\s+([-\d.]+)\s+wait\s+([-\d.]+)\s+exit\s+([-\d.]+)", line)
    if m: rows.append((m.group(1).strip(), float(m.group(2)), float(m.group(3)), float(m.group(4))))
starts=[r[2] for r in rows if r[0]=="flow_embed"]
d=[round(b-a,1) for a,b in zip(starts,starts[1:])]
print("frame periods (us):", d)
