import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; data = rows[2:]
col = {h:i for i,h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot = {h:0 for h in stall_cols}; ns = 0; per = []
for r in data:
    try: n = int(r[col['# Samples']])
    except: continue
    ns += n
    for h in stall_cols:
        try: tot[h] += int(r[col[h]])
        except: pass
    per.append((n, r[col['Source']][:80], r[col['Instructions Executed']]))
print("total samples", ns)
for h,v in sorted(tot.items(), key=lambda kv:-kv[1])[:8]: print(f"  {h:28s} {v:7d} {100*v/ns:5.1f}%")
for n,src,ex in sorted(per, key=lambda x:-x[0])[:25]: print(f"  {n:6d} {ex:>9s}  {src}")
