import csv, sys
f=sys.argv[1]
rows=list(csv.reader(open(f)))
hdr=rows[0]; units=rows[1]; data=rows[2:]
pats=sys.argv[2:]
for i,h in enumerate(hdr):
    if any(p in h for p in pats):
        print(h, '['+units[i]+']', [r[i] for r in data])
