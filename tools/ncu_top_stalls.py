import csv, sys, subprocess
rep, pat = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 28
out = subprocess.run(["ncu","-i",rep,"--page","source","--csv","--kernel-name","regex:"+pat],capture_output=True,text=True).stdout
rows = [r for r in csv.reader(out.splitlines())]
hdr = rows[1]
isrc, isamp, iex = hdr.index('Source'), hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Instructions Executed')
data = []
for i, r in enumerate(rows[2:]):
    if len(r) != len(hdr):
        break                      # next section of the page
    data.append((int(r[isamp] or 0), r[isrc].strip(), int(r[iex] or 0), i))
tot = sum(d[0] for d in data)
print('total samples', tot, 'instructions', len(data))
for s, src, ex, i in sorted(data, key=lambda d: -d[0])[:n]:
    print(f'{100*s/tot:5.1f}%  #{i:5d} ex={ex:9d}  {src[:110]}')
if len(sys.argv) > 4:
    lo, hi = map(int, sys.argv[4].split(':'))
    for s, src, ex, i in data[lo:hi]:
        print(f'{s:6d} #{i:5d} ex={ex:9d} {src[:120]}')
