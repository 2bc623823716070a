"""Per-source-line stall samples from `ncu --page source --csv --print-source cuda,sass` output."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, out, tot = None, [], 0
for r in rows:
    if len(r) == 2 and r[0] == 'File Path':
        cur = r[1].split('/')[-1]
        continue
    if len(r) < 8 or not r[0].isdigit():
        continue
    try:
        n = int(r[4])
    except ValueError:
        continue
    out.append((n, cur, int(r[0]), r[1].strip()[:105], int(r[7]) if r[7].isdigit() else 0))
    tot += n
print("total samples", tot)
out.sort(reverse=True)
for n, f, l, src, ins in out[:top]:
    print("%6d %5.1f%% %-16s %5d inst=%-10d %s" % (n, 100 * n / tot, f, l, ins, src))
