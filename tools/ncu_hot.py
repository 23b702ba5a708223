"""Top stall locations of one kernel from an ncu report:  python tools/ncu_hot.py <rep> <kernel-id regex> [n]
Prints the SASS lines with the most warp-stall samples and their dominant stall reasons."""
import csv, subprocess, sys
rep, kid = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-id', kid], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[hi]
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
si = hdr.index('# Samples')
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
total = sum(int(r[si]) for r in body)
print('total samples', total, 'lines', len(body))
order = sorted(range(len(body)), key=lambda i: -int(body[i][si]))[:n]
for i in sorted(order):
    r = body[i]
    st = sorted(((int(r[c]), hdr[c][6:]) for c in stall_cols), reverse=True)[:3]
    print(f'{i:5d} {int(r[si]):7d} {100.0*int(r[si])/total:5.1f}%  {r[1].strip()[:70]:70s} ', ' '.join(f'{b}:{a}' for a, b in st if a))
