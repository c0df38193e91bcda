"""Per-source-line stall summary of an .ncu-rep captured with --import-source on (kernels built with -lineinfo):
    ncu -i X.ncu-rep --page source --print-source cuda,sass --csv > /tmp/x.csv ; python tools/ncu_lines.py /tmp/x.csv [N]"""
import csv, sys
def num(s):
    try: return int(s)
    except ValueError: return 0
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out, cur_file, hdr = [], None, None
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur_file = r[1]; continue
    if r[0] == 'Function Name': continue
    if r[0] == 'Line No': hdr = r; continue
    if hdr and r[0] != '':
        i_s = hdr.index('# Samples')
        stall_names = [(i, h) for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
        st = sorted(((num(r[i]), h) for i, h in stall_names), reverse=True)[:3]
        out.append((cur_file.split('/')[-1], int(r[0]), r[1].strip()[:100], num(r[i_s]), num(r[hdr.index('Instructions Executed')]), st))
tot = sum(o[3] for o in out)
ti = sum(o[4] for o in out)
print('samples', tot, 'warp instructions', ti)
for o in sorted(out, key=lambda o: -o[3])[:top]:
    print(f'{o[0]}:{o[1]:5d} {100*o[3]/tot:5.1f}% inst={100*o[4]/ti:5.1f}% {o[2]}  {[(a, b[6:]) for a, b in o[5] if a]}')
