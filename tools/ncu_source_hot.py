"""Aggregate an `ncu --page source --csv --print-source cuda,sass` export per CUDA source line."""
import csv, collections, sys
cur = None; hdr = None
inst = collections.Counter(); stall = collections.Counter(); thr = collections.Counter(); text = {}
for row in csv.reader(open(sys.argv[1])):
    if not row: continue
    if row[0] == 'File Path': cur = row[1].split('/')[-1]; hdr = None; continue
    if row[0] == 'Function Name': continue
    if row[0] == 'Line No':
        hdr = row; ix = hdr.index('Instructions Executed'); sx = hdr.index('# Samples'); tx = hdr.index('Thread Instructions Executed'); continue
    if hdr is None or not row[0].isdigit() or row[2] != '-': continue
    try: n = int(row[ix]); s = int(row[sx]); t = int(row[tx])
    except ValueError: continue
    k = (cur, int(row[0])); inst[k] += n; stall[k] += s; thr[k] += t; text[k] = row[1].strip()[:90]
tot = sum(inst.values()); ts = sum(stall.values())
print("total warp-inst %d, samples %d, avg lanes %.1f" % (tot, ts, sum(thr.values()) / max(tot, 1)))
byfile = collections.Counter()
for k, v in inst.items(): byfile[k[0]] += v
print({k: "%.1f%%" % (100 * v / tot) for k, v in byfile.most_common()})
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
for k, v in inst.most_common(top):
    print("%-13s %4d %5.2f%% inst %5.2f%% stall lanes %4.1f | %s" % (k[0], k[1], 100 * v / tot, 100 * stall[k] / max(ts, 1), thr[k] / max(v, 1), text[k]))
