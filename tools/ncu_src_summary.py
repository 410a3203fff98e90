"""Summarise an `ncu --page source --csv` export: instruction-class shares of stall samples and the hottest SASS lines."""
import csv, sys, re, collections
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path)))
hdr = rows[1]
# several launches in one export: sections repeat the two header rows; keep section `sec` (third argument, default 0)
sec = int(sys.argv[3]) if len(sys.argv) > 3 else 0
starts = [i for i, r in enumerate(rows) if r == hdr]
a = starts[sec]; b = starts[sec + 1] - 1 if sec + 1 < len(starts) else len(rows)
body = rows[a + 1:b]; rows = [rows[a - 1]] + rows[a:]
ci = {h: i for i, h in enumerate(hdr)}
S, NI, EX = ci['Warp Stall Sampling (All Samples)'], ci['Warp Stall Sampling (Not-issued Samples)'], ci['Instructions Executed']
tot = sum(int(r[S] or 0) for r in body); totex = sum(int(r[EX] or 0) for r in body)
print("kernel:", rows[0][1][:100]); print("SASS lines %d, stall samples %d, warp instructions executed %d" % (len(body), tot, totex))
cls = collections.Counter(); clsex = collections.Counter()
for r in body:
    m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)', r[1]); op = m.group(2) if m else '?'
    cls[op] += int(r[S] or 0); clsex[op] += int(r[EX] or 0)
print("-- by opcode (stall-sample share | executed share)")
for op, v in cls.most_common(14): print("  %-10s %5.1f%% | %5.1f%%" % (op, 100.0 * v / tot, 100.0 * clsex[op] / totex))
print("-- hottest lines")
for i in sorted(range(len(body)), key=lambda i: -int(body[i][S] or 0))[:top]:
    r = body[i]; print("  #%-5d %5.1f%%  ex=%-10s %s" % (i, 100.0 * int(r[S] or 0) / tot, r[EX], r[1].strip()[:90]))
