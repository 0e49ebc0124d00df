"""Per-source-line instruction and stall-sample shares of one kernel launch of an ncu report.

  ncu -i rep --page source --csv --kernel-name regex:K --launch-skip S --launch-count 1 > sass.csv
  cuobjdump -xelf all rt_core.o ; nvdisasm -g -c rt_core.sm_100a.cubin > all.dis
  python tools/ncu_lines.py sass.csv all.dis <mangled kernel prefix> [top]
The SASS rows of the report and of the disassembly are the same instructions in the same order; the disassembly carries
the `//## File ..., line N` markers (-lineinfo), including the inlined-at chain, of which the innermost line is used.
"""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
hdr = rows[h]
ie, st, te = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
sass = [(r[1].strip(), int(r[ie]), int(r[st]), int(r[te])) for r in rows[h + 1:] if len(r) > te and r[ie].isdigit()]
lines = open(sys.argv[2]).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith(".text." + sys.argv[3]))
cur, seq = None, []
for l in lines[start + 1:]:
    if l.startswith("//-----"):
        break
    m = re.match(r'\s*//## File "(.*)", line (\d+)(.*)', l)
    if m:
        if "inlined at" not in l or cur is None or True:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4}\*/", l):
        seq.append((cur, l.strip()))
n = min(len(seq), len(sass))
agg, samp, thr = collections.Counter(), collections.Counter(), collections.Counter()
for (c, l), (s, e, sm, t) in zip(seq[:n], sass[:n]):
    agg[c] += e; samp[c] += sm; thr[c] += t
tot, ts = sum(agg.values()), sum(samp.values())
print(f"sass rows {len(sass)} / {len(seq)}; warp instructions {tot}, stall samples {ts}")
for k, v in agg.most_common(int(sys.argv[4]) if len(sys.argv) > 4 else 40):
    print(f"{str(k):34s} {100 * v / tot:5.1f} % inst  {100 * samp[k] / ts:5.1f} % samples  lanes {thr[k] / max(1, v):4.1f}")
