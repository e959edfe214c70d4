"""Per-opcode instruction mix of the hot loop from an `ncu --page source --csv --print-source sass` dump."""
import csv, collections, re, sys
path, warps, iters = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
rows = list(csv.reader(open(path)))
hdr = rows[1]
iS, iE, iSamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
tot = 0; hot = collections.Counter(); samp = collections.Counter()
for r in rows[2:]:
    if r and r[0] == "Kernel Name":
        break
    if len(r) <= iE or r[0] == "Address":
        continue
    try:
        n = int(r[iE])
    except ValueError:
        continue
    src = r[iS].strip()
    op = re.sub(r'^@!?U?P\d+\s+', '', src).split()[0].split('.')[0]
    if n >= warps * iters * 0.5:
        hot[op] += n; samp[op] += int(r[iSamp])
    tot += n
loop = sum(hot.values())
print("total warp insts", tot, "in-loop", loop, "per step", loop / (warps * iters))
for op, n in hot.most_common(45):
    print(f"{op:10s} {n / (warps * iters):7.1f}  samples {samp[op]}")
