"""Static instruction mix of the innermost backward-branch loop of a kernel in libshipsim.so."""
import re, subprocess, sys, collections
pat = sys.argv[1]
txt = subprocess.run(["cuobjdump", "-sass", "sac_maritime_ast_b200/libshipsim.so"], capture_output=True, text=True).stdout
blocks = re.split(r'\n\s*Function : ', txt)
for b in blocks[1:]:
    name = b.split('\n')[0]
    if not re.search(pat, name):
        continue
    ins = re.findall(r'/\*([0-9a-f]{4})\*/\s+((?:@!?U?P\d\s+)?[A-Z0-9_.]+)([^;]*);', b)
    addr = [int(a, 16) for a, _, _ in ins]
    # find backward branches; take the largest loop body
    best = None
    for idx, (a, op, rest) in enumerate(ins):
        if 'BRA' in op:
            m = re.search(r'0x([0-9a-f]+)', rest)
            if m:
                tgt = int(m.group(1), 16)
                if tgt < addr[idx]:
                    span = (tgt, addr[idx])
                    if best is None or span[1] - span[0] > best[1] - best[0]:
                        best = span
    print(name[:90], "total", len(ins))
    if best:
        body = [(op, rest) for (a, op, rest), ad in zip(ins, addr) if best[0] <= ad <= best[1]]
        c = collections.Counter(re.sub(r'^@!?U?P\d\s+', '', op).split('.')[0] for op, _ in body)
        print(" loop body", len(body), dict(c.most_common(30)))
        if len(sys.argv) > 2:
            for op, rest in body:
                print("   ", op, rest)
