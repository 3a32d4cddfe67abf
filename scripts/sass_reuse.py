"""Development helper: estimate register-file read pressure of FP64 instructions in a
SASS dump (cuobjdump -sass).  Model: each DFMA/DMUL/DADD issues every 2 cycles and
each fresh 64-bit vector-register source operand costs 1 cycle of register-file
bandwidth; an operand hits the reuse cache if the previous instruction had the same
register in the same slot flagged .reuse.  Prints per-loop statistics."""
import re
import sys

def parse(path, func_pat):
    cur = None
    out = []
    for line in open(path):
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        if cur is None or func_pat not in cur:
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m:
            out.append((int(m.group(1), 16), m.group(2).strip()))
    return out

def analyse(instrs, lo=None, hi=None):
    prev = {}
    n = fresh_tot = 0
    hist = {}
    other = 0
    for addr, txt in instrs:
        if lo is not None and not (lo <= addr <= hi):
            continue
        txt = re.sub(r"^@!?U?P\d+\s+", "", txt)
        op = txt.split()[0]
        if not op.startswith(("DFMA", "DMUL", "DADD")):
            other += 1
            prev = {}  # conservative: any other instruction clears the chain
            continue
        ops = [o.strip() for o in txt[len(op):].split(",")][1:]
        cur = {}
        fresh = set()
        for slot, o in enumerate(ops):
            m = re.match(r"[-|~]*\|?(R\d+)\|?(\.reuse)?", o)
            if not m:
                continue
            reg, reuse = m.group(1), bool(m.group(2))
            if prev.get(slot) != reg:
                fresh.add(reg)
            if reuse:
                cur[slot] = reg
        prev = cur
        k = len(fresh)
        hist[k] = hist.get(k, 0) + 1
        n += 1
        fresh_tot += max(2, k)
    return n, fresh_tot, hist, other

if __name__ == "__main__":
    path, pat = sys.argv[1], sys.argv[2]
    ins = parse(path, pat)
    lo = int(sys.argv[3], 16) if len(sys.argv) > 3 else None
    hi = int(sys.argv[4], 16) if len(sys.argv) > 4 else None
    n, cyc, hist, other = analyse(ins, lo, hi)
    print(f"fp64 instr {n}, other {other}, fresh-operand histogram {dict(sorted(hist.items()))}, "
          f"model cycles {cyc} vs ideal {2 * n} -> {2 * n / max(cyc, 1):.3f}")
