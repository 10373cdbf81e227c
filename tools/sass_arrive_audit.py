"""Audit: is any shared-memory load still outstanding when a stage-release arrive is issued?

For every plain mbarrier arrive (SYNCS.ARRIVE...A1T0) in a cubin's SASS, walk back through the straight-line code before it
and report LDS destinations that no instruction between the load and the arrive has read.  An arrive that releases a
TMA-filled stage must not be issued before the loads from that stage have returned (measured on B200: a refill overtook them,
em_wide3.cu); a register that was consumed proves the load has returned.
usage: python tools/sass_arrive_audit.py file.cubin [...]
"""
import re, subprocess, sys

def regs_of(tok, width):
    m = re.match(r"R(\d+)", tok)
    if not m: return []
    b = int(m.group(1))
    return [b + i for i in range(width)]

def audit(path):
    sass = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout.splitlines()
    func = None; ins = []; out = []
    def flush():
        nonlocal ins
        for i, (txt) in enumerate(ins):
            if "SYNCS.ARRIVE" in txt and "A1T0" in txt:
                pend = {}
                for j in range(max(0, i - 200), i):
                    t = ins[j]
                    if re.search(r"\b(BRA|EXIT|BSYNC|BSSY|RET)\b", t) and "LDS" not in t: pend = {} if "BRA" in t and "@" not in t else pend
                    ops = re.sub(r"^@!?U?P\d+\s+", "", t).split(None, 1)
                    if len(ops) < 2: continue
                    op, rest = ops
                    toks = [x.strip() for x in rest.rstrip(" ;").split(",")]
                    srcs = toks[1:] if not op.startswith(("ST", "RED", "ATOM", "SYNCS", "BAR")) else toks
                    used = set()
                    for s_ in srcs:
                        for m in re.finditer(r"R(\d+)", s_):
                            w = 2 if ("DMMA" in op or ".64" in op or "DADD" in op or "DFMA" in op or "DMUL" in op) else 1
                            if "DMMA" in op and s_ is srcs[-1]: w = 4
                            for k in range(w): used.add(int(m.group(1)) + k)
                    for r in list(pend):
                        if r in used: del pend[r]
                    if op.startswith("LDS"):
                        w = 4 if ".128" in op else 2 if ".64" in op else 1
                        for r in regs_of(toks[0], w): pend[r] = j
                if pend:
                    out.append((func, txt.strip()[:60], sorted(pend)))
        ins = []
    for l in sass:
        m = re.search(r"Function : (\S+)", l)
        if m: flush(); func = m.group(1); continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?);", l)
        if m: ins.append(m.group(1))
    flush()
    return out

for p in sys.argv[1:]:
    res = audit(p)
    print(p, ":", "no outstanding loads at any arrive" if not res else "")
    for f, t, r in res: print("   ", f[:70], "|", t, "| unconsumed LDS destinations:", r)
