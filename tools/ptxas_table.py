"""Parse `nvcc -Xptxas -v` output (stdin) into one line per kernel: registers, spills, stack."""
import re, sys
cur = None
rows = []
for line in sys.stdin:
    m = re.search(r"Compiling entry function '(\S+)'", line)
    if m:
        cur = {"name": m.group(1), "regs": None, "spill": None}
        rows.append(cur)
        continue
    if cur is None:
        continue
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
    if m and cur["spill"] is None:
        cur["spill"] = (int(m.group(1)), int(m.group(2)), int(m.group(3)))
    m = re.search(r"Used (\d+) registers", line)
    if m and cur["regs"] is None:
        cur["regs"] = int(m.group(1))
import subprocess
for r in rows:
    try:
        name = subprocess.run(["c++filt", r["name"]], capture_output=True, text=True).stdout.strip()
    except Exception:
        name = r["name"]
    print(f"{r['regs']:>4} regs  stack/spill {r['spill']}  {name[:110]}")
