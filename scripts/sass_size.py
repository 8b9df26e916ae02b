"""SASS bytes per kernel of the built library (cuobjdump -sass): the fused wf1024_kernel must stay
below the 64 KB instruction-cache cliff (profiles/README.md)."""
import re, subprocess, sys, os
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "vector_b200", "csrc", "libb200vec.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
name, last, rows = None, None, []
for line in sass.splitlines():
    if "Function :" in line:
        if name is not None and last is not None:
            rows.append((last + 16, name))
        name, last = line.split("Function :")[1].strip(), None
    else:
        m = re.match(r"\s*/\*([0-9a-f]{4,6})\*/", line)
        if m:
            last = int(m.group(1), 16)
if name is not None and last is not None:
    rows.append((last + 16, name))
for size, nm in sorted(rows, reverse=True):
    dem = subprocess.run(["c++filt", nm], capture_output=True, text=True).stdout.strip()
    print(f"{size:8d}  {dem[:150]}")
