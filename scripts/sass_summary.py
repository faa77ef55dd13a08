"""opcode histogram per kernel of the shipped library (profiles/r02_sass_summary.txt): python scripts/sass_summary.py > profiles/..."""
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "htool_python_b200/libhmb200.so"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
cols = [("UBLKCP", r"\bUBLKCP"), ("SYNCS", r"\bSYNCS"), ("DMMA", r"\bDMMA"), ("DFMA", r"\bDFMA"), ("LDS128", r"\bLDS\.128"), ("LDS64", r"\bLDS\.64"),
        ("WSYNC", r"\bWARPSYNC"), ("MBSYS", r"\bMEMBAR\.ALL\.SYS"), ("LDL/STL", r"\b(LDL|STL)\b")]
rows = []
for name, body in zip(names, re.split(r"Function : \S+", sass)[1:]):
    ins = [l for l in body.split("\n") if re.search(r"/\*[0-9a-f]{4}\*/", l)]
    text = "\n".join(ins)
    rows.append((re.sub(r"hmb200::", "", name), len(ins), [len(re.findall(rx, text)) for _, rx in cols]))
rows.sort()
w = max(len(r[0]) for r in rows)
print(f"{'kernel':<{w}} {'instr':>6} " + " ".join(f"{c:>6}" for c, _ in cols))
for name, n, counts in rows:
    print(f"{name:<{w}} {n:>6} " + " ".join(f"{c:>6}" for c in counts))
