"""Instruction count of every kernel in a library (code size matters: k_build stalls on instruction fetch):
    python scripts/sass_size.py duckdb_vss_b200/libgvss.so [substring]"""
import re, subprocess, sys
out = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
cur, n = None, {}
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1); n[cur] = 0; continue
    if cur and re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+\S", line):
        n[cur] += 1
for k, v in sorted(n.items(), key=lambda kv: -kv[1]):
    if len(sys.argv) < 3 or sys.argv[2] in k:
        print(f"{v:8d} instr {v * 16 / 1024:8.1f} KB  {k}")
