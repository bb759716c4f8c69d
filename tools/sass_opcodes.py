"""Opcode histogram of the shipped library's SASS (evidence for which data paths the kernels use).
    python tools/sass_opcodes.py > profiles/r2_sass_opcodes.txt"""
import collections, re, subprocess, sys
from pathlib import Path
so = Path(__file__).resolve().parent.parent / "tensorrt-dft-plugins_b200" / "trt_dft_plugins" / "libtrt_dft_plugins.so"
out = subprocess.run(["cuobjdump", "-sass", str(so)], capture_output=True, text=True).stdout
total = collections.Counter()
per_fn = collections.defaultdict(collections.Counter)
fn = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and fn:
        op = m.group(1)
        total[op] += 1
        per_fn[fn][op] += 1
watch = ["UBLKCP", "UTMALDG", "UTMASTG", "LDGSTS", "SYNCS", "BAR", "LDG", "STG", "LDS", "STS", "SHFL", "LDL", "STL"]
print("# data-path opcodes over the whole library (prefix match)")
for w in watch:
    print(f"{w:8s} {sum(c for o, c in total.items() if o.startswith(w))}")
print("\n# per kernel: TMA / async-copy / local-memory opcodes")
for f, c in sorted(per_fn.items()):
    row = {w: sum(n for o, n in c.items() if o.startswith(w)) for w in ("UBLKCP", "UTMALDG", "LDGSTS", "LDL", "STL")}
    if any(row.values()):
        print(f[:150], row)
print("\n# top 40 opcodes")
for op, n in total.most_common(40):
    print(f"{op:28s} {n}")
