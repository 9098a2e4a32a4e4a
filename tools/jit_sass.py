"""SASS of the run-time specialised trace kernel for a set of -D switches (no GPU needed):
python tools/jit_sass.py [Shape..] [-D...] > file   -- lines "addr instruction" like tools/sass_dump.sh"""
import re, subprocess, sys, tempfile
from jit_compile_check import compile_shape
args = sys.argv[1:]
shape = args.pop(0) if args and args[0].startswith("Shape") else "ShapeFixed<3, 6, 2, 0xa05u, 2, 0x0u, 0x4deu>"
cubin, _ = compile_shape(shape, args)
with tempfile.NamedTemporaryFile(suffix=".cubin") as f:
    f.write(cubin); f.flush()
    out = subprocess.run(["cuobjdump", "-sass", f.name], capture_output=True, text=True).stdout
seen = set()
for l in out.replace("\x00", " ").splitlines():
    m = re.match(r"\s+/\*([0-9a-f]{4})\*/\s+(.*?;)", l)
    if m and m.group(1) not in seen:
        seen.add(m.group(1))
        print(m.group(1), m.group(2))
