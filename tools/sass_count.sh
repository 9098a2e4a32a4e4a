#!/bin/bash
# registers / stack / shared memory and static SASS size of the production trace kernels in an object file or library:
#   tools/sass_count.sh terminal-raytracer_b200/csrc/build/trt_native.o
for f in "$@"; do
  echo "== $f"
  cuobjdump -res-usage "$f" 2>/dev/null | grep -A1 "trace_native_kernel" | grep -o "trace_native_kernel[A-Za-z0-9_]*\|REG:[0-9]* STACK:[0-9]* SHARED:[0-9]*" | paste - - | cut -c1-160
  cuobjdump -sass "$f" | awk '/Function :/ {name=$3} /^ +\/\*[0-9a-f][0-9a-f][0-9a-f][0-9a-f]\*\// {n[name]++} END {for (k in n) if (k ~ /trace_native/) print n[k], substr(k, 1, 100)}' | sort -k2
done
