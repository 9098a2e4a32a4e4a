#!/bin/bash
# static SASS size / registers of the trace kernels in an object file: tools/sass_count.sh build/trt_native.o
for f in "$@"; do
  echo "== $f"
  cuobjdump -res-usage "$f" 2>/dev/null | grep -A1 "trace_native_kernel" | grep -E "Function|REG" | sed -E 's/.*Function (_ZN3trt19trace_native_kernel[A-Za-z0-9_]*):/\1/; s/ +SHARED.*//' | paste - - | sed -E 's/_ZN3trt19trace_native_kernelINS_//'
  cuobjdump -sass "$f" | awk '/Function :/ {name=$3} /^\s+\/\*[0-9a-f]{4}\*\// {n[name]++} END {for (k in n) if (k ~ /trace_native/) print n[k], k}' | sort -k2
done
