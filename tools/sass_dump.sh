#!/bin/bash
# SASS of one kernel of an object file as "addr instruction": tools/sass_dump.sh build/trt_native.o ShapeFixed
cuobjdump -sass "$1" | awk -v pat="$2" '/Function :/ {on = ($3 ~ pat) && ($3 ~ /trace_native/)} on && /^ +\/\*[0-9a-f][0-9a-f][0-9a-f][0-9a-f]\*\// {a=$1; gsub(/[\/\*]/,"",a); $1=""; sub(/ *\/\*.*$/,""); print a $0}'
