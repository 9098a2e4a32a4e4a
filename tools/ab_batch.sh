#!/bin/bash
# A/B of guarded kernel variants through the run-time compiler on one GPU box:
#   tools/ab_batch.sh variants.txt [scene] [spp] > log      (lines: "[ENV=val ...] :: -DDEFS ..." ; '#' comments)
# The first line's frame becomes the reference the later lines report their RMSE against (tools/ab_jit.py).
F=${1:-tools/ab_variants.txt}; SCENE=${2:-cornell}; SPP=${3:-256}
export AB_REF=/tmp/ab_ref_$$_$SCENE.npy
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
while IFS= read -r line; do
  [[ -z "$line" || "$line" == \#* ]] && continue
  envs="${line%%::*}"; defs="${line#*::}"; defs="${defs# }"
  echo "### $line"
  env $envs TRT_B200_JIT=force TRT_B200_JIT_DEFS="$defs" timeout 120 python tools/ab_jit.py $SCENE $SPP 2>&1
done < "$F"
