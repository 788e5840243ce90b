#!/bin/bash
# Interleaved A/B of an environment toggle of the library (the toggles are read once per process, so every sample is
# its own process; ABAB order cancels thermal / power-cap drift).  On the GPU box:
#   tools/env_ab.sh CXL_K1_PACK 0 1 4 -- python tools/kbench.py --B 8192 --only k1
# prints the first output line of every run, tagged with the value used.
var=$1; a=$2; b=$3; rounds=$4; shift 5
for i in $(seq 1 "$rounds"); do
  for v in "$a" "$b"; do
    printf '%s=%s  ' "$var" "$v"
    env "$var=$v" "$@" 2>/dev/null | head -1
  done
done
