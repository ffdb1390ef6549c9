#!/bin/bash
# compute-sanitizer passes over every kernel family at small shapes (run on the GPU box):
#   tools/sanitize.sh [memcheck|racecheck|synccheck|initcheck ...]   -> gpurun_out/sanitize_<tool>_<group>.log
# Each group is a pytest selection whose kernels are small enough for the sanitizer's serialised replay.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TOOLS="${*:-memcheck racecheck}"
declare -A SELS=(
  [selftest]="tests/test_gpu_selftest.py"
  [k1_pair]="tests/test_gpu_mlp.py -k 255-npar0 or 1000-npar3 or rays_mode"
  [k1_single]="tests/test_gpu_mlp.py -k 255-npar0 or 1000-npar3"
  [train]="tests/test_gpu_train.py -k saves_tile or native_backward"
  [k3]="tests/test_gpu_sampling.py -k 100-63-128 or 37-64 or hierarchical"
  [k2]="tests/test_gpu_composite.py -k 300-256 or 7-33 or instanced or backward"
  [instancer]="tests/test_gpu_instancer.py"
)
for tool in $TOOLS; do
  for g in selftest k1_pair k1_single train k3 k2 instancer; do
    sel="${SELS[$g]}"
    file="${sel%% -k *}"
    expr=""
    [[ "$sel" == *" -k "* ]] && expr="${sel#* -k }"
    log="gpurun_out/sanitize_${tool}_${g}.log"
    env_extra=""
    [[ "$g" == "k1_single" ]] && export NTX_MLP_CTAS=1 || unset NTX_MLP_CTAS
    echo "== $tool $g: $sel" | tee "$log"
    if [[ -n "$expr" ]]; then
      timeout 900 compute-sanitizer --tool "$tool" --print-limit 20 python -m pytest "$file" -x -q -m gpu -k "$expr" >> "$log" 2>&1
    else
      timeout 900 compute-sanitizer --tool "$tool" --print-limit 20 python -m pytest "$file" -x -q -m gpu >> "$log" 2>&1
    fi
    echo "exit $?" >> "$log"
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|exit " "$log" | tail -4
  done
done
