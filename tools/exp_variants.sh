#!/bin/bash
# One GPU call for kernel variants built by tools/variant.sh: a pytest selection (parity vs the oracle) and the 4K per-task
# times, per variant.   tools/exp_variants.sh "<pytest -k expression>" "<task name regex to print>" name1 name2 ...   ("base" = default library)
cd "$(dirname "$0")/.."
kexpr=$1; task=$2; shift 2
out=gpurun_out/exp_variants; mkdir -p $out
for v in "$@"; do
  if [ "$v" = base ]; then unset VKR_B200_LIB; else export VKR_B200_LIB=$PWD/vk-renderer_b200/libvkr_b200_$v.so; fi
  echo "=== $v"
  timeout 600 python -m pytest tests/test_parity_gpu.py -q -x -k "$kexpr" > $out/pytest_$v.txt 2>&1; echo "pytest rc=$? $(tail -1 $out/pytest_$v.txt)"
  timeout 300 python tools/time_passes.py 3840 2160 10 > $out/time_$v.txt 2>&1; echo "time rc=$?"
  grep -E "$task|chain" $out/time_$v.txt
done
