#!/bin/bash
# One GPU call: parity (1280x720 pass-level, vs the oracle) and 4K per-task times of the SSR-trace variants built by tools/variant.sh.
#   tools/exp_trace.sh name1 name2 ...      ("base" = the default library); results under gpurun_out/exp_trace/
cd "$(dirname "$0")/.."
out=gpurun_out/exp_trace; mkdir -p $out
for v in "$@"; do
  if [ "$v" = base ]; then unset VKR_B200_LIB; else export VKR_B200_LIB=$PWD/vk-renderer_b200/libvkr_b200_$v.so; fi
  echo "=== $v"
  timeout 600 python -m pytest tests/test_parity_gpu.py -q -x -k "pass_level_parity and 1280" > $out/pytest_$v.txt 2>&1; echo "pytest rc=$?" | tee -a $out/pytest_$v.txt
  cp gpurun_out/parity_pass_1280x720.json $out/parity_$v.json 2>/dev/null
  timeout 300 python tools/time_passes.py 3840 2160 10 > $out/time_$v.txt 2>&1; echo "time rc=$?"
  grep -E "SSSR_trace|chain" $out/time_$v.txt
  python - <<PY
import json
r=[x for x in json.load(open("$out/parity_$v.json")) if x["task"]=="SSSR_trace"]
for x in r: print({k:(round(v,6) if isinstance(v,float) else v) for k,v in x.items() if k not in ("task","kind")})
PY
done
