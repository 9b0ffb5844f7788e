set -x
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "^E  |FAILED|passed|failed|Error" gpurun_out/pytest_gpu.log | head -30
JB_IMC_VALIDATE=1 python -m pytest tests/test_parity_gpu.py -m gpu -q -k "synthetic or inf_deck_counts or smr_two or outflow or large_run" > gpurun_out/pytest_validate.log 2>&1; tail -3 gpurun_out/pytest_validate.log
B="python bench.py --cells 128 --steps 2 --warmup 2 --no-cpu --no-e2e"
for cfg in "0 4 4" "1 4 1" "1 4 4" "1 4 8" "1 4 16" "1 3 4" "1 5 4" "1 3 8"; do
  set -- $cfg
  echo "== variant=$1 ctas=$2 refill=$3"
  JB_TRANSPORT_VARIANT=$1 JB_IMC_CTAS_PER_SM=$2 JB_IMC_REFILL_MIN=$3 $B 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('value %.4g seg/s  step %.2f ms  transport %.2f ms' % (d['value'], d['ms_per_step'], d['transport_ms_per_step']))"
done
