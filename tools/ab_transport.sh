set -x
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "^E  |FAILED|passed|failed|Error" gpurun_out/pytest_gpu.log | head -30
JB_IMC_VALIDATE=1 python -m pytest tests/test_parity_gpu.py -m gpu -q -k "synthetic or inf_deck_counts or smr_two or outflow or large_run" > gpurun_out/pytest_validate.log 2>&1; tail -3 gpurun_out/pytest_validate.log
B="python bench.py --cells 128 --steps 2 --warmup 2 --no-cpu --no-e2e"
for cfg in ${AB_CONFIGS:-"0:4:4" "1:3:4" "1:4:4" "1:5:4" "1:6:4" "1:8:4" "1:6:8" "1:6:1"}; do
  IFS=: read v c r <<< "$cfg"
  echo "== variant=$v ctas=$c refill=$r"
  JB_TRANSPORT_VARIANT=$v JB_IMC_CTAS_PER_SM=$c JB_IMC_REFILL_MIN=$r $B 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('value %.4g seg/s  step %.2f ms  transport %.2f ms' % (d['value'], d['ms_per_step'], d['transport_ms_per_step']))"
done
