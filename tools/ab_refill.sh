# A/B of the IMC kernel's launch tuning at the headline size: resident CTAs per SM x refill threshold
B="python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e"
for cfg in ${AB_CONFIGS:-"8:4" "8:2" "8:8" "8:16" "7:4" "6:4"}; do
  IFS=: read c r <<< "$cfg"
  echo "== ctas=$c refill=$r"
  JB_IMC_CTAS_PER_SM=$c JB_IMC_REFILL_MIN=$r timeout 300 $B 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('value %.4g seg/s  step %.2f ms  transport %.2f ms' % (d['value'], d['ms_per_step'], d['transport_ms_per_step']))"
done
