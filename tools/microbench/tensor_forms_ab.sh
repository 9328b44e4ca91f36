# A/B of the two K2 forms on one box (boxes differ by several % under the power cap)
for rep in 1 2; do for form in single pair; do for shape in "64 1024 40000000" "128 4096 8000000"; do set -- $shape
  timeout -k 10 120 python bench.py --tensor-form $form --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --d $1 --k $2 --points $3 2>/dev/null > gpurun_out/ab.json
  python -c "import json; j=json.load(open('gpurun_out/ab.json')); print('rep$rep $form d=$1', round(j['ms_per_step'],3), 'ms', j['clocks']['sm_mhz'], j['clocks']['reasons'])"
done; done; done
