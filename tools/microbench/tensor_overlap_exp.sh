# needs the experiment build: make -C hadoop-1.0.3-gpu_b200 clean && make -C hadoop-1.0.3-gpu_b200 EXPERIMENTS=1 (bench.py's count check fails by design: pass over it)
# K2 alone under HGM_TENSOR_DEBUG: 0 = normal, 1 = no MMA issued, 2 = no TMEM loads
for dbg in 1 2; do for shape in "64 1024 40000000" "128 4096 8000000"; do set -- $shape
  HGM_TENSOR_DEBUG=$dbg timeout -k 10 120 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --d $1 --k $2 --points $3 > gpurun_out/exp_${dbg}_$1.json 2> gpurun_out/exp_${dbg}_$1.err
  echo "debug=$dbg d=$1 rc=$? $(cut -c1-170 gpurun_out/exp_${dbg}_$1.json) $(tail -2 gpurun_out/exp_${dbg}_$1.err | cut -c1-200)"
done; done
