# usage: bash profiles/gpu_sweep.sh "ENV1=a ENV2=b" "ENV1=c" ...   (one bench run per argument)
for cfg in "$@"; do
  echo "== $cfg"
  env $cfg python bench.py --steps 10 --warmup 3 > gpurun_out/sweep.json 2> gpurun_out/sweep.err || tail -3 gpurun_out/sweep.err
  python profiles/show_bench.py gpurun_out/sweep.json
done
