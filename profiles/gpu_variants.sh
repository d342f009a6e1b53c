# usage: bash profiles/gpu_variants.sh file.cu "EXTRA flags A" "EXTRA flags B" ...  — rebuilds one source with each flag
# set on the GPU box and prints the per-stage times (A/B runs of kernel variants; the last build is NOT kept in-tree)
F=$1; shift
C=3d-gaussian-splatting_b200/csrc
for ex in "$@"; do
  touch $C/$F && make -s -C $C EXTRA="$ex" > /dev/null 2>&1 || { echo "build failed: $ex"; continue; }
  TAG="[$ex]" python profiles/stage_times.py 40 ${CFG:-A} 2>&1 | tail -1
done
