# usage (under gpurun): bash profiles/micro/gpu_proj_probe.sh "flags A" "flags B" ... — forward projection alone, one build per flag set
C=3d-gaussian-splatting_b200/csrc
for ex in "$@"; do
  touch $C/project_sh.cu && make -s -C $C EXTRA="$ex" > /dev/null 2>&1 || { echo "build failed: $ex"; continue; }
  TAG="[$ex]" python profiles/micro/proj_probe.py 50 2>&1 | tail -1
done
