#!/bin/bash
# Assembles profiles/ncu_r02_summary.md from the committed captures under profiles/r02/ (exported on the build box with
# `ncu -i X.ncu-rep --page raw --csv`; the .ncu-rep itself stays in the untracked gpurun_out/).
cd "$(dirname "$0")/.."
R=profiles/r02
{
echo "# ncu summary, round 2 (config A = 1 M Gaussians SH3, 1920x1080, N_vis 857 449, M 6 500 218; B200, clocks 1965 MHz, no throttle)"
echo
echo "Command: \`bash profiles/gpu_r2_final.sh\` under gpurun (one B200).  ncu times are cold-cache and serialised; the bench's CUDA-event"
echo "stage times (\`profiles/r02/bench_A.json\`, \`roofline.stages_one_stream_ms\`) are the authoritative durations — the SHARE of each"
echo "kernel agrees (blend bwd ~0.43, blend fwd ~0.22, projection bwd ~0.08, depth sort ~0.06, two-level binning ~0.09, projection fwd ~0.04)."
echo "Raw export: \`profiles/r02/ncu_full_configA_raw.csv\`; DRAM bytes per bench stage: \`profiles/ncu_traffic_configA.json\` (\`profiles/make_traffic.py\`);"
echo "per-source-line tables of the blend kernels: \`profiles/r02/raster_bwd_lines.md\`, \`raster_fwd_lines.md\` (\`profiles/line_table.py\`)."
echo
python profiles/make_summary.py $R/ncu_full_configA_raw.csv $R/launches_configA.csv
echo
echo "## launch list of the bench command itself, \`python bench.py --steps 2 --warmup 3\` (first 400 launches: the device leg, two view streams serialised by ncu)"
echo
python profiles/summarize.py launches $R/launches_bench.csv
echo
echo "## launch list, two config-C views (6 M Gaussians, 3840x2160, M = 104.8 M)"
echo
python profiles/summarize.py launches $R/launches_configC.csv
echo
cat $R/notes.md
} > profiles/ncu_r02_summary.md
wc -l profiles/ncu_r02_summary.md
