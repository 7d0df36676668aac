#!/bin/bash
# One gpurun call: GPU tests, both bench arms, ncu launch list, ncu --set full of the hot kernels, summaries.
# usage: tools/gpu_evidence.sh <tag>      (writes gpurun_out/<tag>_*; copy what should be judged into profiles/)
T=${1:-r2x}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_tests.log 2>&1; echo "tests rc=$?"; tail -3 $O/${T}_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/${T}_smoke.log
timeout 400 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"
timeout 400 python bench.py --impl reference --steps 1 --warmup 1 > $O/${T}_ref.json 2> $O/${T}_ref.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/${T}_launches.csv \
  python bench.py --eager --steps 2 --warmup 3 --no-cpu-baseline --no-roofline > $O/${T}_ncu_launch.log 2>&1; echo "launchlist rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:"ssim_fwd_kernel|ssim_bwd_kernel|mlp_fwd_kernel|mlp_dgrad_kernel|mlp_wgrad_kernel|render_fwd_kernel|render_bwd_kernel|preprocess_fwd_kernel|preprocess_bwd_kernel|tile_count_kernel|tile_scatter_kernel|tile_sort_kernel" \
  --launch-skip 48 --launch-count 12 -o $O/${T}_top -f python bench.py --eager --steps 1 --warmup 3 --no-cpu-baseline --no-roofline > $O/${T}_ncu_full.log 2>&1; echo "full rc=$?"
ncu -i $O/${T}_top.ncu-rep --page raw --csv > $O/${T}_raw.csv 2>/dev/null
python tools/summarize_ncu.py launches $O/${T}_launches.csv $O/${T}_launches.md "bench.py --steps 2 --warmup 3 (C2), $T" || true
python tools/summarize_ncu.py raw $O/${T}_raw.csv $O/${T}_top_kernels.md "hot kernels of one C2 frame, $T" || true
python tools/summarize_ncu.py traffic $O/${T}_raw.csv $O/${T}_traffic_C2.json "$T ncu --set full" || true
head -c 1500 $O/${T}_bench.json
