#!/bin/bash
cd "$(dirname "$0")/.."
tag=${1:-var1}
VARIANTS="base ld default" bash scripts/gpu_variants.sh $tag
VARIANTS="default" GROUPS_LIST="2" bash scripts/gpu_variants.sh $tag
VARIANTS="default" RESORTS="100 160" bash scripts/gpu_variants.sh $tag
WORKLOAD=default_scene_1Mi_waters_per_gpu VARIANTS="default" bash scripts/gpu_variants.sh $tag
echo "== staged" | tee -a gpurun_out/$tag/variants.log
MRISIM_STAGE_MNP=1 WORKLOAD=default_scene_1Mi_waters_per_gpu VARIANTS="default" bash scripts/gpu_variants.sh $tag
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "traced_sprint or cfg4b or dense" 2>&1 | tail -3 | tee -a gpurun_out/$tag/variants.log
