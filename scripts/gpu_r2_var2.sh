#!/bin/bash
cd "$(dirname "$0")/.."
tag=${1:-var2}
VARIANTS="default b64 b32" bash scripts/gpu_variants.sh $tag
WORKLOAD=cfg4b_dense8000cells_1Mi_waters_per_gpu VARIANTS="default b64 b32" bash scripts/gpu_variants.sh $tag
WORKLOAD=default_scene_1Mi_waters_per_gpu VARIANTS="default b32" bash scripts/gpu_variants.sh $tag
