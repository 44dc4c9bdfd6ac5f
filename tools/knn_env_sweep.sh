#!/bin/bash
# usage: tools/knn_env_sweep.sh N "rho list"   (warm estimate_normals timing of the shipped library per staging radius)
N=$1; RHOS=$2
for r in $RHOS; do
  echo "== rho=$r"
  O3DPP_KNN_RHO=$r O3DPP_STATS=1 python tools/prof_knn.py $N 2>&1 | grep -v "o3dpp stats" | tail -2 | cut -c1-330
done
