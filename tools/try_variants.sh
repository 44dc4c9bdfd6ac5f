#!/bin/bash
# usage: tools/try_variants.sh N "rho list" name [name...]   (warm estimate_normals timing per variant; name "main" = the shipped library)
N=$1; RHOS=$2; shift; shift
for v in "$@"; do
  for r in $RHOS; do
  echo "== $v rho=$r"
  L=$PWD/open3d-py-extension_b200/variants/libo3dpp_$v.so; [ "$v" = main ] && L=$PWD/open3d-py-extension_b200/libo3dpp_b200.so
  O3DPP_KNN_RHO=$r O3DPP_LIB=$L O3DPP_STATS=1 python tools/prof_knn.py $N 2>&1 | grep -v "o3dpp stats" | tail -2 | cut -c1-400
  done
done
