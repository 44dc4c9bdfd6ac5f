#!/bin/bash
# usage: tools/knn_env_sweep.sh N "ENV=V ENV=V" "ENV=V ..." ...   (warm estimate_normals timing per environment setting;
# LIBV=<name> in a setting selects the tuning build open3d-py-extension_b200/variants/libo3dpp_<name>.so)
N=$1; shift
for cfg in "$@"; do
  echo "== $cfg"
  extra=""
  for kv in $cfg; do case $kv in LIBV=*) extra="O3DPP_LIB=$PWD/open3d-py-extension_b200/variants/libo3dpp_${kv#LIBV=}.so";; esac; done
  env $cfg $extra O3DPP_STATS=1 python tools/prof_knn.py $N 2>&1 | grep "lane\]" | tail -1 | cut -c1-330
done
