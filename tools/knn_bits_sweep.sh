for b in 13 12 11 10; do echo "== bits=$b"; O3DPP_KNN_BITS=$b O3DPP_STATS=1 python tools/prof_knn.py 10000000 2>&1 | grep -E "o3dpp lane|^ms" | tail -2 | cut -c1-230; done
