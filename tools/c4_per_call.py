"""C4 (frames through the three blocks) against the number of mats handed to one run_once call: python tools/c4_per_call.py"""
import sys, os, json
sys.path.insert(0, '.'); sys.path.insert(0, './open3d-py-extension_b200')
import bench
for pc in (1, 2, 4, 8):
    r = bench.run_c4(0, 1, 0, None, lambda: None, per_call=pc)
    print(pc, round(r['frames_per_s'], 1), round(r['ms_per_frame_per_gpu'], 3))
