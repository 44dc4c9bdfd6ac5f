"""Tuning builds of the library: python tools/variants.py name:DEF=V,DEF=V ...  -> open3d-py-extension_b200/variants/libo3dpp_<name>.so
Run one with O3DPP_LIB=<path>.  (Built files are git-ignored; they travel to the GPU box with the snapshot.)"""
import importlib.util, os, sys
from concurrent.futures import ThreadPoolExecutor
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "open3d-py-extension_b200")
spec = importlib.util.spec_from_file_location("_b", os.path.join(PKG, "build.py"))
b = importlib.util.module_from_spec(spec); spec.loader.exec_module(b)
os.makedirs(os.path.join(PKG, "variants"), exist_ok=True)
def one(arg):
    name, _, defs = arg.partition(":")
    out = os.path.join(PKG, "variants", f"libo3dpp_{name}.so")
    b.build(out=out, defines=tuple(d for d in defs.split(",") if d))
    return out
with ThreadPoolExecutor(8) as ex:
    for o in ex.map(one, sys.argv[1:]): print(o)
