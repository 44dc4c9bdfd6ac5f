import cProfile, pstats, sys, os
sys.argv = ["bench_frames.py", "--frames", "48"]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pr = cProfile.Profile(); pr.enable()
exec(open(os.path.join(ROOT, "tools", "bench_frames.py")).read())
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
