"""cProfile of the host side of a grid fit outside the EM loop (construction, MAP decode, result assembly):
    python scripts/prof_host.py [workload]"""
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else 'config2'
    from remixt_b200 import pipeline
    exp, inits = bench.load_workload(name)
    cfg = {'max_copy_number': 6, 'num_em_iter': 0, 'num_update_iter': 5}
    pipeline.fit_grid_batched(exp, dict(enumerate(inits)), cfg)
    pr = cProfile.Profile()
    prof = {}
    pr.enable()
    pipeline.fit_grid_batched(exp, dict(enumerate(inits)), cfg, profile=prof)
    pr.disable()
    print(prof)
    st = pstats.Stats(pr)
    st.sort_stats('cumulative').print_stats(40)
    st.sort_stats('tottime').print_stats(20)


if __name__ == '__main__':
    main()
