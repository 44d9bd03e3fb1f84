"""cProfile of a complete grid fit (host side of the lock-step M-step).  python scripts/prof_mstep.py [workload]"""
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
    cfg = {'max_copy_number': 6, 'num_em_iter': 5, 'num_update_iter': 5}
    pipeline.fit_grid_batched(exp, dict(enumerate(inits)), cfg)
    pr = cProfile.Profile()
    pr.enable()
    pipeline.fit_grid_batched(exp, dict(enumerate(inits)), cfg)
    pr.disable()
    st = pstats.Stats(pr)
    st.sort_stats('cumulative').print_stats(45)
    st.sort_stats('tottime').print_stats(25)
    st.sort_stats('cumulative').print_stats('pipeline.py|cn_model.py|bpmodel.py|results.py|_lib.py')


if __name__ == '__main__':
    main()
