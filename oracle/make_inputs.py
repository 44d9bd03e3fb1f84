"""Generate simulated experiments with the UNMODIFIED reference simulator (this container only).

TEST INFRASTRUCTURE ONLY.  Writes plain-array ``.npz`` experiment files (x, l, adjacencies,
breakpoints, init params) that travel to the GPU box; the reference simulator itself cannot.

Sampler chain as in reference ``simulations/pipeline.py:144-155``; parameters per SURVEY.md 8(d).

Usage: python oracle/make_inputs.py NAME OUT.npz      (NAME in CONFIGS below)
"""
import os
import sys
import warnings

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import refenv  # noqa: E402

CHROM_LENGTHS_HG19 = {
    '1': 249250621, '2': 243199373, '3': 198022430, '4': 191154276, '5': 180915260,
    '6': 171115067, '7': 159138663, '8': 146364022, '9': 141213431, '10': 135534747,
    '11': 135006516, '12': 133851895, '13': 115169878, '14': 107349540, '15': 102531392,
    '16': 90354753, '17': 81195210, '18': 78077248, '19': 59128983, '20': 63025520,
    '21': 48129895, '22': 51304566, 'X': 155270560,
}

CONFIGS = {
    # config 1 of BASELINE.json: single chromosome, ~1k segments, ~50 breakpoints
    'config1': dict(N=1000, M=3, chromosome_lengths={'15': 101991189}, h_total=0.1,
                    frac_normal=0.4, frac_clone_1=0.4, num_ancestral_events=25,
                    num_descendent_events=10, num_false_breakpoints=5,
                    proportion_breakpoints_detected=0.9, random_seed=999),
    # small multi-chromosome case for fast parity tests
    'small': dict(N=240, M=3, chromosome_lengths={'20': 63025520, '21': 48129895, '22': 51304566},
                  h_total=0.1, frac_normal=0.4, frac_clone_1=0.4, num_ancestral_events=10,
                  num_descendent_events=5, num_false_breakpoints=3,
                  proportion_breakpoints_detected=0.9, random_seed=2001),
    # config 2: whole genome ~30k segments, ~300 breakpoints
    'config2': dict(N=30000, M=3, chromosome_lengths=CHROM_LENGTHS_HG19, h_total=0.1,
                    frac_normal=0.4, frac_clone_1=0.4, num_ancestral_events=100,
                    num_descendent_events=50, num_false_breakpoints=50,
                    proportion_breakpoints_detected=0.9, random_seed=999),
    # mid-size whole genome (fast to simulate)
    'genome5k': dict(N=5000, M=3, chromosome_lengths=CHROM_LENGTHS_HG19, h_total=0.1,
                     frac_normal=0.4, frac_clone_1=0.4, num_ancestral_events=100,
                     num_descendent_events=50, num_false_breakpoints=50,
                     proportion_breakpoints_detected=0.9, random_seed=999),
}


def simulate(params):
    refenv.install()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import remixt.simulations.experiment as sim
    params = dict(params)
    hs = sim.RearrangementHistorySampler(params)
    gs = sim.GenomeCollectionSampler(hs, params)
    ms = sim.GenomeMixtureSampler(params)
    es = sim.ExperimentSampler(params)
    np.random.seed(params['random_seed'])
    genomes = gs.sample_genome_collection()
    mixture = ms.sample_genome_mixture(genomes)
    return es.sample_experiment(mixture)


def experiment_arrays(experiment):
    """Flatten the reference Experiment object (analysis/experiment.py:244) to plain arrays."""
    adj = np.array(sorted(experiment.adjacencies), dtype=np.int64).reshape(-1, 2)
    bp_ids, bps = [], []
    for bp_id, bp in experiment.breakpoints.items():
        (n1, s1), (n2, s2) = tuple(bp)
        bp_ids.append(str(bp_id))
        bps.append((n1, s1, n2, s2))
    out = dict(
        x=np.asarray(experiment.x, dtype=np.float64),
        l=np.asarray(experiment.l, dtype=np.float64),
        adjacencies=adj,
        breakpoints=np.array(bps, dtype=np.int64).reshape(-1, 4),
        breakpoint_ids=np.array(bp_ids),
        chromosome=np.asarray(experiment.segment_chromosome_id).astype(str),
        true_h=np.asarray(experiment.h, dtype=np.float64),
    )
    return out


def main():
    name, out = sys.argv[1], sys.argv[2]
    exp = simulate(CONFIGS[name])
    arrs = experiment_arrays(exp)
    np.savez_compressed(out, **arrs)
    print(name, 'N=%d adjacencies=%d breakpoints=%d' % (
        arrs['x'].shape[0], arrs['adjacencies'].shape[0], arrs['breakpoints'].shape[0]))


if __name__ == '__main__':
    main()
