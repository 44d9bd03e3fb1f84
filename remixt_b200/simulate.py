"""Synthetic experiments: the count-level sampler of the reference's simulator and a fast generator of the clone
genomes it mixes.

* :func:`sample_experiment` is ``simulations/experiment.py:ExperimentSampler.sample_experiment`` (``:1243-1400``) for
  the ``negbin_betabin`` emission model (the default, ``:1232``): negative-binomial total counts and beta-binomial
  allele counts, each a two-component mixture with a 1 % heavy-dispersion component, major / minor reordering.  It
  consumes the global ``numpy.random`` stream exactly as the reference does, so given the same genome mixture and
  seed it returns the reference's counts bit for bit (``tests/test_simulate.py``).  The reference draws the negative
  binomials one Python call per segment (``:1193-1197``); numpy's legacy generators fill arrays element by element
  with the same scalar routine, so the vectorised calls here visit the stream identically.

* :func:`sample_genome_mixture_fast` replaces ``RearrangementHistorySampler`` + ``GenomeCollectionSampler`` +
  ``GenomeMixtureSampler`` (``:628-1127``).  The reference evolves a swarm of 100 explicit rearranged genomes per event
  and re-scores each (O(events x swarm x N) Python; ~30 s at N = 1 k, hours at 30 k) to steer ploidy, LOH
  and divergence to targets.  The fits only consume what comes out at the end -- per-clone allele-specific copy number
  on a segmentation, wild-type adjacencies, and breakpoints at copy-number changes -- so this generator writes that
  down directly: segmental gains / losses on one allele of an ancestral tumour clone, a few more on each descendant,
  one breakpoint per event joining the segment ends the event makes adjacent, false breakpoints and a detection rate
  as in ``GenomeMixtureSampler`` (``:1066-1125``).  It is NOT stream-compatible with the reference (different
  algorithm); it is the input generator of the aggregate-throughput workloads (BASELINE config 5: 64 genomes).
"""
import numpy as np

from .pipeline import ArrayExperiment

CHROMOSOME_LENGTHS_HG19 = {
    '1': 249250621, '2': 243199373, '3': 198022430, '4': 191154276, '5': 180915260, '6': 171115067, '7': 159138663,
    '8': 146364022, '9': 141213431, '10': 135534747, '11': 135006516, '12': 133851895, '13': 115169878,
    '14': 107349540, '15': 102531392, '16': 90354753, '17': 81195210, '18': 78077248, '19': 59128983, '20': 63025520,
    '21': 48129895, '22': 51304566, 'X': 155270560,
}

_ALLELE_MEASUREMENT = np.array([[1, 0, 1], [0, 1, 1]])       # likelihood.py:68


class GenomeMixture(object):
    """What ``sample_experiment`` reads of the reference's ``GenomeMixture`` (``simulations/experiment.py:965-1033``)."""

    def __init__(self, l, cn, frac, adjacencies, breakpoints, chromosome, breakpoint_cn=None):
        self.l = np.asarray(l, dtype=float)
        self.cn = np.asarray(cn)
        self.frac = np.asarray(frac, dtype=float)
        self.N = self.l.shape[0]
        self.M = self.cn.shape[1]
        self.adjacencies = adjacencies
        self.detected_breakpoints = breakpoints
        self.segment_chromosome_id = np.asarray(chromosome)
        self.breakpoint_cn = breakpoint_cn or {}


def expected_read_count(l, cn, h, phi):
    """Expected major, minor and total read counts (``likelihood.py:102-134``)."""
    p = np.vstack([phi, phi, np.ones(phi.shape)]).T
    gamma = np.sum(cn * np.vstack([h, h]).T, axis=-2)
    x1 = np.dot(_ALLELE_MEASUREMENT.T, gamma.T).T
    x3 = ((x1 * p).T * l.T).T
    x3 += 1e-16
    if np.any(x3 <= 0) or np.any(np.isnan(x3)):
        raise ValueError('mu <= 0 or nan in expected_read_count')
    return x3


def _sample_negbin(mu, r):
    mu += 1e-16                                     # in place, as simulations/experiment.py:1194
    return np.random.negative_binomial(r, r / (r + mu)).reshape(mu.shape)


def _sample_negbin_mix(mu, r_0, r_1, mix):
    x_0 = _sample_negbin(mu, r_0)
    x_1 = _sample_negbin(mu, r_1)
    is_0 = np.random.random(size=x_0.shape) > mix
    return np.where(is_0, x_0, x_1), is_0


def _sample_betabin(n, p, M):
    return np.random.binomial(n, np.random.beta(M * p, M * (1 - p)))


def _sample_betabin_mix(n, p, M_0, M_1, mix):
    x_0 = _sample_betabin(n, p, M_0)
    x_1 = _sample_betabin(n, p, M_1)
    is_0 = np.random.random(size=x_0.shape) > mix
    return np.where(is_0, x_0, x_1), is_0


def sample_experiment(mixture, params=None):
    """Read counts of a sequencing experiment on ``mixture`` (``ExperimentSampler.sample_experiment``, emission model
    ``negbin_betabin``).  Returns ``dict(x (N,3) major/minor/total, h, phi, major_is_allele_a, is_outlier_total,
    is_outlier_allele)``."""
    params = params or {}
    h_total = params.get('h_total', 0.1)
    N, l, cn = mixture.N, mixture.l, mixture.cn
    h = mixture.frac * h_total
    phi = np.random.uniform(low=params.get('phi_min', 0.05), high=params.get('phi_max', 0.2), size=N)
    mu = expected_read_count(l, cn, h, phi)
    x = np.zeros(mu.shape)
    x_total, total_is_0 = _sample_negbin_mix(mu[:, 2] + 1e-16, params.get('negbin_r_0', 1000.),
                                             params.get('negbin_r_1', 10.), params.get('negbin_mix', 0.01))
    allele_total = (phi * x_total).astype(int)
    p_true = mu[:, 0] / (mu[:, 0:2].sum(axis=1) + 1e-16)
    x_allele_1, allele_is_0 = _sample_betabin_mix(allele_total, p_true, params.get('betabin_M_0', 2000.),
                                                  params.get('betabin_M_1', 10.), params.get('betabin_mix', 0.01))
    x[:, 2] = x_total
    x[:, 0] = x_allele_1
    x[:, 1] = allele_total - x_allele_1
    major_is_allele_a = x[:, 0] > x[:, 1]
    swap = ~major_is_allele_a
    x[swap, 0], x[swap, 1] = x[swap, 1].copy(), x[swap, 0].copy()
    return dict(x=x, h=h, phi=phi, major_is_allele_a=major_is_allele_a * 1, is_outlier_total=~total_is_0,
                is_outlier_allele=~allele_is_0)


def sample_random_breakpoints(N, num_breakpoints, adjacencies, excluded_breakpoints=None):
    """False-positive breakpoints (``simulations/experiment.py:1035-1063``), same draws in the same order."""
    breakpoints = set()
    while len(breakpoints) < num_breakpoints:
        n_1 = np.random.randint(N)
        n_2 = np.random.randint(N)
        side_1 = np.random.randint(2)
        side_2 = np.random.randint(2)
        if (n_1, n_2) in adjacencies and side_1 == 1 and side_2 == 0:
            continue
        if (n_2, n_1) in adjacencies and side_2 == 1 and side_1 == 0:
            continue
        if (n_1, side_1) == (n_2, side_2):
            continue
        breakpoint = frozenset([(n_1, side_1), (n_2, side_2)])
        if excluded_breakpoints is not None and breakpoint in excluded_breakpoints:
            continue
        breakpoints.add(breakpoint)
    return breakpoints


def sample_genome_mixture_fast(params, rng=None):
    """A clonal genome mixture of ``params['N']`` segments over ``params['chromosome_lengths']`` without simulating a
    rearrangement history (see the module docstring).  ``M`` = 3 (normal + two tumour clones that share
    ``num_ancestral_events`` events and differ by ``num_descendent_events`` each, per-allele clone difference at most
    one, as the model's state space assumes).  Draws from ``rng`` (a ``numpy.random.RandomState``; default: the global
    stream).  Returns a :class:`GenomeMixture`."""
    rs = rng if rng is not None else np.random
    N = int(params.get('N', 1000))
    lengths = params.get('chromosome_lengths', CHROMOSOME_LENGTHS_HG19)
    chroms = list(lengths.keys())
    total = float(sum(lengths.values()))
    # segments per chromosome in proportion to its length (at least 2), random segment lengths within
    per = np.maximum(2, np.round(np.array([lengths[c] for c in chroms]) / total * N).astype(int))
    while per.sum() > N:
        per[np.argmax(per)] -= 1
    while per.sum() < N:
        per[np.argmax([lengths[c] / per[k] for k, c in enumerate(chroms)])] += 1
    l, chromosome, chrom_start = [], [], []
    for k, c in enumerate(chroms):
        w = rs.dirichlet(np.full(per[k], 5.))
        chrom_start.append(len(l))
        l.extend(list(np.maximum(1., np.floor(w * lengths[c]))))
        chromosome.extend([c] * per[k])
    l = np.array(l)
    adjacencies = set()
    for k in range(len(chroms)):
        for n in range(chrom_start[k], chrom_start[k] + per[k] - 1):
            adjacencies.add((n, n + 1))

    M = 3
    max_cn = int(params.get('max_copy_number', 6))
    cn = np.ones((N, M, 2), dtype=int)
    breakpoint_cn = {}

    def event(clones, ancestral):
        # a gain or loss of a run of segments on one allele of the given tumour clones, and the breakpoint it makes
        for _ in range(200):
            k = rs.randint(len(chroms))
            span = max(1, int(rs.exponential(per[k] / 6.)))
            a = chrom_start[k] + rs.randint(per[k])
            b = min(a + span, chrom_start[k] + per[k] - 1)
            allele = rs.randint(2)
            gain = rs.random_sample() < params.get('proportion_gain', 0.55)
            new = cn[a:b + 1][:, clones, :].copy()
            new[:, :, allele] += 1 if gain else -1
            full = cn[a:b + 1].copy()
            full[:, clones, :] = new
            tum = full[:, 1:, :]
            if (new < 0).any() or (tum.sum(axis=2) > max_cn).any() or (np.abs(tum[:, 0, :] - tum[:, 1, :]) > 1).any():
                continue
            if gain:      # tandem duplication: right end of b joins left end of a
                bp = frozenset([(b, 1), (a, 0)]) if a != b or True else None
            else:         # deletion: right end of a - 1 joins left end of b + 1 (inside the chromosome only)
                if a == chrom_start[k] or b == chrom_start[k] + per[k] - 1:
                    continue
                bp = frozenset([(a - 1, 1), (b + 1, 0)])
            if bp in breakpoint_cn or len(bp) < 2:
                continue
            cn[a:b + 1] = full
            brk = np.zeros(M, dtype=int)
            brk[clones] = 1
            breakpoint_cn[bp] = brk
            return
        raise ValueError('unable to place a copy number event')

    for _ in range(int(params.get('num_ancestral_events', 25))):
        event([1, 2], True)
    for m in (1, 2):
        for _ in range(int(params.get('num_descendent_events', 10))):
            event([m], False)

    frac_normal = params.get('frac_normal', 0.4)
    frac_clone_1 = params.get('frac_clone_1', None)
    if frac_clone_1 is None:
        frac = np.concatenate([[frac_normal], rs.dirichlet([params.get('frac_clone_concentration', 1.)] * 2) * (1 - frac_normal)])
    else:
        frac = np.array([frac_normal, frac_clone_1, 1. - frac_normal - frac_clone_1])

    true_breakpoints = list(breakpoint_cn.keys())
    detected = list(true_breakpoints)
    rs.shuffle(detected)
    detected = detected[:int(params.get('proportion_breakpoints_detected', 0.9) * len(detected))]
    state = None
    if rng is not None:         # sample_random_breakpoints draws from the global stream
        state = np.random.get_state()
        np.random.seed(rs.randint(2 ** 31 - 1))
    false = sample_random_breakpoints(N, int(params.get('num_false_breakpoints', 50)), adjacencies,
                                      excluded_breakpoints=set(true_breakpoints))
    if state is not None:
        np.random.set_state(state)
    detected.extend(sorted(false, key=lambda b: sorted(b)))
    return GenomeMixture(l, cn, frac, adjacencies, dict(enumerate(detected)), np.array(chromosome), breakpoint_cn)


def simulate_experiment(params, seed=None):
    """A complete simulated experiment in the form the fits take (:class:`pipeline.ArrayExperiment`), with the truth
    attached (``true_cn``, ``true_h``, ``true_brk_cn``): :func:`sample_genome_mixture_fast` then
    :func:`sample_experiment`, both from ``RandomState(seed)`` (``params['random_seed']`` by default, the seed style of
    ``benchmark/sim_defs.yaml``)."""
    seed = params.get('random_seed', 999) if seed is None else seed
    rs = np.random.RandomState(seed)
    mixture = sample_genome_mixture_fast(params, rng=rs)
    state = np.random.get_state()
    np.random.seed(rs.randint(2 ** 31 - 1))
    try:
        counts = sample_experiment(mixture, params)
    finally:
        np.random.set_state(state)
    ids = sorted(mixture.detected_breakpoints.keys())
    bps = np.array([[v for end in sorted(mixture.detected_breakpoints[i]) for v in end] for i in ids],
                   dtype=np.int64).reshape(-1, 4)
    exp = ArrayExperiment(counts['x'], mixture.l, np.array(sorted(mixture.adjacencies), dtype=np.int64), bps,
                          breakpoint_ids=[str(i) for i in ids], chromosome=mixture.segment_chromosome_id,
                          major_is_allele_a=counts['major_is_allele_a'])
    exp.true_cn = mixture.cn
    exp.true_h = counts['h']
    exp.true_brk_cn = dict((str(i), mixture.breakpoint_cn.get(mixture.detected_breakpoints[i], np.zeros(mixture.M, dtype=int)))
                           for i in ids)
    return exp
