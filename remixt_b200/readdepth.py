"""Initialisation candidates from read depth: the mirror of ``remixt/analysis/readdepth.py``.

These are the host steps every "full fit" starts with: modes of the minor-allele depth
(k-means, k = 5, on 10 000 length-weighted resamples), candidate ``(h_normal, h_tumour)`` pairs from those modes,
and the ploidy each candidate implies.  The arithmetic is a few vector operations over N segments and one tiny
k-means, so it stays on the host, with the same library calls in the same order as the reference
(``sklearn.cluster.KMeans`` seeded through the global ``numpy.random`` stream, ``readdepth.py:77``;
``numpy.random.multinomial`` under a state-restoring guard, ``utils.py:24-29``): given the same experiment and seed
the candidates equal the reference's bit for bit (``tests/test_init_results.py``).
"""
import numpy as np
import pandas as pd


def _attr(experiment, name, default):
    v = getattr(experiment, name, None)
    return default if v is None else v


def create_segment_table(experiment):
    """Per-segment table of counts, lengths and depths (``analysis/experiment.py:323-351``)."""
    n = len(experiment.l)
    l = np.asarray(experiment.l)
    x = np.asarray(experiment.x)
    data = pd.DataFrame({
        'chromosome': np.asarray(experiment.segment_chromosome_id),
        'start': _attr(experiment, 'segment_start', np.zeros(n, dtype=int)),
        'end': _attr(experiment, 'segment_end', np.maximum(l.astype(int) - 1, 0)),
        'major_is_allele_a': _attr(experiment, 'segment_major_is_allele_a', np.ones(n, dtype=int)),
        'length': l,
        'major_readcount': x[:, 0],
        'minor_readcount': x[:, 1],
        'readcount': x[:, 2],
    })
    data['allele_ratio'] = data['minor_readcount'] / (data['major_readcount'] + data['minor_readcount'])
    data['allele_ratio'] = data['allele_ratio'].fillna(0)
    data['major_depth'] = data['readcount'] * (1. - data['allele_ratio']) / data['length']
    data['minor_depth'] = data['readcount'] * data['allele_ratio'] / data['length']
    data['total_depth'] = data['readcount'] / data['length']
    return data


def estimate_phi(x):
    """Proportion of genotypable reads per segment (``likelihood.py:71-84``)."""
    return x[:, 0:2].sum(axis=1).astype(float) / (x[:, 2].astype(float) + 1.0)


def calculate_depth(experiment):
    """Major, minor and total depth per segment (``readdepth.py:12-58``); segments of zero length or without
    genotypable reads are dropped."""
    data = create_segment_table(experiment)
    data['segment_length'] = data['end'] - data['start'] + 1
    data['length_ratio'] = data['length'] / data['segment_length']
    data['allele_readcount'] = data['minor_readcount'] + data['major_readcount']
    data['high_quality'] = (
        (data['length'] > np.percentile(data['length'].values, 10)) &
        (data['allele_readcount'] > np.percentile(data['allele_readcount'].values, 10)) &
        (data['length_ratio'] > np.percentile(data['length_ratio'].values, 10)))
    phi = estimate_phi(np.asarray(experiment.x))
    p = np.vstack([phi, phi, np.ones(phi.shape)]).T          # likelihood.py:88-99
    data = data[(data['length'] > 0) & np.all(p > 0, axis=1)]
    data = data.rename(columns={'major_depth': 'major', 'minor_depth': 'minor', 'total_depth': 'total'})
    return data[['chromosome', 'start', 'end', 'length', 'major', 'minor', 'total', 'high_quality']]


def weighted_resample(data, weights, num_samples=10000):
    """``utils.py:24-29``: multinomial resampling that leaves the global random stream where it found it."""
    norm_weights = weights.astype(float) / float(weights.sum())
    state = np.random.get_state()
    counts = np.random.multinomial(num_samples, norm_weights)
    np.random.set_state(state)
    return np.repeat(data, counts)


def calculate_minor_modes(read_depth):
    """Modes of the minor-allele depth distribution (``readdepth.py:60-89``)."""
    import sklearn.cluster
    amp_rd = np.percentile(read_depth['minor'], 95)
    read_depth = read_depth[read_depth['minor'] < amp_rd]
    rd_samples = weighted_resample(read_depth['minor'].values, read_depth['length'].values)
    kmm = sklearn.cluster.KMeans(n_clusters=5)
    kmm.fit(rd_samples.reshape((rd_samples.size, 1)))
    means = kmm.cluster_centers_[:, 0]
    cluster_idx = kmm.predict(rd_samples.reshape((rd_samples.size, 1)))
    cluster_counts = np.bincount(cluster_idx)
    cluster_prop = cluster_counts.astype(float) / cluster_counts.sum()
    return means[cluster_prop >= 0.01]


def calculate_candidate_h_monoclonal(minor_modes, h_normal=None, h_tumour=None):
    """Candidate ``[h_normal, h_tumour]`` pairs (``readdepth.py:92-128``): the lowest mode is the normal
    contamination; every higher mode is one or two tumour copies above it."""
    if h_normal is None:
        h_normal = minor_modes.min()
    if h_tumour is not None:
        return np.array([[h_normal, h_tumour]])
    candidates = []
    for mode in minor_modes:
        if mode <= h_normal:
            continue
        tumour = mode - h_normal
        for scale in (1., 0.5):
            candidates.append(np.array([h_normal, tumour * scale]))
    return candidates


def estimate_ploidy(h, experiment, read_depth=None):
    """Ploidy a haploid-depth candidate implies (``readdepth.py:131-153``).  ``read_depth`` (extension) takes the
    table ``calculate_depth`` already produced instead of recomputing it per candidate."""
    read_depth = calculate_depth(experiment) if read_depth is None else read_depth.copy()
    read_depth['major_raw'] = (read_depth['major'] - h[0]) / h[1:].sum()
    read_depth['minor_raw'] = (read_depth['minor'] - h[0]) / h[1:].sum()
    major, minor, length = read_depth.replace(np.inf, np.nan).dropna()[['major_raw', 'minor_raw', 'length']].values.T
    return ((major + minor) * length).sum() / length.sum()
