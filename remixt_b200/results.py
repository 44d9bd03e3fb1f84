"""Result tables and the on-disk layout of a collated fit: the mirror of ``remixt/analysis/experiment.py``
(``create_cn_table`` ``:354``, ``create_brk_cn_table`` ``:397``), of the store layout of ``analysis/pipeline.py``
(``store_fit_results`` ``:231-250``, ``store_optimal_solution`` ``:253-264``, ``collate`` ``:267-293``) and of
``ui/write_results.py`` (``write_results_tables`` ``:8-49``).

Keys and columns are the reference's: ``stats``; ``solutions/solution_{id}/{h,cn,mix,brk_cn}``; ``read_depth`` and
``minor_modes`` from the init step; and ``/cn``, ``/mix``, ``/brk_cn`` of the selected solution.  The reference writes
a PyTables HDF5 file through ``pandas.HDFStore``; :func:`open_store` returns exactly that when PyTables is
importable, and otherwise a :class:`TableStore` with the same mapping interface that keeps one parquet file per key
under a directory (pyarrow), so that the rest of the code does not care which of the two it holds.
"""
import json
import os

import numpy as np
import pandas as pd

from . import readdepth
from .pipeline import get_param


# ---------------------------------------------------------------------------------------------------------------
# tables
# ---------------------------------------------------------------------------------------------------------------
def create_cn_table(experiment, cn, h, phi=None):
    """Per-segment table of observed depths, MAP copy number and the depths it implies (``experiment.py:354-394``)."""
    data = readdepth.create_segment_table(experiment)
    for m in range(0, cn.shape[1]):
        data['major_{0}'.format(m)] = cn[:, m, 0]
        data['minor_{0}'.format(m)] = cn[:, m, 1]
    data['major_raw'] = (data['major_depth'] - data['major_0'] * h[0]) / h[1:].sum()
    data['minor_raw'] = (data['minor_depth'] - data['minor_0'] * h[0]) / h[1:].sum()
    data['major_depth_e'] = (cn[:, :, 0] * h[np.newaxis, :]).sum(axis=-1)
    data['minor_depth_e'] = (cn[:, :, 1] * h[np.newaxis, :]).sum(axis=-1)
    data['total_depth_e'] = (cn.sum(axis=-1) * h[np.newaxis, :]).sum(axis=-1)
    data['major_e'] = data['major_depth_e'] * experiment.l
    data['minor_e'] = data['minor_depth_e'] * experiment.l
    data['total_e'] = data['total_depth_e'] * experiment.l
    data['major_raw_e'] = (data['major_depth_e'] - data['major_0'] * h[0]) / h[1:].sum()
    data['minor_raw_e'] = (data['minor_depth_e'] - data['minor_0'] * h[0]) / h[1:].sum()
    if 'major_2' in data:
        data['major_diff'] = np.absolute(data['major_1'] - data['major_2'])
        data['minor_diff'] = np.absolute(data['minor_1'] - data['minor_2'])
    return data


def breakpoint_segment_table(experiment):
    """``breakpoint_segment_data`` of an experiment: the reference ``Experiment`` carries it
    (``experiment.py:281-282``); for array experiments it is rebuilt from the breakpoint dict with the columns
    ``convert_breakpoints_to_dict`` reads (``:219-225``)."""
    table = getattr(experiment, 'breakpoint_segment_data', None)
    if table is not None:
        return table
    rows = []
    for prediction_id, bp in experiment.breakpoints.items():
        (n_1, side_1), (n_2, side_2) = sorted(bp)
        rows.append(dict(prediction_id=prediction_id, n_1=n_1, side_1=side_1, n_2=n_2, side_2=side_2))
    return pd.DataFrame(rows, columns=['prediction_id', 'n_1', 'side_1', 'n_2', 'side_2'])


def create_brk_cn_table(brk_cn, breakpoint_segment_data):
    """Per-breakpoint copy number table (``experiment.py:397-424``)."""
    if len(brk_cn) == 0:
        return pd.DataFrame(columns=['prediction_id'])
    table = pd.DataFrame(brk_cn.values(), index=brk_cn.keys())
    table.columns = ['cn_{}'.format(m) for m in table.columns]
    table.index.name = 'prediction_id'
    table = table.reset_index()
    return table.merge(breakpoint_segment_data, on='prediction_id').fillna(0.)


# ---------------------------------------------------------------------------------------------------------------
# store
# ---------------------------------------------------------------------------------------------------------------
class TableStore(object):
    """``pandas.HDFStore``'s mapping interface (``store[key] = frame_or_series``, ``store[key]``, ``keys()``,
    context manager) over a directory of parquet files, one per key."""

    def __init__(self, path, mode='r'):
        self.path = path
        if mode == 'w':
            os.makedirs(path, exist_ok=True)
            for f in os.listdir(path):
                if f.endswith('.parquet') or f == 'index.json':
                    os.remove(os.path.join(path, f))
            self.index = {}
        else:
            with open(os.path.join(path, 'index.json')) as f:
                self.index = json.load(f)

    @staticmethod
    def _norm(key):
        return '/' + key.strip('/')

    def _file(self, key):
        return os.path.join(self.path, self._norm(key).strip('/').replace('/', '__') + '.parquet')

    def __setitem__(self, key, value):
        key = self._norm(key)
        if isinstance(value, pd.Series):
            frame = value.to_frame(name='value')
            kind = 'series'
        else:
            frame, kind = value, 'frame'
        frame = frame.copy()
        frame.columns = [str(c) for c in frame.columns]
        frame.to_parquet(self._file(key), index=True)
        self.index[key] = kind
        with open(os.path.join(self.path, 'index.json'), 'w') as f:
            json.dump(self.index, f)

    def __getitem__(self, key):
        key = self._norm(key)
        if key not in self.index:
            raise KeyError(key)
        frame = pd.read_parquet(self._file(key))
        return frame['value'] if self.index[key] == 'series' else frame

    def __contains__(self, key):
        return self._norm(key) in self.index

    def keys(self):
        return list(self.index.keys())

    def items(self):
        return [(k, self[k]) for k in self.keys()]

    def close(self):
        pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


def open_store(path, mode='r', prefer_hdf5=True):
    """The collated-results store: ``pandas.HDFStore`` (the reference's format) when PyTables is available and
    ``path`` is not a directory, else a :class:`TableStore`."""
    if prefer_hdf5 and not os.path.isdir(path):
        try:
            import tables  # noqa: F401
            return pd.HDFStore(path, mode)
        except ImportError:
            pass
    return TableStore(path, mode)


def store_fit_results(store, experiment, fit_results, key_prefix):
    """One solution under ``key_prefix`` (``analysis/pipeline.py:231-250``)."""
    h = fit_results['h']
    cn_table = create_cn_table(experiment, fit_results['cn'], h)
    cn_table['prob_is_outlier_total'] = fit_results['p_outlier_total'][:, 1]
    cn_table['prob_is_outlier_allele'] = fit_results['p_outlier_allele'][:, 1]
    cn_table['total_likelihood_mask'] = fit_results['total_likelihood_mask']
    cn_table['allele_likelihood_mask'] = fit_results['allele_likelihood_mask']
    brk_cn_table = create_brk_cn_table(fit_results['brk_cn'], breakpoint_segment_table(experiment))
    store[key_prefix + '/h'] = pd.Series(h, index=range(len(h)))
    store[key_prefix + '/cn'] = cn_table
    store[key_prefix + '/mix'] = pd.Series(h / h.sum(), index=range(len(h)))
    store[key_prefix + '/brk_cn'] = brk_cn_table


def optimal_solution(stats, config):
    """``init_id`` of the solution ``store_optimal_solution`` copies to the top level (``analysis/pipeline.py:253-259``):
    the best ELBO among the solutions with ``proportion_divergent < max_prop_diverge`` (all if none qualifies)."""
    max_prop_diverge = get_param(config, 'max_prop_diverge')
    if (stats['proportion_divergent'] < max_prop_diverge).any():
        stats = stats[stats['proportion_divergent'] < max_prop_diverge].copy()
    stats = stats.sort_values('elbo', ascending=False, kind='stable')
    return stats.loc[stats.index[0], 'init_id']


def store_optimal_solution(stats, store, config):
    solution_idx = optimal_solution(stats, config)
    key_prefix = '/solutions/solution_{}'.format(solution_idx)
    store['/cn'] = store[key_prefix + '/cn']
    store['/mix'] = store[key_prefix + '/mix']
    store['/brk_cn'] = store[key_prefix + '/brk_cn']
    return solution_idx


def collate(store, experiment, fit_results_by_init, config, init_store=None):
    """The collated results of a grid of fits (``analysis/pipeline.py:267-293``): the ``stats`` table, what the init
    step stored, every solution, and the selected one at the top level.  ``fit_results_by_init`` is the ``results``
    dict of ``pipeline.fit_grid_batched``; returns the selected ``init_id``."""
    rows = []
    for init_id, results in fit_results_by_init.items():
        row = dict(results['stats'])
        row['init_id'] = init_id
        rows.append(row)
    stats_table = pd.DataFrame(rows)
    store['stats'] = stats_table
    if init_store is not None:
        for key, value in init_store.items():
            store[key] = value
    for init_id, results in fit_results_by_init.items():
        store_fit_results(store, experiment, results, 'solutions/solution_{0}'.format(init_id))
    return store_optimal_solution(stats_table, store, config)


def write_results_tables(store, cn_filename, brk_cn_filename, meta_filename, max_proportion_divergent=0.5,
                         max_ploidy=None, min_ploidy=None):
    """The user-facing outputs of ``ui/write_results.py:8-49``: the copy-number and breakpoint tables of the best
    solution that passes the filters (tab separated) and its statistics as YAML."""
    import yaml
    stats = store['stats']
    stats = stats[stats['proportion_divergent'] <= max_proportion_divergent]
    if max_ploidy is not None:
        stats = stats[stats['ploidy'] < max_ploidy]
    if min_ploidy is not None:
        stats = stats[stats['ploidy'] > min_ploidy]
    if stats.empty:
        raise ValueError('filters to restrictive, no solutions')
    best = stats.sort_values('elbo', ascending=False).iloc[0]
    solution = best['init_id']
    prefix = 'solutions/solution_{0}'.format(solution)
    store[prefix + '/cn'].to_csv(cn_filename, sep='\t', index=False)
    store[prefix + '/brk_cn'].to_csv(brk_cn_filename, sep='\t', index=False)
    metadata = dict()
    for key, value in best.items():
        metadata[key] = value.item() if hasattr(value, 'item') else value
    metadata['h'] = list(store[prefix + '/h'].tolist())
    metadata['mix'] = list(store[prefix + '/mix'].tolist())
    with open(meta_filename, 'w') as f:
        yaml.dump(metadata, f, default_flow_style=False)
    return solution
