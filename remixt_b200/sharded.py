"""A ``RemixtModel`` whose chromosomes are sharded over several models / GPUs.

Chromosome chains are exactly independent given the frame log-probabilities and ``p_breakpoint``
(telomere rows of ``log_transmat`` are zero, ``bpmodel.pyx:649-650``), so a fit over a large genome
splits into contiguous chromosome groups, one :class:`remixt_b200.bpmodel.RemixtModel` per group
(constructed with ``shard=True``: all K breakpoints, the breakends of its own chromosomes).  What
couples the shards is small and explicit:

  * ``update_p_breakpoint`` (``bpmodel.pyx:964-985``): the ``(K, B)`` ``log q`` is a sum over breakend
    rows; every shard contributes the rows it owns (``breakpoint_logq``), the partial sums are added
    (one allreduce of ``K * B`` doubles) and every shard normalises the same total
    (``apply_breakpoint_logq``);
  * scalars: ELBO terms, ``hmm_log_norm_const``, the M-step objective and its ``M``-vector gradient are
    sums over segments / chains, reduced the same way.  The entropy of ``p_breakpoint`` is a global term
    every shard would count; it is taken out of the shards' entropies and added back once;
  * whole-genome arrays the host driver reads (posteriors, outlier factors, MAP copy number) are
    gathered in segment order.

The class keeps the reference's constructor signature and method names, so
``cn_model.BreakpointModel(..., model_class=ShardedRemixtModel, comm=...)`` drives it unchanged: every
rank runs the same host code (same RNG stream, same M-step decisions) on its own shards.

``comm`` moves the reductions: :class:`LocalComm` (one process; several shards on one GPU),
:class:`TorchDistComm` (``torch.distributed``, operands staged by the host: gloo in the CPU tests) or
:class:`NcclLibComm` (one shard per GPU, ``ncclAllReduce`` on the model's own stream inside the library).
"""
import numpy as np


class LocalComm(object):
    """One process owns every shard."""
    rank = 0
    size = 1

    def allreduce_sum(self, a):
        return np.asarray(a, dtype=np.float64)

    def allgather(self, obj):
        return [obj]


class TorchDistComm(object):
    """Reductions over ``torch.distributed`` (process group already initialised).  With the NCCL backend the
    operands are staged through this rank's GPU: they are ``K * B`` doubles and a handful of scalars."""

    def __init__(self, device=None, group=None):
        import torch
        import torch.distributed as dist
        self._torch, self._dist, self._group = torch, dist, group
        self.rank = dist.get_rank(group)
        self.size = dist.get_world_size(group)
        self._cuda = dist.get_backend(group) == 'nccl'
        self._device = torch.device('cuda', torch.cuda.current_device() if device is None else int(device)) \
            if self._cuda else torch.device('cpu')

    def allreduce_sum(self, a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        t = self._torch.from_numpy(a.copy()).to(self._device)
        self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self._group)
        return t.cpu().numpy().reshape(a.shape)

    def allgather(self, obj):
        out = [None] * self.size
        self._dist.all_gather_object(out, obj, group=self._group)
        return out


class NcclLibComm(TorchDistComm):
    """One shard per rank with the reductions INSIDE ``libremixt_b200.so`` (``rmx_comm_init``): the ``(K, B)``
    ``log q`` allreduce of ``update_p_breakpoint`` and the scalar reductions run as ``ncclAllReduce`` on the model's
    CUDA stream, so a round of the five updates (or ``variational_update(n)``, fused) needs no host round trip.
    ``torch.distributed`` only carries the communicator's unique id and the occasional gather of whole-genome arrays
    for the host driver."""
    in_library = True

    def attach(self, model):
        from remixt_b200 import bpmodel
        ids = [bpmodel.nccl_unique_id() if self.rank == 0 else None]
        self._dist.broadcast_object_list(ids, src=0, group=self._group)
        model.comm_init(ids[0], self.rank, self.size)


def plan_shards(is_telomere, num_shards, breakpoint_idx=None):
    """Contiguous chromosome groups balanced by segment count: ``[(start, end), ...]`` (``end`` exclusive).

    A chain ends at every ``is_telomere`` segment and at the last segment.  A group boundary is a telomere
    row, whose transition is identically zero; it must not also be a breakend row (its joint would be
    needed across the boundary)."""
    is_telomere = np.asarray(is_telomere)
    N = is_telomere.shape[0]
    ends = [n + 1 for n in range(N - 1) if is_telomere[n] > 0] + [N]
    if num_shards < 1 or num_shards > len(ends):
        raise ValueError('cannot split %d chromosome chains into %d shards' % (len(ends), num_shards))
    bounds, start, k = [], 0, 0
    for c, e in enumerate(ends):
        remaining_chains = len(ends) - (c + 1)
        remaining_shards = num_shards - (k + 1)
        target = N * (k + 1) / float(num_shards)
        # close the group once it reaches its share, or when every remaining chain is needed by a remaining shard
        if k < num_shards - 1 and (e >= target or remaining_chains == remaining_shards):
            bounds.append((start, e))
            start, k = e, k + 1
    bounds.append((start, N))
    assert len(bounds) == num_shards and all(b > a for a, b in bounds)
    if breakpoint_idx is not None:
        for a, b in bounds[:-1]:
            if np.asarray(breakpoint_idx)[b - 1] >= 0:
                raise ValueError('segment %d ends a chromosome group and carries a breakend; the chromosome-sharded '
                                 'fit needs breakends inside chromosomes' % (b - 1))
    return bounds


def _xlogx_sum(p):
    """``_entropy`` of ``bpmodel.pyx:110-115``: sum of x log x over x > 0."""
    p = np.asarray(p, dtype=np.float64).ravel()
    pos = p > 0.
    return float(np.sum(p[pos] * np.log(p[pos])))


_SCALARS = ('negbin_r_0', 'negbin_r_1', 'negbin_hdel_mu', 'negbin_hdel_r_0', 'negbin_hdel_r_1',
            'betabin_M_0', 'betabin_M_1', 'betabin_loh_p', 'betabin_loh_M_0', 'betabin_loh_M_1',
            'prior_outlier_total', 'prior_outlier_allele', 'transition_model', 'divergence_weight',
            'transition_penalty')
_SEGMENT_ARRAYS = ('posterior_marginals', 'framelogprob', 'p_allele_swap', 'p_outlier_total', 'p_outlier_allele',
                   'total_likelihood_mask', 'allele_likelihood_mask', 'is_hdel', 'is_loh')


class ShardedRemixtModel(object):
    """``RemixtModel`` (``bpmodel.pyx:397``) over chromosome shards; see the module docstring."""

    def __init__(self, num_clones, num_segments, num_breakpoints, normal_contamination,
                 cn_states, brk_states, h_init, l, x, y, is_telomere, breakpoint_idx,
                 breakpoint_orient, transition_penalty, divergence_weight,
                 comm=None, num_shards=None, model_class=None, local_kwargs=None):
        d = self.__dict__
        d['comm'] = comm if comm is not None else LocalComm()
        if model_class is None:
            from remixt_b200 import bpmodel
            model_class = bpmodel.RemixtModel
        if local_kwargs is None:
            local_kwargs = dict(shard=True)
        num_shards = self.comm.size if num_shards is None else int(num_shards)
        if num_shards < self.comm.size:
            raise ValueError('every rank needs a shard: num_shards %d < %d ranks' % (num_shards, self.comm.size))
        cn_states = np.asarray(cn_states)
        d['num_clones'], d['num_segments'], d['num_breakpoints'] = int(num_clones), int(num_segments), int(num_breakpoints)
        d['normal_contamination'] = bool(normal_contamination)
        d['num_alleles'] = 2
        d['cn_states'], d['brk_states'] = cn_states, np.asarray(brk_states)
        d['num_cn_states'], d['num_brk_states'] = cn_states.shape[1], d['brk_states'].shape[0]
        d['is_telomere'] = np.asarray(is_telomere)
        d['breakpoint_idx'], d['breakpoint_orient'] = np.asarray(breakpoint_idx), np.asarray(breakpoint_orient)
        if d['breakpoint_idx'].shape[0] != num_segments or d['breakpoint_idx'].max() + 1 != num_breakpoints:
            raise ValueError('breakpoint_idx must have maximum of num_breakpoints positive indices')
        d['bounds'] = plan_shards(is_telomere, num_shards, breakpoint_idx)
        d['shard_ids'] = [k for k in range(num_shards) if k % self.comm.size == self.comm.rank]
        d['shards'] = []
        l, x, y = np.asarray(l, dtype=np.float64), np.asarray(x, dtype=np.float64), np.asarray(y, dtype=np.float64)
        for k in self.shard_ids:
            a, b = self.bounds[k]
            self.shards.append(model_class(
                num_clones, b - a, num_breakpoints, normal_contamination,
                np.ascontiguousarray(cn_states[a:b]), brk_states, h_init,
                np.ascontiguousarray(l[a:b]), np.ascontiguousarray(x[a:b]), np.ascontiguousarray(y[a:b]),
                np.ascontiguousarray(self.is_telomere[a:b]), np.ascontiguousarray(self.breakpoint_idx[a:b]),
                np.ascontiguousarray(self.breakpoint_orient[a:b]), transition_penalty, divergence_weight,
                **local_kwargs))
        d['in_library'] = bool(getattr(self.comm, 'in_library', False))
        if self.in_library:
            if len(self.shards) != 1 or num_shards != self.comm.size:
                raise ValueError('in-library reductions need exactly one shard per rank')
            self.comm.attach(self.shards[0])

    # ------------------------------------------------------------------ plumbing
    def _local(self):
        return zip(self.shard_ids, self.shards)

    def _gather_segments(self, parts):
        """``parts``: this rank's per-shard arrays (leading dimension = the shard's segments)."""
        got = {}
        for piece in self.comm.allgather(dict(zip(self.shard_ids, parts))):
            got.update(piece)
        return np.concatenate([got[k] for k in range(len(self.bounds))], axis=0)

    def __getattr__(self, name):
        if name in _SCALARS or name in ('h', 'p_breakpoint', 'cn_max'):
            return getattr(self.shards[0], name)
        if name in _SEGMENT_ARRAYS:
            return self._gather_segments([np.asarray(getattr(m, name)) for m in self.shards])
        if name == 'hmm_log_norm_const' and self.__dict__.get('in_library'):
            return float(self.shards[0].hmm_log_norm_const)
        if name == 'hmm_log_norm_const':
            return float(self.comm.allreduce_sum(np.array([sum(float(m.hmm_log_norm_const) for m in self.shards)]))[0])
        if name in ('l', 'x', 'y'):
            return self._gather_segments([np.asarray(getattr(m, name)) for m in self.shards])
        raise AttributeError(name)

    def __setattr__(self, name, value):
        if name in _SCALARS or name in ('h', 'p_breakpoint'):
            for m in self.shards:
                setattr(m, name, value)
        elif name in ('total_likelihood_mask', 'allele_likelihood_mask'):
            value = np.asarray(value)
            if value.shape[0] != self.num_segments:
                raise ValueError('%s must have num_segments entries' % name)
            for k, m in self._local():
                a, b = self.bounds[k]
                setattr(m, name, np.ascontiguousarray(value[a:b]))
        else:
            self.__dict__[name] = value

    # ------------------------------------------------------------------ coordinate updates (cn_model.py:444-460)
    def update_p_allele_swap(self):
        for m in self.shards:
            m.update_p_allele_swap()

    def update_p_cn(self):
        for m in self.shards:
            m.update_p_cn()

    def update_p_breakpoint(self):
        """``bpmodel.pyx:964-985``: partial sums per shard, one allreduce of ``(K, B)`` doubles, the same
        normalisation everywhere."""
        if self.in_library:
            self.shards[0].update_p_breakpoint()        # partial sums -> ncclAllReduce on the stream -> normalise
            return
        partial = np.zeros((self.num_breakpoints, self.num_brk_states))
        for m in self.shards:
            partial += m.breakpoint_logq()
        total = self.comm.allreduce_sum(partial)
        for m in self.shards:
            m.apply_breakpoint_logq(total)

    def update_p_outlier_total(self):
        for m in self.shards:
            m.update_p_outlier_total()

    def update_p_outlier_allele(self):
        for m in self.shards:
            m.update_p_outlier_allele()

    def update_framelogprob(self):
        for m in self.shards:
            m.update_framelogprob()

    def variational_update(self, n_iter=1):
        if self.in_library:
            self.shards[0].variational_update(n_iter)   # the fused loop, collectives in stream
            return
        for _ in range(int(n_iter)):
            self.update_p_allele_swap()
            self.update_p_cn()
            self.update_p_breakpoint()
            self.update_p_outlier_total()
            self.update_p_outlier_allele()

    # ------------------------------------------------------------------ objective
    def _energy_entropy(self):
        if self.in_library:
            m = self.shards[0]
            return float(m.calculate_variational_energy()), float(m.calculate_variational_entropy())
        hb = _xlogx_sum(self.shards[0].p_breakpoint)     # the one global entropy term, identical on every shard
        loc = np.zeros(2)
        for m in self.shards:
            loc[0] += m.calculate_variational_energy()
            loc[1] += m.calculate_variational_entropy() - hb
        tot = self.comm.allreduce_sum(loc)
        return float(tot[0]), float(tot[1]) + hb

    def calculate_variational_energy(self):
        return self._energy_entropy()[0]

    def calculate_variational_entropy(self):
        return self._energy_entropy()[1]

    def calculate_elbo(self):
        energy, entropy = self._energy_entropy()
        return energy - entropy

    def calculate_expected_log_likelihood(self, sample):
        sample = np.asarray(sample)
        if self.in_library:
            a, b = self.bounds[self.shard_ids[0]]
            return float(self.shards[0].calculate_expected_log_likelihood(np.ascontiguousarray(sample[a:b])))
        loc = 0.
        for k, m in self._local():
            a, b = self.bounds[k]
            loc += m.calculate_expected_log_likelihood(np.ascontiguousarray(sample[a:b]))
        return float(self.comm.allreduce_sum(np.array([loc]))[0])

    def calculate_expected_log_likelihood_partial_h(self, sample, partial_h):
        sample = np.asarray(sample)
        if self.in_library:
            a, b = self.bounds[self.shard_ids[0]]
            self.shards[0].calculate_expected_log_likelihood_partial_h(np.ascontiguousarray(sample[a:b]), partial_h)
            return
        loc = np.zeros(self.num_clones)
        g = np.zeros(self.num_clones)
        for k, m in self._local():
            a, b = self.bounds[k]
            m.calculate_expected_log_likelihood_partial_h(np.ascontiguousarray(sample[a:b]), g)
            loc += g
        partial_h[:] = self.comm.allreduce_sum(loc)

    # ------------------------------------------------------------------ MAP decode (bpmodel.pyx:1197-1210)
    def infer_cn(self, cn):
        parts = []
        for k, m in self._local():
            a, b = self.bounds[k]
            c = np.zeros((b - a, self.num_clones, self.num_alleles), dtype=np.int64)
            m.infer_cn(c)
            parts.append(c)
        cn[:] = self._gather_segments(parts)
