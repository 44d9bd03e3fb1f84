"""rmx_max_product_noalloc: the Viterbi decode with its lattice in the library's file-scope device array.

The entry point has no counterpart in the reference (include/remixt_b200.h says why it exists).  It is checked here,
through the C-ABI, against the existing entry point rmx_max_product on the same inputs: same kernel, so the state
sequence is bit-exact and the log probability is held to the tight tolerance of tests/test_parity_gpu.py.  Two calls
in a row share the static array (the second one is smaller, so the array still holds cells of the first), and the
free device memory must not move across a call.  The file name sorts last on purpose: pytest -x stops at the first
failure, and a failure here must not hide another test.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL_TIGHT = 1e-7
MAX_CELLS = 1 << 21          # RMX_NOALLOC_MAX_CELLS of include/remixt_b200.h


class _Cuda(object):
    """The few CUDA runtime calls the TEST needs for the caller-owned buffers (the entry point itself allocates
    nothing): the library links libcudart.so.12 dynamically, so this is the runtime instance it uses."""

    def __init__(self):
        self.rt = C.CDLL('libcudart.so.12')
        self.rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
        self.rt.cudaFree.argtypes = [C.c_void_p]
        self.rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        self.rt.cudaMemGetInfo.argtypes = [C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]
        self.rt.cudaStreamCreate.argtypes = [C.POINTER(C.c_void_p)]
        self.rt.cudaStreamSynchronize.argtypes = [C.c_void_p]
        self.rt.cudaStreamDestroy.argtypes = [C.c_void_p]
        self.owned = []

    def ok(self, rc):
        assert rc == 0, 'CUDA runtime error %d' % rc

    def alloc(self, nbytes):
        p = C.c_void_p()
        self.ok(self.rt.cudaMalloc(C.byref(p), max(int(nbytes), 8)))
        self.owned.append(p)
        return p

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        p = self.alloc(arr.nbytes)
        if arr.nbytes:
            self.ok(self.rt.cudaMemcpy(p, arr.ctypes.data_as(C.c_void_p), arr.nbytes, 1))
        return p

    def download(self, p, shape, dtype):
        out = np.empty(shape, dtype=dtype)
        self.ok(self.rt.cudaMemcpy(out.ctypes.data_as(C.c_void_p), p, out.nbytes, 2))
        return out

    def free_bytes(self):
        free, total = C.c_size_t(), C.c_size_t()
        self.ok(self.rt.cudaMemGetInfo(C.byref(free), C.byref(total)))
        return free.value

    def close(self):
        for p in self.owned:
            self.rt.cudaFree(p)
        self.owned = []


def _inputs(seed, n_obs, n_comp):
    rng = np.random.RandomState(seed)
    F = rng.normal(size=(n_obs, n_comp)) * 3.
    T = np.log(rng.dirichlet(np.ones(n_comp), size=(max(n_obs - 1, 0), n_comp)))
    # a few forbidden transitions and ties, as the model's own matrices have
    T[rng.uniform(size=T.shape) < 0.05] = -np.inf
    if n_obs > 2:
        F[n_obs // 2, :] = F[n_obs // 2, 0]
    return np.ascontiguousarray(F), np.ascontiguousarray(T)


def _existing(lib, check, F, T):
    n_obs, n_comp = F.shape
    path = np.empty(n_obs, dtype=np.int64)
    lp = C.c_double()
    check(lib.rmx_max_product(0, n_obs, n_comp, F.ctypes.data_as(C.c_void_p), T.ctypes.data_as(C.c_void_p),
                              path.ctypes.data_as(C.c_void_p), C.byref(lp)), None)
    return path, lp.value


def test_no_alloc_max_product_matches_existing_entry_point():
    from remixt_b200 import _lib
    lib, check = _lib.lib, _lib.check
    assert _lib.device_count() > 0, 'no CUDA device: the GPU tests must not silently skip'
    cuda = _Cuda()
    stream = C.c_void_p()
    cuda.ok(cuda.rt.cudaStreamCreate(C.byref(stream)))
    try:
        # calls in a row on the shared array: the later ones are smaller and of another shape, so the array still holds
        # cells of the earlier ones (sizes bounded by the (n_obs - 1, n_comp, n_comp) matrix the caller supplies)
        cases = [(1, 1500, 97, None), (2, 400, 165, stream), (3, 1, 5, stream)]
        assert all(n * s <= MAX_CELLS for _, n, s, _ in cases)
        staged = []
        for seed, n_obs, n_comp, st in cases:
            F, T = _inputs(seed, n_obs, n_comp)
            staged.append((F, T, cuda.upload(F), cuda.upload(T), cuda.alloc(8 * n_obs), cuda.alloc(8), st))
        for k, (F, T, dF, dT, dpath, dlp, st) in enumerate(staged):
            n_obs, n_comp = F.shape
            free_before = cuda.free_bytes()
            check(lib.rmx_max_product_noalloc(0, st, n_obs, n_comp, dF, dT, dpath, dlp), None)
            cuda.ok(cuda.rt.cudaStreamSynchronize(st))
            if k > 0:       # the first call may load the module (and with it the static array)
                assert cuda.free_bytes() == free_before, 'the allocation-free entry point changed the free device memory'
            path = cuda.download(dpath, (n_obs,), np.int64)
            lp = float(cuda.download(dlp, (1,), np.float64)[0])
            path_ref, lp_ref = _existing(lib, check, F, T)
            assert np.array_equal(path, path_ref), 'case %d: state sequence differs' % k
            assert abs(lp - lp_ref) <= RTOL_TIGHT * abs(lp_ref), (k, lp, lp_ref)
            assert path.min() >= 0 and path.max() < n_comp
    finally:
        cuda.rt.cudaStreamDestroy(stream)
        cuda.close()


def test_no_alloc_max_product_rejects_sizes_beyond_the_static_array():
    from remixt_b200 import _lib
    lib = _lib.lib
    dummy = C.c_void_p(256)
    n_comp = 97
    n_obs = MAX_CELLS // n_comp + 1
    rc = lib.rmx_max_product_noalloc(0, None, n_obs, n_comp, dummy, dummy, dummy, dummy)
    assert rc == _lib.RMX_ERR_ARG
    msg = _lib.last_error(None)
    assert 'RMX_NOALLOC_MAX_CELLS' in msg and str(n_obs * n_comp) in msg, msg
    with pytest.raises(TypeError):
        _lib.check(rc, None)
    assert lib.rmx_max_product_noalloc(0, None, 10, 5000, dummy, dummy, dummy, dummy) == _lib.RMX_ERR_ARG
