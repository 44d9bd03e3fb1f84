"""Import harness for the UNMODIFIED reference Python modules (this container only).

TEST INFRASTRUCTURE ONLY.  Nothing under ``remixt_b200/`` may import this.

``install()`` makes ``import remixt.cn_model`` etc. work without copying any reference
source: a synthetic ``remixt`` package whose ``__path__`` lists ``oracle/_ref/remixt`` (the
compiled ``bpmodel`` extension, see build_ref.py) and ``/root/reference/remixt`` (the .py
files, read in place).  Absent third-party modules the reference imports but never uses on
the fit path are stubbed: pypeliner, blossomv, statsmodels;
``scipy.misc.logsumexp`` is aliased to ``scipy.special.logsumexp``
(``simulations/experiment.py:698``).

``install(bpmodel=module)`` aliases ``remixt.bpmodel`` to another module (the B200 drop-in)
so the reference's own ``cn_model.py`` drives it unmodified.

``/root/reference`` does not exist on the GPU box: ``available()`` is False there and the
callers fall back to the committed golden fixtures.
"""
import importlib
import importlib.machinery
import os
import sys
import types
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("REMIXT_REFERENCE", "/root/reference")
REF_PKG = os.path.join(REF_ROOT, "remixt")
BIN_PKG = os.path.join(HERE, "_ref", "remixt")


def _staged():
    # the reference's fit-path modules staged next to the compiled extension by build_ref.stage_py
    return os.path.exists(os.path.join(BIN_PKG, "cn_model.py"))


def available():
    """True when the reference's Python modules can be imported: from the reference tree (this container) or from
    the copies build_ref staged under oracle/_ref (the GPU box)."""
    return os.path.isdir(REF_PKG) or _staged()


def _stub(name, **attrs):
    mod = types.ModuleType(name)
    mod.__dict__.update(attrs)
    mod.__path__ = []
    sys.modules[name] = mod
    return mod


def _install_stubs():
    if "pypeliner" not in sys.modules:
        p = _stub("pypeliner")
        p.commandline = _stub("pypeliner.commandline")
    if "blossomv" not in sys.modules:
        b = _stub("blossomv")
        b.blossomv = _stub("blossomv.blossomv")
    try:
        import statsmodels.tools.numdiff  # noqa: F401
    except Exception:
        def approx_fprime(x, f, epsilon=None, args=(), kwargs={}, centered=False):
            import numpy as np
            x = np.asarray(x, dtype=float)
            f0 = f(x, *args, **kwargs)
            grad = np.zeros_like(x)
            for i in range(x.shape[0]):
                h = 1.4901161193847656e-08 * max(abs(x[i]), 0.1)
                xh = x.copy()
                xh[i] += h
                grad[i] = (f(xh, *args, **kwargs) - f0) / h
            return grad
        s = _stub("statsmodels")
        s.tools = _stub("statsmodels.tools")
        s.tools.numdiff = _stub("statsmodels.tools.numdiff", approx_fprime=approx_fprime)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        try:
            import scipy.misc
        except Exception:
            import scipy
            scipy.misc = _stub("scipy.misc")
        import scipy.special
        if not hasattr(scipy.misc, "logsumexp"):
            scipy.misc.logsumexp = scipy.special.logsumexp


def install(bpmodel=None):
    """Make the reference package importable; optionally alias remixt.bpmodel."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    _install_stubs()
    if "remixt" not in sys.modules or getattr(sys.modules["remixt"], "_oracle_env", False) is False:
        pkg = types.ModuleType("remixt")
        pkg.__path__ = [BIN_PKG, REF_PKG] if os.path.isdir(REF_PKG) else [BIN_PKG]
        pkg.__version__ = "0.5.20-oracle"
        pkg._oracle_env = True
        pkg.__spec__ = importlib.machinery.ModuleSpec("remixt", None, is_package=True)
        pkg.__spec__.submodule_search_locations = pkg.__path__
        sys.modules["remixt"] = pkg
    if bpmodel is not None:
        sys.modules["remixt.bpmodel"] = bpmodel
        sys.modules["remixt"].bpmodel = bpmodel
    return sys.modules["remixt"]


def uninstall():
    for k in [k for k in sys.modules if k == "remixt" or k.startswith("remixt.")]:
        del sys.modules[k]


def load_ref_bpmodel():
    """Import ONLY the compiled reference extension (works on the GPU box too: the built
    .so under oracle/_ref travels, the reference .py files do not)."""
    import importlib.util
    import sysconfig
    so = os.path.join(BIN_PKG, "bpmodel" + sysconfig.get_config_var("EXT_SUFFIX"))
    if not os.path.exists(so):
        return None
    name = "remixt.bpmodel"
    if name in sys.modules and getattr(sys.modules[name], "__file__", None) == so:
        return sys.modules[name]
    had_pkg = "remixt" in sys.modules
    if not had_pkg:
        pkg = types.ModuleType("remixt")
        pkg.__path__ = [BIN_PKG]
        sys.modules["remixt"] = pkg
    saved = sys.modules.pop(name, None)
    try:
        spec = importlib.util.spec_from_file_location(name, so)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
    finally:
        if saved is not None:
            sys.modules[name] = saved
        else:
            sys.modules.pop(name, None)
        if not had_pkg:
            sys.modules.pop("remixt", None)
    return mod
