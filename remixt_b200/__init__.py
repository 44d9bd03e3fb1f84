"""remixt_b200: B200-native implementation of ReMixT's variational-inference hot path.

``remixt_b200.bpmodel``  drop-in for the reference's Cython module ``remixt.bpmodel``
``remixt_b200.cn_model`` host driver mirroring ``remixt.cn_model.BreakpointModel``
``remixt_b200.pipeline`` fit / init-grid orchestration mirroring ``remixt.analysis.pipeline``
``remixt_b200.mstep``    lock-step M-step of an initialisation grid (batched objective evaluations)
``remixt_b200.readdepth`` initialisation candidates mirroring ``remixt.analysis.readdepth``
``remixt_b200.results``  result tables / collated store mirroring ``remixt.analysis.experiment`` and ``pipeline.collate``

Importing the package loads ``libremixt_b200.so`` and fails if it has not been built.
"""
from . import _lib          # noqa: F401  (fails loudly when the CUDA library is missing)
from . import bpmodel       # noqa: F401
from . import cn_model      # noqa: F401
from . import pipeline      # noqa: F401
from . import mstep         # noqa: F401
from . import readdepth     # noqa: F401
from . import results       # noqa: F401

__version__ = '0.1.0'


def install_as_remixt_bpmodel():
    """Alias this package's ``bpmodel`` as ``remixt.bpmodel`` so that an installed reference
    ``remixt.cn_model`` / ``remixt.analysis.pipeline`` runs on the B200 kernels unmodified.
    Call before ``import remixt.cn_model``."""
    import sys
    sys.modules['remixt.bpmodel'] = bpmodel
    if 'remixt' in sys.modules:
        sys.modules['remixt'].bpmodel = bpmodel
    return bpmodel
