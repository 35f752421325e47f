"""Drop-in for the part of the reference's `cim_qn_train` package that sits on the denoising hot path: the Enhance
wrappers of layers_enhance.py on the sm_100a kernels. Everything else of that package (the
ProgressiveTrain orchestration, analyzers) resolves to the reference's own files when its checkout is on sys.path too."""
import os
import sys

for _p in sys.path:
    _cand = os.path.join(_p, "cim_qn_train")
    if os.path.isdir(_cand) and os.path.abspath(_cand) != os.path.dirname(os.path.abspath(__file__)) \
            and _cand not in __path__:
        __path__.append(_cand)
