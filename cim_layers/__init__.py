"""Drop-in for the reference's `cim_layers` package: the LSQ quant-noise layers (layers_qn_lsq.py) and the ADDA
compute-in-memory layers (layers_qn_lsq_adda_cim.py) on the sm_100a kernels, and the CIM trainers' optimizer
(IBA_optimizer.PercentOptimizerFP). Modules this repo does not replace (the sibling experimental variants,
register_dict, the other optimisers) resolve to the reference's own files when its checkout is on sys.path as well."""
import os
import sys

for _p in sys.path:
    _cand = os.path.join(_p, "cim_layers")
    if os.path.isdir(_cand) and os.path.abspath(_cand) != os.path.dirname(os.path.abspath(__file__)) \
            and _cand not in __path__:
        __path__.append(_cand)
