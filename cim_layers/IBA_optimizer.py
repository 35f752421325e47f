"""Drop-in for `cim_layers/IBA_optimizer.py` of the reference: `PercentOptimizerFP`, the optimizer of the CIM trainers
(Model_DiT_9L_train.py:17,266), as two table-driven sm_100a launches per step (sdb200.optim.PercentOptimizerFP,
csrc/optim.cu). The other names of the reference's file (AdamOptimizerINT, DDFP_scheduler: chip-emulation
experiments, out of scope here) resolve to its own copy when that checkout is on sys.path as well."""
import importlib.util
import os
import sys

from sdb200.optim import PercentOptimizerFP  # noqa: F401

_here = os.path.abspath(__file__)
for _p in sys.path:
    _cand = os.path.join(_p or ".", "cim_layers", "IBA_optimizer.py")
    if os.path.isfile(_cand) and os.path.abspath(_cand) != _here:
        _spec = importlib.util.spec_from_file_location("_reference_IBA_optimizer", _cand)
        _mod = importlib.util.module_from_spec(_spec)
        _spec.loader.exec_module(_mod)
        for _k, _v in vars(_mod).items():
            if not _k.startswith("_") and _k not in globals():
                globals()[_k] = _v
        break
