"""Reference-surface model package (models.unet_cond_base, models.unet_base, models.blocks, models.vqvae,
models.transformer, ...). Hot-path modules are implemented here; when the reference checkout is also on sys.path,
its remaining `models.*` modules (lpips, discriminator, vae, ...) stay importable through the extended path."""
import os
import sys

for _p in sys.path:
    _cand = os.path.join(_p, "models")
    if os.path.isdir(_cand) and os.path.abspath(_cand) != os.path.dirname(os.path.abspath(__file__)) \
            and os.path.exists(os.path.join(_cand, "unet_cond_base.py")):
        __path__.append(_cand)
