"""B200-native DDPM denoising hot path: CUDA kernels (csrc/), the C-ABI library (lib/libb200sd.so), its ctypes
binding (runtime.py) and the fused execution engine the reference-surface modules (models/, scheduler/,
cim_layers/ at the repo root) call into."""
__all__ = ["runtime", "build"]
