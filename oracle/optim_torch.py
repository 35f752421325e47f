"""TEST INFRASTRUCTURE (oracle): CPU restatement of the reference's PercentOptimizerFP.step
(cim_layers/IBA_optimizer.py:80-143) as a pure function over lists of tensors. Pinned bit-for-bit against the real
reference class by oracle/gen_golden.py (tests/golden/percent_opt.pt). Only tests/, __graft_entry__.smoke() and
bench.py's CPU legs may import this package."""
import torch


def percent_step(params, grads, exp_avg, exp_avg_sq, step, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
    """One optimizer step, in place on `params`, `exp_avg`, `exp_avg_sq` (lists of fp32 tensors); `step` is the 1-based
    count of this step (IBA_optimizer.py:105). Tensors whose gradient is None are skipped (:93-94)."""
    b1, b2 = betas
    for p, g, m, v in zip(params, grads, exp_avg, exp_avg_sq):
        if g is None:
            continue
        m.copy_(b1 * m + (1 - b1) * g)                                   # :108-109
        v.copy_(b2 * v + (1 - b2) * (g * g))                             # :112-113
        bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step                        # :116-117
        adjusted = (bc2 ** 0.5) / bc1                                    # :120
        update = (adjusted * m) / ((v ** 0.5) / (bc2 ** 0.5) + eps)      # :123-124
        max_u, max_w = update.abs().max(), p.abs().max()                 # :129-130
        scale = max(lr * max_w / (max_u + eps), lr)                      # :132 (Python max: lr only if strictly larger)
        update *= scale                                                  # :133
        p.copy_(p - update)                                              # :137-138
