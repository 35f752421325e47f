"""Host-side launch planning that needs no GPU: the one-pass GroupNorm backward's cluster / slab plan
(b200sd_groupnorm_bwd_plan, csrc/norm.cu gn_fused_plan)."""
import itertools

import pytest

import sdb200  # noqa: F401
from sdb200 import runtime as rt

SMEM_MAX = 200 * 1024


@pytest.mark.parametrize("with_add", [False, True])
def test_groupnorm_backward_plan_invariants(with_add, monkeypatch):
    """Every shape either gets a plan that tiles it exactly — whole groups per cluster in 16-byte vectors, every CTA of
    the cluster owning at least one row, the slab (x, dy and the optional extra gradient) within the shared-memory
    budget, at most 256 threads — or falls back to the two-pass kernels."""
    monkeypatch.delenv("B200SD_GN_BWD_FUSED", raising=False)
    monkeypatch.delenv("B200SD_GN_MIN_CH", raising=False)
    seen_fused = seen_fallback = 0
    for n, hw, (c, groups) in itertools.product(
            (1, 2, 16, 64), (1, 16, 64, 100, 256, 1024, 4096, 16384),
            ((24, 4), (48, 8), (80, 8), (96, 8), (128, 32), (256, 32), (320, 32), (384, 32), (512, 32), (640, 32),
             (768, 32), (960, 32), (1024, 32), (1536, 32), (2048, 32))):
        p = rt.groupnorm_bwd_plan(n, hw, c, groups, with_add)
        if p is None:
            seen_fallback += 1
            continue
        seen_fused += 1
        G, K, rows, rpp = p["groups_per_cluster"], p["ctas_per_cluster"], p["rows_per_cta"], p["rows_per_pass"]
        cpg = c // groups
        cc = G * cpg
        assert groups % G == 0 and cc % 8 == 0 and groups // G <= 64, (n, hw, c, p)
        assert cc >= 64 or G == groups, (n, hw, c, p)
        assert K in (1, 2, 4, 8) and rows * K >= hw and rows * (K - 1) < hw, (n, hw, c, p)
        cv = cc // 8
        assert 1 <= rpp <= rows and cv * rpp <= 256 and p["threads"] % 32 == 0 and cv * rpp <= p["threads"] <= 256
        slabs = 3 if with_add else 2
        assert p["smem"] >= slabs * rows * cc * 2 and p["smem"] <= SMEM_MAX, (n, hw, c, p)
    assert seen_fused > 300 and seen_fallback > 10


def test_groupnorm_backward_plan_of_the_training_shapes(monkeypatch):
    """The batch-16 shapes of the conditioned UNet: 8 CTAs per cluster at 32 x 32, slabs that leave room for a second
    CTA on the SM; and the switch that forces the two-pass route."""
    monkeypatch.delenv("B200SD_GN_BWD_FUSED", raising=False)
    monkeypatch.delenv("B200SD_GN_MIN_CH", raising=False)
    p = rt.groupnorm_bwd_plan(16, 1024, 384, 32, True)
    assert (p["groups_per_cluster"], p["ctas_per_cluster"], p["rows_per_cta"]) == (8, 8, 128) and p["smem"] < 108 * 1024
    p = rt.groupnorm_bwd_plan(16, 64, 1536, 32, False)
    assert p["groups_per_cluster"] * (1536 // 32) >= 64 and p["ctas_per_cluster"] == 2
    assert rt.groupnorm_bwd_plan(1, 16384, 256, 32, False) is None           # 2048 rows x 64 channels per CTA: no fit
    assert rt.groupnorm_bwd_plan(16, 1024, 380, 32, False) is None           # channels not a multiple of 8 / of groups
    monkeypatch.setenv("B200SD_GN_BWD_FUSED", "0")
    assert rt.groupnorm_bwd_plan(16, 1024, 384, 32, True) is None
