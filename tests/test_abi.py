"""The C-ABI library loads without a GPU and exports every symbol include/b200sd.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b200sd.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200sd_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(rt):
    names = _declared()
    assert len(names) >= 15
    L = rt.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/b200sd.h but not exported"
    assert sorted(rt.EXPORTS) == names, "runtime.EXPORTS is out of sync with the header"
    assert L.b200sd_version() == 100
    assert isinstance(rt.launch_count(), int)


def test_param_struct_layout_matches_header(rt):
    """ctypes mirror of b200sd_igemm_params must have the C layout (sizes from the header's field list)."""
    assert ctypes.sizeof(rt.Src) == 8 + 16 + 24
    assert ctypes.sizeof(rt.Seg) == 20
    # 8 (n_src,n_seg) + 4*48 + 40*20 + 8 + 4 + 12 (+pad) ...: just check alignment-sensitive offsets
    assert rt.IgemmParams.srcs.offset == 8
    assert rt.IgemmParams.segs.offset == 8 + 4 * 48
    assert rt.IgemmParams.weight.offset == 8 + 4 * 48 + 40 * 20
    assert rt.IgemmParams.weight.offset % 8 == 0
    assert rt.IgemmParams.splitk_ws.offset == 1176 and rt.IgemmParams.splitk_max.offset == 1204   # gcc's offsetof
    assert ctypes.sizeof(rt.IgemmParams) == 1208


def test_no_cpu_fallback(rt):
    import pytest
    import torch
    with pytest.raises(rt.B200Error):
        rt.sched_add_noise(torch.zeros(2, 4), torch.zeros(2, 4), torch.zeros(2, dtype=torch.long),
                           torch.zeros(10), torch.zeros(10))
    with pytest.raises(rt.B200Error):
        rt.linear(torch.zeros(4, 64, dtype=torch.bfloat16), torch.zeros(8, 64, dtype=torch.bfloat16))
