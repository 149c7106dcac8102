"""Level table of the multiresolution hash grid (row f2: a reference checkpoint's flat `params`
vector is laid out by it).  tinycudann evaluates scale_l = exp2f(l * log2f(b)) * N_min - 1.0f and
res_l = (uint32)ceilf(scale_l) + 1 in fp32 on the host (grid.h, called from the encoding's constructor); the
product (radnerf_b200.tcnn.grid_meta) and the oracle (oracle/tcnn_cpu.GridMeta) both follow that rule by default and
keep round 1's rule (fp64, snapped) behind rule="fp64_snap" / RNB_GRID_LEVEL_RULE=fp64_snap.

The two differ only where the real-valued scale is an integer: scene scale 0.5 (b = 64^(1/15)) at levels 5, 10, 15,
where fp32(b) rounds up, log2f returns 0.40000004 and the level comes out one finer (65 / 257 / 1025) -- which moves
every later level's offset.  SURVEY 2.4's 5 710 032 entries is the real-valued count; tinycudann's arithmetic gives
5 722 520.  PARITY UNPINNED: tinycudann is absent here, so which one a reference checkpoint carries cannot be checked
on this box; the fp32 rule is the one its source states.
"""
import numpy as np
import pytest

import radnerf_b200  # noqa: F401
from oracle import tcnn_cpu as T
from radnerf_b200 import tcnn


def per_level_scale(scale):
    return float(np.exp(np.log(2048 * scale / 16) / 15))      # models/networks.py:38-40, 229-231


@pytest.mark.parametrize("scale,n_fp32,n_real", [(0.5, 5722520, 5710032), (4.0, 6537456, 6537456), (16.0, 6811592, 6811592)])
def test_level_tables(scale, n_fp32, n_real):
    b = per_level_scale(scale)
    m32, n32 = tcnn.grid_meta(16, 2, 19, 16, b, rule="tcnn_fp32")
    m64, n64 = tcnn.grid_meta(16, 2, 19, 16, b, rule="fp64_snap")
    print(f"\nscale {scale}: b = {b!r}")
    print("  level  res(fp32)  offset(fp32)   res(real)  offset(real)")
    for l in range(16):
        print(f"  {l:5d}  {m32.res[l]:9d}  {m32.offset[l]:12d}   {m64.res[l]:9d}  {m64.offset[l]:12d}")
    print(f"  entries: tinycudann fp32 rule {n32}, real-valued rule {n64}")
    assert n32 == n_fp32 and n64 == n_real
    # product and oracle agree on both rules, entry for entry
    for rule, m, n in (("tcnn_fp32", m32, n32), ("fp64_snap", m64, n64)):
        om = T.GridMeta(per_level_scale=b, rule=rule)
        assert om.n_entries == n and list(m.res[:16]) == om.res and list(m.offset[:17]) == om.offsets
        assert [np.float32(x) for x in m.scale[:16]] == om.scales
    # the rules agree except on integer-valued levels
    diff = [l for l in range(16) if m32.res[l] != m64.res[l]]
    assert diff == ([5, 10, 15] if scale == 0.5 else [])


def test_default_rule_is_tinycudann_fp32(monkeypatch):
    monkeypatch.delenv("RNB_GRID_LEVEL_RULE", raising=False)
    assert tcnn.default_level_rule() == "tcnn_fp32"
    m, n = tcnn.grid_meta(16, 2, 19, 16, per_level_scale(0.5))
    assert n == 5722520 and m.res[5] == 65
    monkeypatch.setenv("RNB_GRID_LEVEL_RULE", "fp64_snap")
    m, n = tcnn.grid_meta(16, 2, 19, 16, per_level_scale(0.5))
    assert n == 5710032 and m.res[5] == 64
