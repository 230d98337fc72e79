"""The library's native descriptor builder (qeb_desc_*, host code, no GPU) against the oracle's restatement of
fft_type_init / ggen / fft_set_nl / gk_sort and of the slab distribution: every array must be identical."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import qe_b200  # noqa: E402
from oracle import descriptor as D  # noqa: E402
from oracle import dist as X  # noqa: E402

SMALL = ["si2_small", "si2_small_gamma", "ortho_small", "ortho_small_gamma", "si2", "al4"]


def native(name):
    c = D.CONFIGS[name]
    return qe_b200.descriptor.Descriptor(c["at"], c["alat"], c["ecutwfc"], c["ecutrho"] or 0.0, c["xk"], c["gamma"])


def test_good_fft_order():
    for n in range(1, 400):
        assert qe_b200.descriptor.good_fft_order(n) == D.good_fft_order(n)


@pytest.mark.parametrize("name", SMALL + ["ausurf112"])
def test_serial_descriptor_matches_oracle(name):
    o = D.config_descriptor(name)
    d = native(name)
    assert (d.nr1, d.nr2, d.nr3, d.ngm, d.npw, d.nst, d.gstart) == (o.nr1, o.nr2, o.nr3, o.ngm, o.npw, o.nst, o.gstart)
    assert np.array_equal(d.mill, o.mill)
    assert np.array_equal(d.nl, o.nl)
    if o.gamma_only:
        assert np.array_equal(d.nlm, o.nlm)
    assert np.array_equal(d.igk, o.igk)
    assert np.array_equal(d.g2kin, o.g2kin)
    assert np.array_equal(d.gg, o.gg)
    assert np.array_equal(d.isind, o.isind) and np.array_equal(d.iplw, o.iplw)


def test_workload_table_matches_survey_sizes():
    """SURVEY section 8 size table (grid, npw, wave sticks) for the BASELINE configs that build in seconds."""
    for name, grid, npw, nst in (("si2", (25, 25, 25), 734, 109), ("al4", (27, 27, 27), 1246, None),
                                 ("ausurf112", (125, 64, 200), 95394, None)):
        d, w = qe_b200.descriptor.workload_descriptor(name)
        assert (d.nr1, d.nr2, d.nr3) == grid
        assert abs(d.npw - npw) <= 0.01 * npw and (nst is None or d.nst == nst)


@pytest.mark.parametrize("name", ["si2_small", "si2_small_gamma", "ortho_small", "ortho_small_gamma"])
@pytest.mark.parametrize("nproc", [1, 2, 3, 4, 8])
def test_distribution_matches_oracle(name, nproc):
    o = D.config_descriptor(name)
    if o.nr3 < nproc:
        pytest.skip("fewer planes than ranks")
    od = X.distribute(o, nproc)
    d = native(name).distribute(nproc)
    for a in ("nr3p", "nsw", "nsp", "iss", "ismap"):
        assert np.array_equal(getattr(d, a), getattr(od, a)), a
    for r in range(nproc):
        a, b = d.ranks[r], od.ranks[r]
        assert (a.ngm, a.npw, a.gstart) == (b.ngm, b.npw, b.gstart)
        for f in ("nl", "igk", "g2kin", "ig_l2g", "pw_l2g"):
            assert np.array_equal(getattr(a, f), getattr(b, f)), (r, f)
        if o.gamma_only:
            assert np.array_equal(a.nlm, b.nlm)


def test_errors():
    with pytest.raises(qe_b200.QebError):
        qe_b200.descriptor.Descriptor(np.eye(3), 10.0, 10.0, 0.0, (0.1, 0, 0), True)  # gamma_only needs k = 0
    d = native("si2_small")
    with pytest.raises(qe_b200.QebError):
        d.distribute(d.nr3 + 1)  # a rank without planes
