"""viekf_batch_get_state_noalloc (include/viekf_batch.h; no counterpart in the reference):
the read-back through the fixed file-scope device scratch array gives exactly what viekf_batch_get_state gives,
on two calls in a row (the array keeps the previous call's contents), and rejects a request beyond the fixed size.

Collected last on purpose; needs nothing from the other tests and no oracle."""
import numpy as np
import pytest

from tests.common import abi

pytestmark = pytest.mark.gpu


def _imu(t):
    return np.array([0.4 * np.sin(1.3 * t), 0.3 * np.cos(0.9 * t), -9.80665 + 0.5 * np.sin(0.7 * t), 0.15 * np.sin(1.1 * t), 0.12 * np.cos(0.8 * t),
                     0.1 * np.sin(0.5 * t)])


def test_get_state_noalloc_matches_get_state_twice(pkg):
    cfg = abi.default_config()
    B, N = 64, 6
    bk = pkg.BatchVIEKF(cfg, B=B, N=N)
    bk.perturb_x0(7)
    bk.propagate_state(_imu(0.0), 0.0)
    rng = np.random.default_rng(3)
    for i in range(4):
        assert bk.init_feature(rng.uniform(100, 400, size=(B, 2)), i, depth=3.0 + i).all()
    for k in range(1, 6):
        bk.propagate_state(_imu(0.005 * k), 0.005 * k)
    # first call: the whole batch (leaves 64 x 47 doubles in the shared array)
    a = bk.get_state_noalloc()
    ref = bk.get_state()
    assert a.shape == (B, 17 + 5 * N) and np.array_equal(a, ref), "bit-identical: same kernel, same copy"
    assert np.abs(a[:, 17:17 + 5 * 4]).max() > 0 and (a[:, 17 + 5 * 4:] == 0).all()
    # second call in a row, after the state moved on, on a sub-range: nothing of the first call's contents leaks through
    bk.keep_only_features([0, 2])
    bk.propagate_state(_imu(0.03), 0.03)
    b = bk.get_state_noalloc(3, 5)
    ref2 = bk.get_state(3, 5)
    assert np.array_equal(b, ref2) and not np.array_equal(b, ref[3:8])
    assert (b[:, 17 + 5 * 2:] == 0).all()
    assert np.array_equal(bk.get_state_noalloc(), bk.get_state())
    bk.close()


def test_get_state_noalloc_rejects_oversize_and_bad_ranges(pkg):
    cfg = abi.default_config()
    N = 25
    per = 17 + 5 * N
    fit = abi.NOALLOC_SCRATCH_DOUBLES // per
    bk = pkg.BatchVIEKF(cfg, B=fit + 1, N=N)
    with pytest.raises(pkg.batch.ViekfError) as e:
        bk.get_state_noalloc()  # one instance too many
    assert "exceed the fixed scratch" in str(e.value) and str(fit) in str(e.value)
    x = bk.get_state_noalloc(1, fit)  # exactly the limit
    assert np.array_equal(x, bk.get_state(1, fit))
    with pytest.raises(pkg.batch.ViekfError):
        bk.get_state_noalloc(2, fit)  # beyond the batch
    bk.close()
