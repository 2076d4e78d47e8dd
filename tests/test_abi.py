"""Host-side checks that need no GPU: the C-ABI library loads, exports every symbol include/viekf_batch.h declares,
struct layouts agree between C and the ctypes / numpy mirrors, and the product path fails loudly without a device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import __graft_entry__ as g

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib(pkg):
    g.build()
    return pkg.abi.load_library()


def test_exports_every_declared_symbol(lib, pkg):
    hdr = open(os.path.join(ROOT, "include", "viekf_batch.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(viekf_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/viekf_batch.h but not exported"
    assert declared == set(lib._declared), declared ^ set(lib._declared)


def test_struct_layouts(lib, pkg, oracle_mod):
    assert oracle_mod.lib().vo_sizeof_event() == pkg.abi.EVENT_DTYPE.itemsize
    assert oracle_mod.lib().vo_sizeof_config() == C.sizeof(pkg.abi.Config)
    c = pkg.abi.Config()
    lib.viekf_config_default(C.byref(c))
    d = pkg.abi.default_config()
    assert bytes(c) == bytes(d), "viekf_config_default (C) and abi.default_config (Python) disagree on params/ekf.yaml"


def test_no_cpu_fallback(lib, pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.batch.ViekfError) as e:
        pkg.BatchVIEKF(B=4, N=2)
    assert "no CUDA device" in str(e.value)


def test_tape_builder_roundtrip(pkg):
    tb = pkg.TapeBuilder()
    tb.propagate(np.arange(6.0), 0.5, sigma=[1, 2, 3, 4, 5, 6])
    tb.update(pkg.abi.FEAT, [1.0, 2.0], np.diag([3.0, 4.0]), id=7, depth=2.5)
    tb.keep_only_features([4, 5, 6])
    tb.stat(0, np.arange(17.0))
    ev = tb.events()
    assert ev.dtype.itemsize == 248 and len(ev) == 4
    assert ev[0]["kind"] == pkg.abi.EV_PROP and ev[0]["t"] == 0.5 and list(ev[0]["sigma"][:6]) == [1, 2, 3, 4, 5, 6]
    assert ev[1]["zdim"] == 2 and ev[1]["rdim"] == 2 and ev[1]["z"][4] == 2.5 and list(ev[1]["R"][:4]) == [3.0, 0.0, 0.0, 4.0]
    assert ev[2]["n_ids"] == 3 and list(tb.id_pool()) == [4, 5, 6]
    assert tb.n_imu_steps() == 1


def test_config_load_yaml(lib, pkg):
    """VIEKF(param_file) / load (reference vi_ekf.cpp:101-155): the C++ reader of params/ekf.yaml-style files."""
    data = os.path.join(ROOT, "tests", "data")
    c = pkg.abi.load_config(os.path.join(data, "ekf_default.yaml"))
    assert bytes(c) == bytes(pkg.abi.default_config()), "the fixture carries the defaults: the loaded configuration must equal viekf_config_default"
    o = pkg.abi.load_config(os.path.join(data, "ekf_override.yaml"))
    assert list(o.Qu) == [2.0, 2.0, 2.0, 1e-3, 1e-3, 1e-3] and o.use_drag_term == 0 and o.use_keyframe_reset == 0 and o.use_partial_update == 1
    assert o.min_depth == 0.75 and list(o.lambda_feat) == [0.5, 0.5, 0.25] and list(o.P0) == list(c.P0)
    with pytest.raises(RuntimeError) as e:
        pkg.abi.load_config(os.path.join(data, "ekf_bad.yaml"))
    assert "P0_feat" in str(e.value)
    with pytest.raises(RuntimeError):
        pkg.abi.load_config(os.path.join(data, "does_not_exist.yaml"))
