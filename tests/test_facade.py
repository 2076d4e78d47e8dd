"""include/viekf_facade.hpp: the header-only C++ facade with the reference's method names compiles with plain g++ against
the C ABI and links the engine; without a GPU it fails loudly (no CPU fallback); on the GPU it reproduces the ctypes path."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def demo(pkg, tmp_path_factory):
    pkg.abi.lib_path()
    out = str(tmp_path_factory.mktemp("facade") / "facade_demo")
    csrc = os.path.join(ROOT, "vi-ekf_b200", "csrc")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "facade_demo.cpp"),
                           "-o", out, "-L", csrc, "-lviekf_b200", f"-Wl,-rpath,{csrc}"])
    return out


def test_facade_compiles_and_fails_loudly_without_gpu(demo):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([demo], capture_output=True, text=True)
    assert r.returncode == 3 and "no CUDA device" in r.stderr


@pytest.mark.gpu
def test_facade_matches_ctypes_path(demo, pkg):
    r = subprocess.run([demo], capture_output=True, text=True, check=True)
    lines = r.stdout.strip().splitlines()
    head = lines[0].split()
    x_cpp = np.array([float(v) for v in lines[1].split()[1:]])
    abi = pkg.abi
    cfg = abi.default_config()
    bk = pkg.BatchVIEKF(cfg, B=1, N=4)
    for k in range(41):
        t = 0.005 * k
        u = np.array([0.1, -0.05, -9.80665, 0.01, 0.02, -0.01])
        bk.propagate_state(u, t)
        bk.update(abi.ACC, u[:2], np.eye(2), active=True)
        if k % 8 == 4:
            for fid in range(3):
                bk.update(abi.FEAT, np.array([300.0 + 20 * fid, 220.0 + 10 * fid]), 0.25 * np.eye(2), active=True, id=fid, depth=4.0)
            bk.keep_only_features([0, 1, 2])
    x = bk.get_state()[0]
    assert int(head[1]) == 3 and int(head[3]) == 3
    assert np.array_equal(x_cpp, x)


@pytest.mark.gpu
def test_facade_queue_mode_matches_ctypes_path(demo, pkg):
    """enable_queue(): add_measurement / handle_measurements with the reference's rewind-replay semantics (vi_ekf_meas.cpp:6-194)."""
    r = subprocess.run([demo, "queue"], capture_output=True, text=True, check=True)
    lines = r.stdout.strip().splitlines()
    head = lines[0].split()
    x_cpp = np.array([float(v) for v in lines[1].split()[1:]])
    abi = pkg.abi
    cfg = abi.default_config()
    cfg.use_keyframe_reset = 0
    bk = pkg.BatchVIEKF(cfg, B=1, N=4)
    bk.enable_queue()
    n_new = 0
    for k in range(41):
        t = 0.005 * (k + 1)
        u = np.array([0.1, -0.05, -9.80665, 0.01, 0.02, -0.01])
        bk.queue_propagate_state(u, t)
        if k > 0:
            bk.add_measurement(t, u[:2], abi.ACC, np.eye(2), True)
        if k % 8 == 0:
            for fid in range(3):
                r_ = bk.add_measurement(t if k == 0 else t - 0.012, np.array([300.0 + 20 * fid + 0.1 * k, 220.0 + 10 * fid]), abi.FEAT,
                                        0.25 * np.eye(2), True, fid, 4.0)
                n_new += r_ == abi.MEAS_NEW_FEATURE
        bk.handle_measurements()
    assert int(head[1]) == 3 and int(head[3]) == n_new == 3
    assert np.array_equal(x_cpp, bk.get_state()[0])
