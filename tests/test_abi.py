"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every symbol the header declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, 'include', 'ekfslam_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(ekfb?_[a-zA-Z0-9_]+)\s*\(', src)))


def test_header_declares_the_boundary():
    names = _declared()
    for must in ('ekf_create', 'ekf_destroy', 'ekf_predict', 'ekf_update', 'ekf_append', 'ekf_associate',
                 'ekf_measure_batch', 'ekf_get_x', 'ekf_get_P', 'ekf_get_s', 'ekf_set_state', 'ekf_sync',
                 'ekf_last_error'):
        assert must in names


def test_library_exports_every_declared_symbol(lib):
    for name in _declared():
        assert hasattr(lib, name), name


def test_python_binding_covers_header(lib):
    from ekf_slam_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared()


def test_params_default_match_reference_constants(lib):
    from ekf_slam_b200 import _lib
    p = _lib.EkfParams()
    assert lib.ekf_params_default(ctypes.byref(p), 0) == 0          # EKF_SLAM.m:12-16
    assert (p.C, p.Rc[0], p.Rc[1], p.s_cost, p.s_thresh) == (0.2, 0.01, 5.0, 1e-11, 1e9)
    assert lib.ekf_params_default(ctypes.byref(p), 1) == 0          # EKF_SLAM_UC.m:13,16
    assert (p.Rc[0], p.Rc[1]) == (0.1, 5.0)


def test_no_cpu_fallback_without_device(lib):
    """Without a CUDA device the product path must fail loudly (skipped where a GPU exists)."""
    import pytest
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip('GPU present')
    except ImportError:
        pass
    import ekf_slam_b200 as E
    with pytest.raises(E.EkfError):
        E.EKF_SLAM()


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, 'ekf_slam_b200')
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith(('.py', '.cu', '.cuh', '.h', '.cpp')):
                txt = open(os.path.join(dp, fn)).read()
                assert not re.search(r'^\s*(from|import)\s+\.*oracle', txt, flags=re.M), fn
                assert 'oracle/' not in txt, fn
