"""ctypes binding of libekfslam_b200.so (C ABI declared in include/ekfslam_b200.h).

There is no CPU fallback: if the CUDA library is missing this module raises at load
time, and ekf_create fails on a machine without a CUDA device.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('EKFSLAM_B200_LIB') or os.path.join(_HERE, 'libekfslam_b200.so')      # override: A/B of two builds on one box

EKF_OK, EKF_EINVAL, EKF_ECAPACITY, EKF_ECUDA, EKF_ENCCL, EKF_ENONFINITE = 0, -1, -2, -3, -4, -5
VARIANT_KNOWN, VARIANT_UC = 0, 1
ASSOC_SIGNATURE, ASSOC_POSITION = 0, 1
STORAGE_FULL, STORAGE_SYM = 0, 1


class EkfParams(C.Structure):
    _fields_ = [('C', C.c_double), ('Rc', C.c_double * 2), ('s_cost', C.c_double),
                ('s_thresh', C.c_double), ('variant', C.c_int32), ('assoc_mode', C.c_int32),
                ('storage', C.c_int32), ('capacity_landmarks', C.c_int32), ('device', C.c_int32),
                ('defer_m', C.c_int32), ('stream', C.c_uint64)]


class EkfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__('ekfslam_b200 error %d: %s' % (code, msg))
        self.code = code


_DP = C.POINTER(C.c_double)
_IP = C.POINTER(C.c_int32)
_H = C.c_void_p

# name -> argtypes ; every function returns int32 unless listed in _RESTYPE
SIGNATURES = {
    'ekf_params_default': [C.POINTER(EkfParams), C.c_int32],
    'ekf_create': [C.POINTER(EkfParams), C.POINTER(_H)],
    'ekf_destroy': [_H],
    'ekf_nccl_unique_id': [C.POINTER(C.c_uint8)],
    'ekf_create_sharded': [C.POINTER(EkfParams), C.c_int32, C.c_int32, C.POINTER(C.c_uint8), C.POINTER(_H)],
    'ekf_create_multi': [C.POINTER(EkfParams), C.c_int32, _IP, C.POINTER(_H)],
    'ekf_shard_info': [_H, _IP, _IP, _IP, _IP],
    'ekf_slab_range': [C.c_int32, C.c_int32, C.c_int32, C.c_int32, _IP, _IP],
    'ekf_update_batch': [_H, C.c_int32, _DP, _DP, _IP],
    'ekf_predict': [_H, _DP],
    'ekf_update': [_H, _DP, _DP, C.c_int32],
    'ekf_append': [_H, _DP, _DP, _DP, C.c_double],
    'ekf_associate': [_H, _DP, _DP, _IP, _IP],
    'ekf_associate_static': [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, _DP, _DP, C.c_double, C.c_double,
                             C.c_int32, C.c_int32, C.c_void_p, C.c_uint64, C.c_char_p, C.c_int32],
    'ekf_measure_batch': [_H, _DP, C.c_int32, _DP, _DP, _IP, C.c_int32, _IP],
    'ekf_measure_batch_async': [_H, _DP, C.c_int32, _DP, _DP, _IP, C.c_int32, _IP],
    'ekf_measure_wait': [_H],
    'ekf_n_landmarks': [_H, _IP],
    'ekf_get_x': [_H, _DP, C.c_int32],
    'ekf_get_s': [_H, _DP, C.c_int32],
    'ekf_get_P': [_H, _DP, C.c_int32],
    'ekf_get_P_block': [_H, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _DP, C.c_int32],
    'ekf_get_ellipses': [_H, _DP, C.c_int32],
    'ekf_set_state': [_H, _DP, _DP, C.c_int32, _DP, C.c_int32],
    'ekf_set_x': [_H, _DP, C.c_int32],
    'ekf_set_P': [_H, _DP, C.c_int32],
    'ekf_set_s': [_H, _DP, C.c_int32],
    'ekf_set_state_lowrank': [_H, _DP, _DP, C.c_int32, _DP, C.c_int32, _DP],
    'ekf_set_noise': [_H, C.c_double, _DP, C.c_double, C.c_double],
    'ekf_defer_info': [_H, _IP, _IP],
    'ekf_flush_info': [_H, _IP, _IP],
    'ekf_flush': [_H],
    'ekf_sync': [_H],
    'ekf_last_error': [_H],
    'ekf_profile_enable': [_H, C.c_int32],
    'ekf_profile_read': [_H, _DP, C.POINTER(C.c_int64), C.POINTER(C.c_int64)],
    'ekf_profile_exchange': [_H, _DP, C.POINTER(C.c_int64), _IP],
    'ekf_downdate_bytes': [_H, _DP],
    'ekf_version': [],
    # Monte-Carlo batch
    'ekfb_create': [C.POINTER(EkfParams), C.c_int32, C.c_int32, C.c_int32, C.POINTER(_H)],
    'ekfb_destroy': [_H],
    'ekfb_step': [_H, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                  C.c_int32, C.c_int32, _IP],
    'ekfb_sync': [_H],
    'ekfb_n_landmarks': [_H, _IP],
    'ekfb_get_x': [_H, C.c_int32, _DP, C.c_int32],
    'ekfb_get_s': [_H, C.c_int32, _DP, C.c_int32],
    'ekfb_get_P': [_H, C.c_int32, _DP, C.c_int32],
    'ekfb_set_state': [_H, C.c_int32, _DP, _DP, C.c_int32, _DP, C.c_int32],
    'ekfb_get_poses': [_H, _DP],
    'ekfb_downdate_bytes': [_H, _DP],
    'ekfb_profile': [_H, C.c_int32, _DP, C.POINTER(C.c_int64), _DP],
    'ekfb_last_error': [_H],
}
_RESTYPE = {'ekf_last_error': C.c_char_p, 'ekf_version': C.c_char_p, 'ekfb_last_error': C.c_char_p}

_lib = None


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError('%s not found: build it with `python -c "import __graft_entry__ as g; g.build()"` '
                          '(or make -C ekf_slam_b200/csrc); there is no CPU fallback' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name)           # AttributeError if the .so lacks a declared symbol
        fn.argtypes = args
        fn.restype = _RESTYPE.get(name, C.c_int32)
    _lib = lib
    return lib


def check(h, rc):
    if rc != EKF_OK:
        msg = load().ekf_last_error(h) if h else b''
        raise EkfError(rc, (msg or b'').decode('utf-8', 'replace'))


def dptr(a):
    return a.ctypes.data_as(_DP)


def iptr(a):
    return a.ctypes.data_as(_IP)
