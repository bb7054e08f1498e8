"""ctypes binding of libudp_codec.so (the C ABI declared in include/udp_codec.h).

There is no CPU fallback: if the shared library is missing or a call fails, this raises.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# UDP_CODEC_LIB: another build of the same library (kernel tuning experiments); no fallback either way
LIB_PATH = os.environ.get("UDP_CODEC_LIB") or os.path.join(HERE, "libudp_codec.so")

_vp, _i, _d, _f = C.c_void_p, C.c_int, C.c_double, C.c_float

# name -> argtypes, exactly the prototypes of include/udp_codec.h
SIGNATURES = {
    "udp_codec_version": [],
    "udp_codec_sm_arch": [],
    "udp_warp_matrices": [_vp, _vp, _i, _vp, _i, _i, _vp, _i, _vp],
    "udp_warp_affine_u8": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp],
    "udp_warp_affine_launches": [_i, _i],
    "udp_accuracy_pck": [_vp, _vp, _i, _i, _i, _i, _i, C.c_double, _vp, _vp, _vp, _vp, _vp, _vp],
    "udp_accuracy_pck_noalloc": [_vp, _vp, _i, _i, _i, _i, _i, C.c_double, _vp, _vp, _vp, _vp, _vp],
    "udp_pack_all_boxes": [_vp, _vp, _i, _vp, _vp, _i, _vp],
    "udp_decode_quarter": [_vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp],
    "udp_rsn_get_results": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _d, _vp, _vp, _i, _vp],
    "udp_aid_erase": [_vp, _i, _i, _i, _vp, _vp, _i, _vp, _vp, _i, _vp],
    "udp_coco_rescore": [_vp, _vp, _i, _i, _d, _vp, _vp],
    "udp_oks_nms": [_vp, _vp, _vp, _vp, _i, _i, _i, _d, _vp, _i, _d, _i, _i, _vp, _vp, _vp, _vp],
    "udp_warp_affine_normalize_f32": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp],
    "udp_rotate_points": [_vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _i, _i, _i, _vp],
    "udp_rsn_transform_joints": [_vp, _i, _vp, _i, _vp, _vp, _vp, _i, _i, _vp, _i, _i, _vp],
    "udp_encode_gaussian": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _d, _vp, _vp, _i, _vp],
    "udp_encode_offset": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _d, _vp, _vp, _i, _vp],
    "udp_encode_rsn": [_vp, _i, _vp, _i, _i, _i, _i, _i, _vp, _i, _vp, _i, _vp],
    "udp_flip_back": [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _i, _vp],
    "udp_decode_gaussian": [_vp, _vp, _vp, _vp, _vp, _i, _f, _i, _i, _i, _i, _vp, _vp, _vp, _vp,
                            _vp, _vp, _i, _vp],
    "udp_decode_offset": [_vp, _vp, _vp, _d, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp,
                          _i, _vp],
    "udp_decode_offset_split": [_vp, _vp, _vp, _vp, _vp, _d, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp,
                                _i, _vp],
    "udp_copy2d_async": [_vp, C.c_size_t, _vp, C.c_size_t, C.c_size_t, C.c_size_t, _i, _vp],
    "udp_host_device_pointer": [_vp, _vp],
    "udp_fliplr_joints": [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp],
    "udp_transform_preds": [_vp, _i, _vp, _vp, _i, _i, _i, _vp, _i, _i, _vp],
}

_lib = None


class UdpCodecError(RuntimeError):
    pass


def load():
    """Load the library (once).  Raises if it has not been built -- never falls back."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise UdpCodecError(
            "%s not found: build it with `python -m udp_pose_b200.build` "
            "(there is no CPU fallback)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so is stale
        fn.argtypes = args
        fn.restype = C.c_int
    _lib = lib
    return lib


def check(rc, name):
    if rc == 0:
        return
    if rc < 0:
        raise UdpCodecError("%s: invalid argument (UDP_EINVAL)" % name)
    raise UdpCodecError("%s: CUDA error %d" % (name, rc))


def call(name, *args):
    check(getattr(load(), name)(*args), name)
