"""CPU: the C-ABI library loads and exports every symbol include/udp_codec.h declares, with
the argument counts the ctypes binding uses.  No compute calls (no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "udp_codec.h")


def declared():
    """{name: n_args} parsed from the header's prototypes."""
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"\bint\s+(udp_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else len(args.split(","))
    return out


@pytest.fixture(scope="module")
def lib():
    from udp_pose_b200 import _lib, build
    build.build()
    return ctypes.CDLL(_lib.LIB_PATH)


def test_header_and_binding_agree():
    from udp_pose_b200 import _lib
    d = declared()
    assert len(d) >= 14
    assert set(d) == set(_lib.SIGNATURES), set(d) ^ set(_lib.SIGNATURES)
    for name, n in d.items():
        assert len(_lib.SIGNATURES[name]) == n, name


def test_every_declared_symbol_is_exported(lib):
    for name in declared():
        assert getattr(lib, name) is not None


def test_version_and_arch(lib):
    assert lib.udp_codec_version() >= 100
    assert lib.udp_codec_sm_arch() == 100  # compiled for sm_100a only


def test_invalid_arguments_are_rejected_without_a_gpu(lib):
    """Argument validation happens before any CUDA call, so it can be exercised here."""
    from udp_pose_b200 import _lib
    L = _lib.load()
    assert L.udp_encode_gaussian(None, None, None, 17, 192, 256, 48, 64, 2.0, None, None, 4, None) == -1
    assert L.udp_warp_affine_u8(None, None, None, None, None, None, 191, 256, 1, None) == -1  # odd W
    assert L.udp_decode_offset(None, None, None, 3.5, None, None, 0, 17, 48, 64, None, None, None, None, None, -1, None) == -1
    # N == 0 is a no-op, not an error
    assert L.udp_encode_offset(None, None, None, 17, 192, 256, 48, 64, 3.5, None, None, 0, None) == 0


def test_no_cpu_fallback_in_product_package():
    """The product package must not import the oracle (or cv2) anywhere."""
    pkg = os.path.join(ROOT, "udp_pose_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), fn
            if fn not in ("reference_api.py",):  # reference_api reads image files with cv2.imread only
                assert "import cv2" not in src, fn
