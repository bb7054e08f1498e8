"""Import the UNMODIFIED reference hot path from /root/reference (this container only).

TEST INFRASTRUCTURE -- never imported by the product package.  `/root/reference` does not
exist on the GPU box, so this module is used only by `tests/golden/make_golden.py` (to
generate the committed fixtures) and by `tests/test_oracle_vs_reference.py` (skipped when
the reference tree is absent).

Two harness-side shims, nothing in the reference is edited:
  1. `np.float = float`  -- deep_hrnet/lib/core/inference.py:136 uses the removed alias.
  2. deep_hrnet/lib/dataset/JointsDataset.py is loaded by file path because the package
     `dataset/__init__.py` pulls in pycocotools / json_tricks, which are not installed.
"""
import importlib.util
import os
import sys
import types

import numpy as np

REF_ROOT = os.environ.get("UDP_REFERENCE_ROOT", "/root/reference")
_HR_LIB = os.path.join(REF_ROOT, "deep_hrnet", "lib")
_RSN = os.path.join(REF_ROOT, "RSN")


def available():
    return os.path.isfile(os.path.join(_HR_LIB, "core", "inference.py"))


_cache = {}


def _load_by_path(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def hrnet():
    """Returns a namespace with the reference modules: inference, transforms, joints_dataset."""
    if "hr" in _cache:
        return _cache["hr"]
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    if not hasattr(np, "float"):
        np.float = float  # shim 1
    if _HR_LIB not in sys.path:
        sys.path.insert(0, _HR_LIB)
    import core.inference as inference  # noqa: E402
    import utils.transforms as transforms  # noqa: E402
    jd = _load_by_path("_ref_JointsDataset",
                       os.path.join(_HR_LIB, "dataset", "JointsDataset.py"))  # shim 2
    ns = types.SimpleNamespace(inference=inference, transforms=transforms, joints_dataset=jd)
    _cache["hr"] = ns
    return ns


def rsn():
    """RSN twin: dataset/JointsDataset.py (generate_heatmap, get_warpmatrix_inverse) and
    lib/utils/transforms.py (flip_back, affine_transform)."""
    if "rsn" in _cache:
        return _cache["rsn"]
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    # RSN/dataset/JointsDataset.py does `from lib.utils.transforms import ...`; deep_hrnet/lib
    # is itself importable as top-level packages, so give RSN's `lib` a private name space.
    saved = {k: sys.modules.get(k) for k in ("lib", "lib.utils", "lib.utils.transforms")}
    tr = _load_by_path("_rsn_transforms", os.path.join(_RSN, "lib", "utils", "transforms.py"))
    lib = types.ModuleType("lib")
    lib_utils = types.ModuleType("lib.utils")
    lib.utils = lib_utils
    lib_utils.transforms = tr
    sys.modules["lib"], sys.modules["lib.utils"], sys.modules["lib.utils.transforms"] = lib, lib_utils, tr
    try:
        jd = _load_by_path("_rsn_JointsDataset", os.path.join(_RSN, "dataset", "JointsDataset.py"))
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    ns = types.SimpleNamespace(transforms=tr, joints_dataset=jd)
    _cache["rsn"] = ns
    return ns


def make_cfg(target_type, image_size, heatmap_size, sigma, kpd=4.0, post_process=True,
             flip_test=True):
    """A SimpleNamespace config carrying the fields the codec reads
    (deep_hrnet/lib/dataset/JointsDataset.py:86-114, core/inference.py:152-159)."""
    S = types.SimpleNamespace
    return S(
        OUTPUT_DIR="",
        DATASET=S(DATA_FORMAT="jpg", SCALE_FACTOR=0.35, ROT_FACTOR=45, FLIP=True,
                  NUM_JOINTS_HALF_BODY=8, PROB_HALF_BODY=0.0, COLOR_RGB=True,
                  CUTOUT=None, HIDE_AND_SEEK=None),
        MODEL=S(TARGET_TYPE=target_type, IMAGE_SIZE=list(image_size),
                HEATMAP_SIZE=list(heatmap_size), SIGMA=sigma),
        LOSS=S(USE_DIFFERENT_JOINTS_WEIGHT=False, KPD=kpd),
        TEST=S(POST_PROCESS=post_process, FLIP_TEST=flip_test),
    )


def make_dataset(cfg, num_joints):
    """JointsDataset(cfg, '', 'val', False) with num_joints set (the subclass normally does)."""
    ds = hrnet().joints_dataset.JointsDataset(cfg, "", "val", False)
    ds.num_joints = num_joints
    return ds


def make_rsn_dataset(keypoint_num, input_shape, output_shape):
    """Object carrying what RSN generate_heatmap reads (RSN/dataset/JointsDataset.py:276-291);
    input_shape/output_shape are (h, w) as in RSN/dataset/attribute.py."""
    jd = rsn().joints_dataset
    ds = object.__new__(jd.JointsDataset)
    ds.keypoint_num = keypoint_num
    ds.input_shape = tuple(input_shape)
    ds.output_shape = tuple(output_shape)
    fx, fy = np.meshgrid(np.arange(0, output_shape[1]), np.arange(0, output_shape[0]))
    ds.feat_x_int = fx.reshape((-1,))
    ds.feat_y_int = fy.reshape((-1,))
    ds.gaussian_kernels = [(15, 15), (11, 11), (9, 9), (7, 7), (5, 5)]
    return ds
