"""The five BASELINE.json workload configurations (SURVEY.md section 8 header table).

Only codec parameters live here -- the fields the reference codec reads from its yacs tree
(deep_hrnet/lib/dataset/JointsDataset.py:86-114, lib/core/inference.py:152-159).
"""
from dataclasses import dataclass, field
from types import SimpleNamespace
from typing import List, Tuple

COCO_FLIP_PAIRS = [[1, 2], [3, 4], [5, 6], [7, 8], [9, 10], [11, 12], [13, 14], [15, 16]]  # coco.py:91-92
MPII_FLIP_PAIRS = [[0, 5], [1, 4], [2, 3], [10, 15], [11, 14], [12, 13]]  # mpii.py:30
RSN_SIGMAS = (2.6, 2.0, 1.7, 1.4105, 1.1105)  # RSN/dataset/JointsDataset.py:19-25


@dataclass(frozen=True)
class CodecConfig:
    name: str
    num_joints: int
    image_size: Tuple[int, int]      # (W, H)
    heatmap_size: Tuple[int, int]    # (w, h)
    target_type: str                 # 'gaussian' | 'offset' | 'rsn'
    sigma: float = 2
    kpd: float = 4.0
    post_process: bool = True
    flip_pairs: List[List[int]] = field(default_factory=list)
    cs_dtype: str = "float32"        # dtype of db center/scale (coco.py:223-227 / mpii.py:55-56)
    rot_factor: float = 45
    pre_div: float = 0.0             # RSN prm test.py:185 divides the maps by 255.0

    @property
    def channels(self):
        return self.num_joints * (3 if self.target_type == "offset" else 1)

    def as_cfg(self):
        """An attribute tree shaped like the reference's yacs config (any object works there)."""
        S = SimpleNamespace
        return S(
            MODEL=S(TARGET_TYPE="offset" if self.target_type == "offset" else "gaussian",
                    IMAGE_SIZE=list(self.image_size), HEATMAP_SIZE=list(self.heatmap_size),
                    SIGMA=self.sigma, NUM_JOINTS=self.num_joints),
            LOSS=S(KPD=self.kpd, USE_DIFFERENT_JOINTS_WEIGHT=False),
            TEST=S(POST_PROCESS=self.post_process, FLIP_TEST=True),
            DATASET=S(SCALE_FACTOR=0.35, ROT_FACTOR=self.rot_factor, FLIP=True,
                      NUM_JOINTS_HALF_BODY=8, PROB_HALF_BODY=0.0, COLOR_RGB=True),
        )


CONFIGS = {
    1: CodecConfig("coco_w32_256x192_udp", 17, (192, 256), (48, 64), "gaussian", sigma=2,
                   flip_pairs=COCO_FLIP_PAIRS),
    2: CodecConfig("coco_w32_256x192_udpv1", 17, (192, 256), (48, 64), "offset", kpd=3.5,
                   flip_pairs=COCO_FLIP_PAIRS),
    3: CodecConfig("coco_w48_384x288_udp", 17, (288, 384), (72, 96), "gaussian", sigma=3,
                   flip_pairs=COCO_FLIP_PAIRS),
    4: CodecConfig("mpii_w32_256x256_udp", 16, (256, 256), (64, 64), "gaussian", sigma=2,
                   flip_pairs=MPII_FLIP_PAIRS, cs_dtype="float64", rot_factor=30),
    5: CodecConfig("rsn18_256x192_udp", 17, (192, 256), (48, 64), "rsn", sigma=2,
                   flip_pairs=COCO_FLIP_PAIRS, pre_div=255.0),
}
