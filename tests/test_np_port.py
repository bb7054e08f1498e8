"""CPU: the numpy/cv2 port that bench.py times as the CPU baseline, against the golden
vectors of the unmodified reference (so the timed baseline computes the reference's results)."""
import numpy as np
import pytest

pytest.importorskip("cv2")
import parity  # noqa: E402
from oracle import np_port as P  # noqa: E402
from udp_pose_b200.configs import CONFIGS  # noqa: E402


class NpPortImpl:
    def warp_matrices(self, rot, c, s, W, H):
        return np.stack([P.unbiased_matrix(float(rot[i]), c[i], s[i], W, H) for i in range(len(rot))])

    def rotate_points(self, jxy, rot, c, s, W, H):
        return np.stack([P.map_joints(jxy[i].copy(), float(rot[i]), c[i], s[i], W, H) for i in range(len(rot))])

    def warp_affine(self, images, index, M, W, H):
        import cv2
        return np.stack([cv2.warpAffine(images[index[i]], M[i], (W, H), flags=cv2.WARP_INVERSE_MAP | cv2.INTER_LINEAR)
                         for i in range(len(M))])

    def encode_gaussian(self, j, v, W, H, w, h, sigma, jw=None):
        r = [P.target_gaussian(j[i], v[i], W, H, w, h, sigma) for i in range(len(j))]
        t, tw = np.stack([a for a, _ in r]), np.stack([b for _, b in r])
        return t, (tw * np.asarray(jw, np.float32).reshape(-1, 1) if jw is not None else tw)

    def encode_offset(self, j, v, W, H, w, h, kpd, jw=None):
        r = [P.target_offset(j[i], v[i], W, H, w, h, kpd) for i in range(len(j))]
        t, tw = np.stack([a for a, _ in r]), np.stack([b for _, b in r])
        return t, (tw * np.asarray(jw, np.float32).reshape(-1, 1) if jw is not None else tw)

    def encode_rsn(self, j, valid, W, H, w, h):
        return np.stack([P.labels_rsn(j[i], valid[i], W, H, w, h) for i in range(len(j))])

    def rsn_transform_joints(self, j, v, rot, c, s, W, H):
        return parity.OracleImpl().rsn_transform_joints(j, v, rot, c, s, W, H)

    def decode(self, cfg, out, flipped, c, s, post_process=True):
        trip = cfg.target_type == "offset"
        hm = out.copy() if flipped is None else P.unflip_average(out, flipped, cfg.flip_pairs, trip)
        with np.errstate(all="ignore"):
            p, mv, pin = P.final_preds(hm, c, s, "offset" if trip else "gaussian", cfg.kpd, post_process, cfg.pre_div)
        return dict(preds=p.astype(np.float64), maxvals=mv, preds_in=pin.astype(np.float64))


@pytest.mark.parametrize("cid", sorted(CONFIGS))
def test_np_port_golden(cid):
    impl = NpPortImpl()
    parity.check_crop_side(impl, cid)
    parity.check_encode(impl, cid)
    parity.check_decode(impl, cid)
