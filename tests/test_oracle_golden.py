"""CPU: the C oracle against the committed golden vectors (outputs of the unmodified reference)."""
import pytest

import parity
from udp_pose_b200.configs import CONFIGS


@pytest.fixture(scope="module")
def impl():
    return parity.OracleImpl()


@pytest.mark.parametrize("cid", sorted(CONFIGS))
def test_crop_side(impl, cid):
    r = parity.check_crop_side(impl, cid)
    assert r["matrix_bit_exact"]  # same glibc sin/cos as the fixture generator


@pytest.mark.parametrize("cid", sorted(CONFIGS))
def test_encode(impl, cid):
    parity.check_encode(impl, cid)


@pytest.mark.parametrize("cid", sorted(CONFIGS))
def test_decode(impl, cid):
    parity.check_decode(impl, cid)


def test_adversarial(impl):
    parity.check_adversarial(impl)
