"""Known-answer vectors for the hand-frame transforms: the 7 cases the reference's own
utils/test_coordinate_transforms.py:22-237 pins (identity, translation, 90 deg Z, T+R, 90 deg Y,
round trip with euler (0.3, 0.4, 0.5), batched mixed rotations) plus the documented ARR-alignment
identity (docs/DESIGN_DECISIONS.md:52-68, observation_encoder.py:1487-1488).  Run against the CPU
oracle and the isaacgym stub; tests/test_gpu_parity.py runs the same vectors through the CUDA
kernel."""
import math
import os
import sys

import pytest
import torch

from oracle import dexstep_oracle as O

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "stubs"))
from isaacgym import torch_utils as stub  # noqa: E402

S = 0.7071067811865476

CASES = [  # hand_pos, hand_quat(xyzw), point_world, expected_local, atol
    ([0, 0, 0], [0, 0, 0, 1], [1, 0, 0], [1, 0, 0], 1e-6),
    ([1, 1, 1], [0, 0, 0, 1], [2, 1, 1], [1, 0, 0], 1e-6),
    ([0, 0, 0], [0, 0, S, S], [0, 1, 0], [1, 0, 0], 1e-5),
    ([1, 1, 1], [0, 0, S, S], [1, 2, 1], [1, 0, 0], 1e-5),
    ([0, 0, 0], [0, S, 0, S], [1, 0, 0], [0, 0, 1], 1e-5),
    ([0, 0, 0], [0, S, 0, S], [0, 1, 0], [0, 1, 0], 1e-5),
    ([0, 0, 0], [0, S, 0, S], [0, 0, 1], [-1, 0, 0], 1e-5),
    ([1, 0, 0], [0, S, 0, S], [2, 0, 0], [0, 0, 1], 1e-5),
]


def known_answer_tensors():
    hp = torch.tensor([c[0] for c in CASES], dtype=torch.float32)
    hq = torch.tensor([c[1] for c in CASES], dtype=torch.float32)
    pw = torch.tensor([c[2] for c in CASES], dtype=torch.float32)
    ex = torch.tensor([c[3] for c in CASES], dtype=torch.float32)
    return hp, hq, pw, ex


def test_point_in_hand_frame_known_answers():
    hp, hq, pw, ex = known_answer_tensors()
    got = O.point_in_hand_frame(pw, hp, hq)
    assert torch.allclose(got, ex, atol=1e-5)
    got_stub = stub.quat_rotate_inverse(hq, pw - hp)
    assert torch.equal(got, got_stub)


def test_round_trip_euler():
    e = torch.tensor([[0.3, 0.4, 0.5]]).repeat(3, 1)
    q = stub.quat_from_euler_xyz(e[:, 0], e[:, 1], e[:, 2])
    hp = torch.tensor([[1.0, 2.0, 3.0]]).repeat(3, 1)
    p = torch.tensor([[4.0, 5.0, 6.0]]).repeat(3, 1)
    back = O.point_in_world_frame(O.point_in_hand_frame(p, hp, q), hp, q)
    assert torch.allclose(back, p, atol=1e-5)
    assert torch.allclose(q.norm(dim=1), torch.ones(3), atol=1e-6)


def test_arr_alignment_identity():
    """ARR = 0 => hand quat = [0, s, 0, s] => aligned quat = identity [0, 0, 0, 1]."""
    q_hand = torch.tensor([[0.0, S, 0.0, S]])
    inv = O.quat_conjugate(torch.tensor([0.0, S, 0.0, S])).unsqueeze(0)
    q = O.quat_mul(q_hand, inv)
    assert torch.allclose(q, torch.tensor([[0.0, 0.0, 0.0, 1.0]]), atol=1e-6)
    assert torch.equal(q, stub.quat_mul(q_hand, stub.quat_conjugate(torch.tensor([[0.0, S, 0.0, S]]))))


def test_quat_mul_matches_hamilton_product():
    g = torch.Generator().manual_seed(0)
    a = torch.randn(64, 4, generator=g, dtype=torch.float64)
    b = torch.randn(64, 4, generator=g, dtype=torch.float64)
    x1, y1, z1, w1 = a.unbind(-1)
    x2, y2, z2, w2 = b.unbind(-1)
    ref = torch.stack([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                       w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2], dim=-1)
    assert torch.allclose(O.quat_mul(a, b), ref, atol=1e-12)


def test_quaternion_algebra_against_scipy():
    """isaacgym.torch_utils is a binary-only dependency that is absent here; its functions are restated in
    oracle/stubs/isaacgym/torch_utils.py (and again in the oracle / the CUDA kernels).  Independent pin of the convention
    (scalar-last x, y, z, w; Hamilton product; active rotation): scipy.spatial.transform.Rotation uses the same convention,
    so  quat_mul(a, b) == (R(a) * R(b)).as_quat(),  quat_conjugate == inverse,  quat_rotate == apply,
    quat_rotate_inverse == inverse().apply,  quat_from_euler_xyz == from_euler("xyz") (extrinsic), on 256 random unit
    quaternions, to float32 accuracy and up to the quaternion double cover (q ~ -q)."""
    R = pytest.importorskip("scipy.spatial.transform").Rotation
    g = torch.Generator().manual_seed(3)
    a = torch.randn(256, 4, generator=g)
    b = torch.randn(256, 4, generator=g)
    a, b = a / a.norm(dim=1, keepdim=True), b / b.norm(dim=1, keepdim=True)
    v = torch.randn(256, 3, generator=g)
    Ra, Rb = R.from_quat(a.double().numpy()), R.from_quat(b.double().numpy())

    def same_rotation(q_got, q_want):
        q_want = torch.from_numpy(q_want).float()
        sign = torch.sign((q_got * q_want).sum(dim=1, keepdim=True))
        return torch.allclose(q_got, q_want * sign, atol=2e-6)

    for mul, conj in ((stub.quat_mul, stub.quat_conjugate), (O.quat_mul, O.quat_conjugate)):
        assert same_rotation(mul(a, b), (Ra * Rb).as_quat())
        assert same_rotation(conj(a), Ra.inv().as_quat())
    assert torch.allclose(stub.quat_rotate(a, v), torch.from_numpy(Ra.apply(v.double().numpy())).float(), atol=3e-6)
    assert torch.allclose(stub.quat_rotate_inverse(a, v), torch.from_numpy(Ra.inv().apply(v.double().numpy())).float(), atol=3e-6)
    assert torch.allclose(O.point_in_hand_frame(v, torch.zeros(256, 3), a),
                          torch.from_numpy(Ra.inv().apply(v.double().numpy())).float(), atol=3e-6)
    e = torch.rand(256, 3, generator=g) * 2.0 - 1.0
    got = stub.quat_from_euler_xyz(e[:, 0], e[:, 1], e[:, 2])
    assert same_rotation(got, R.from_euler("xyz", e.double().numpy()).as_quat())
