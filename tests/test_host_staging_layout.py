"""CPU tests of the host-staging helpers (no GPU, no kernel call): the block carver DexStepEnv.alloc_host_buffers / step_host
share between the pinned and the device side, and the compact / packed body layouts the kernels are pointed at."""
import torch

from dexrobot_isaac_b200.params import BodyLayout
from dexrobot_isaac_b200.pipeline import carve_block, result_segments


def test_carve_block_offsets_and_views():
    segs = [("a", (100, 26, 2), torch.float32), ("b", (100, 12, 7), torch.float32), ("c", (100,), torch.bool),
            ("d", (100, 18), torch.float32)]
    block, views = carve_block(segs)
    block2, views2 = carve_block(segs)
    assert block.dtype == torch.uint8 and block.numel() % 256 == 0 and int(block.sum()) == 0
    end = 0
    for name, shape, dtype in segs:
        v = views[name]
        assert tuple(v.shape) == shape and v.dtype == dtype and v.is_contiguous()
        off = v.data_ptr() - block.data_ptr()
        assert off % 256 == 0 and off >= end  # aligned, in order, no overlap
        assert off == views2[name].data_ptr() - block2.data_ptr()  # same list -> same offsets (host block == device block)
        end = off + v.numel() * v.element_size()
    assert end <= block.numel() < end + 256
    views["b"].fill_(1.0)
    views["c"].fill_(True)
    assert float(views["a"].sum()) == 0.0 and float(views["d"].sum()) == 0.0  # the views do not alias each other
    assert int(block.ne(0).sum()) == 100 * 12 * 7 * 2 + 100  # 0x3f800000: two non-zero bytes per float


def test_result_segments_are_what_a_step_hands_back():
    names = [n for n, _, _ in result_segments(64, 158)]
    assert names == ["obs_buf", "rew_buf", "reset_buf"]
    _, v = carve_block(result_segments(64, 158))
    assert tuple(v["obs_buf"].shape) == (64, 158) and v["reset_buf"].dtype == torch.bool
    assert v["obs_buf"].data_ptr() % 16 == 0  # the kernels store a full obs tile with one 16-byte aligned bulk copy


def test_compact_and_packed_layouts():
    full = BodyLayout.default(with_box=True)
    rigid, contact, actors = full.compact_gather_indices()
    assert rigid == [full.hand_body] + list(full.tip_bodies) + list(full.pad_bodies)
    assert contact == list(full.contact_bodies) + [full.box_body] and actors == [full.box_actor]
    c = full.compact()
    assert (c.num_bodies, c.rigid_row_floats, c.hand_body) == (11, 13, 0)
    assert list(c.tip_bodies) == [1, 2, 3, 4, 5] and list(c.pad_bodies) == [6, 7, 8, 9, 10]
    assert (c.contact_rows, c.box_body, c.num_actors, c.box_actor) == (6, 5, 1, 0)
    p = full.compact(packed=True)
    assert (p.num_bodies, p.rigid_row_floats, p.hand_body) == (12, 7, 0)
    # the 13-float hand row occupies slots 0-1; no pose slot overlaps it or another one, all inside the env's 84 floats
    spans = [(p.hand_body * 7, p.hand_body * 7 + 13)] + [(b * 7, b * 7 + 7) for b in list(p.tip_bodies) + list(p.pad_bodies)]
    spans.sort()
    assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:])) and spans[-1][1] == p.num_bodies * 7
    nobox = BodyLayout.default(with_box=False).compact(packed=True)
    assert (nobox.contact_rows, nobox.box_body, nobox.box_actor) == (5, -1, -1)
