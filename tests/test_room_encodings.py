"""Host-side encodings around the Sokoban step (CPU): the nibble rooms ragen_sokoban_step_compact keeps in HBM and the
three-bit / nibble glyph rows it and its siblings write.  The kernels are checked against the oracle in test_gpu_sokoban.py;
here the host converters are checked against plain numpy restatements of the layouts include/ragen_b200.h describes."""
import numpy as np
import pytest
import torch

from oracle.sokoban_np import SokobanBatchOracle, synthetic_rooms
from ragen_b200.env.sokoban import pack_rooms, packed3_bytes, unpack3_glyphs, unpack_glyphs, unpack_rooms


@pytest.mark.parametrize("H,W,nb,n", [(6, 6, 1, 33), (8, 8, 2, 17), (10, 10, 2, 65)])
def test_nibble_rooms_round_trip(H, W, nb, n):
    _, state = synthetic_rooms(n, H, W, nb, seed=n)
    st = torch.from_numpy(state.reshape(n, -1).copy())
    s4 = pack_rooms(st)
    assert s4.shape == (n, H * W // 2) and s4.dtype == torch.uint8
    ref = state.reshape(n, -1)
    assert np.array_equal(s4.numpy(), ref[:, 0::2] | (ref[:, 1::2] << 4))  # cell 2i low nibble, cell 2i+1 high nibble
    assert torch.equal(unpack_rooms(s4), st)
    out = torch.zeros_like(st)
    assert unpack_rooms(s4, out=out) is out and torch.equal(out, st)
    buf = torch.zeros_like(s4)
    pack_rooms(st, out=buf)
    assert torch.equal(buf, s4)


@pytest.mark.parametrize("H,W,nb,n", [(6, 6, 1, 9), (8, 8, 2, 5), (10, 10, 2, 7)])
def test_three_bit_and_nibble_rows_unpack_to_the_glyph_rows(H, W, nb, n):
    fixed, state = synthetic_rooms(n, H, W, nb, seed=n + 3)
    o = SokobanBatchOracle(fixed, state, H, W, nb, 50)
    rng = np.random.default_rng(n)
    for _ in range(6):
        o.step_turn(rng.integers(1, 5, (n, 2)).astype(np.uint8), None)
    g = o.glyphs()  # [n, H*(W+1)] codes, 7 = row separator, 6 = player on a target
    cells = g.reshape(n, H, W + 1)[:, :, :W].reshape(n, -1)
    # three bits per cell, little-endian bit stream, padded to an even byte count
    B = packed3_bytes(H, W)
    assert B == ((3 * H * W + 15) // 16) * 2
    bits = np.zeros((n, B * 8), dtype=np.uint8)
    for k in range(3):
        bits[:, k: 3 * H * W: 3] = (cells >> k) & 1
    rows3 = np.packbits(bits, axis=1, bitorder="little")
    assert np.array_equal(unpack3_glyphs(rows3, H, W), g)
    # two codes per byte, low nibble first
    rows4 = (g[:, 0::2] | (g[:, 1::2] << 4)).astype(np.uint8)
    assert np.array_equal(unpack_glyphs(rows4), g)
