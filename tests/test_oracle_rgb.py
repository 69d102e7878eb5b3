"""The oracle's room_to_rgb restatement on a hand-checked room (CPU)."""
import numpy as np

from oracle import sokoban as S


def test_room_to_rgb_places_sprites():
    surfaces = np.zeros((7, 2, 2, 3), dtype=np.uint8)
    for k in range(7):
        surfaces[k] = 10 * k + np.arange(12, dtype=np.uint8).reshape(2, 2, 3)
    state = np.array([[0, 0, 0], [1, 5, 4], [3, 2, 0]])
    fixed = np.array([[0, 0, 0], [1, 2, 1], [2, 2, 0]])
    img = S.room_to_rgb(state, fixed, surfaces)
    assert img.shape == (6, 6, 3) and img.dtype == np.uint8
    want_ids = np.array([[0, 0, 0], [1, 6, 4], [3, 2, 0]])  # the player stands on a target
    for i in range(3):
        for j in range(3):
            assert np.array_equal(img[2 * i:2 * i + 2, 2 * j:2 * j + 2], surfaces[want_ids[i, j]])
    # without the structure the player keeps surface 5 (room_to_rgb(room, None))
    assert np.array_equal(S.room_to_rgb(state, None, surfaces)[2:4, 2:4], surfaces[5])
    assert state[1, 1] == 5  # the input is not modified
