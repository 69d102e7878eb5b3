"""Sokoban on the GPU: device-resident batch + per-env views with the reference API.

`SokobanBatch` holds n rooms as struct-of-arrays uint8 in HBM and advances all of them with one
kernel launch (`ragen_sokoban_step`).  `SokobanEnv` keeps the constructor / reset / step / render /
get_all_actions contract of ragen/env/sokoban/env.py:10-67 and is a view onto one slot of a batch,
so it can be registered under REGISTERED_ENVS['sokoban'] unchanged.
"""
from dataclasses import dataclass, field
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from .. import _lib
from .base import BaseDiscreteActionEnv

NEWLINE = 7


@dataclass
class SokobanEnvConfig:
    """Field-for-field the reference dataclass (ragen/env/sokoban/config.py:4-21)."""
    dim_room: Tuple[int, int] = (6, 6)
    max_steps: int = 100
    num_boxes: int = 3
    search_depth: int = 300
    grid_lookup: Optional[Dict[int, str]] = field(
        default_factory=lambda: {0: "#", 1: "_", 2: "O", 3: "√", 4: "X", 5: "P", 6: "S"})
    grid_vocab: Optional[Dict[str, str]] = field(
        default_factory=lambda: {"#": "wall", "_": "empty", "O": "target", "√": "box on target", "X": "box",
                                 "P": "player", "S": "player on target"})
    action_lookup: Optional[Dict[int, str]] = field(default_factory=lambda: {1: "Up", 2: "Down", 3: "Left", 4: "Right"})
    dim_x: Optional[int] = None
    dim_y: Optional[int] = None
    render_mode: str = "text"

    def __post_init__(self):
        if self.dim_x is not None and self.dim_y is not None:
            self.dim_room = (self.dim_x, self.dim_y)
            delattr(self, "dim_x")
            delattr(self, "dim_y")


SURFACE_FILES = ("wall", "floor", "box_target", "box_on_target", "box", "player", "player_on_target")


def load_sprite_atlas(directory=None):
    """uint8 [7, 16, 16, 3]: the PNG sprites room_to_rgb loads (gym_sokoban/envs/surface/*.png).  gym_sokoban is a third-party
    package the reference imports; when it is not installed the sprites do not exist and the caller must supply an atlas."""
    import importlib.util
    import os

    if directory is None:
        spec = importlib.util.find_spec("gym_sokoban")
        if spec is None or not spec.submodule_search_locations:
            raise RuntimeError("rgb_array rendering needs gym_sokoban's sprite PNGs (package not installed): "
                               "pass them with SokobanBatch.set_atlas(uint8[7, S, S, 3]) or load_sprite_atlas(directory)")
        directory = os.path.join(list(spec.submodule_search_locations)[0], "envs", "surface")
    from PIL import Image

    return np.stack([np.asarray(Image.open(os.path.join(directory, f + ".png")).convert("RGB"), dtype=np.uint8)
                     for f in SURFACE_FILES])


@dataclass
class StepOut:
    """Device tensors written by one step launch (views of the batch's output buffers)."""
    reward: torch.Tensor      # float64 [n] summed step rewards
    done: torch.Tensor        # uint8 [n]
    flags: torch.Tensor       # uint8 [n] bit0 effective, bit1 success, bit2 executed
    n_executed: torch.Tensor  # uint8 [n]
    obs: torch.Tensor         # uint8 [n, H*(W+1)] glyph rows


def glyph_table(grid_lookup):
    """index 0..7 -> str; 7 is the row separator, missing codes render as '?' (env.py:56)."""
    tab = [grid_lookup.get(i, "?") for i in range(7)] + ["\n"]
    return tab


def glyph_pair_table(grid_lookup):
    """For packed glyph rows (two 4-bit codes per byte, low nibble first): byte value -> the two glyph strings, as a
    str.translate table.  `packed_row.tobytes().decode("latin1").translate(table)[:-1]` is the env's text."""
    tab = glyph_table(grid_lookup) + ["?"] * 8
    return {b: tab[b & 15] + tab[b >> 4] for b in range(256)}


def packed_to_text(packed_row, pair_table):
    """packed_row: 1-D uint8 numpy array of length H*(W+1)/2; the trailing separator is dropped."""
    return packed_row.tobytes().decode("latin1").translate(pair_table)[:-1]


def unpack_glyphs(packed):
    """[n, G/2] packed rows -> [n, G] glyph codes (numpy; tests and debugging)."""
    p = np.asarray(packed)
    out = np.empty((p.shape[0], p.shape[1] * 2), dtype=np.uint8)
    out[:, 0::2] = p & 15
    out[:, 1::2] = p >> 4
    return out


def packed3_bytes(H, W):
    """Bytes per env of a three-bit glyph row (include/ragen_b200.h: ragen_sokoban_step_packed3)."""
    return ((3 * H * W + 15) // 16) * 2


def unpack3_glyphs(packed3, H, W):
    """[n, packed3_bytes] three-bit rows -> [n, H*(W+1)] glyph codes with the row separators (code 7) put back: the same array
    the byte and nibble formats unpack to (numpy)."""
    p = np.ascontiguousarray(np.asarray(packed3), dtype=np.uint8)
    n = p.shape[0]
    bits = np.unpackbits(p, axis=1, bitorder="little")[:, : 3 * H * W].reshape(n, H * W, 3)
    cells = (bits[..., 0] | (bits[..., 1] << 1) | (bits[..., 2] << 2)).astype(np.uint8).reshape(n, H, W)
    out = np.full((n, H, W + 1), 7, dtype=np.uint8)
    out[:, :, :W] = cells
    return out.reshape(n, H * (W + 1))


def pack_rooms(state, out=None):
    """room_state [n, H*W] (uint8 tensor, codes 0..5, H*W even) -> the nibble rooms ragen_sokoban_step_compact keeps in HBM:
    [n, H*W/2], cell 2i in the low nibble of byte i."""
    return torch.bitwise_or(state[:, 0::2], state[:, 1::2] << 4, out=out)


def unpack_rooms(state4, out=None):
    """Inverse of pack_rooms."""
    if out is None:
        out = torch.empty((state4.shape[0], 2 * state4.shape[1]), dtype=torch.uint8, device=state4.device)
    out[:, 0::2] = state4 & 15
    out[:, 1::2] = state4 >> 4
    return out


REWARD_CODES = (-0.1, -0.1 + 1, -0.1 + -1, -0.1 + 1 + 10, -0.1 + 10)  # gym_sokoban _calc_reward's sums, in its addition order


def decode_delta(delta):
    """delta uint8 [n, K, 4] (ragen_sokoban_step_delta) or [n, K] (the one-byte records of ragen_sokoban_step_delta1) ->
    (acc_reward f64 [n], done u8 [n], flags u8 [n], n_executed u8 [n]) as the other step entry points report them: rewards summed
    left to right in doubles, done / flags of the last executed action."""
    d = np.asarray(delta)
    n, K = d.shape[0], d.shape[1]
    info = d[:, :, 3] if d.ndim == 3 else d
    executed = (info >> 6) & 1
    table = np.array(REWARD_CODES + (0.0, 0.0, 0.0), dtype=np.float64)
    acc = np.zeros(n, dtype=np.float64)
    for k in range(K):
        acc = np.where(executed[:, k] == 1, acc + table[info[:, k] & 7], acc)
    nexec = executed.sum(axis=1).astype(np.uint8)
    last = info[np.arange(n), np.maximum(nexec.astype(np.int64) - 1, 0)]
    has = nexec > 0
    done = np.where(has, (last >> 3) & 1, 0).astype(np.uint8)
    flags = np.where(has, ((last >> 4) & 1) | (((last >> 5) & 1) << 1) | 4, 0).astype(np.uint8)
    return acc, done, flags, nexec


class HostMirror:
    """The rooms of a batch mirrored in host memory and kept current from step deltas: 4 bytes per action cross the
    device->host link instead of a re-rendered room.  `state` is room_state [n, H*W] (codes 0..5), `fixed` room_fixed."""

    def __init__(self, fixed, state, H, W):
        self.H, self.W = int(H), int(W)
        self.fixed = np.ascontiguousarray(fixed, dtype=np.uint8).reshape(-1, H * W).copy()
        self.state = np.ascontiguousarray(state, dtype=np.uint8).reshape(-1, H * W).copy()
        self.n = self.state.shape[0]

    def apply(self, delta, actions=None):
        """Patches the mirror from a turn's records: [n, K, 4] records carry the cells; the one-byte records [n, K] need the
        turn's `actions` [n, K] (the mirror knows where every player stands)."""
        d = np.asarray(delta)
        if d.ndim == 2:
            return self._apply1(d, np.asarray(actions).reshape(d.shape))
        for k in range(d.shape[1]):
            rec = d[:, k]
            rows = np.flatnonzero((rec[:, 3] >> 4) & 1)  # the player moved (a push moves it too)
            if rows.size == 0:
                continue
            before, after, box = rec[rows, 0].astype(np.int64), rec[rows, 1].astype(np.int64), rec[rows, 2].astype(np.int64)
            self.state[rows, before] = self.fixed[rows, before]
            self.state[rows, after] = 5
            if getattr(self, "_player", None) is not None:
                self._player[rows] = after
            pushed = box != 0xff
            pr, pc = rows[pushed], box[pushed]
            self.state[pr, pc] = np.where(self.fixed[pr, pc] == 2, 3, 4)

    def players(self):
        """Cell of every room's player (from the mirrored state; -1 where a room has none)."""
        has = (self.state == 5).any(axis=1)
        return np.where(has, (self.state == 5).argmax(axis=1), -1).astype(np.int64)

    def _apply1(self, d, actions):
        step = np.array([0, -self.W, self.W, -1, 1, 0, 0, 0], dtype=np.int64)  # config.py:12: 1 up, 2 down, 3 left, 4 right
        player = getattr(self, "_player", None)
        if player is None:
            player = self._player = self.players()
        for k in range(d.shape[1]):
            rec = d[:, k]
            rows = np.flatnonzero((rec >> 4) & 1)  # the player moved (a push moves it too)
            if rows.size == 0:
                continue
            dlt = step[actions[rows, k] & 7]
            before = player[rows]
            after = before + dlt
            self.state[rows, before] = self.fixed[rows, before]
            self.state[rows, after] = 5
            player[rows] = after
            pushed = (rec[rows] >> 7) & 1 == 1
            pr, pc = rows[pushed], after[pushed] + dlt[pushed]
            self.state[pr, pc] = np.where(self.fixed[pr, pc] == 2, 3, 4)

    def refresh(self, rows, fixed, state):
        """Rooms that were reset: take their new fixed / state rows."""
        self.fixed[rows], self.state[rows] = fixed, state
        if getattr(self, "_player", None) is not None:
            self._player = self.players()

    def glyphs(self, rows=None):
        """[m, H*(W+1)] glyph codes with separators: what the render entry points write."""
        s = self.state if rows is None else self.state[rows]
        f = self.fixed if rows is None else self.fixed[rows]
        g = np.where((s == 5) & (f == 2), 6, s).reshape(-1, self.H, self.W)
        out = np.full((g.shape[0], self.H, self.W + 1), 7, dtype=np.uint8)
        out[:, :, : self.W] = g
        return out.reshape(g.shape[0], -1)


class HostStepper:
    """SokobanBatch.host_stepper: one CUDA graph = copy the pinned actions up, step (delta records), copy the records down."""

    def __init__(self, batch, h_actions, record_bytes=1):
        assert isinstance(h_actions, torch.Tensor) and h_actions.is_pinned() and h_actions.dtype == torch.uint8
        assert record_bytes in (1, 4)
        self.batch = batch
        self.actions = h_actions if h_actions.dim() == 2 else h_actions.unsqueeze(1)
        n, K = self.actions.shape
        assert n == batch.n
        self.K = K
        self.stream = torch.cuda.Stream(batch.device)
        self._d_act = torch.empty((n, K), dtype=torch.uint8, device=batch.device)
        shape = (n, K) if record_bytes == 1 else (n, K, 4)
        self._h_rec = torch.empty(shape, dtype=torch.uint8, pin_memory=True)
        self.records = self._h_rec.numpy()
        self.record_bytes = record_bytes
        self.stream.wait_stream(torch.cuda.current_stream(batch.device))
        with torch.cuda.stream(self.stream):
            # once outright, executing NO action (n_actions = 0 leaves rooms and counters as they are): the batch's record buffer
            # and the library's state exist before the capture
            self._body(torch.zeros(n, dtype=torch.uint8, device=batch.device))
            self.stream.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph, stream=self.stream):
                self._body(None)

    def _body(self, n_actions):
        self._d_act.copy_(self.actions, non_blocking=True)
        rec = self.batch.step_delta(self._d_act, n_actions, record_bytes=self.record_bytes)  # the batch's own (persistent) buffer
        self._h_rec.copy_(rec, non_blocking=True)

    def launch(self):
        with torch.cuda.stream(self.stream):
            self.graph.replay()
        self.batch._state_stale = True
        self.batch.obs_stale = True

    def wait(self):
        self.stream.synchronize()
        return self.records


def unpack3_rooms(state3, HW, out=None):
    """Glyph rooms [n, B3] (three bits per cell, codes 0..6) -> room_state [n, H*W] uint8 on the same device (6 read as 5)."""
    n = state3.shape[0]
    bits = (state3.unsqueeze(-1) >> torch.arange(8, device=state3.device, dtype=torch.uint8)) & 1  # [n, B3, 8], little-endian
    cells = bits.reshape(n, -1)[:, : 3 * HW].reshape(n, HW, 3)
    codes = cells[..., 0] | (cells[..., 1] << 1) | (cells[..., 2] << 2)
    codes = torch.where(codes == 6, torch.full_like(codes, 5), codes)
    if out is not None:
        out.copy_(codes)
        return out
    return codes


def codes_to_text(codes_row, table):
    """codes_row: 1-D uint8 numpy array of length H*(W+1); the trailing separator is dropped."""
    return "".join(table[c] for c in codes_row[:-1].tolist())


class SokobanBatch:
    def __init__(self, n, dim_room=(6, 6), num_boxes=1, max_steps=100, search_depth=300, device=0, compact=False, rooms=None):
        self.ctx = _lib.Context.get(device)
        self.lib = self.ctx.lib
        self.device = torch.device("cuda", int(device))
        self.n = int(n)
        self.H, self.W = int(dim_room[0]), int(dim_room[1])
        self.num_boxes, self.max_steps, self.search_depth = int(num_boxes), int(max_steps), int(search_depth)
        HW, G = self.H * self.W, self.H * (self.W + 1)
        dev = self.device
        # compact=True: the rooms live in HBM as nibbles (state4 [n, HW/2], cell 2i in the low nibble of byte i) and are stepped by
        # ragen_sokoban_step_compact - half the bytes of a step's dominant read-modify-write.  `state` stays the byte view every
        # other consumer sees; it is refreshed from state4 when read after a step and packed into state4 by adopt().
        # rooms="glyph3": the rooms live in HBM as THREE-BIT GLYPH CODES (state3 [n, packed3_bytes]; 0..5 as room_state, 6 = the player
        # on a target) and are stepped by ragen_sokoban_step_glyph3: the room is its own rendered observation (StepOut.obs IS
        # state3), a step moves 38 B in and 38 B out per 10x10 env and nothing is rendered.
        if rooms not in (None, "bytes", "nibbles", "glyph3"):
            raise ValueError(f"rooms must be 'bytes', 'nibbles' or 'glyph3', got {rooms!r}")
        self.glyph3 = rooms == "glyph3"
        self.compact = bool(compact) or rooms == "nibbles"
        if (self.compact or self.glyph3) and not (self.H == self.W and self.H in (6, 8, 10)):
            raise ValueError(f"compact rooms are built for 6x6, 8x8 and 10x10; room is {self.H}x{self.W}")
        self._state = torch.zeros((self.n, HW), dtype=torch.uint8, device=dev)
        self.state4 = torch.zeros((self.n, HW // 2), dtype=torch.uint8, device=dev) if self.compact else None
        self.state3 = torch.zeros((self.n, packed3_bytes(self.H, self.W)), dtype=torch.uint8, device=dev) if self.glyph3 else None
        self._state_stale = False
        self.fixed = torch.zeros((self.n, HW), dtype=torch.uint8, device=dev)
        # one 8-byte record per env (include/ragen_b200.h): player, num_env_steps, boxes_on_target|under, actual
        self.meta = torch.zeros((self.n, 4), dtype=torch.int16, device=dev)
        # per-step results live in ONE slab [three-bit rows | reward f64, done, flags, n_executed | nibble rows] so that the
        # host-buffer path (step_host) brings a step back with a single device->host copy in either packed format
        n_ = self.n
        self._b3 = packed3_bytes(self.H, self.W) if (self.H == self.W and self.H in (6, 8, 10)) else 0
        self._off_sc = (n_ * self._b3 + 15) & ~15
        self._off_obs = (self._off_sc + 11 * n_ + 15) & ~15
        self._slab_bytes = self._off_obs + (n_ * (G // 2) if G % 2 == 0 else 0)
        self._slab = torch.zeros(max(self._slab_bytes, 16), dtype=torch.uint8, device=dev)
        o = self._off_sc
        self.reward = self._slab[o: o + 8 * n_].view(torch.float64)
        self.done = self._slab[o + 8 * n_: o + 9 * n_]
        self.flags = self._slab[o + 9 * n_: o + 10 * n_]
        self.n_executed = self._slab[o + 10 * n_: o + 11 * n_]
        self.obs = torch.zeros((self.n, G), dtype=torch.uint8, device=dev)
        # [n, G/2] two 4-bit codes per byte (needs an even G)
        self.obs_packed = self._slab[self._off_obs: self._off_obs + n_ * (G // 2)].view(n_, G // 2) if G % 2 == 0 else None
        # [n, packed3_bytes] three bits per cell, no separators (6x6 / 8x8 / 10x10)
        self.obs_packed3 = self._slab[: n_ * self._b3].view(n_, self._b3) if self._b3 else None
        self.obs_is_packed = False  # which of the two buffers the latest render wrote
        self.status = torch.zeros(self.n, dtype=torch.int32, device=dev)
        self._pinned = {}

    @property
    def state(self):
        """room_state as uint8 [n, H*W] (the reference's cell codes)."""
        if self._state_stale:
            if self.glyph3:
                unpack3_rooms(self.state3, self.H * self.W, out=self._state)
            else:
                unpack_rooms(self.state4, out=self._state)
            self._state_stale = False
        return self._state

    # views of the per-env record (reference attribute names)
    @property
    def player(self):
        return self.meta[:, 0]

    @property
    def env_steps(self):
        return self.meta[:, 1].view(torch.uint16) if hasattr(torch, "uint16") else self.meta[:, 1]

    @property
    def on_target(self):
        return self.meta.view(torch.int8)[:, 4]

    # -- reset ------------------------------------------------------------------------------
    def load_rooms(self, fixed, state):
        """Adopt rooms given as arrays [n, H*W] (or [n, H, W]); the bookkeeping of
        SokobanEnv.reset (env.py:37-38) runs on the device."""
        f = torch.as_tensor(np.ascontiguousarray(fixed), dtype=torch.uint8).reshape(self.n, -1)
        s = torch.as_tensor(np.ascontiguousarray(state), dtype=torch.uint8).reshape(self.n, -1)
        self.fixed.copy_(f, non_blocking=False)
        self.state.copy_(s, non_blocking=False)
        self.adopt()

    def generate(self, seeds):
        """Rooms for `seeds` (one per seed, any count) on the device.  Returns (fixed u8 [k, H*W], state u8 [k, H*W],
        status i32 [k], explored i32 [k]) CUDA tensors; status as in include/ragen_b200.h."""
        s = torch.as_tensor(np.asarray(seeds, dtype=np.uint64).astype(np.int64)).to(self.device)
        k, HW = s.numel(), self.H * self.W
        fixed = torch.empty((k, HW), dtype=torch.uint8, device=self.device)
        state = torch.empty((k, HW), dtype=torch.uint8, device=self.device)
        status = torch.empty(k, dtype=torch.int32, device=self.device)
        explored = torch.empty(k, dtype=torch.int32, device=self.device)
        _lib.check(self.lib.ragen_sokoban_generate(
            self.ctx.handle, k, self.H, self.W, self.num_boxes, self.search_depth, int(1.7 * (self.H + self.W)), _lib.ptr(s),
            _lib.ptr(fixed), _lib.ptr(state), _lib.ptr(status), _lib.ptr(explored), self.ctx.stream()))
        return fixed, state, status, explored

    def reset(self, seeds, slots=None):
        """SokobanEnv.reset for many envs (env.py:28-42): env slots[i] (default: all envs) gets the room of seeds[i].
        Envs sharing a seed (every env of a group does, es_manager.py:80-82) share one generated room.  Seeds whose
        generation fails are re-seeded exactly as the reference does (abs(hash(str(seed))) % 2**32, env.py:41).
        Returns the list of seeds finally used per env."""
        seeds = [int(x) for x in seeds]
        slots = list(range(self.n)) if slots is None else [int(x) for x in slots]
        assert len(seeds) == len(slots)
        if not seeds:
            return []
        rooms = {}
        todo = sorted(set(seeds))
        alias = {}
        if todo:
            # common case (every seed generates): no per-room Python work, one gather on the device
            fixed, state, status, _ = self.generate(todo)
            if not bool(status.any()):
                src = torch.from_numpy(np.searchsorted(np.asarray(todo, dtype=np.uint64), np.asarray(seeds, dtype=np.uint64))).to(self.device)
                dst = torch.tensor(slots, dtype=torch.long, device=self.device)
                self.fixed[dst] = fixed[src]
                self.state[dst] = state[src]
                self.adopt(slots=dst)
                return seeds
            st = status.cpu().tolist()
            nxt = []
            for j, sd in enumerate(todo):
                if st[j] == 0:
                    rooms[sd] = (fixed[j], state[j])
                else:
                    alias[sd] = abs(hash(str(sd))) % (2 ** 32)
                    if alias[sd] not in rooms and alias[sd] not in nxt:
                        nxt.append(alias[sd])
            todo = nxt
        while todo:
            fixed, state, status, _ = self.generate(todo)
            st = status.cpu().tolist()
            nxt = []
            for j, sd in enumerate(todo):
                if st[j] == 0:
                    rooms[sd] = (fixed[j], state[j])
                else:
                    alias[sd] = abs(hash(str(sd))) % (2 ** 32)
                    if alias[sd] not in rooms and alias[sd] not in nxt:
                        nxt.append(alias[sd])
            todo = nxt

        def resolve(sd):
            while sd not in rooms:
                sd = alias[sd]
            return sd

        used = [resolve(sd) for sd in seeds]
        order = sorted(set(used))
        pos = {sd: i for i, sd in enumerate(order)}
        f_all = torch.stack([rooms[sd][0] for sd in order])
        s_all = torch.stack([rooms[sd][1] for sd in order])
        src = torch.tensor([pos[sd] for sd in used], dtype=torch.long, device=self.device)
        dst = torch.tensor(slots, dtype=torch.long, device=self.device)
        self.fixed[dst] = f_all[src]
        self.state[dst] = s_all[src]
        self.adopt(slots=dst)
        return used

    def solve(self, rows=None, max_depth=100, want_path=True):
        """Shortest push solutions of the current rooms (all envs, or `rows`): see solve_rooms."""
        f = self.fixed if rows is None else self.fixed[rows]
        s = self.state if rows is None else self.state[rows]
        return solve_rooms(f, s, (self.H, self.W), self.num_boxes, max_depth, self.device.index, want_path)

    def adopt(self, check=True, slots=None):
        """Post-reset bookkeeping (env.py:37-38) of all envs, or of `slots` only (a device index tensor / list): the other envs
        keep their record - an env in mid-episode must not lose its step count or its stored box count because a neighbour in
        the batch was reset.  The nibble rooms of a compact batch are written by the same launch."""
        only = None
        if slots is not None:
            idx = torch.as_tensor(slots, dtype=torch.long, device=self.device)
            if idx.numel() != self.n or bool((idx != torch.arange(self.n, device=self.device)).any()):
                only = torch.zeros(self.n, dtype=torch.uint8, device=self.device)
                only[idx] = 1
        state = self.state  # refreshes the byte view of a compact batch before the untouched rooms are read
        _lib.check(self.lib.ragen_sokoban_adopt_slots(
            self.ctx.handle, self.n, self.H, self.W, self.num_boxes, _lib.ptr(state), _lib.ptr(self.fixed),
            _lib.ptr(self.meta), _lib.ptr(self.status), _lib.ptr(only), _lib.ptr(self.state4) if self.compact else None,
            _lib.ptr(self.state3) if self.glyph3 else None, self.ctx.stream()))
        if check:
            bad = torch.nonzero(self.status if only is None else self.status * only).flatten()
            if bad.numel():
                i = int(bad[0])
                raise ValueError(f"room {i} rejected (status {int(self.status[i])}: 1 = no player, 2 = border not wall, "
                                 "3 = room_fixed inconsistent with room_state)")
        self.render()

    # -- step -------------------------------------------------------------------------------


    @property
    def can_pack(self):
        return (self.H * (self.W + 1)) % 2 == 0

    def _packed3_buffer(self):
        if self.glyph3:
            return self.state3  # glyph rooms are their own three-bit rows
        if self.obs_packed3 is None:
            raise ValueError(f"three-bit glyph rows are built for 6x6, 8x8 and 10x10 rooms; room is {self.H}x{self.W}")
        return self.obs_packed3

    def _packed_buffer(self):
        if self.obs_packed is None:  # lives in the result slab whenever the row stream is even
            raise ValueError(f"packed glyph rows need an even H*(W+1); room is {self.H}x{self.W}")
        return self.obs_packed

    def step(self, actions, n_actions=None, render=True, packed=False) -> StepOut:
        """actions: uint8 device tensor [n] or [n, K]; n_actions: optional uint8 [n].  packed=True renders the glyph rows
        as two 4-bit codes per byte into `obs_packed` (half the bytes; StepOut.obs is then that tensor)."""
        if actions.dim() == 1:
            actions = actions.unsqueeze(1)
        assert actions.dtype == torch.uint8 and actions.is_cuda and actions.is_contiguous() and actions.shape[0] == self.n
        K = actions.shape[1]
        fmt = 3 if (packed == 3 and packed is not True) else (1 if packed else 0)  # 0 bytes, 1 nibbles, 3 three-bit rows
        if self.glyph3:
            if render and fmt != 3:
                raise ValueError("glyph rooms ARE three-bit glyph rows (packed=3); use render() for the other formats")
            self.obs_is_packed = 3
            self.obs_stale = False
            _lib.check(self.lib.ragen_sokoban_step_glyph3(
                self.ctx.handle, self.n, self.H, self.W, self.num_boxes, self.max_steps, _lib.ptr(actions), K, _lib.ptr(n_actions),
                _lib.ptr(self.state3), _lib.ptr(self.meta), _lib.ptr(self.reward), _lib.ptr(self.done), _lib.ptr(self.flags),
                _lib.ptr(self.n_executed), None, self.ctx.stream()))
            self._state_stale = True
            return StepOut(self.reward, self.done, self.flags, self.n_executed, self.state3)
        if self.compact:
            if render and fmt != 3:
                raise ValueError("a compact batch renders three-bit rows while stepping (packed=3); use render() for the other formats")
            obs = self._packed3_buffer() if render else None
            if render:
                self.obs_is_packed = 3
            self.obs_stale = not render
            _lib.check(self.lib.ragen_sokoban_step_compact(
                self.ctx.handle, self.n, self.H, self.W, self.num_boxes, self.max_steps, _lib.ptr(actions), K, _lib.ptr(n_actions),
                _lib.ptr(self.state4), _lib.ptr(self.meta), _lib.ptr(self.reward), _lib.ptr(self.done), _lib.ptr(self.flags),
                _lib.ptr(self.n_executed), _lib.ptr(obs), self.ctx.stream()))
            self._state_stale = True
            return StepOut(self.reward, self.done, self.flags, self.n_executed, obs if render else self.obs)
        obs = (self._packed3_buffer() if fmt == 3 else self._packed_buffer() if fmt == 1 else self.obs) if render else None
        if render:
            self.obs_is_packed = fmt if fmt else False
        self.obs_stale = not render  # texts() renders again before reading rows a step did not write
        fn = (self.lib.ragen_sokoban_step_packed3 if fmt == 3 else self.lib.ragen_sokoban_step_packed if fmt == 1
              else self.lib.ragen_sokoban_step)
        _lib.check(fn(self.ctx.handle, self.n, self.H, self.W, self.num_boxes, self.max_steps, _lib.ptr(actions), K,
                      _lib.ptr(n_actions), _lib.ptr(self.state), _lib.ptr(self.meta), _lib.ptr(self.reward), _lib.ptr(self.done),
                      _lib.ptr(self.flags), _lib.ptr(self.n_executed), _lib.ptr(obs), self.ctx.stream()))
        return StepOut(self.reward, self.done, self.flags, self.n_executed, obs if render else self.obs)

    def step_delta(self, actions, n_actions=None, record_bytes=4):
        """The compact step reporting DELTAS (ragen_sokoban_step_delta): returns the device tensor uint8 [n, K, 4] - or, with
        record_bytes=1, the one-byte records [n, K] of ragen_sokoban_step_delta1; nothing is rendered and the result arrays are
        not written (decode_delta / HostMirror turn deltas into them on the host)."""
        if not (self.compact or self.glyph3):
            raise ValueError("step_delta steps nibble or glyph rooms: construct the batch with compact=True or rooms='glyph3'")
        if actions.dim() == 1:
            actions = actions.unsqueeze(1)
        assert actions.dtype == torch.uint8 and actions.is_cuda and actions.is_contiguous() and actions.shape[0] == self.n
        K = actions.shape[1]
        if record_bytes == 1:
            buf = getattr(self, "_delta1", None)
            if buf is None or buf.shape[1] != K:
                buf = self._delta1 = torch.zeros((self.n, K), dtype=torch.uint8, device=self.device)
            rooms = self.state3 if self.glyph3 else self.state4
            _lib.check(self.lib.ragen_sokoban_step_delta1(
                self.ctx.handle, self.n, self.H, self.W, self.num_boxes, self.max_steps, _lib.ptr(actions), K, _lib.ptr(n_actions),
                4 if self.glyph3 else 3, _lib.ptr(rooms), _lib.ptr(self.meta), _lib.ptr(buf), self.ctx.stream()))
            self._state_stale = True
            self.obs_stale = True
            return buf
        buf = getattr(self, "_delta", None)
        if buf is None or buf.shape[1] != K:
            buf = self._delta = torch.zeros((self.n, K, 4), dtype=torch.uint8, device=self.device)
        if self.glyph3:
            _lib.check(self.lib.ragen_sokoban_step_glyph3(
                self.ctx.handle, self.n, self.H, self.W, self.num_boxes, self.max_steps, _lib.ptr(actions), K, _lib.ptr(n_actions),
                _lib.ptr(self.state3), _lib.ptr(self.meta), None, None, None, None, _lib.ptr(buf), self.ctx.stream()))
        else:
            _lib.check(self.lib.ragen_sokoban_step_delta(
                self.ctx.handle, self.n, self.H, self.W, self.num_boxes, self.max_steps, _lib.ptr(actions), K, _lib.ptr(n_actions),
                _lib.ptr(self.state4), _lib.ptr(self.meta), _lib.ptr(buf), self.ctx.stream()))
        self._state_stale = True
        self.obs_stale = True
        return buf

    def text_rows(self, grid_lookup, out=None, fmt=None):
        """The latest render as UTF-8 TEXT on the device (ragen_glyph_utf8): uint8 [n, stride] NUL-padded rows (the env's string
        plus a final newline) and int32 [n] lengths.  `np.asarray(rows.cpu()).view(f"S{stride}")` is an array of byte strings."""
        fmt = self.obs_is_packed if fmt is None else fmt
        fmt = 3 if (fmt == 3 and fmt is not True) else (1 if fmt else 0)
        src = self._packed3_buffer() if fmt == 3 else self._packed_buffer() if fmt == 1 else self.obs
        G = self.H * (self.W + 1)
        tab = np.zeros((8, 4), dtype=np.int32)
        ln = np.zeros(8, dtype=np.int32)
        for c, g in enumerate(glyph_table(grid_lookup)):
            b = g.encode("utf-8")
            if len(b) > 4:
                raise ValueError(f"glyph {g!r} is longer than 4 UTF-8 bytes")
            tab[c, : len(b)] = list(b)
            ln[c] = len(b)
        worst = int(ln.max())
        if int(np.delete(ln, 3).max()) <= 1:  # only "box on target" is multi-byte (the default vocabulary): at most num_boxes of them
            stride = G + (int(ln[3]) - 1) * self.num_boxes
        else:
            stride = G * worst
        stride = (stride + 3) & ~3
        if out is None:
            out = (torch.empty((self.n, stride), dtype=torch.uint8, device=self.device),
                   torch.empty(self.n, dtype=torch.int32, device=self.device))
        _lib.check(self.lib.ragen_glyph_utf8(self.ctx.handle, self.n, G, fmt, self.W, _lib.ptr(src), int(src.shape[1]),
                                             tab.ctypes.data_as(_lib._P), ln.ctypes.data_as(_lib._P), int(out[0].shape[1]),
                                             _lib.ptr(out[0]), _lib.ptr(out[1]), self.ctx.stream()))
        return out

    def host_mirror(self):
        """A HostMirror of the batch's current rooms (one device->host copy of state and fixed)."""
        return HostMirror(self.fixed.cpu().numpy(), self.state.cpu().numpy(), self.H, self.W)

    # -- pipelined host-buffer stepping: launch step i+1 while the results of step i drain ------------------------------------
    def step_host_begin(self, h_actions, fmt="delta"):
        """h_actions: PINNED uint8 host tensor [n] or [n, K].  Queues host->device copy + step on the compute stream and the
        device->host copy of the results on a copy stream; returns a ticket for step_host_end.  fmt: "delta" (4 B per action),
        "delta1" (1 B per action) or 3 (three-bit rows + scalars, 49 B per 10x10 env).  Up to two tickets may be open per batch."""
        assert isinstance(h_actions, torch.Tensor) and h_actions.is_pinned() and h_actions.dtype == torch.uint8
        h_act = h_actions if h_actions.dim() == 2 else h_actions.unsqueeze(1)
        pipe = getattr(self, "_pipe", None)
        if pipe is None:
            pipe = self._pipe = {"copy": torch.cuda.Stream(self.device), "slot": 0, "d_act": [None, None]}
        j = pipe["slot"]
        pipe["slot"] = j ^ 1
        main = torch.cuda.current_stream(self.device)
        if pipe.get("last_done") is not None:
            main.wait_event(pipe["last_done"])  # the step below overwrites the result buffer the previous copy reads
        if pipe["d_act"][j] is None or pipe["d_act"][j].shape != h_act.shape:
            pipe["d_act"][j] = torch.empty(h_act.shape, dtype=torch.uint8, device=self.device)
        d_act = pipe["d_act"][j]
        d_act.copy_(h_act, non_blocking=True)
        if fmt == "delta":
            src = self.step_delta(d_act).view(-1)
        elif fmt == "delta1":
            src = self.step_delta(d_act, record_bytes=1).view(-1)
        else:
            if self.glyph3:
                raise ValueError("glyph rooms: pipelined host stepping carries deltas (fmt='delta' / 'delta1'); full rows come from step_host")
            self.step(d_act, packed=3)
            src = self._slab[: self._off_sc + 11 * self.n]
        h = self._pin(("pipe", fmt, j), (src.numel(),), torch.uint8)
        ready = torch.cuda.Event()
        ready.record(main)
        with torch.cuda.stream(pipe["copy"]):
            pipe["copy"].wait_event(ready)
            h.copy_(src, non_blocking=True)
            done = torch.cuda.Event()
            done.record(pipe["copy"])
        pipe["last_done"] = done
        return (fmt, h, done, h_act.shape[1])

    def step_host_end(self, ticket):
        """Waits for the ticket's results.  "delta": numpy uint8 [n, K, 4]; "delta1": [n, K]; 3: (rows, reward, done, flags, n_executed)."""
        fmt, h, done, K = ticket
        done.synchronize()
        hn = h.numpy()
        if fmt == "delta":
            return hn.reshape(self.n, K, 4)
        if fmt == "delta1":
            return hn.reshape(self.n, K)
        n_, o = self.n, self._off_sc
        return (hn[: n_ * self._b3].reshape(n_, -1), hn[o: o + 8 * n_].view(np.float64), hn[o + 8 * n_: o + 9 * n_],
                hn[o + 9 * n_: o + 10 * n_], hn[o + 10 * n_: o + 11 * n_])

    def host_stepper(self, h_actions, record_bytes=1):
        """A captured host-buffer step for callers that step the same batch many times: returns a HostStepper whose launch() replays
        [actions host->device | step with delta records | records device->host] as ONE CUDA graph on the stepper's own stream -
        a few microseconds of host work per step instead of a dozen framework calls - and whose wait() returns the records
        (numpy view of pinned memory, decode_delta / HostMirror.apply take them).  h_actions: the PINNED uint8 tensor [n] or
        [n, K] the caller fills before every launch()."""
        return HostStepper(self, h_actions, record_bytes)

    def _pin(self, name, shape, dtype):
        key = (name, tuple(shape), dtype)
        if key not in self._pinned:
            self._pinned[key] = torch.empty(shape, dtype=dtype, pin_memory=True)
        return self._pinned[key]

    def step_host(self, actions_np, n_actions_np=None, want_obs=True, packed=False):
        """End-to-end call with HOST buffers: actions go host->device, the step runs, and reward /
        done / flags / n_executed (and the glyph rows; packed=True: two codes per byte, see glyph_pair_table) come back
        device->host.  Returns numpy views of pinned staging buffers (valid until the next call)."""
        if isinstance(actions_np, torch.Tensor) and actions_np.dtype == torch.uint8 and actions_np.is_contiguous() and actions_np.is_pinned():
            h_act = actions_np if actions_np.dim() == 2 else actions_np.unsqueeze(1)  # already pinned host memory: no staging copy
        else:
            a = np.ascontiguousarray(actions_np, dtype=np.uint8)
            if a.ndim == 1:
                a = a[:, None]
            h_act = self._pin("act", a.shape, torch.uint8)
            h_act.numpy()[...] = a
        d_act = h_act.to(self.device, non_blocking=True)
        d_na = None
        if n_actions_np is not None:
            h_na = self._pin("na", (self.n,), torch.uint8)
            h_na.numpy()[...] = n_actions_np
            d_na = h_na.to(self.device, non_blocking=True)
        out = self.step(d_act, d_na, render=want_obs, packed=(packed if want_obs else False))
        if self.glyph3:  # the rows are the rooms themselves: one copy for them, one for the scalars
            n_, o = self.n, self._off_sc
            h_rows = self._pin("rows3", tuple(self.state3.shape), torch.uint8)
            h_sc = self._pin("sc", (11 * n_,), torch.uint8)
            if want_obs:
                h_rows.copy_(self.state3, non_blocking=True)
            h_sc.copy_(self._slab[o: o + 11 * n_], non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            hn = h_sc.numpy()
            return (h_rows.numpy() if want_obs else None, hn[: 8 * n_].view(np.float64), hn[8 * n_: 9 * n_], hn[9 * n_: 10 * n_],
                    hn[10 * n_: 11 * n_])
        if packed and want_obs:  # one copy: the packed rows and the scalars are neighbours in the result slab
            n_ = self.n
            three = packed == 3 and packed is not True
            lo, hi = (0, self._off_sc + 11 * n_) if three else (self._off_sc, self._slab_bytes)
            h = self._pin("slab3" if three else "slab", (hi - lo,), torch.uint8)
            h.copy_(self._slab[lo:hi], non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            hn = h.numpy()
            o = self._off_sc - lo
            rows = hn[: n_ * self._b3].reshape(n_, -1) if three else hn[self._off_obs - lo:].reshape(n_, -1)
            return (rows, hn[o: o + 8 * n_].view(np.float64), hn[o + 8 * n_: o + 9 * n_], hn[o + 9 * n_: o + 10 * n_],
                    hn[o + 10 * n_: o + 11 * n_])
        h_rew = self._pin("rew", (self.n,), torch.float64)
        h_done = self._pin("done", (self.n,), torch.uint8)
        h_flags = self._pin("flags", (self.n,), torch.uint8)
        h_nex = self._pin("nex", (self.n,), torch.uint8)
        h_rew.copy_(out.reward, non_blocking=True)
        h_done.copy_(out.done, non_blocking=True)
        h_flags.copy_(out.flags, non_blocking=True)
        h_nex.copy_(out.n_executed, non_blocking=True)
        h_obs = None
        if want_obs:
            h_obs = self._pin("obs_packed" if packed else "obs", tuple(out.obs.shape), torch.uint8)
            h_obs.copy_(out.obs, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return (h_obs.numpy() if want_obs else None, h_rew.numpy(), h_done.numpy(), h_flags.numpy(), h_nex.numpy())

    # -- render -----------------------------------------------------------------------------
    def render(self, packed=False):
        fmt = 3 if (packed == 3 and packed is not True) else (1 if packed else 0)
        if self.glyph3 and fmt == 3:  # nothing to do: the rooms are the rows
            self.obs_is_packed, self.obs_stale = 3, False
            return self.state3
        obs = self._packed3_buffer() if fmt == 3 else self._packed_buffer() if fmt == 1 else self.obs
        self.obs_is_packed = fmt if fmt else False
        self.obs_stale = False
        fn = (self.lib.ragen_sokoban_render_packed3 if fmt == 3 else self.lib.ragen_sokoban_render_packed if fmt == 1
              else self.lib.ragen_sokoban_render)
        _lib.check(fn(self.ctx.handle, self.n, self.H, self.W, _lib.ptr(self.state), _lib.ptr(self.meta), _lib.ptr(obs),
                      self.ctx.stream()))
        return obs

    def set_atlas(self, surfaces):
        """Sprites for render_rgb: uint8 [7, S, S, 3] in gym_sokoban's surface order (wall, floor, box_target, box_on_target,
        box, player, player_on_target)."""
        a = np.ascontiguousarray(surfaces, dtype=np.uint8)
        if a.ndim != 4 or a.shape[0] != 7 or a.shape[1] != a.shape[2] or a.shape[3] != 3:
            raise ValueError(f"atlas must be uint8 [7, S, S, 3], got {a.shape}")
        self.atlas = torch.from_numpy(a).to(self.device)
        return self.atlas

    def render_rgb(self, rows=None):
        """rgb_array observations (env.py:58-59) of all envs or of `rows`: uint8 device tensor [m, H*S, W*S, 3]."""
        if getattr(self, "atlas", None) is None:
            self.set_atlas(load_sprite_atlas())
        S = int(self.atlas.shape[1])
        d_rows, m = None, self.n
        if rows is not None:
            d_rows = torch.as_tensor(rows, dtype=torch.int64).to(self.device)
            m = int(d_rows.numel())
            if m and (int(d_rows.min()) < 0 or int(d_rows.max()) >= self.n):
                raise IndexError("render_rgb: env row out of range")
        out = torch.empty((m, self.H * S, self.W * S, 3), dtype=torch.uint8, device=self.device)
        if m == 0:
            return out
        _lib.check(self.lib.ragen_sokoban_render_rgb(self.ctx.handle, self.n, self.H, self.W, S, _lib.ptr(self.state),
                                                     _lib.ptr(self.meta), _lib.ptr(self.atlas), _lib.ptr(d_rows), m if d_rows is not None else 0,
                                                     _lib.ptr(out), self.ctx.stream()))
        return out

    def texts(self, grid_lookup, rows=None):
        """Text observations of the latest render (all envs, or `rows`): one device->host copy + one translate per env."""
        if getattr(self, "obs_stale", False):
            self.render(packed=self.obs_is_packed)
        if self.obs_is_packed == 3:
            tr = {i: g for i, g in enumerate(glyph_table(grid_lookup))}
            buf3 = self._packed3_buffer()
            rows3 = (buf3 if rows is None else buf3[rows]).cpu().numpy()
            codes = unpack3_glyphs(rows3, self.H, self.W)  # vectorised; then one C-level translate per env
            return [bytes(r[:-1]).decode("latin1").translate(tr) for r in codes]
        if self.obs_is_packed:
            pairs = glyph_pair_table(grid_lookup)
            codes = (self.obs_packed if rows is None else self.obs_packed[rows]).cpu().numpy()
            return [packed_to_text(r, pairs) for r in codes]
        tab = glyph_table(grid_lookup)
        codes = (self.obs if rows is None else self.obs[rows]).cpu().numpy()
        return [codes_to_text(r, tab) for r in codes]


def _state_bound(H, W, num_boxes):
    """Upper bound on distinct (player, boxes) states of an H x W room with a wall border; the solver's queue size."""
    import math

    cells = max(1, (H - 2) * (W - 2))
    bound = cells * math.comb(max(cells - 1, 0), max(int(num_boxes), 0)) + 1
    return int(min(max(bound, 64), 1 << 20))


def solve_rooms(fixed, state, dim_room, num_boxes=None, max_depth=100, device=0, want_path=True):
    """get_shortest_action_path (ragen/env/sokoban/utils.py:11-101) for a batch of rooms on the device.
    fixed / state: uint8 [k, H*W] (CUDA tensors or anything np.asarray takes).  Returns (status i32 [k], length i32 [k],
    path u8 [k, max_depth] or None, explored i32 [k]) CUDA tensors; status as in include/ragen_b200.h (0 solved, 1 none within
    max_depth, 2 state budget exceeded, 3 malformed room).  The path is the reference's own (same BFS discovery order)."""
    H, W = int(dim_room[0]), int(dim_room[1])
    dev = torch.device("cuda", int(device))

    def dev_u8(x):
        t = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(x), dtype=np.uint8))
        return t.to(dev, torch.uint8).reshape(-1, H * W).contiguous()

    f, s = dev_u8(fixed), dev_u8(state)
    assert f.shape == s.shape
    k = f.shape[0]
    if num_boxes is None:
        num_boxes = int(((s == 3) | (s == 4)).sum(dim=1).max()) if k else 0
    ctx = _lib.Context.get(int(device))
    status = torch.empty(k, dtype=torch.int32, device=dev)
    length = torch.empty(k, dtype=torch.int32, device=dev)
    explored = torch.empty(k, dtype=torch.int32, device=dev)
    path = torch.zeros((k, max(int(max_depth), 1)), dtype=torch.uint8, device=dev) if want_path else None
    _lib.check(ctx.lib.ragen_sokoban_solve(ctx.handle, k, H, W, _lib.ptr(f), _lib.ptr(s), int(max_depth), _state_bound(H, W, min(num_boxes, 4)),
                                           _lib.ptr(status), _lib.ptr(length), _lib.ptr(path) if int(max_depth) >= 1 else None,
                                           _lib.ptr(explored), ctx.stream()))
    return status, length, path, explored


def get_shortest_action_path(room_fixed, room_state, MAX_DEPTH=100, device=0):
    """The reference's function (utils.py:11-101), one room: list of actions 1..4, or [] (already solved / no solution)."""
    f = np.asarray(room_fixed)
    H, W = f.shape
    status, length, path, _ = solve_rooms(f.reshape(1, -1), np.asarray(room_state).reshape(1, -1), (H, W), max_depth=MAX_DEPTH, device=device)
    st = int(status[0])
    if st >= 2:
        raise RuntimeError(f"sokoban solver: status {st} (2 = state budget exceeded, 3 = malformed room)")
    return path[0, : int(length[0])].cpu().tolist() if st == 0 else []


class SokobanEnv(BaseDiscreteActionEnv):
    """One Sokoban env with the reference interface (ragen/env/sokoban/env.py:10-67), backed by a
    slot of a SokobanBatch.  Standalone construction creates a private batch of one env."""

    def __init__(self, config=None, batch: Optional[SokobanBatch] = None, slot: int = 0, **kwargs):
        BaseDiscreteActionEnv.__init__(self)
        self.config = config or SokobanEnvConfig()
        self.GRID_LOOKUP = self.config.grid_lookup
        self.ACTION_LOOKUP = self.config.action_lookup
        self.search_depth = self.config.search_depth
        self.render_mode = self.config.render_mode
        self.dim_room = tuple(self.config.dim_room)
        self.num_boxes = self.config.num_boxes
        self.max_steps = self.config.max_steps
        if batch is None:
            batch = SokobanBatch(1, self.dim_room, self.num_boxes, self.max_steps, self.search_depth,
                                 device=kwargs.get("device", 0))
            slot = 0
        self.batch, self.slot = batch, slot
        self._act = self._na = None  # staging tensors of the single-env step path, made on first use

    # the reference exposes these as attributes; here they read the device state of the slot
    @property
    def room_state(self):
        return self.batch.state[self.slot].cpu().numpy().reshape(self.dim_room).astype(np.int64)

    @property
    def room_fixed(self):
        return self.batch.fixed[self.slot].cpu().numpy().reshape(self.dim_room).astype(np.int64)

    @property
    def player_position(self):
        p = int(self.batch.player[self.slot])
        return np.array([p // self.dim_room[1], p % self.dim_room[1]])

    @property
    def num_env_steps(self):
        return int(self.batch.env_steps[self.slot])

    @property
    def boxes_on_target(self):
        return int(self.batch.on_target[self.slot])

    def load_room(self, fixed, state):
        """Start from a given room instead of the generator (used by parity tests)."""
        b = self.batch
        b.fixed[self.slot].copy_(torch.as_tensor(np.asarray(fixed, dtype=np.uint8).reshape(-1)))
        b.state[self.slot].copy_(torch.as_tensor(np.asarray(state, dtype=np.uint8).reshape(-1)))
        b.adopt(slots=[self.slot])
        return self.render()

    def reset(self, seed=None, mode=None):
        if seed is None:  # the reference then uses the process's unseeded global RNG state
            seed = int(np.random.SeedSequence().entropy % (2 ** 32))
        self.seed_used = self.batch.reset([seed], [self.slot])[0]
        return self.render()

    def _staging(self):
        if self._act is None:
            batch = self.batch
            self._act = torch.zeros((batch.n, 1), dtype=torch.uint8, device=batch.device)
            self._na = torch.zeros(batch.n, dtype=torch.uint8, device=batch.device)

    def step(self, action: int):
        b = self.batch
        self._staging()
        self._act.zero_()
        self._na.zero_()
        self._act[self.slot, 0] = int(action)
        self._na[self.slot] = 1
        out = b.step(self._act, self._na)
        reward = float(out.reward[self.slot])
        fl = int(out.flags[self.slot])
        done = bool(out.done[self.slot])
        info = {"action_is_effective": bool(fl & 1), "action_is_valid": True, "success": bool(fl & 2)}
        return self.render(), reward, done, info

    def render(self, mode=None):
        render_mode = mode if mode is not None else self.render_mode
        if render_mode == "text":
            return self.batch.texts(self.GRID_LOOKUP, rows=[self.slot])[0]
        if render_mode == "rgb_array":
            return self.batch.render_rgb(rows=[self.slot])[0].cpu().numpy()
        raise ValueError(f"Invalid mode: {render_mode}")

    def get_all_actions(self):
        return list(self.ACTION_LOOKUP.keys())

    def close(self):
        self.render_cache = None
