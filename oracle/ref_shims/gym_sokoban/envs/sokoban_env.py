"""Shim of gym_sokoban.envs.sokoban_env.SokobanEnv (test infrastructure).

Restates the published behaviour of mpSchrader/gym-sokoban's ``SokobanEnv`` for exactly the
members RAGEN touches (ragen/env/sokoban/env.py:20-26, 31, 37, 46, 49, 59).  The upstream
package is un-vendored and un-pinned (reference requirements.txt:4), so this restatement is the
stated source of truth for push/move/reward/done semantics.

Cell codes: 0 wall, 1 floor, 2 target, 3 box on target, 4 box, 5 player.
Actions: 0 no-op, 1-4 push (falls back to move) up/down/left/right, 5-8 plain move.
"""
import numpy as np

_DELTA = ((-1, 0), (1, 0), (0, -1), (0, 1))
_N_ACTIONS = 9


class SokobanEnv:
    metadata = {"render.modes": ["rgb_array"]}

    def __init__(self, dim_room=(10, 10), max_steps=120, num_boxes=4, num_gen_steps=None, reset=True):
        self.dim_room = dim_room
        self.num_gen_steps = int(1.7 * (dim_room[0] + dim_room[1])) if num_gen_steps is None else num_gen_steps
        self.num_boxes = num_boxes
        self.boxes_on_target = 0
        self.penalty_for_step = -0.1
        self.penalty_box_off_target = -1
        self.reward_box_on_target = 1
        self.reward_finished = 10
        self.reward_last = 0
        self.viewer = None
        self.max_steps = max_steps
        if reset:
            self.reset()

    # -- dynamics ---------------------------------------------------------------------------
    def step(self, action, observation_mode="rgb_array"):
        assert 0 <= action < _N_ACTIONS
        self.num_env_steps += 1
        self.new_box_position = None
        self.old_box_position = None
        moved_box = False
        if action == 0:
            moved_player = False
        elif action < 5:
            moved_player, moved_box = self._push(action)
        else:
            moved_player = self._move(action)
        self._calc_reward()
        done = self._check_if_done()
        observation = self.render(mode=observation_mode)
        info = {"action.moved_player": moved_player, "action.moved_box": moved_box}
        if done:
            info["maxsteps_used"] = self._check_if_maxsteps()
            info["all_boxes_on_target"] = self._check_if_all_boxes_on_target()
        return observation, self.reward_last, done, info

    def _push(self, action):
        d = np.array(_DELTA[(action - 1) % 4])
        here = self.player_position.copy()
        ahead = self.player_position + d
        beyond = ahead + d
        if beyond[0] >= self.room_state.shape[0] or beyond[1] >= self.room_state.shape[1]:
            return False, False
        pushable = self.room_state[ahead[0], ahead[1]] in (3, 4)
        pushable = pushable and self.room_state[beyond[0], beyond[1]] in (1, 2)
        if not pushable:
            return self._move(action), False
        self.new_box_position = tuple(beyond)
        self.old_box_position = tuple(ahead)
        self.player_position = ahead
        self.room_state[ahead[0], ahead[1]] = 5
        self.room_state[here[0], here[1]] = self.room_fixed[here[0], here[1]]
        self.room_state[beyond[0], beyond[1]] = 3 if self.room_fixed[beyond[0], beyond[1]] == 2 else 4
        return True, True

    def _move(self, action):
        d = np.array(_DELTA[(action - 1) % 4])
        here = self.player_position.copy()
        ahead = self.player_position + d
        if self.room_state[ahead[0], ahead[1]] in (1, 2):
            self.player_position = ahead
            self.room_state[ahead[0], ahead[1]] = 5
            self.room_state[here[0], here[1]] = self.room_fixed[here[0], here[1]]
            return True
        return False

    def _open_targets(self):
        return int(np.count_nonzero((self.room_state == 2) | ((self.room_fixed == 2) & (self.room_state == 5))))

    def _calc_reward(self):
        self.reward_last = self.penalty_for_step
        now_on = self.num_boxes - self._open_targets()
        if now_on > self.boxes_on_target:
            self.reward_last += self.reward_box_on_target
        elif now_on < self.boxes_on_target:
            self.reward_last += self.penalty_box_off_target
        if self._check_if_all_boxes_on_target():
            self.reward_last += self.reward_finished
        self.boxes_on_target = now_on

    def _check_if_done(self):
        return self._check_if_all_boxes_on_target() or self._check_if_maxsteps()

    def _check_if_all_boxes_on_target(self):
        return self._open_targets() == 0

    def _check_if_maxsteps(self):
        return self.max_steps == self.num_env_steps

    # -- members RAGEN overrides (reset/render) or never reaches in text mode ------------------
    def reset(self, *a, **k):
        # RAGEN's subclass overrides reset (ragen/env/sokoban/env.py:28); the upstream ctor calls
        # self.reset() once, which therefore resolves to the subclass method.
        raise NotImplementedError("upstream room generation is not restated; RAGEN overrides reset")

    def render(self, mode="rgb_array", **k):
        return None

    def get_image(self, mode="rgb_array", scale=1):
        # Upstream builds an RGB frame from sprite surfaces here.  step() calls it once per step
        # (through RAGEN's render(mode='rgb_array'), ragen/env/sokoban/env.py:58-59) and the
        # caller discards the result (env.py:46), so the shim returns a placeholder.
        return None

    def close(self):
        pass
