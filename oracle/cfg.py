"""TEST INFRASTRUCTURE — attribute+item access config tree standing in for OmegaConf objects
(the reference only uses attribute and [] access, `.get`, `.items`: es_manager.py:36-38, 59-67;
ctx_manager.py:93-129)."""


class AttrDict(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


def to_attr(x):
    if isinstance(x, dict):
        return AttrDict({k: to_attr(v) for k, v in x.items()})
    if isinstance(x, (list, tuple)):
        return [to_attr(v) for v in x]
    return x


SOKOBAN_INSTRUCTION = (
    "You are solving the Sokoban puzzle. You are the player and you need to push all boxes to targets. "
    "When you are right next to a box, you can push it by moving in the same direction. You cannot push a box "
    "through a wall, and you cannot pull a box. The answer should be a sequence of actions, like "
    "<answer>Right || Right || Up</answer>")


def base_config(tags=("SimpleSokoban",), n_groups=(8,), group_size=16, mode_both=True, **agent_proxy):
    """The slice of config/base.yaml + config/envs.yaml the hot path reads."""
    custom_envs = {
        "SimpleSokoban": {"env_type": "sokoban", "max_actions_per_traj": 10, "env_instruction": SOKOBAN_INSTRUCTION,
                          "max_tokens": 100, "env_config": {"dim_x": 6, "dim_y": 6, "num_boxes": 1, "max_steps": 100}},
        "LargerSokoban": {"env_type": "sokoban", "max_actions_per_traj": 10, "env_instruction": SOKOBAN_INSTRUCTION,
                          "max_tokens": 100,
                          "env_config": {"dim_x": 8, "dim_y": 8, "num_boxes": 2, "max_steps": 100, "search_depth": 10}},
        "FrozenLake": {"env_type": "frozen_lake", "max_actions_per_traj": 10,
                       "env_instruction": "You are solving the FrozenLake puzzle. Forbid the whole and go to the target. "
                                          "You may move to the unintended direction due to the slippery ice. Example answer "
                                          "format: <think>To forbid the hole and go to the target, I should go left then go "
                                          "up.</think><answer>Left || Up</answer>",
                       "max_tokens": 100, "env_config": None},
        "Bandit": {"env_type": "bandit", "max_actions_per_traj": 1, "env_instruction": "", "max_tokens": 100,
                   "env_config": {"lo_arm_name": "Phoenix", "hi_arm_name": "Dragon"}},
    }
    es = {"env_groups": int(sum(n_groups)), "group_size": group_size,
          "env_configs": {"tags": list(tags), "n_groups": list(n_groups)}}
    ap = {"max_context_window": -1, "max_turn": 5, "action_sep": "||", "max_actions_per_turn": 5,
          "use_turn_scores": False, "enable_think": True,
          "reward_normalization": {"grouping": "state", "method": "identity"}}
    ap.update(agent_proxy)
    cfg = {"custom_envs": custom_envs, "agent_proxy": ap, "enable_response_mask": True,
           "es_manager": {"format_penalty": -0.1, "train": es, "val": dict(es)},
           "actor_rollout_ref": {"rollout": {"response_length": 400}}}
    return to_attr(cfg)
