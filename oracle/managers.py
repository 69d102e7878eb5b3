"""ORACLE (test infrastructure, CPU) — EnvStateManager and ContextManager passes restated.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU legs may import this.  Validated in
this container against the unmodified reference and pinned by tests/golden/.

Follows
  * ragen/llm_agent/es_manager.py:75-103   reset (group seeding)
  * ragen/llm_agent/es_manager.py:105-171  step (action lookup, execute, penalty, truncation)
  * ragen/llm_agent/es_manager.py:173-207  get_rollout_states (metrics)
  * ragen/llm_agent/ctx_manager.py:148-173 _parse_response
  * ragen/llm_agent/ctx_manager.py:35-70   get_masks_and_scores
  * ragen/llm_agent/ctx_manager.py:175-226 _normalize_score_tensor
  * ragen/llm_agent/ctx_manager.py:308-329 metric means
"""
import re
from itertools import zip_longest

import numpy as np
import torch

SPECIAL_STRINGS = ["<think>", "</think>", "<answer>", "</answer>", "<|im_start|>", "<|im_end|>"]  # ctx_manager.py:94


# ---------------------------------------------------------------------------------------------
# EnvStateManager
# ---------------------------------------------------------------------------------------------
class ManagerOracle:
    """CPU restatement of EnvStateManager over a list of oracle envs.

    envs: list of dicts {tag, env, max_actions_per_traj, action_lookup (dict int->str or None)};
    env_id is the list index, group_id = env_id // group_size.
    """

    def __init__(self, envs, group_size, format_penalty=-0.1):
        self.envs = envs
        self.group_size = group_size
        self.format_penalty = format_penalty
        self.status = None
        self.cache = None

    def group_seeds(self, seed):
        return [seed + i // self.group_size for i in range(len(self.envs))]

    def reset(self, seed, reset_fn=None):
        seeds = self.group_seeds(seed)
        self.status = []
        self.cache = []
        for i, (e, s) in enumerate(zip(self.envs, seeds)):
            obs = reset_fn(i, e, s) if reset_fn is not None else e["env"].reset(seed=s)
            self.status.append({"truncated": False, "terminated": False, "num_actions": 0, "rewards": [], "seed": s})
            self.cache.append({"env_id": i, "group_id": i // self.group_size, "tag": e["tag"], "penalty": 0,
                               "history": [{"state": obs, "actions_left": e["max_actions_per_traj"]}]})
        return self.cache

    @staticmethod
    def map_actions(action_lookup, actions):
        if action_lookup is None:
            return list(actions)
        rev = {v.lower(): k for k, v in action_lookup.items()}
        return [rev[a.lower()] for a in actions if a.lower() in rev]

    def step(self, env_inputs):
        alive = []
        for inp in env_inputs:
            i = inp["env_id"]
            e, st, cache = self.envs[i], self.status[i], self.cache[i]
            left = e["max_actions_per_traj"] - st["num_actions"]
            lookup = getattr(e["env"], "action_lookup", e.get("action_lookup"))
            valid = self.map_actions(lookup, inp["actions"])
            acc, info, done, executed = 0, {}, False, []
            for a in valid[:left]:
                _, r, d, step_info = e["env"].step(a)
                acc += r
                info.update(step_info)
                executed.append(a)
                if d:
                    done = True
                    break
            if len(valid) != len(inp["actions"]) or not valid:
                cache["penalty"] += self.format_penalty
            st["num_actions"] += len(executed)
            st["rewards"].append(acc)
            if done:
                st["terminated"] = True
                st["truncated"] = not info.get("success", False)
            cache["history"][-1].update({"actions": executed, "reward": acc, "info": info,
                                         "llm_response": inp["llm_response"], "llm_raw_response": inp["llm_raw_response"]})
            cache["history"].append({"state": e["env"].render(), "actions_left": e["max_actions_per_traj"] - st["num_actions"]})
            if st["num_actions"] >= e["max_actions_per_traj"] and not done:
                st["truncated"] = True
                st["terminated"] = True
                done = True
            if not done:
                alive.append(cache)
        return alive

    def rollout_states(self):
        for e, st, cache in zip(self.envs, self.status, self.cache):
            m = {"success": float(st["terminated"] and not st["truncated"]), "num_actions": st["num_actions"]}
            per_key = {}
            for turn in cache["history"]:
                for k, v in turn.get("info", {}).items():
                    if k != "success":
                        per_key.setdefault(k, []).append(float(v))
            for k, v in per_key.items():
                m[k] = np.sum(v) / (len(cache["history"]) - 1)
            cache["history"][-1]["metrics"] = per_key
            cache["metrics"] = {f"{e['tag']}/{k}": v for k, v in m.items()}
        return self.cache


def mean_metrics(env_outputs, env_nums):
    """ctx_manager.py:308-327."""
    acc = {}
    for out in env_outputs:
        for k, v in out["metrics"].items():
            acc.setdefault(k, []).append(v)
    res = {k: np.sum(v) / env_nums[k.split("/")[0]] for k, v in acc.items()}
    for k, v in acc.items():
        nz = [x for x in v if x != 0]
        if nz:
            head, tail = k.split("/", 1)
            res[f"{head}/non-zero/{tail}"] = np.mean(nz)
    return res


# ---------------------------------------------------------------------------------------------
# ContextManager passes
# ---------------------------------------------------------------------------------------------
def parse_response(response: str, enable_think=True, action_sep="||", max_actions=5):
    """_parse_response (ctx_manager.py:148-173): returns (llm_response, actions)."""
    rx = r"<think>(.*?)</think>\s*<answer>(.*?)</answer>" if enable_think else r"<answer>(.*?)</answer>"
    m = re.search(rx, response, re.DOTALL)
    if m is None:
        return response, []
    think, answer = (m.group(1), m.group(2)) if enable_think else ("", m.group(1))
    for tok in SPECIAL_STRINGS:
        answer = answer.replace(tok, "").strip()
        think = think.replace(tok, "").strip()
    actions = [a.strip() for a in answer.split(action_sep) if a.strip()]
    if len(actions) > max_actions:
        actions = actions[:max_actions]
        answer = (" " + action_sep + " ").join(actions)
    text = f"<think>{think}</think><answer>{answer}</answer>" if enable_think else f"<answer>{answer}</answer>"
    return text, actions


def masks_and_scores(input_ids, all_scores, im_start, im_end, use_turn_scores=False, enable_response_mask=True,
                     roll_scores=True):
    """get_masks_and_scores (ctx_manager.py:35-70) with the token ids given explicitly.
    roll_scores=True is the Qwen branch (:60-62)."""
    turn = torch.cumsum((input_ids == im_start).long(), dim=-1)
    response_mask = (turn % 2 == 1) & (turn > 1)
    loss_mask = response_mask if enable_response_mask else (turn > 1)
    score = torch.zeros_like(input_ids, dtype=torch.float32)
    if use_turn_scores:
        for k, col in enumerate(zip_longest(*all_scores, fillvalue=0)):
            where = (input_ids == im_end) & (turn == 2 * k + 3)
            where[~where.any(dim=-1), -1] = True
            score[where] = torch.tensor(col, dtype=torch.float32)
        if roll_scores:
            score = score.roll(shifts=1, dims=-1)
    else:
        score[:, -1] = torch.tensor([sum(s) for s in all_scores], dtype=torch.float32)
    return score[:, 1:], loss_mask[:, :-1], response_mask[:, :-1]


def normalize_scores(last_scores, penalties, group_keys, method):
    """_normalize_score_tensor (ctx_manager.py:175-226) on the last column only.
    last_scores, penalties: float32 [B]; group_keys: list of hashables; returns float32 [B]."""
    x = last_scores.clone().float() + penalties.float()
    groups = {}
    for i, g in enumerate(group_keys):
        groups.setdefault(g, []).append(i)
    if len(groups) >= x.shape[0]:
        return x

    def whiten(v):
        sd = v.std(dim=-1, keepdim=True)
        return (v - v.mean(dim=-1, keepdim=True)) / (sd + 1e-6) if sd.abs().max() > 1e-6 else torch.zeros_like(v)

    fn = {"identity": lambda v: v, "mean": lambda v: v - v.mean(dim=-1, keepdim=True), "mean_std": whiten,
          "asym_clip": lambda v: whiten(v).clamp(min=-1, max=3)}[method]
    for idx in groups.values():
        idx = torch.tensor(idx)
        x[idx] = fn(x[idx])
    return x
