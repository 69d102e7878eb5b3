"""ORACLE (test infrastructure, CPU) — the reference's rollout loop restated end to end, for the manager-level CPU baseline.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU legs (`bench.py --rows`) may import this.

Follows
  * ragen/llm_agent/agent_proxy.py:143-159   LLMAgentProxy.rollout (reset -> [lm inputs -> generate -> env inputs -> step] x max_turn
                                             -> rollout states -> formulate_rollouts)
  * ragen/llm_agent/es_manager.py:75-207     EnvStateManager (oracle/managers.py: ManagerOracle over per-env Python oracle envs)
  * ragen/llm_agent/ctx_manager.py:106-146   instruction prefix
  * ragen/llm_agent/ctx_manager.py:228-330   get_lm_inputs: messages -> chat template -> tokenizer (whole history, every turn),
                                             masks / scores / normalisation / metrics when preparing the update
  * ragen/llm_agent/ctx_manager.py:332-352   get_env_inputs: decode -> regex parse

This is the execution model the GPU path is measured against: one Python env object per env, a serial step loop, and the whole
conversation rebuilt as strings and re-tokenised for every env on every turn.  The Sokoban rooms come from the caller (the
reference's generator is ~4 ms per 6x6 room and dominates a reset; the rollout rows time the turn loop)."""
import numpy as np
import torch

from . import managers as M
from . import sokoban as S


def instruction_prefix(env_instruction, grid_vocab, action_lookup, max_actions_per_traj, action_sep):
    """ctx_manager.py:117-131."""
    text = env_instruction
    if grid_vocab:
        text += "\nThe meaning of each symbol in the state is:\n" + ", ".join(f"{k}: {v}" for k, v in grid_vocab.items())
    if action_lookup:
        text += "\nYour available actions are:\n" + ", ".join(f"{v}" for v in action_lookup.values())
        text += f"\nYou can make up to {max_actions_per_traj} actions, separated by the action separator \" " + action_sep + " \"\n"
    return text


class RolloutPort:
    """SimpleSokoban-style tags only (what C1 / C5 use): n_groups x group_size Sokoban envs of one shape."""

    def __init__(self, tokenizer, n_groups, group_size, dim_room=(6, 6), num_boxes=1, max_steps=100, max_actions_per_traj=10,
                 max_turn=5, max_actions_per_turn=5, enable_think=True, action_sep="||", env_instruction="", max_tokens=100,
                 norm_method="identity", tag="SimpleSokoban", format_penalty=-0.1):
        self.tok = tokenizer
        self.n, self.group_size = n_groups * group_size, group_size
        self.H, self.W = dim_room
        self.max_turn, self.max_actions_per_turn = max_turn, max_actions_per_turn
        self.enable_think, self.action_sep, self.max_tokens = enable_think, action_sep, max_tokens
        self.norm_method, self.tag = norm_method, tag
        self.lookup = {1: "Up", 2: "Down", 3: "Left", 4: "Right"}
        self.envs = [{"tag": tag, "env": S.SokobanOracle(dim_room, num_boxes, max_steps), "max_actions_per_traj": max_actions_per_traj,
                      "action_lookup": self.lookup} for _ in range(self.n)]
        self.mgr = M.ManagerOracle(self.envs, group_size, format_penalty)
        vocab = {"#": "wall", "_": "empty", "O": "target", "√": "box on target", "X": "box", "P": "player", "S": "player on target"}
        self.prefix = instruction_prefix(env_instruction, vocab, self.lookup, max_actions_per_traj, action_sep)
        name = tokenizer.name_or_path.lower()
        assert "qwen" in name
        self.im_start, self.im_end = tokenizer.encode("<|im_start|>")[0], tokenizer.encode("<|im_end|>")[0]

    def reset(self, fixed, state):
        """rooms [n, H*W] (the caller generated them once per group, as es_manager.py:80-95 does per env)."""
        def load(i, e, seed):
            return e["env"].load(fixed[i], state[i])
        return self.mgr.reset(0, reset_fn=load)

    def lm_inputs(self, env_outputs, prepare_for_update):
        texts = []
        fmt = "<think> [Your thoughts] </think> <answer> [your answer] </answer>" if self.enable_think else "<answer> [your answer] </answer>"
        length = f"Max response length: {self.max_tokens} words (tokens)."
        for eo in env_outputs:
            if "state" in eo["history"][-1] and prepare_for_update:
                eo["history"] = eo["history"][:-1]
            messages = [{"role": "system", "content": "You're a helpful assistant. "}, {"role": "user", "content": self.prefix}]
            hist = eo["history"]
            for idx, c in enumerate(hist):
                messages[-1]["content"] += f"\nTurn {idx + 1}:\n"
                if "state" in c:
                    messages[-1]["content"] += (f"State:\n{c['state']}\nYou have {c['actions_left']} actions left. Always output: {fmt} "
                                                f"with no extra text. Strictly follow this format. {length}\n")
                if "llm_response" in c:
                    messages.append({"role": "assistant", "content": c["llm_response"]})
                if "reward" in c and not (prepare_for_update and idx == len(hist) - 1):
                    messages.append({"role": "user", "content": f"Reward:\n{c['reward']}\n"})
            text = self.tok.apply_chat_template(messages, add_generation_prompt=(not prepare_for_update), tokenize=False)
            if not prepare_for_update:
                text += "<think>" if self.enable_think else "<answer>"
            texts.append(text)
        enc = self.tok(texts, return_tensors="pt", padding=True, padding_side="left", truncation=False)
        ids, att = enc.input_ids, enc.attention_mask
        out = {"input_ids": ids, "attention_mask": att, "position_ids": att.cumsum(dim=-1), "env_ids": [eo["env_id"] for eo in env_outputs]}
        if prepare_for_update:
            scores = [[h.get("reward", 0.0) for h in eo["history"]] for eo in env_outputs]
            sc, lm, rm = M.masks_and_scores(ids, scores, self.im_start, self.im_end, False, True, True)
            pen = torch.tensor([eo.get("penalty", 0) for eo in env_outputs], dtype=torch.float32)
            sc[:, -1] = M.normalize_scores(sc[:, -1], pen, [eo["group_id"] for eo in env_outputs], self.norm_method)
            metrics = M.mean_metrics(env_outputs, {self.tag: self.n})
            metrics["response_length"] = rm.sum(dim=-1).float().mean().item()
            out.update({"loss_mask": lm, "rm_scores": sc, "metrics": metrics})
        return out

    def env_inputs(self, env_ids, responses):
        head = "<think>" if self.enable_think else "<answer>"
        out = []
        for env_id, r in zip(env_ids, responses):
            r = head + r
            llm_response, actions = M.parse_response(r, self.enable_think, self.action_sep, self.max_actions_per_turn)
            out.append({"env_id": env_id, "llm_raw_response": r, "llm_response": llm_response, "actions": actions})
        return out

    def rollout(self, fixed, state, policy):
        """policy(lm_inputs dict) -> list of response strings (one per row)."""
        env_outputs = self.reset(fixed, state)
        for _ in range(self.max_turn):
            lm = self.lm_inputs(env_outputs, False)
            env_outputs = self.mgr.step(self.env_inputs(lm["env_ids"], policy(lm)))
            if not env_outputs:
                break
        return self.lm_inputs(self.mgr.rollout_states(), True)
