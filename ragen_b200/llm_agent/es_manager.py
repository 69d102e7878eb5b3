"""EnvStateManager with the reference's interface (ragen/llm_agent/es_manager.py:28-258), backed by device batches.

Same constructor, `reset(seed)`, `step(all_env_inputs)`, `get_rollout_states()`, `render()`, `close()`; same
`envs` entries and `rollout_cache` dict shapes, so ContextManager / LLMAgentProxy / the trainer use it unchanged.
What differs is where the work runs: all envs of a tag live in ONE device batch and a manager call is one kernel
launch per tag (reset = room/map generation, step = up to K actions per env + reward + done + render), followed by one
device->host copy of the per-env results; the Python loop that remains only files those results into the history
dicts the API exposes.
"""
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np
import torch

from ..env import BATCH_TYPES, REGISTERED_ENV_CONFIGS, REGISTERED_ENVS
from ..env.frozen_lake import fl_glyph_table
from ..env.sokoban import glyph_table, unpack_glyphs
from .assembly import ObsLog


@dataclass
class EnvStatus:
    """Status of an environment (es_manager.py:17-24)."""
    truncated: bool = False
    terminated: bool = False
    num_actions: int = 0
    rewards: List[float] = field(default_factory=list)
    seed: Optional[int] = None


def _translate_rows(codes: np.ndarray, table) -> List[str]:
    """uint8 glyph rows [k, G] -> strings (last separator dropped); one C-level translate per row."""
    tr = {i: g for i, g in enumerate(table)}
    return [bytes(r[:-1]).decode("latin1").translate(tr) for r in codes]


class _Block:
    """All envs of one tag: a device batch + the entries that view it."""

    def __init__(self, tag, env_type, batch, first, count):
        self.tag, self.env_type, self.batch, self.first, self.count = tag, env_type, batch, first, count
        self.K = 1
        self.act = self.nact = None
        self.log = None  # ObsLog of a grid block in text mode: the glyph rows of every turn stay on the device
        self.stage = None     # pinned + device staging of a turn's actions
        self.host_obs = None  # glyph rows the last step copied back together with its results


class EnvStateManager:
    def __init__(self, config, mode: str = "train", device: int = 0, rank: int = 0, world_size: int = 1):
        """rank / world_size: this process keeps only the env groups with group_id % world_size == rank (env_id and
        group_id stay global, so the union over ranks is the single-process manager; ragen_b200/sharding.py)."""
        self.sys_config = config
        self.mode = mode
        self.device = int(device)
        self.rank, self.world_size = int(rank), int(world_size)
        self.config = getattr(self.sys_config.es_manager, mode)
        self.env_groups = int(self.config.env_groups)
        self.group_size = self.config.group_size
        self._init_envs()
        self.rollout_cache = None

    # ---- construction (es_manager.py:42-73) -----------------------------------------------------------------
    def _init_envs(self):
        cfg = self.config
        assert sum(cfg.env_configs.n_groups) == self.env_groups, \
            f"Sum of n_groups must equal env_groups. Got sum({cfg.env_configs.n_groups}) != {self.env_groups}"
        assert len(cfg.env_configs.tags) == len(cfg.env_configs.n_groups), \
            f"Number of tags must equal number of n_groups. Got {len(cfg.env_configs.tags)} != {len(cfg.env_configs.n_groups)}"
        self.envs, self.blocks = [], []
        K = int(getattr(self.sys_config.agent_proxy, "max_actions_per_turn", 5))
        done_groups = 0
        for tag, n_group in zip(cfg.env_configs.tags, cfg.env_configs.n_groups):
            tpl = self.sys_config.custom_envs[tag]
            env_type = tpl.env_type
            groups = [g for g in range(done_groups, done_groups + n_group) if g % self.world_size == self.rank]
            done_groups += n_group
            if not groups:
                continue
            first, count = len(self.envs), len(groups) * self.group_size
            make_cfg = (lambda: REGISTERED_ENV_CONFIGS[env_type]()) if tpl.env_config is None else \
                (lambda: REGISTERED_ENV_CONFIGS[env_type](**tpl.env_config))
            c0 = make_cfg()
            if env_type == "sokoban":
                batch = BATCH_TYPES[env_type](count, tuple(c0.dim_room), c0.num_boxes, c0.max_steps, c0.search_depth, device=self.device)
            elif env_type == "frozen_lake":
                batch = BATCH_TYPES[env_type](count, c0.size, c0.p, c0.is_slippery, device=self.device)
            else:
                batch = BATCH_TYPES[env_type](count, c0, device=self.device)
            block = _Block(tag, env_type, batch, first, count)
            block.K = max(K, 1)
            extra = {}
            if env_type == "countdown":
                from ..env.countdown import load_problems
                extra["data"] = load_problems(c0)
            for i in range(count):
                env_config = make_cfg()
                env_obj = REGISTERED_ENVS[env_type](env_config, batch=batch, slot=i, **extra)
                group_id = groups[i // self.group_size]
                self.envs.append({"tag": tag, "group_id": group_id, "env_id": group_id * self.group_size + i % self.group_size,
                                  "env": env_obj, "config": env_config, "status": EnvStatus(),
                                  "max_actions_per_traj": tpl.max_actions_per_traj})
            self.blocks.append(block)
        # env_id -> position in self.envs / self.rollout_cache (the identity when nothing is sharded away)
        self._pos = {e["env_id"]: i for i, e in enumerate(self.envs)}
        self._block_of = {}
        for b in self.blocks:
            for i in range(b.count):
                self._block_of[self.envs[b.first + i]["env_id"]] = b

    # ---- observations ----------------------------------------------------------------------------------------
    def _block_texts(self, block, record=False) -> List:
        """Current observation of every env of the block: text, or a list of images (one device->host copy for grid envs)."""
        envs = self.envs[block.first:block.first + block.count]
        if block.env_type == "sokoban" and envs[0]["env"].render_mode == "rgb_array":
            # one blit launch + one device->host copy for the whole block, then _handle_mm_state (:242-250) per env
            import PIL.Image

            frames = block.batch.render_rgb().cpu().numpy()
            return [[PIL.Image.fromarray(f, mode="RGB")] for f in frames]
        if block.env_type in ("sokoban", "frozen_lake"):
            b = block.batch
            if block.env_type == "sokoban":  # packed glyph rows (two codes per byte) when the room shape allows: half the copy
                if getattr(b, "obs_stale", False):
                    b.render(packed=b.obs_is_packed)
                if b.obs_is_packed == 3:
                    b.render(packed=b.can_pack)
                packed = bool(b.obs_is_packed)
                dev = b.obs_packed if packed else b.obs
                table, H, W = glyph_table(envs[0]["env"].GRID_LOOKUP), b.H, b.W
            else:
                packed, dev = False, b.obs
                table, H, W = fl_glyph_table(envs[0]["env"].GRID_LOOKUP), b.size, b.size
            fresh = getattr(block, "host_obs", None)  # the step just brought these rows over with its results
            block.host_obs = None
            if fresh is not None and fresh[0] is dev and fresh[1] == packed:
                host = fresh[2]
            else:
                host = dev.cpu().numpy()
            codes = unpack_glyphs(host) if packed else host
            texts = _translate_rows(codes, table)
            if record:
                # the device keeps the glyph rows of the turn and the strings carry a reference to them: the ContextManager
                # assembles token trajectories from the codes without reading the characters (llm_agent/assembly.py)
                if block.log is None:
                    block.log = ObsLog(block.count, H, W, table, dev.device)
                    block.log.begin()
                texts = block.log.wrap(texts, block.log.record(dev, packed, codes))
            return texts
        return [e["env"].render() for e in envs]

    # ---- reset (es_manager.py:75-103) ------------------------------------------------------------------------
    def reset(self, seed: Optional[int] = None):
        import random

        envs = self.envs
        rollout_cache = [{"env_id": e["env_id"], "history": [], "group_id": e["group_id"], "tag": e["tag"], "penalty": 0}
                         for e in envs]
        if self.mode == "train":
            seed = random.randint(0, 1000000) if seed is None else seed
        else:
            seed = 123
        seeds = [seed + e["group_id"] for e in envs]  # every env of a group shares the seed (:80-82)
        for block in self.blocks:
            bs = seeds[block.first:block.first + block.count]
            entries = envs[block.first:block.first + block.count]
            if block.env_type == "sokoban":
                used = block.batch.reset(bs)  # one room per distinct seed, generated on the device
                for e, u in zip(entries, used):
                    e["env"].seed_used = u
            elif block.env_type == "frozen_lake":
                block.batch.reset(bs)
                for e, s in zip(entries, bs):
                    e["config"].map_seed = s
            elif block.env_type == "bandit":
                block.batch.reset(bs)
                lo_first = block.batch.lo_first.cpu().numpy()
                for e, lf in zip(entries, lo_first):
                    e["env"].adopt_reset(bool(lf))
            else:
                for e, s in zip(entries, bs):
                    e["env"].reset(seed=s, mode=self.mode)
        for s, e in zip(seeds, envs):
            e["status"] = EnvStatus(seed=s)
        for block in self.blocks:
            if block.log is not None:
                block.log.begin()  # a new rollout: references into the previous one go stale
            texts = self._block_texts(block, record=True)
            for i, t in enumerate(texts):
                cache, e = rollout_cache[block.first + i], envs[block.first + i]
                cache["history"] = self._update_cache_history(cache["history"], next_state=t,
                                                              actions_left=e["max_actions_per_traj"], num_actions_info=None)
        self.rollout_cache = rollout_cache
        return rollout_cache

    # ---- step (es_manager.py:105-171) --------------------------------------------------------------------------
    def step(self, all_env_inputs: List[Dict]):
        envs = self.envs
        # 1. valid actions per input (:230-240); ContextManager.get_env_inputs already mapped them on the device
        plans = []
        per_block = {}
        for env_input in all_env_inputs:
            entry = envs[self._pos[env_input["env_id"]]]
            left = entry["max_actions_per_traj"] - entry["status"].num_actions
            valid = self._extract_map_valid_actions(entry, env_input["actions"], env_input.get("action_codes"))
            todo = valid[:left] if left > 0 else []
            plans.append((env_input, entry, valid, todo))
            per_block.setdefault(id(self._block_of[entry["env_id"]]), []).append(len(plans) - 1)
        # 2. one launch per tag
        results = {}
        for block in self.blocks:
            idxs = per_block.get(id(block))
            if idxs:
                self._run_block(block, [plans[i] for i in idxs], idxs, results)
        # 3. file the results into status / history, in input order
        env_outputs = []
        texts = {id(b): self._block_texts(b, record=True) for b in self.blocks if id(b) in per_block}
        for i, (env_input, entry, valid, todo) in enumerate(plans):
            block = self._block_of[entry["env_id"]]
            pos = self._pos[entry["env_id"]]  # position in self.envs / self.rollout_cache
            acc_reward, turn_info, turn_done, n_exec = results[i]
            executed = todo[:n_exec]
            if len(valid) != len(env_input["actions"]) or not valid:
                self.rollout_cache[pos]["penalty"] += self.sys_config.es_manager.format_penalty
            status = entry["status"]
            status.num_actions += len(executed)
            status.rewards.append(acc_reward)
            actions_left = entry["max_actions_per_traj"] - status.num_actions
            if turn_done:
                status.terminated = True
                status.truncated = not turn_info.get("success", False)
            history = self._update_cache_history(
                self.rollout_cache[pos]["history"], next_state=texts[id(block)][pos - block.first], actions_left=actions_left,
                num_actions_info={"actions": executed, "reward": acc_reward, "info": turn_info,
                                  "llm_response": env_input["llm_response"], "llm_raw_response": env_input["llm_raw_response"]})
            if status.num_actions >= entry["max_actions_per_traj"] and not turn_done:
                status.truncated = True
                status.terminated = True
                turn_done = True
            self.rollout_cache[pos]["history"] = history
            if not turn_done:
                env_outputs.append(self.rollout_cache[pos])
        return env_outputs

    def _run_block(self, block, plans, idxs, results):
        """Executes the planned actions of one tag's envs with one launch and stores, per plan index,
        (acc_reward, turn_info, turn_done, n_executed) - what _execute_actions returns (:116-128)."""
        b, dev = block.batch, block.batch.device
        if block.env_type in ("sokoban", "frozen_lake"):
            K = max(block.K, max(len(p[3]) for p in plans), 1)
            n = block.count
            # actions and per-env action counts travel in ONE pinned staging buffer [n * K | n]: one host->device copy per turn
            stage = getattr(block, "stage", None)
            if stage is None or stage[0].numel() != n * (K + 1):
                pinned = torch.zeros(n * (K + 1), dtype=torch.uint8, pin_memory=True)
                stage = block.stage = (pinned, torch.empty(n * (K + 1), dtype=torch.uint8, device=dev), K)
            pinned, d_stage, _ = stage
            host_stage = pinned.numpy()
            host_stage[:] = 0
            act, nact = host_stage[: n * K].reshape(n, K), host_stage[n * K:]
            for env_input, entry, valid, todo in plans:
                s = self._pos[entry["env_id"]] - block.first
                codes = [entry["env"].action_map[a] + 1 for a in todo] if block.env_type == "frozen_lake" else todo
                act[s, :len(codes)] = codes
                nact[s] = len(codes)
            d_stage.copy_(pinned, non_blocking=True)
            d_act, d_nact = d_stage[: n * K].view(n, K), d_stage[n * K:]
            kw = {}
            if block.env_type == "sokoban":
                text = self.envs[block.first]["env"].render_mode == "text"
                kw = {"packed": b.can_pack and text, "render": text}
            out = b.step(d_act, d_nact, **kw)
            if block.env_type == "sokoban" and kw["render"] and kw["packed"]:
                # reward / done / flags / n_executed and the nibble glyph rows are neighbours in the batch's result slab:
                # ONE device->host copy brings the whole turn back (the texts below are built from this copy)
                slab = b._slab[b._off_sc: b._slab_bytes].cpu().numpy()
                rew = slab[: 8 * n].view(np.float64)
                done_h, flags_h, nexec_h = slab[8 * n: 9 * n], slab[9 * n: 10 * n], slab[10 * n: 11 * n]
                block.host_obs = (b.obs_packed, True, slab[b._off_obs - b._off_sc:].reshape(n, -1))
            else:
                host = torch.stack([out.done, out.flags, out.n_executed]).cpu().numpy()
                done_h, flags_h, nexec_h = host[0], host[1], host[2]
                rew = out.reward.cpu().numpy()
                block.host_obs = None
            for i, (env_input, entry, valid, todo) in zip(idxs, plans):
                s = self._pos[entry["env_id"]] - block.first
                n_exec, fl = int(nexec_h[s]), int(flags_h[s])
                info = {"action_is_effective": bool(fl & 1), "action_is_valid": True, "success": bool(fl & 2)} if n_exec else {}
                results[i] = (float(rew[s]) if n_exec else 0, info, bool(done_h[s]), n_exec)
        elif block.env_type == "bandit":
            act = np.zeros(block.count, dtype=np.uint8)
            on = np.zeros(block.count, dtype=np.uint8)
            for env_input, entry, valid, todo in plans:
                if todo:
                    s = self._pos[entry["env_id"]] - block.first
                    assert todo[0] in entry["env"].ACTION_LOOKUP, f"Invalid action: {todo[0]}"
                    act[s], on[s] = todo[0], 1
            rew, fl = b.step(torch.from_numpy(act).to(dev), torch.from_numpy(on).to(dev))
            rew, fl = rew.cpu().numpy(), fl.cpu().numpy()
            for i, (env_input, entry, valid, todo) in zip(idxs, plans):
                s = self._pos[entry["env_id"]] - block.first
                if todo:
                    _, r, done, info = entry["env"].finish_step(todo[0], float(rew[s]), int(fl[s]))
                    results[i] = (0 + r, info, done, 1)
                else:
                    results[i] = (0, {}, False, 0)
        elif block.env_type == "metamathqa":
            # free-text answers; a wrong answer does not end the episode, so action r of every env still going is scored in
            # round r - one launch per round instead of one regex pass per (env, action)
            state = {i: [0, {}, False, 0] for i in idxs}  # acc_reward, info, done, executed
            for r in range(max((len(p[3]) for p in plans), default=0)):
                live = [(i, p) for i, p in zip(idxs, plans) if len(p[3]) > r and not state[i][2]]
                if not live:
                    break
                bits = b.score([p[3][r] for _, p in live], [p[1]["env"].correct_answer or "" for _, p in live], "metamathqa")
                for (i, p), bt in zip(live, bits):
                    _, reward, done, info = p[1]["env"].finish_step(bool(bt & 1), bool(bt & 2))
                    st = state[i]
                    st[0] += reward
                    st[1].update(info)
                    st[2], st[3] = done, st[3] + 1
            for i in idxs:
                results[i] = tuple(state[i])
        else:  # countdown: actions are free text, the first one ends the episode
            todo_plans = [(i, p) for i, p in zip(idxs, plans) if p[3]]
            if todo_plans:
                gts = [p[1]["env"].data[p[1]["env"].index] for _, p in todo_plans]
                res, uns = b.score([p[3][0] for _, p in todo_plans], [list(g["nums"]) for g in gts], [g["target"] for g in gts])
                for (i, p), r, u in zip(todo_plans, res, uns):
                    env = p[1]["env"]
                    env.note_unsupported(p[3][0], bool(u))  # warns: scored as an eval() exception, never re-evaluated on the host
                    _, reward, done, info = env.finish_step(env.reward_from_result(int(r)))
                    results[i] = (0 + reward, info, done, 1)
            for i, p in zip(idxs, plans):
                if not p[3]:
                    results[i] = (0, {}, False, 0)

    def _extract_map_valid_actions(self, entry: Dict, actions: List[str], codes=None):
        """es_manager.py:230-240.  `codes` are the ids ragen_parse_actions already produced for these actions (set by
        ContextManager.get_env_inputs); inputs built elsewhere are mapped here with the same rule."""
        action_lookup = getattr(entry["env"].config, "action_lookup", None)
        if action_lookup is None:
            return actions
        if codes is not None:
            return list(codes)
        cached = entry.get("_rev")
        if cached is None or cached[0] is not action_lookup:  # Bandit installs a new lookup dict at every reset
            cached = entry["_rev"] = (action_lookup, {v.lower(): k for k, v in action_lookup.items()})
        rev = cached[1]
        return [rev[a] for a in (x.lower() for x in actions) if a in rev]

    # ---- final states (es_manager.py:173-207) --------------------------------------------------------------------
    def get_rollout_states(self):
        TURN_LVL_METRICS = ["action_is_effective", "action_is_valid", "end_of_page"]
        for entry, cache in zip(self.envs, self.rollout_cache):
            status = entry["status"]
            env_metric = {"success": float(status.terminated and (not status.truncated)), "num_actions": status.num_actions}
            custom_metric = {}
            for turn in cache["history"]:
                for k, v in turn.get("info", {}).items():
                    if k == "success":
                        continue
                    custom_metric.setdefault(k, []).append(float(v))
            for k, v in custom_metric.items():
                if "Webshop" not in k or ("Webshop" in k and k in TURN_LVL_METRICS):
                    env_metric[k] = np.sum(v) / (len(cache["history"]) - 1)
                else:
                    env_metric[k] = np.sum(v)
            cache["history"][-1]["metrics"] = custom_metric
            cache["metrics"] = {f"{entry['tag']}/{k}": v for k, v in env_metric.items()}
            if entry["tag"] == "MetamathQA":  # es_manager.py:205-206
                cache["correct_answer"] = entry["env"].correct_answer
        return self.rollout_cache

    def _update_cache_history(self, history: List[Dict], next_state, actions_left, num_actions_info: Optional[Dict] = None):
        """es_manager.py:212-228."""
        if num_actions_info is not None:
            assert len(history), "History should not be empty"
            history[-1].update(num_actions_info)
        if isinstance(next_state, str):
            history.append({"state": next_state, "actions_left": actions_left})
        else:  # list of images
            history.append({"state": "<images>" * len(next_state), "images": next_state, "actions_left": actions_left})
        return history

    def render(self):
        out = []
        for block in self.blocks:
            out.extend(self._block_texts(block))
        return out

    def close(self):
        for entry in self.envs:
            entry["env"].close()
