"""ContextManager with the reference's interface (ragen/llm_agent/ctx_manager.py:74-356).

The token trajectories are assembled ON THE DEVICE (assembly="device", the default; ragen_b200/llm_agent/assembly.py): every
text chunk of a prompt is tokenised once into a token pool in HBM, grid observations go from the env kernels' glyph codes to
token ids without becoming text, and one launch gathers each row's segments into the left-padded input_ids / attention_mask /
position_ids the reference gets from "messages -> chat template -> tokenizer" (:248-280).  The reference's host path is kept as
assembly="host" (explicit choice, and the parity oracle of the tests).  The other tensor passes run on the device through
libragen_b200 as well: response parsing + action lookup (ragen_parse_actions), loss / response masks and score placement
(ragen_masks_scores), group reward normalisation (ragen_group_norm).  Outputs have the reference's DataProto layout
(batch / non_tensor_batch / meta_info).
"""
from dataclasses import asdict
from typing import Dict, List

import numpy as np
import torch

from .. import ops
from ..env import REGISTERED_ENV_CONFIGS

try:  # the real container when verl is installed (ragen/llm_agent/ctx_manager.py:14)
    from verl import DataProto  # type: ignore
    from tensordict import TensorDict  # type: ignore
except Exception:  # stand-in with the members this path touches (verl/verl/protocol.py:200)
    class TensorDict(dict):
        def __init__(self, source=None, batch_size=None):
            super().__init__(source or {})
            self.batch_size = batch_size

    class DataProto:
        def __init__(self, batch=None, non_tensor_batch=None, meta_info=None):
            self.batch = batch
            self.non_tensor_batch = non_tensor_batch if non_tensor_batch is not None else {}
            self.meta_info = meta_info if meta_info is not None else {}


class LazyBatch(dict):
    """non_tensor_batch whose `messages_list` is built when somebody reads it: on the device path no message string exists
    until then.  Reads through [], get, items, values and iteration over items all see the real value."""

    def __init__(self, base, key, thunk):
        super().__init__(base)
        self._lazy_key, self._thunk = key, thunk
        super().__setitem__(key, None)

    def _force(self):
        if self._thunk is not None:
            thunk, self._thunk = self._thunk, None
            super().__setitem__(self._lazy_key, thunk())

    def __getitem__(self, k):
        if k == self._lazy_key:
            self._force()
        return super().__getitem__(k)

    def get(self, k, default=None):
        if k == self._lazy_key:
            self._force()
        return super().get(k, default)

    def items(self):
        self._force()
        return super().items()

    def values(self):
        self._force()
        return super().values()

    def __setitem__(self, k, v):
        if k == self._lazy_key:
            self._thunk = None
        super().__setitem__(k, v)

    def copy(self):
        self._force()
        return dict(self)

    def __reduce__(self):
        self._force()
        return (dict, (dict(self),))


def get_special_tokens(tokenizer):
    """ctx_manager.py:24-33."""
    name = tokenizer.name_or_path.lower()
    if "qwen" in name:
        return tokenizer.encode("<|im_start|>")[0], tokenizer.encode("<|im_end|>")[0]
    if "llama-3" in name:
        return 128006, 128009
    raise ValueError(f"Unsupported model: {tokenizer.name_or_path}")


def get_masks_and_scores(input_ids, tokenizer, all_scores=None, use_turn_scores=False, enable_response_mask=False, device=None):
    """ctx_manager.py:35-70 on the device.  input_ids [B, L] (any device); all_scores: list of per-turn Python floats
    per row.  Returns (score_tensor f32, loss_mask bool, response_mask bool), all [B, L-1], on input_ids' device."""
    special_token, reward_token = get_special_tokens(tokenizer)
    home = input_ids.device
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    ids = input_ids.to(dev)
    B = ids.shape[0]
    K = max((len(s) for s in all_scores), default=0)
    ts = np.zeros((B, max(K, 1)), dtype=np.float64)
    for i, s in enumerate(all_scores):
        ts[i, :len(s)] = s
    nt = torch.tensor([len(s) for s in all_scores], dtype=torch.int32, device=dev)
    sc, lm, rm = ops.masks_scores(ids, torch.from_numpy(ts).to(dev) if K else None, nt, special_token, reward_token,
                                  use_turn_scores, enable_response_mask, "qwen" in tokenizer.name_or_path.lower())
    return sc.to(home), lm.bool().to(home), rm.bool().to(home)


class ContextManager:
    def __init__(self, config, tokenizer, processor=None, mode: str = "train", device: int = 0, assembly: str = "device",
                 output_device="cpu", stable_outputs: bool = False):
        """assembly: "device" (token trajectories gathered on the GPU from pre-tokenised segments; a tokenizer that cannot be
        decomposed that way is refused with TokenizerNotDecomposable) or "host" (the reference's string path).
        output_device: where the tensors of the returned DataProto live ("cpu" as in the reference, or "cuda").
        stable_outputs: with device assembly and CPU outputs, the tensors of a generation-time call (prepare_for_update=False) are
        views of a pinned staging buffer that is reused two calls later; True gives every call memory of its own."""
        if assembly not in ("device", "host"):
            raise ValueError(f"assembly must be 'device' or 'host', got {assembly!r}")
        self.config = config
        self.tokenizer = tokenizer
        self.processor = processor
        self.device = torch.device("cuda", int(device))
        self.assembly = assembly
        self.stable_outputs = bool(stable_outputs)
        self._host_ring, self._ring_next = [None, None], 0
        self.output_device = torch.device(output_device) if output_device != "cuda" else self.device
        self._assembler = None
        self.action_sep = self.config.agent_proxy.action_sep
        self.special_token_list = ["<think>", "</think>", "<answer>", "</answer>", "<|im_start|>", "<|im_end|>"]
        self.es_cfg = self.config.es_manager[mode]
        self.env_nums = {env_tag: n_group * self.es_cfg.group_size
                         for n_group, env_tag in zip(self.es_cfg.env_configs.n_groups, self.es_cfg.env_configs.tags)}
        self._init_prefix_lookup()
        self._tables = None        # action lookup tables on the device, built from the envs at first use
        self._env_table = None
        self.exchange = None       # set by shard(): batch-global metrics when env groups are spread over ranks

    def shard(self, world_size: int, group=None, extra_keys=()):
        """This manager sees only its rank's env groups (EnvStateManager(rank=, world_size=)); the rollout metrics of
        formulate_rollouts are then all-reduced so every rank reports the whole batch (ragen_b200/sharding.py)."""
        from ..sharding import RolloutExchange

        self.world_size = int(world_size)
        self.exchange = RolloutExchange(self.es_cfg.env_configs.tags, self.env_nums, extra_keys, group) if world_size > 1 else None
        return self

    def _check_env_installed(self, env_type: str):
        if env_type not in REGISTERED_ENV_CONFIGS:
            raise ValueError(f"Environment {env_type} is not installed. Please install it using the scripts/setup_{env_type}.sh script.")

    def _init_prefix_lookup(self):
        """ctx_manager.py:106-146."""
        prefix_lookup, prefixes, env_config_lookup = {}, {}, {}
        for env_tag, env_config in self.config.custom_envs.items():
            if env_tag not in self.es_cfg.env_configs.tags:
                continue
            self._check_env_installed(env_config.env_type)
            env_config_new = asdict(REGISTERED_ENV_CONFIGS[env_config.env_type]())
            for k, v in env_config.items():
                env_config_new[k] = v
            env_instruction = env_config_new.get("env_instruction", "")
            if env_config_new.get("grid_vocab", False):
                env_instruction += "\nThe meaning of each symbol in the state is:\n" + \
                    ", ".join([f"{k}: {v}" for k, v in env_config_new["grid_vocab"].items()])
            if env_config_new.get("action_lookup", False):
                s = "\nYour available actions are:\n" + ", ".join([f"{v}" for k, v in env_config_new["action_lookup"].items()])
                s += f"\nYou can make up to {env_config_new['max_actions_per_traj']} actions, separated by the action separator \" " + \
                    self.action_sep + " \"\n"
                env_instruction += s
            prefixes[env_tag] = env_instruction
            env_config_lookup[env_tag] = {"max_tokens": env_config.get("max_tokens", self.config.actor_rollout_ref.rollout.response_length)}
        cur_group = 0
        group_size = self.es_cfg.group_size
        for env_tag, n_group in zip(self.es_cfg.env_configs.tags, self.es_cfg.env_configs.n_groups):
            for i in range(cur_group * group_size, (cur_group + n_group) * group_size):
                prefix_lookup[i] = prefixes[env_tag]
                env_config_lookup[i] = env_config_lookup[env_tag]
            cur_group += n_group
        self.prefix_lookup = prefix_lookup
        self.env_config_lookup = env_config_lookup

    # ---- response parsing (ctx_manager.py:148-173, 332-352) ------------------------------------------------------
    def bind_envs(self, envs: List[Dict]):
        """Gives the parser the envs' `config.action_lookup` dicts (EnvStateManager.envs entries) so that the action-id
        lookup of es_manager.py:230-240 happens in the same kernel as the parse.  Call after every reset: Bandit
        re-draws its arm order then."""
        uniq, ids, env_table = [], {}, {}
        for e in envs:
            lk = getattr(e["env"].config, "action_lookup", None)
            if lk is None:
                env_table[e["env_id"]] = -1
                continue
            key = tuple(lk.items())
            if key not in ids:
                ids[key] = len(uniq)
                uniq.append(dict(lk))
            env_table[e["env_id"]] = ids[key]
        self._tables = ops.ActionTables(uniq, self.device) if uniq else None
        self._env_table = env_table
        if self._assembler is not None:
            self._assembler.new_rollout()

    def _parse_batch(self, responses: List[str], env_ids=None):
        tid = None
        if self._tables is not None and env_ids is not None:
            tid = [self._env_table.get(int(e), -1) for e in env_ids]
        res = ops.parse_actions(responses, self._tables if tid is not None else None, tid,
                                bool(self.config.agent_proxy.enable_think), self.action_sep,
                                int(self.config.agent_proxy.max_actions_per_turn), device=self.device)
        host = res.to_host()
        flags = getattr(res, "flags_host", None)
        return res, host, flags if flags is not None else res.flags.cpu().numpy()

    def _parse_response(self, response: str) -> List:
        _, host, _ = self._parse_batch([response])
        return host[0][0], host[0][1]

    def get_env_inputs(self, lm_outputs) -> List[Dict]:
        if lm_outputs.batch is not None and "responses" in lm_outputs.batch.keys():
            responses = self.tokenizer.batch_decode(lm_outputs.batch["responses"], skip_special_tokens=True)
        else:
            responses = lm_outputs.non_tensor_batch["response_texts"]
        head = "<think>" if self.config.agent_proxy.enable_think else "<answer>"
        responses = [head + r for r in responses]  # generation does not include the opening tag (:340)
        env_ids = lm_outputs.non_tensor_batch["env_ids"]
        _, host, flags = self._parse_batch(responses, env_ids)
        env_inputs = []
        for env_id, response, (llm_response, actions, codes), fl in zip(env_ids, responses, host, flags):
            item = {"env_id": env_id, "llm_raw_response": response, "llm_response": llm_response, "actions": actions}
            if fl & 16 or (self._env_table is not None and self._env_table.get(int(env_id), -1) >= 0):
                item["action_codes"] = codes  # ids of the valid actions, mapped on the device
            env_inputs.append(item)
        return env_inputs

    # ---- reward normalisation (ctx_manager.py:175-226) -------------------------------------------------------------
    def _normalize_score_tensor(self, score_tensor: torch.Tensor, env_outputs: List[Dict]) -> torch.Tensor:
        assert self.config.agent_proxy.use_turn_scores == False, \
            "Reward normalization is not supported for use_turn_scores == True"  # noqa: E712
        rn_cfg = self.config.agent_proxy.reward_normalization
        grouping, method = rn_cfg.grouping, rn_cfg.method
        if grouping == "state":
            group_tags = [env_output["group_id"] for env_output in env_outputs]
        elif grouping == "inductive":
            group_tags = [env_output["tag"] for env_output in env_outputs]
        elif grouping == "batch":
            group_tags = [1] * len(env_outputs)
        else:
            raise ValueError(f"Invalid grouping: {grouping}")
        if method not in ("mean_std", "mean", "asym_clip", "identity"):
            raise ValueError(f"Invalid normalization method: {method}")
        keys = self._cross_rank_norm_keys()
        if keys is not None:
            # sharded rows and groups that span ranks (per tag / whole batch): this rank's partial {count, sum, sum of squares}
            # per group ride in the rollout's one all-reduce (_mean_metrics), the normalisation is applied after it
            dev = self.device
            gi = torch.tensor([keys.index(t) for t in group_tags], dtype=torch.int32, device=dev)
            penalty = torch.tensor([env_output.get("penalty", 0) for env_output in env_outputs], dtype=torch.float32, device=dev)
            x, gstats = ops.group_norm_stats(score_tensor[:, -1].to(dev), penalty, gi, len(keys))
            self._pending_norm = (x, gi, gstats, method, score_tensor)
            return score_tensor
        dense = {}
        gi = [dense.setdefault(t, len(dense)) for t in group_tags]
        dev = self.device
        penalty = torch.tensor([env_output.get("penalty", 0) for env_output in env_outputs], dtype=torch.float32, device=dev)
        out = ops.group_norm(score_tensor[:, -1].to(dev), penalty, torch.tensor(gi, dtype=torch.int32, device=dev), len(dense), method)
        score_tensor[:, -1] = out.to(score_tensor.device)  # in place, as the reference (rm_scores aliases original_rm_scores)
        return score_tensor

    def _cross_rank_norm_keys(self):
        """The group keys of a normalisation whose groups span ranks (sharded rollout, grouping by tag or whole batch, a method
        other than identity), in an order every rank derives from the config alone; None when the normalisation is rank-local."""
        rn_cfg = self.config.agent_proxy.reward_normalization
        if self.exchange is None or rn_cfg.grouping == "state" or rn_cfg.method == "identity" or self.config.agent_proxy.use_turn_scores:
            return None
        return list(dict.fromkeys(self.es_cfg.env_configs.tags)) if rn_cfg.grouping == "inductive" else [1]

    # ---- LM inputs (ctx_manager.py:228-330) ---------------------------------------------------------------------------
    def _trim_history(self, env_output: Dict, prepare_for_update: bool):
        """In place, as the reference does (:241-246)."""
        if "state" in env_output["history"][-1] and prepare_for_update:
            env_output["history"] = env_output["history"][:-1]  # the state of turn n+1 is not part of the trajectory
        max_k = getattr(self.config.agent_proxy, "max_context_window", None)
        if max_k is not None and isinstance(max_k, int) and max_k > 0:
            env_output["history"] = env_output["history"][-max_k:]

    def _messages(self, env_output: Dict, prepare_for_update: bool) -> List[Dict]:
        """Chat messages of one trajectory (:248-267); trims `history` in place the way the reference does."""
        self._trim_history(env_output, prepare_for_update)
        return self._messages_of(env_output, prepare_for_update)

    def _render_text(self, env_output: Dict, prepare_for_update: bool) -> str:
        """The prompt text of an already trimmed trajectory (:269-274)."""
        text = self.tokenizer.apply_chat_template(self._messages_of(env_output, prepare_for_update),
                                                  add_generation_prompt=(not prepare_for_update), tokenize=False)
        if not prepare_for_update:
            text += "<think>" if self.config.agent_proxy.enable_think else "<answer>"  # the opening tag is forced
        return text

    def _messages_of(self, env_output: Dict, prepare_for_update: bool) -> List[Dict]:
        enable_think = self.config.agent_proxy.enable_think
        fmt = "<think> [Your thoughts] </think> <answer> [your answer] </answer>" if enable_think else "<answer> [your answer] </answer>"
        length = f"Max response length: {self.env_config_lookup[env_output['env_id']]['max_tokens']} words (tokens)."
        messages = [{"role": "system", "content": "You're a helpful assistant. "},
                    {"role": "user", "content": self.prefix_lookup[env_output["env_id"]]}]
        history = env_output["history"]
        for idx, content in enumerate(history):
            messages[-1]["content"] += f"\nTurn {idx + 1}:\n"
            if "state" in content:
                messages[-1]["content"] += (f"State:\n{content['state']}\nYou have {content['actions_left']} actions left. "
                                            f"Always output: {fmt} with no extra text. Strictly follow this format. {length}\n")
            if "llm_response" in content:
                messages.append({"role": "assistant", "content": content["llm_response"]})
            if "reward" in content and not (prepare_for_update and idx == len(history) - 1):
                messages.append({"role": "user", "content": f"Reward:\n{content['reward']}\n"})
        assert all(msg["role"] == "assistant" for msg in messages[2::2])  # the loss mask counts on this alternation
        return messages

    def _empty_lm_inputs(self, prepare_for_update: bool) -> DataProto:
        """A rank that owns no env group (fewer groups than ranks) still takes part in the metrics exchange."""
        ids = torch.zeros((0, 1), dtype=torch.long)
        batch = {"input_ids": ids, "attention_mask": ids.clone(), "position_ids": ids.clone(), "responses": ids[:, 1:]}
        out = DataProto()
        if prepare_for_update:
            score = torch.zeros((0, 0), dtype=torch.float32)
            batch.update({"loss_mask": torch.zeros((0, 0), dtype=torch.bool), "rm_scores": score, "original_rm_scores": score})
            self._response_tokens = (0.0, 0)
            out.meta_info = {"metrics": self._mean_metrics([], 0.0)}
        out.batch = TensorDict(batch, batch_size=0)
        out.non_tensor_batch = {"env_ids": np.array([], dtype=object), "group_ids": np.array([], dtype=object),
                                "messages_list": np.array([], dtype=object)}
        return out

    def get_lm_inputs(self, env_outputs: List[Dict], prepare_for_update: bool) -> DataProto:
        if len(env_outputs) == 0:
            return self._empty_lm_inputs(prepare_for_update)
        if self.assembly == "device":
            # segments of the token pool gathered on the GPU; no prompt string is built, nothing is re-tokenised
            if self._assembler is None:
                from .assembly import TrajectoryAssembler

                self._assembler = TrajectoryAssembler(self)
            for env_output in env_outputs:
                self._trim_history(env_output, prepare_for_update)
            input_ids, attention_mask, position_ids = self._assembler.assemble(env_outputs, prepare_for_update)
            messages_list = None  # built on first access (only API-calling worker groups read it, agent_proxy.py:96)
        else:
            llm_input_texts, messages_list = [], []
            for env_output in env_outputs:
                messages = self._messages(env_output, prepare_for_update)
                text = self.tokenizer.apply_chat_template(messages, add_generation_prompt=(not prepare_for_update), tokenize=False)
                if not prepare_for_update:
                    text += "<think>" if self.config.agent_proxy.enable_think else "<answer>"  # the opening tag is forced
                llm_input_texts.append(text)
                messages_list.append(messages)
            inputs = self.tokenizer(llm_input_texts, return_tensors="pt", padding=True, padding_side="left", truncation=False)
            input_ids, attention_mask = inputs.input_ids, inputs.attention_mask
            position_ids = attention_mask.cumsum(dim=-1)
        if prepare_for_update:
            scores = [[i.get("reward", 0.0) for i in env_output["history"]] for env_output in env_outputs]
            score_tensor, loss_mask, response_mask = get_masks_and_scores(
                input_ids, self.tokenizer, scores, use_turn_scores=self.config.agent_proxy.use_turn_scores,
                enable_response_mask=self.config.enable_response_mask, device=self.device)
            normalized_score_tensor = score_tensor
            if not self.config.agent_proxy.use_turn_scores:
                normalized_score_tensor = self._normalize_score_tensor(score_tensor, env_outputs)
            per_row = response_mask.sum(dim=-1).cpu()  # B integers; the mean is taken where the reference takes it (CPU fp32, :288)
            response_length = per_row.float().mean().item()
            self._response_tokens = (float(per_row.sum()), int(response_mask.shape[0]))
            metrics = self._mean_metrics(env_outputs, response_length)  # sharded: the rollout's one all-reduce happens here

        if self.assembly == "device" and self.output_device.type == "cpu":
            # the reference hands CPU tensors to the trainer: asynchronous copies into pinned memory, one synchronisation
            aliased = prepare_for_update and normalized_score_tensor is score_tensor
            dev_t = [input_ids, attention_mask, position_ids] + ([loss_mask, score_tensor] if prepare_for_update else [])
            if prepare_for_update and not aliased:
                dev_t.append(normalized_score_tensor)
            # Pinned staging: two grow-only buffers used in turn.  The per-turn prompts (prepare_for_update=False) are VIEWS of a
            # staging buffer - they are consumed by generate_sequences at once and stay valid until the second next call; the final
            # batch of a rollout is copied out into memory of its own.  stable_outputs=True copies every call's outputs.
            sizes = [(t.numel() * t.element_size() + 15) & ~15 for t in dev_t]
            ring = self._host_ring
            j = self._ring_next
            self._ring_next = j ^ 1
            if ring[j] is None or ring[j].numel() < sum(sizes):
                ring[j] = torch.empty(max(sum(sizes) * 5 // 4, 1 << 20), dtype=torch.uint8, pin_memory=True)
            host_t, at = [], 0
            for t, nb in zip(dev_t, sizes):
                raw = ring[j][at: at + t.numel() * t.element_size()]
                host_t.append(raw.view(t.dtype).view(t.shape))
                at += nb
            for h, t in zip(host_t, dev_t):
                h.copy_(t, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            if prepare_for_update or self.stable_outputs:
                host_t = [h.clone() for h in host_t]
            input_ids, attention_mask, position_ids = host_t[:3]
            if prepare_for_update:
                loss_mask, score_tensor = host_t[3], host_t[4]
                normalized_score_tensor = score_tensor if aliased else host_t[5]
        llm_inputs = DataProto()
        llm_inputs.batch = TensorDict({"input_ids": input_ids, "attention_mask": attention_mask, "position_ids": position_ids,
                                       "responses": input_ids[:, 1:]}, batch_size=input_ids.shape[0])
        if prepare_for_update:
            llm_inputs.batch["loss_mask"] = loss_mask
            llm_inputs.batch["rm_scores"] = normalized_score_tensor
            llm_inputs.batch["original_rm_scores"] = score_tensor  # the SAME tensor when normalised: the reference aliases them
        non_tensor = {
            "env_ids": np.array([env_output["env_id"] for env_output in env_outputs], dtype=object),
            "group_ids": np.array([env_output["group_id"] for env_output in env_outputs], dtype=object),
        }
        if messages_list is None:
            # the managers keep mutating these dicts (es_manager.py:212-228): the messages are those of THIS call
            # (only the LAST entry of a history is ever updated in place; the finished ones are shared, the list itself is copied)
            snapshot = [{"env_id": eo["env_id"], "history": eo["history"][:-1] + [dict(eo["history"][-1])]} for eo in env_outputs]
            non_tensor = LazyBatch(non_tensor, "messages_list", lambda: np.array(
                [self._messages_of(eo, prepare_for_update) for eo in snapshot], dtype=object))
        else:
            non_tensor["messages_list"] = np.array(messages_list, dtype=object)
        llm_inputs.non_tensor_batch = non_tensor
        if prepare_for_update:
            llm_inputs.meta_info = {"metrics": metrics}
        return llm_inputs

    def _mean_metrics(self, env_outputs: List[Dict], response_length: float) -> Dict:
        """Per-tag means over ALL envs of the tag plus the non-zero means (:309-328)."""
        if self.exchange is not None:  # one all-reduce of the fixed-layout vector, then the same arithmetic on global sums
            ex = self.exchange
            on_dev = self.device if torch.distributed.get_backend(ex.group) == "nccl" else None
            vec = ex.pack(env_outputs, *self._response_tokens, device=on_dev)
            keys = self._cross_rank_norm_keys()
            pending, self._pending_norm = getattr(self, "_pending_norm", None), None
            if keys is not None:  # tail: per-key {count, sum, sum of squares} of this rank's rows, then its row count
                tail = torch.zeros(3 * len(keys) + 1, dtype=torch.float64, device=vec.device)
                if pending is not None:
                    tail[:-1] = pending[2].flatten().to(vec.device)
                    tail[-1] = float(pending[0].numel())
                vec = torch.cat([vec, tail])
            vec = ex.all_reduce(vec)
            if keys is not None and pending is not None:
                x, gi, _, method, score_tensor = pending
                gstats, rows = vec[ex.size:-1].reshape(len(keys), 3), float(vec[-1])
                if len(keys) < rows:  # "if len(group2index) < acc_scores.shape[0]" (:219), on the whole batch
                    ops.group_norm_apply(x, gi, len(keys), gstats, method)
                score_tensor[:, -1] = x.to(score_tensor.device)
            return ex.unpack(vec[: ex.size])[0]
        metrics = {}
        for env_output in env_outputs:
            for key, value in env_output["metrics"].items():
                metrics.setdefault(key, []).append(value)
        mean_metrics = {key: np.sum(value) / self.env_nums[key.split("/")[0]] for key, value in metrics.items()}
        for key, values in metrics.items():
            prefix, suffix = key.split("/", 1)
            non_zero_values = [v for v in values if v != 0]
            if non_zero_values:
                mean_metrics[f"{prefix}/non-zero/{suffix}"] = np.mean(non_zero_values)
        mean_metrics["response_length"] = response_length
        return mean_metrics

    def formulate_rollouts(self, env_outputs: List[Dict]) -> DataProto:
        return self.get_lm_inputs(env_outputs, prepare_for_update=True)
