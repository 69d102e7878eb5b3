"""LLMAgentProxy with the reference's interface (ragen/llm_agent/agent_proxy.py:118-159).

The rollout loop is the reference's: reset -> [get_lm_inputs -> generate_sequences -> get_env_inputs -> step] x max_turn ->
get_rollout_states -> formulate_rollouts.  Both managers are the device-backed ones of this package; the policy LLM is
not part of this path: `actor_rollout_wg` is whatever the trainer hands in (a verl RayWorkerGroup, the reference's
VllmWrapperWg / ApiCallingWrapperWg, a scripted stand-in in the tests) and only its `generate_sequences` is called.
"""
from typing import Dict, List

from .ctx_manager import ContextManager, DataProto
from .es_manager import EnvStateManager


class LLMAgentProxy:
    def __init__(self, config, actor_rollout_wg, tokenizer, device: int = 0, rank: int = 0, world_size: int = 1, group=None,
                 assembly: str = "device"):
        """rank / world_size / group (absent in the reference, which is single-process): one proxy per GPU, each owning the
        env groups with group_id % world_size == rank; rollout metrics come back batch-global (ragen_b200/sharding.py).
        assembly: "device" (token trajectories gathered on the GPU, llm_agent/assembly.py) or "host" (the reference's path)."""
        self.config = config
        self.train_ctx_manager = ContextManager(config, tokenizer, mode="train", device=device, assembly=assembly).shard(world_size, group)
        self.train_es_manager = EnvStateManager(config, mode="train", device=device, rank=rank, world_size=world_size)
        self.val_ctx_manager = ContextManager(config, tokenizer, mode="val", device=device, assembly=assembly).shard(world_size, group)
        self.val_es_manager = EnvStateManager(config, mode="val", device=device, rank=rank, world_size=world_size)
        self.actor_wg = actor_rollout_wg
        self.tokenizer = tokenizer

    def generate_sequences(self, lm_inputs: DataProto):
        """agent_proxy.py:129-141.  A worker group that shards the batch (it has `world_size`, as RayWorkerGroup) gets it
        padded to a multiple of its size with verl's own helpers; wrappers are called as they are."""
        wg = self.actor_wg
        if not hasattr(wg, "generate_sequences"):
            raise ValueError(f"Unsupported actor worker type: {type(wg)}")
        if any(c.__name__ == "RayWorkerGroup" for c in type(wg).__mro__):
            from verl.protocol import pad_dataproto_to_divisor, unpad_dataproto  # present whenever a RayWorkerGroup is

            padded_lm_inputs, pad_size = pad_dataproto_to_divisor(lm_inputs, wg.world_size)
            lm_outputs = unpad_dataproto(wg.generate_sequences(padded_lm_inputs), pad_size=pad_size)
            lm_outputs.meta_info = lm_inputs.meta_info
            lm_outputs.non_tensor_batch = lm_inputs.non_tensor_batch
            return lm_outputs
        return wg.generate_sequences(lm_inputs)

    def rollout(self, dataproto: DataProto, val: bool = False, seed=None):
        """agent_proxy.py:143-159.  `seed` (absent in the reference, which always draws one) pins the train reset."""
        es_manager = self.val_es_manager if val else self.train_es_manager
        ctx_manager = self.val_ctx_manager if val else self.train_ctx_manager
        env_outputs = es_manager.reset() if seed is None else es_manager.reset(seed=seed)
        ctx_manager.bind_envs(es_manager.envs)  # action-id lookup joins the parse kernel (Bandit re-draws names per reset)
        for _ in range(self.config.agent_proxy.max_turn):
            if len(env_outputs) == 0:  # a rank without env groups (more ranks than groups) has nothing to generate
                break
            lm_inputs: DataProto = ctx_manager.get_lm_inputs(env_outputs, prepare_for_update=False)
            lm_inputs.meta_info = dataproto.meta_info
            lm_outputs: DataProto = self.generate_sequences(lm_inputs)
            env_inputs: List[Dict] = ctx_manager.get_env_inputs(lm_outputs)
            env_outputs: List[Dict] = es_manager.step(env_inputs)
            if len(env_outputs) == 0:  # all finished
                break
        rollout_states = es_manager.get_rollout_states()
        return ctx_manager.formulate_rollouts(rollout_states)
