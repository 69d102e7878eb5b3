"""Rollout orchestration with the reference's interfaces (ragen/llm_agent/): EnvStateManager, ContextManager,
LLMAgentProxy - backed by the kernels of libragen_b200."""
