"""Wall-clock share of each phase of LLMAgentProxy.rollout (C5 shape, scripted policy), without a profiler's overhead."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from fake_tokenizer import FakeQwenTokenizer, ScriptedWg  # noqa: E402
from oracle import cfg as C  # noqa: E402  (config builder only)
from ragen_b200.llm_agent.agent_proxy import LLMAgentProxy  # noqa: E402
from ragen_b200.llm_agent.ctx_manager import DataProto  # noqa: E402

cfg = C.base_config(tags=("SimpleSokoban",), n_groups=(8,), group_size=16, reward_normalization={"grouping": "state", "method": "mean_std"})
tok = FakeQwenTokenizer()
proxy = LLMAgentProxy(cfg, ScriptedWg(tok, DataProto), tok)
spent = {}


def timed(obj, name, label):
    fn = getattr(obj, name)

    def wrap(*a, **k):
        t0 = time.perf_counter()
        out = fn(*a, **k)
        spent[label] = spent.get(label, 0.0) + time.perf_counter() - t0
        return out
    setattr(obj, name, wrap)


es, cm = proxy.train_es_manager, proxy.train_ctx_manager
timed(es, "reset", "es.reset")
timed(es, "step", "es.step")
timed(es, "get_rollout_states", "es.get_rollout_states")
timed(cm, "get_lm_inputs", "ctx.get_lm_inputs (incl. formulate)")
timed(cm, "get_env_inputs", "ctx.get_env_inputs")
timed(cm, "bind_envs", "ctx.bind_envs")
timed(proxy.actor_wg, "generate_sequences", "policy stub")
from ragen_b200 import ops  # noqa: E402
timed(ops, "stage_responses", "  ops.stage_responses")
timed(ops, "parse_actions_device", "  ops.parse_actions_device")
timed(ops.ParseResult, "to_host", "  ParseResult.to_host")
timed(cm, "_parse_batch", "  ctx._parse_batch")
from ragen_b200.llm_agent import assembly as _A  # noqa: E402
timed(_A.TrajectoryAssembler, "assemble", "  assembler.assemble")
timed(_A.TrajectoryAssembler, "row_elements", "    assembler.row_elements")
timed(_A.TrajectoryAssembler, "_obs_segments", "    assembler._obs_segments")
timed(_A.TokenPool, "put", "    pool.put")
timed(_A, "encode_many", "    encode_many")
timed(es, "_block_texts", "  es._block_texts")
timed(es, "_run_block", "  es._run_block")
for rep in range(2):
    proxy.rollout(DataProto(batch=None, non_tensor_batch=None, meta_info={}), val=False, seed=1 + rep)
spent.clear()
import gc
if os.environ.get("NOGC"):
    gc.collect()
    gc.freeze()
    gc.disable()
R = 5
t0 = time.perf_counter()
for rep in range(R):
    proxy.rollout(DataProto(batch=None, non_tensor_batch=None, meta_info={}), val=False, seed=123 + rep)
torch.cuda.synchronize()
total = time.perf_counter() - t0
print(f"rollout {1e3 * total / R:.2f} ms")
for k, v in sorted(spent.items(), key=lambda kv: -kv[1]):
    print(f"  {k:40s} {1e3 * v / R:7.2f} ms")
sys.exit(0)
import cProfile, pstats, io
pr = cProfile.Profile()
pr.enable()
for rep in range(3):
    proxy.rollout(DataProto(batch=None, non_tensor_batch=None, meta_info={}), val=False, seed=500 + rep)
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(28)
print(s.getvalue()[:5000])
