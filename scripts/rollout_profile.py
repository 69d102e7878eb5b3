"""cProfile of LLMAgentProxy.rollout (C5 shape, scripted policy) on the GPU box: where this package's host time goes."""
import cProfile
import io
import os
import pstats
import sys

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
proxy.rollout(DataProto(batch=None, non_tensor_batch=None, meta_info={}), val=False, seed=1)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for rep in range(3):
    proxy.rollout(DataProto(batch=None, non_tensor_batch=None, meta_info={}), val=False, seed=123 + rep)
torch.cuda.synchronize()
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45)
print(s.getvalue()[:9000])
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(25)
print(s.getvalue()[:6000])
s = io.StringIO()
pstats.Stats(pr, stream=s).print_callers("inner|fromnumeric.py.*sum|tolist")
print(s.getvalue()[:5000])
