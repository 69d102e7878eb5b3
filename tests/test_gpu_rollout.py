"""GPU parity of the whole seam: LLMAgentProxy.rollout (device-backed managers + scripted policy) against the same loop
over the UNMODIFIED reference managers (tests/golden/rollouts.json), then compute_advantage on the rollout (C5 shape:
rollout -> token-level rewards -> GAE / bi-level GAE / GRPO) against the oracle estimators and the reference goldens."""
import hashlib

import numpy as np
import pytest
import torch

from conftest import golden_json, golden_npz
from fake_tokenizer import FakeQwenTokenizer, ScriptedWg
from oracle import advantage as A
from oracle import cfg as C

pytestmark = pytest.mark.gpu
CASES = golden_json("rollouts.json")


def _digest(t):
    a = np.ascontiguousarray(t.numpy() if hasattr(t, "numpy") else t)
    return hashlib.sha256(a.tobytes()).hexdigest()[:24] + ":" + "x".join(map(str, a.shape)) + ":" + str(a.dtype)


class _Recorder(ScriptedWg):
    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self.turns = []

    def generate_sequences(self, lm_inputs):
        out = super().generate_sequences(lm_inputs)
        self.turns.append({"n": len(out.non_tensor_batch["response_texts"]), "input_ids": _digest(lm_inputs.batch["input_ids"]),
                           "responses": out.non_tensor_batch["response_texts"][:3]})
        return out


def _rollout(case, assembly="device", tok=None):
    from ragen_b200.llm_agent.agent_proxy import LLMAgentProxy
    from ragen_b200.llm_agent.ctx_manager import DataProto

    cfg = C.base_config(tags=tuple(case["tags"]), n_groups=tuple(case["n_groups"]), group_size=case["group_size"], **case["ap"])
    tok = tok or FakeQwenTokenizer()
    wg = _Recorder(tok, DataProto)
    proxy = LLMAgentProxy(cfg, wg, tok, assembly=assembly)
    dp = proxy.rollout(DataProto(batch=None, non_tensor_batch=None, meta_info={}), val=False, seed=case["seed"])
    wg.proxy = proxy
    return dp, wg


@pytest.mark.parametrize("assembly", ["device", "host"])
@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_rollout_matches_reference_managers(case, assembly):
    """assembly="device": every prompt of every turn is gathered on the GPU from token segments, the observations from the
    glyph rows the env kernels wrote (no prompt string is built); "host": the reference's string path.  Both must give the
    unmodified reference's tokens, turn by turn."""
    dp, wg = _rollout(case, assembly)
    if assembly == "device":
        modes = wg.proxy.train_ctx_manager._assembler._modes
        assert modes and set(modes.values()) == {"cell"}  # grid observations went through ragen_glyph_tokens
        msgs = dp.non_tensor_batch["messages_list"]  # built on demand; same shape as the reference's object array
        assert len(msgs) == dp.batch["input_ids"].shape[0] and msgs[0][0]["role"] == "system"
    assert wg.turns == case["turns"]  # every prompt of every turn, token for token
    assert list(dp.batch["input_ids"].shape) == case["shape"]
    for k in ("input_ids", "attention_mask", "position_ids", "responses", "loss_mask"):
        assert _digest(dp.batch[k]) == case["digest"][k], k
    rm = dp.batch["rm_scores"]
    want = torch.tensor(case["rm_last"])
    assert float((rm[:, -1] - want).abs().max()) <= 1e-5 * max(1.0, float(want.abs().max()))
    assert float((rm[:, :-1]).abs().max()) == 0.0
    assert dp.batch["loss_mask"].sum(-1).tolist() == case["loss_mask_sum"]
    assert set(dp.meta_info["metrics"]) == set(case["metrics"])
    for k, w in case["metrics"].items():
        assert float(dp.meta_info["metrics"][k]) == pytest.approx(w, rel=1e-12), k
    if case["ap"].get("reward_normalization", {}).get("method", "identity") == "identity":
        assert _digest(rm) == case["digest"]["rm_scores"]


@pytest.mark.parametrize("case", CASES[:2], ids=[c["name"] for c in CASES[:2]])
def test_rollout_with_a_bpe_tokenizer_device_equals_host(case):
    """A real BPE tokenizer (Qwen's pre-tokenizer, merges across glyphs): the device assembly takes the row-dictionary mode and
    must give the host path's tensors for every turn and for the final batch."""
    from fake_tokenizer import make_bpe_tokenizer

    tok = make_bpe_tokenizer()
    dev, wg_d = _rollout(case, "device", tok)
    host, wg_h = _rollout(case, "host", tok)
    assert wg_d.turns == wg_h.turns and len(wg_d.turns) >= 2
    modes = wg_d.proxy.train_ctx_manager._assembler._modes
    assert set(modes.values()) <= {"row", "text"} and "row" in modes.values()
    for k in ("input_ids", "attention_mask", "position_ids", "responses", "loss_mask", "rm_scores"):
        assert torch.equal(dev.batch[k], host.batch[k]), k
    assert dev.meta_info["metrics"] == host.meta_info["metrics"]


def test_rollout_to_advantage_c5_shape():
    """C5: rollout -> compute_advantage (GAE, bi-level GAE, GRPO multi-turn) with the rollout's own masks and scores."""
    from ragen_b200.trainer import AdvantageEstimator, compute_advantage

    dp, _ = _rollout(CASES[1])
    batch = dp.batch
    B, T = batch["rm_scores"].shape
    gen = torch.Generator().manual_seed(11)
    batch["response_mask"] = batch["loss_mask"].float()
    batch["token_level_rewards"] = batch["rm_scores"].clone()
    batch["values"] = torch.randn(B, T, generator=gen) * batch["response_mask"]
    r, v, m = batch["token_level_rewards"], batch["values"], batch["response_mask"]

    compute_advantage(dp, AdvantageEstimator.GAE, gamma=1.0, lam=1.0)
    wa, wr = A.gae(r, v, m, 1.0, 1.0)
    assert torch.equal(batch["returns"], wr)
    assert float((batch["advantages"] - wa).abs().max()) <= 1e-5 * max(1.0, float(wa.abs().max()))

    # bi-level needs a reward at the last loss-mask position of each row (core_algos.py:79); the rollout puts the whole
    # trajectory reward on the last column, so move it there the way the trainer's multi-turn path sees it
    r2 = torch.zeros_like(r)
    for b in range(B):
        pos = m[b].nonzero()
        if len(pos):
            r2[b, int(pos[-1])] = r[b, -1] if float(r[b, -1]) != 0 else 0.5
    batch["token_level_rewards"] = r2
    compute_advantage(dp, AdvantageEstimator.GAE, gamma=1.0, lam=1.0, bi_level_gae=True, high_level_gamma=0.95)
    wa, wr = A.bilevel_gae(r2, v, m, 1.0, 1.0, 0.95)
    assert torch.equal(batch["returns"], wr)
    assert float((batch["advantages"] - wa).abs().max()) <= 1e-5 * max(1.0, float(wa.abs().max()))

    batch["token_level_rewards"] = r
    dp.non_tensor_batch["uid"] = dp.non_tensor_batch["group_ids"]
    for by_std in (True, False):
        compute_advantage(dp, AdvantageEstimator.GRPO, multi_turn=True, norm_adv_by_std_in_grpo=by_std)
        wa = A.grpo_outcome(r, batch["loss_mask"].float(), [int(g) for g in dp.non_tensor_batch["group_ids"]], norm_by_std=by_std)
        wa = wa[0] if isinstance(wa, tuple) else wa
        assert float((batch["advantages"] - wa).abs().max()) <= 1e-5 * max(1.0, float(wa.abs().max()))
        assert batch["advantages"].data_ptr() == batch["returns"].data_ptr()

    with pytest.raises(NotImplementedError):
        compute_advantage(dp, "no_such_estimator")


def test_reference_named_estimators_on_golden():
    """The reference-named functions (whitening included) against tensors the UNMODIFIED reference produced."""
    from ragen_b200.trainer import core_algos as K

    g = golden_npz("advantage.npz")
    r, v, m = torch.tensor(g["rewards"]), torch.tensor(g["values"]), torch.tensor(g["mask"])
    for tag in ("a", "b"):
        gamma, lam, hg = (float(x) for x in g[f"cfg_{tag}"])
        adv, ret = K.compute_gae_advantage_return(r, v, m, gamma, lam)
        assert not adv.is_cuda and torch.equal(ret, torch.tensor(g[f"gae_{tag}_ret"]))
        assert float((adv - torch.tensor(g[f"gae_{tag}_adv"])).abs().max()) <= 1e-5 * max(1.0, float(np.abs(g[f"gae_{tag}_adv"]).max()))
        adv, ret = K.compute_bi_level_gae_advantage_return(r, v, m, gamma, lam, hg)
        assert torch.equal(ret, torch.tensor(g[f"bi_{tag}_ret"]))
        assert float((adv - torch.tensor(g[f"bi_{tag}_adv"])).abs().max()) <= 1e-5 * max(1.0, float(np.abs(g[f"bi_{tag}_adv"]).max()))
    for by_std, key in ((True, "grpo_std_adv"), (False, "grpo_mean_adv")):
        adv, ret = K.compute_grpo_outcome_advantage(torch.tensor(g["outcome"]), m, g["uid"], norm_adv_by_std_in_grpo=by_std)
        assert adv is ret
        assert float((adv - torch.tensor(g[key])).abs().max()) <= 1e-5 * max(1.0, float(np.abs(g[key]).max()))
    w = K.masked_whiten(v, m)
    ref = A.masked_whiten(v, m)
    assert float((w - ref).abs().max()) <= 1e-5 * max(1.0, float(ref.abs().max()))
    with pytest.raises(ValueError):
        K.masked_whiten(v, torch.zeros_like(m))
