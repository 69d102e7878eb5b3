"""RLOO / REINFORCE++ / REINFORCE++-baseline / ReMax / KL-in-reward / StarPO-S filter.
CPU part: the oracle restatements against tensors produced by the UNMODIFIED verl functions (tests/golden/advantage_more.npz).
GPU part (marked): the kernels, through ragen_b200.trainer's reference-named functions, against the same goldens + the oracle."""
import numpy as np
import pytest
import torch

from conftest import golden_npz
from oracle import advantage as A

G = golden_npz("advantage_more.npz")
T_ = lambda k: torch.tensor(G[k])  # noqa: E731
RTOL = 1e-5


def close(got, want):
    got, want = torch.as_tensor(got).double().cpu(), torch.as_tensor(want).double()
    assert got.shape == want.shape
    err, scale = (got - want).abs(), max(1.0, float(want.abs().max()))
    assert float(err.max()) <= RTOL * scale, (float(err.max()), scale)
    big = want.abs() > 1e-3  # per element: 1e-5 relative (two fp32 ulps of the tensor's scale as cancellation slack)
    if bool(big.any()):
        rel = float(((err[big] - 2.0 ** -22 * scale).clamp(min=0) / want.abs()[big]).max())
        assert rel <= RTOL, rel


def test_oracle_matches_reference_outputs():
    mask, uid = T_("mask"), list(G["uid"])
    close(A.rloo(T_("outcome"), mask, uid), G["rloo"])
    close(A.rfpp_baseline(T_("outcome"), mask, uid), G["rfpp_baseline"])
    for tag, gamma in (("1", 1.0), ("97", 0.97)):
        adv, ret = A.rfpp(T_("dense"), mask, gamma)
        assert torch.equal(ret, T_(f"rfpp_ret_{tag}"))
        close(adv, G[f"rfpp_adv_{tag}"])
    adv, ret = A.remax(T_("dense"), T_("remax_base"), mask)
    assert torch.equal(ret, T_("remax_ret")) and torch.equal(adv, T_("remax_adv"))
    for kind in ("kl", "abs", "mse", "low_var_kl"):
        rew, cur = A.apply_kl_penalty(T_("outcome"), T_("logp"), T_("ref_logp"), mask, kind, 0.05)
        assert torch.equal(rew, T_(f"klrew_{kind}"))
        assert cur == pytest.approx(float(G[f"klcur_{kind}"]), rel=1e-6)
    std, mx, mean = A.group_filter_stats(T_("filter_scores"), 8, 4)
    assert torch.equal(std, T_("filter_std")) and torch.equal(mx, T_("filter_max")) and torch.equal(mean, T_("filter_mean"))
    keep, _ = A.filter_rollout(T_("filter_scores"), 8, 4, 0.25, "std")
    assert keep.view(8, 4)[:, 0].nonzero().flatten().tolist() == sorted(G["filter_top_std"].tolist())


@pytest.mark.gpu
def test_estimators_on_device_match_reference():
    from ragen_b200.trainer import core_algos as K

    mask, uid = T_("mask"), G["uid"]
    adv, ret = K.compute_rloo_outcome_advantage(T_("outcome"), mask, uid)
    assert adv is ret and not adv.is_cuda
    close(adv, G["rloo"])
    adv, ret = K.compute_reinforce_plus_plus_baseline_outcome_advantage(T_("outcome"), mask, uid)
    close(adv, G["rfpp_baseline"])
    for tag, gamma in (("1", 1.0), ("97", 0.97)):
        adv, ret = K.compute_reinforce_plus_plus_outcome_advantage(T_("dense"), mask, gamma)
        assert torch.equal(ret, T_(f"rfpp_ret_{tag}"))  # sequential fp32 recurrence: bit-exact
        close(adv, G[f"rfpp_adv_{tag}"])
        assert float(adv[mask == 0].abs().max()) == 0.0
    adv, ret = K.compute_remax_outcome_advantage(T_("dense"), T_("remax_base"), mask)
    assert torch.equal(ret, T_("remax_ret")) and torch.equal(adv, T_("remax_adv"))  # double accumulation as torch's CPU cumsum


@pytest.mark.gpu
def test_compute_advantage_dispatch_and_kl_penalty():
    from ragen_b200.llm_agent.ctx_manager import DataProto, TensorDict
    from ragen_b200.trainer import AdvantageEstimator, apply_kl_penalty, compute_advantage
    from ragen_b200.trainer.core_algos import AdaptiveKLController, FixedKLController

    mask = T_("mask")
    B, T = mask.shape
    dp = DataProto(batch=TensorDict({"responses": torch.zeros(B, T, dtype=torch.long), "attention_mask": mask.clone(),
                                     "token_level_rewards": T_("outcome"), "reward_baselines": T_("remax_base")}, batch_size=B),
                   non_tensor_batch={"uid": G["uid"]})
    for est, key in ((AdvantageEstimator.RLOO, "rloo"), (AdvantageEstimator.REINFORCE_PLUS_PLUS_BASELINE, "rfpp_baseline")):
        compute_advantage(dp, est)
        close(dp.batch["advantages"], G[key])
    dp.batch["token_level_rewards"] = T_("dense")
    compute_advantage(dp, AdvantageEstimator.REINFORCE_PLUS_PLUS, gamma=0.97)
    close(dp.batch["advantages"], G["rfpp_adv_97"])
    assert torch.equal(dp.batch["returns"], T_("rfpp_ret_97"))
    compute_advantage(dp, AdvantageEstimator.REMAX)
    assert torch.equal(dp.batch["advantages"], T_("remax_adv"))
    with pytest.raises(NotImplementedError):
        compute_advantage(dp, "something_else")
    # KL in reward
    dp.batch["token_level_scores"] = T_("outcome")
    dp.batch["old_log_probs"], dp.batch["ref_log_prob"] = T_("logp"), T_("ref_logp")
    for kind in ("kl", "abs", "mse", "low_var_kl"):
        ctrl = AdaptiveKLController(0.05, 0.1, 100)
        _, metrics = apply_kl_penalty(dp, ctrl, kl_penalty=kind)
        close(dp.batch["token_level_rewards"], G[f"klrew_{kind}"])
        cur = float(G[f"klcur_{kind}"])
        assert metrics["actor/reward_kl_penalty"] == pytest.approx(cur, rel=1e-5, abs=1e-7)
        assert metrics["actor/reward_kl_penalty_coeff"] == 0.05
        want = 0.05 * (1 + np.clip(metrics["actor/reward_kl_penalty"] / 0.1 - 1, -0.2, 0.2) * B / 100)
        assert ctrl.value == pytest.approx(want, rel=1e-12)
    ctrl = FixedKLController(0.05)
    apply_kl_penalty(dp, ctrl, kl_penalty="kl")
    assert ctrl.value == 0.05
    with pytest.raises(NotImplementedError):
        apply_kl_penalty(dp, ctrl, kl_penalty="full")


@pytest.mark.gpu
@pytest.mark.parametrize("kind,ratio", [("std", 0.25), ("std_rev", 0.25), ("std", 1), ("std", 0.5)])
def test_filter_rollout(kind, ratio):
    from ragen_b200 import ops
    from ragen_b200.llm_agent.ctx_manager import DataProto, TensorDict
    from ragen_b200.trainer import filter_rollout

    sc = T_("filter_scores")
    row, std, mx, mean = ops.group_filter_stats(sc.cuda(), 8, 4)
    close(std, G["filter_std"]); close(mx, G["filter_max"]); close(mean, G["filter_mean"]); close(row, sc.sum(-1))
    n = sc.shape[0]
    dp = DataProto(batch=TensorDict({"original_rm_scores": sc.clone(), "input_ids": torch.arange(n).unsqueeze(1)}, batch_size=n),
                   non_tensor_batch={"env_ids": np.arange(n, dtype=object), "tags": [str(i) for i in range(n)]})
    out, metrics = filter_rollout(dp, 8, 4, ratio, kind)
    keep, want = A.filter_rollout(sc, 8, 4, ratio, kind)
    rows = keep.nonzero().flatten().tolist()
    assert out.batch["input_ids"].flatten().tolist() == rows
    assert out.non_tensor_batch["env_ids"].tolist() == rows and out.non_tensor_batch["tags"] == [str(i) for i in rows]
    if ratio == 0.25:
        top = G["filter_top_std" if kind == "std" else "filter_top_std_rev"].tolist()
        assert sorted(set(r // 4 for r in rows)) == sorted(top)
    assert set(metrics) == set(want)
    for k in want:
        assert float(metrics[k]) == pytest.approx(float(want[k]), rel=1e-5, abs=1e-6), k
