"""GPU parity of the text-env answer scoring: ragen_text_score and the MetaMathQA / Static env classes
against outputs of the UNMODIFIED reference functions (tests/golden/text_scores.json, metamath_episodes.json) and the oracle."""
import numpy as np
import pytest

from conftest import golden_json
from oracle import cfg as C
from oracle import toy_envs as TE

pytestmark = pytest.mark.gpu
MODES = ("metamathqa", "exact_match", "numeric", "multiple_choice")


def test_text_score_matches_reference_functions():
    from ragen_b200 import ops

    rows = golden_json("text_scores.json")
    preds, labels = [r["pred"] for r in rows], [r["label"] for r in rows]
    for mode in MODES:
        bits = ops.text_score(preds, labels, mode).cpu().numpy()
        assert [[bool(b & 1), bool(b & 2)] for b in bits] == [r[mode] for r in rows], mode
        assert not (bits & 128).any()


def test_text_score_batch_against_oracle():
    from ragen_b200 import ops

    rng = np.random.default_rng(3)
    alphabet = list("abcdABCD xyzXYZ  0123456789..,;:-_()\t\n")
    preds, labels = [], []
    for _ in range(5000):
        p = "".join(rng.choice(alphabet, int(rng.integers(0, 40))))
        l = p if rng.random() < 0.25 else "".join(rng.choice(alphabet, int(rng.integers(0, 12))))
        if rng.random() < 0.2:
            l = p.upper().replace(" ", "  ")
        preds.append(p)
        labels.append(l)
    for mode in MODES:
        bits = ops.text_score(preds, labels, mode).cpu().numpy()
        for p, l, b in zip(preds, labels, bits):
            if not b & 128:
                assert (bool(b & 1), bool(b & 2)) == tuple(TE.SCORERS[mode](p, l)), (mode, p, l)
        assert (bits & 128).mean() < 0.2
    assert ops.text_score([], [], "numeric").numel() == 0
    flagged = ops.text_score(["École", "1.00000000000000000001"], ["école", "1"], "numeric").cpu().numpy()
    assert (flagged & 128).all()


def test_metamathqa_env_matches_reference_episodes():
    from ragen_b200.env import MetaMathQAEnv, MetaMathQAEnvConfig

    g = golden_json("metamath_episodes.json")
    for ep in g["episodes"]:
        env = MetaMathQAEnv(MetaMathQAEnvConfig(split="train", data={"train": g["rows"]}))
        assert env.reset(seed=ep["seed"]) == ep["question"]
        assert env.current_question_idx == ep["idx"] and env.correct_answer == ep["label"]
        assert env.render() == ep["question"]
        for st in ep["steps"]:
            obs, reward, done, info = env.step(st["answer"])
            assert (obs, reward, done, info) == (st["obs"], st["reward"], st["done"], st["info"])
            assert env.render() == obs


def test_static_env_scorers():
    from ragen_b200.env import StaticEnv, StaticEnvConfig

    data = {"gsm8k": {"test": [{"question": "2+2?", "answer": "Two and two.\n#### 4"}]},
            "mmlu": {"test": [{"question": "Pick", "choices": ["x", "y", "z", "w"], "answer": 2}]},
            "metamathqa": {"train": [{"query": "6*7?", "response": "so\nThe answer is: 42"}]}}
    env = StaticEnv(StaticEnvConfig(dataset_name="gsm8k", data=data["gsm8k"]))
    assert env.reset(seed=0) == "2+2?" and env.correct_answer == "4"
    assert env.step("it is 5") == ("Incorrect. Please think again.", 0.0, False, {"success": False, "is_valid": True})
    assert env.step("4.0 then") == ("Correct!", 0.5, True, {"success": True, "is_valid": True})
    env = StaticEnv(StaticEnvConfig(dataset_name="mmlu", data=data["mmlu"]))
    assert env.reset(seed=1) == "Pick\nA. x\nB. y\nC. z\nD. w" and env.correct_answer == "C"
    assert env.step("no idea")[3] == {"success": False, "is_valid": True}  # "no idea".upper() holds a D
    assert env.step("(c)") == ("Correct!", 0.5, True, {"success": True, "is_valid": True})
    env = StaticEnv(StaticEnvConfig(dataset_name="metamathqa", data=data["metamathqa"]))
    env.reset(seed=2)
    assert env.step(" 42. ") == ("Correct!", 1.0, True, {"success": True, "is_valid": True})


def test_manager_runs_metamathqa_tag():
    """EnvStateManager over a text-QA tag: several answers per turn are scored round by round, with the reference's accounting."""
    from ragen_b200.llm_agent.es_manager import EnvStateManager

    g = golden_json("metamath_episodes.json")
    cfg = C.base_config(tags=("MetaMathQA",), n_groups=(2,), group_size=2)
    cfg.custom_envs["MetaMathQA"] = C.to_attr({"env_type": "metamathqa", "max_actions_per_traj": 3, "env_instruction": "Solve.", "max_tokens": 100,
                                              "env_config": {"split": "train", "data": {"train": g["rows"]}}})
    mgr = EnvStateManager(cfg, mode="train")
    cache = mgr.reset(seed=3)
    labels = [e["env"].correct_answer for e in mgr.envs]
    assert labels[0] == labels[1] and labels[2] == labels[3]  # a group shares the seed, hence the question
    inputs = [{"env_id": 0, "llm_raw_response": "r", "llm_response": "r", "actions": ["wrong", labels[0] or "x", "never reached"]},
              {"env_id": 1, "llm_raw_response": "r", "llm_response": "r", "actions": ["wrong"]},
              {"env_id": 2, "llm_raw_response": "r", "llm_response": "r", "actions": []},
              {"env_id": 3, "llm_raw_response": "r", "llm_response": "r", "actions": [(labels[2] or "x").upper()]}]
    out = mgr.step(inputs)
    states = mgr.get_rollout_states()
    if labels[0]:
        assert states[0]["history"][0]["reward"] == 0.5 and states[0]["history"][0]["actions"] == ["wrong", labels[0]]
        assert mgr.envs[0]["status"].terminated and not mgr.envs[0]["status"].truncated
    assert states[1]["history"][0]["reward"] == 0.0 and states[1]["history"][1]["state"] == "Incorrect. Please think again."
    assert states[2]["penalty"] == cfg.es_manager.format_penalty and states[2]["history"][0]["reward"] == 0
    assert [c["env_id"] for c in out] == [i for i in range(4) if not mgr.envs[i]["status"].terminated]
