"""CPU: oracle/rollout.py (the reference's rollout loop restated end to end: the manager-level CPU baseline of bench.py --rows)
against tests/golden/rollouts.json, which the UNMODIFIED reference managers produced with the same scripted policy: every
turn's prompt tokens and the final batch, by digest."""
import hashlib

import numpy as np

from conftest import golden_json
from fake_tokenizer import FakeQwenTokenizer, ScriptedWg
from oracle import cfg as C
from oracle import sokoban as S
from oracle.rollout import RolloutPort


def _digest(t):
    a = np.ascontiguousarray(t.numpy() if hasattr(t, "numpy") else t)
    return hashlib.sha256(a.tobytes()).hexdigest()[:24] + ":" + "x".join(map(str, a.shape)) + ":" + str(a.dtype)


class _Proto:
    def __init__(self, batch=None, non_tensor_batch=None, meta_info=None):
        self.batch, self.non_tensor_batch, self.meta_info = batch, non_tensor_batch or {}, meta_info or {}


def test_rollout_port_reproduces_the_reference_rollout():
    case = golden_json("rollouts.json")[0]
    assert case["tags"] == ["SimpleSokoban"]
    tok = FakeQwenTokenizer()
    n_groups, gs = case["n_groups"][0], case["group_size"]
    port = RolloutPort(tok, n_groups, gs, dim_room=(6, 6), num_boxes=1, max_steps=100, max_actions_per_traj=10,
                       env_instruction=C.SOKOBAN_INSTRUCTION, max_tokens=100,
                       norm_method=case["ap"]["reward_normalization"]["method"])
    rooms = [S.generate_room(6, 6, 1, 300, case["seed"] + g) for g in range(n_groups)]  # one room per group seed (es_manager.py:80-82)
    fixed = np.stack([rooms[i // gs][0].reshape(-1) for i in range(n_groups * gs)])
    state = np.stack([rooms[i // gs][1].reshape(-1) for i in range(n_groups * gs)])
    wg = ScriptedWg(tok, _Proto)
    turns = []

    def policy(lm):
        ids = np.array(lm["env_ids"], dtype=object)
        out = wg.generate_sequences(_Proto(batch={"input_ids": lm["input_ids"]}, non_tensor_batch={"env_ids": ids, "group_ids": ids}))
        turns.append({"n": len(ids), "input_ids": _digest(lm["input_ids"]), "responses": out.non_tensor_batch["response_texts"][:3]})
        return out.non_tensor_batch["response_texts"]

    final = port.rollout(fixed, state, policy)
    assert turns == case["turns"]
    assert list(final["input_ids"].shape) == case["shape"]
    for k in ("input_ids", "attention_mask", "position_ids", "loss_mask"):
        assert _digest(final[k]) == case["digest"][k], k
    want = np.array(case["rm_last"])
    assert np.abs(final["rm_scores"][:, -1].numpy() - want).max() <= 1e-5 * max(1.0, np.abs(want).max())
    for k, w in case["metrics"].items():
        assert abs(float(final["metrics"][k]) - w) <= 1e-12 * max(1.0, abs(w)), k
