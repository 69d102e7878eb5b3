"""GPU parity of the manager seam: ragen_b200.llm_agent.EnvStateManager (device batches behind the reference API) against
the scripted rollout recorded from the UNMODIFIED reference EnvStateManager (tests/golden/es_manager.json) and against
the oracle manager on mixed tags (Sokoban + FrozenLake + Bandit)."""
import copy
import json

import numpy as np
import pytest

from conftest import golden_json
from oracle import cfg as C
from oracle import managers as M
from oracle import sokoban as S
from oracle import toy_envs as TE

pytestmark = pytest.mark.gpu
HIST_KEYS = ("state", "actions_left", "actions", "reward", "info", "llm_response", "llm_raw_response")


def _same_caches(got, want):
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert (g["env_id"], g["group_id"], g["tag"]) == (w["env_id"], w["group_id"], w["tag"])
        assert g["penalty"] == w["penalty"], g["env_id"]
        assert g.get("metrics") == w.get("metrics"), g["env_id"]
        assert len(g["history"]) == len(w["history"]), g["env_id"]
        for hg, hw in zip(g["history"], w["history"]):
            for k in HIST_KEYS:
                assert hg.get(k) == hw.get(k), (g["env_id"], k, hg.get(k), hw.get(k))
                assert repr(hg.get(k)) == repr(hw.get(k)), (g["env_id"], k)  # 0 vs 0.0 matters: rewards are printed into prompts


def test_reference_script_sokoban():
    from ragen_b200.llm_agent.es_manager import EnvStateManager

    g = golden_json("es_manager.json")
    cfg = C.base_config(tags=("SimpleSokoban",), n_groups=(2,), group_size=4)
    mgr = EnvStateManager(cfg, mode="train")
    cache = mgr.reset(seed=g["seed"])
    batch = mgr.blocks[0].batch
    assert np.array_equal(batch.fixed.cpu().numpy(), np.array(g["rooms_fixed"], dtype=np.uint8))
    assert np.array_equal(batch.state.cpu().numpy(), np.array(g["rooms_state"], dtype=np.uint8))
    for c, w in zip(cache, g["final"]):
        assert c["history"][0]["state"] == w["history"][0]["state"]
    for inputs in g["script"]:
        mgr.step(json.loads(json.dumps(inputs)))
    final = mgr.get_rollout_states()
    _same_caches(final, g["final"])
    for e, want in zip(mgr.envs, g["status"]):
        st = e["status"]
        assert {"truncated": st.truncated, "terminated": st.terminated, "num_actions": st.num_actions, "rewards": st.rewards,
                "seed": st.seed} == want


def _oracle_manager(cfg, group_size):
    envs = []
    for tag, n_group in zip(cfg.es_manager.train.env_configs.tags, cfg.es_manager.train.env_configs.n_groups):
        tpl = cfg.custom_envs[tag]
        for _ in range(n_group * group_size):
            if tpl.env_type == "sokoban":
                ec = tpl.env_config
                env = S.SokobanOracle((ec["dim_x"], ec["dim_y"]), ec["num_boxes"], ec["max_steps"], ec.get("search_depth", 300))
                envs.append({"tag": tag, "env": env, "max_actions_per_traj": tpl.max_actions_per_traj,
                             "action_lookup": {1: "Up", 2: "Down", 3: "Left", 4: "Right"}})
            elif tpl.env_type == "frozen_lake":
                envs.append({"tag": tag, "env": TE.FrozenLakeOracle(), "max_actions_per_traj": tpl.max_actions_per_traj,
                             "action_lookup": {1: "Left", 2: "Down", 3: "Right", 4: "Up"}})
            else:
                env = TE.BanditOracle(tpl.env_config["lo_arm_name"], tpl.env_config["hi_arm_name"])
                envs.append({"tag": tag, "env": env, "max_actions_per_traj": tpl.max_actions_per_traj})
    return M.ManagerOracle(envs, group_size)


@pytest.mark.parametrize("seed", [123, 77])
def test_mixed_tags_against_oracle_manager(seed):
    """C1-style rollout, 5 turns: 8 groups x 16 envs over four tags, random / invalid / surplus actions."""
    from ragen_b200.llm_agent.es_manager import EnvStateManager

    tags, n_groups, gs = ("SimpleSokoban", "FrozenLake", "Bandit", "LargerSokoban"), (3, 2, 2, 1), 16
    cfg = C.base_config(tags=tags, n_groups=n_groups, group_size=gs)
    mgr = EnvStateManager(cfg, mode="train")
    ora = _oracle_manager(cfg, gs)
    got = mgr.reset(seed=seed)
    want = ora.reset(seed)
    _same_caches(got, want)
    rng = np.random.default_rng(seed)
    names = {"sokoban": ["Up", "Down", "Left", "Right"], "frozen_lake": ["Left", "Down", "Right", "Up"], "bandit": ["Phoenix", "Dragon"]}
    alive = [c["env_id"] for c in got]
    for turn in range(5):
        inputs = []
        for env_id in alive:
            kind = cfg.custom_envs[mgr.envs[env_id]["tag"]].env_type
            k = int(rng.integers(0, 7)) if kind != "bandit" else int(rng.integers(0, 3))
            acts = [names[kind][int(rng.integers(0, len(names[kind])))] if rng.random() > 0.15 else "Jump" for _ in range(k)]
            if rng.random() < 0.3:
                acts = [a.upper() if rng.random() < 0.5 else a.lower() for a in acts]
            resp = "<think>t</think><answer>" + " || ".join(acts) + "</answer>"
            inputs.append({"env_id": env_id, "llm_raw_response": resp, "llm_response": resp, "actions": acts[:5]})
        out_g = mgr.step(copy.deepcopy(inputs))
        out_w = ora.step(copy.deepcopy(inputs))
        assert [c["env_id"] for c in out_g] == [c["env_id"] for c in out_w], turn
        alive = [c["env_id"] for c in out_g]
        if not alive:
            break
    _same_caches(mgr.get_rollout_states(), ora.rollout_states())
    assert mgr.render() == [e["env"].render() if hasattr(e["env"], "render") else e["env"].obs for e in ora.envs]
