"""Generates the golden vectors under tests/golden/ by running the UNMODIFIED reference
(/root/reference, loaded through oracle/ref_loader.py + oracle/ref_shims) in this container.

    PYTHONHASHSEED=0 python tests/golden/make_golden.py

The reference ships no golden vectors of its own for this path, so these
files are what pins the oracle and, through it, the CUDA kernels.  Third-party step semantics
(gym_sokoban, gymnasium) come from the shims - see oracle/ref_shims/README.md.
"""
import json
import os
import random
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import cfg as C  # noqa: E402
from oracle import ref_loader as R  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def ref_generate(H, W, nb, depth, seed):
    """generate_room under all_seed(seed); returns (fixed, state) uint8 or None on the
    RuntimeError/RuntimeWarning that SokobanEnv.reset would catch."""
    U = R.sokoban_utils()
    all_seed = R.module("ragen.utils").all_seed
    try:
        with all_seed(seed):
            fixed, state, _, _ = U.generate_room(dim=(H, W), num_steps=int(1.7 * (H + W)), num_boxes=nb, search_depth=depth)
        return fixed.astype(np.uint8).reshape(-1), state.astype(np.uint8).reshape(-1)
    except (RuntimeError, RuntimeWarning):
        return None


def sokoban_rooms():
    out = {}
    for name, (H, W, nb, depth, seeds) in {
        "s6": (6, 6, 1, 300, list(range(0, 96)) + list(range(1010, 1020)) + list(range(123, 131))),
        "s6d30": (6, 6, 1, 30, list(range(0, 32))),
        "s8": (8, 8, 2, 10, list(range(0, 64))),
        "s6b3": (6, 6, 3, 300, list(range(0, 16))),
        "s10": (10, 10, 2, 300, list(range(0, 8))),
        "s7x9": (7, 9, 2, 50, list(range(0, 16))),
    }.items():
        fixed, state, ok = [], [], []
        for s in seeds:
            r = ref_generate(H, W, nb, depth, s)
            ok.append(r is not None)
            fixed.append(r[0] if r else np.zeros(H * W, np.uint8))
            state.append(r[1] if r else np.zeros(H * W, np.uint8))
        out[f"{name}_cfg"] = np.array([H, W, nb, depth])
        out[f"{name}_seeds"] = np.array(seeds)
        out[f"{name}_ok"] = np.array(ok)
        out[f"{name}_fixed"] = np.stack(fixed)
        out[f"{name}_state"] = np.stack(state)
        print(name, "rooms", len(seeds), "failed", len(seeds) - int(np.sum(ok)))
    np.savez_compressed(os.path.join(OUT, "sokoban_rooms.npz"), **out)
    return out


def sokoban_steps(rooms):
    """Step streams through the reference SokobanEnv (RAGEN wrapper over the gym_sokoban shim)."""
    se = R.sokoban_env()
    out = {}
    texts = {}
    for name, T, max_steps in (("s6", 40, 100), ("s8", 40, 100), ("s10", 60, 25), ("s6b3", 30, 100)):
        H, W, nb, depth = rooms[f"{name}_cfg"]
        ok = rooms[f"{name}_ok"]
        idx = np.flatnonzero(ok)[:24]
        rng = np.random.default_rng(7)
        cfg = se.SokobanEnvConfig(dim_room=(int(H), int(W)), num_boxes=int(nb), max_steps=max_steps, search_depth=int(depth))
        acts = rng.integers(1, 5, size=(len(idx), T)).astype(np.uint8)
        acts[rng.random(acts.shape) < 0.03] = 0  # a few gym no-op steps
        # bias half the streams toward the shortest solution so pushes / wins / +1 / +10 rewards occur
        U = R.sokoban_utils()
        states = np.zeros((len(idx), T, H * W), np.uint8)
        rew = np.zeros((len(idx), T)); done = np.zeros((len(idx), T), np.uint8); flags = np.zeros((len(idx), T), np.uint8)
        for j, i in enumerate(idx):
            env = se.SokobanEnv(cfg)
            env.room_fixed = rooms[f"{name}_fixed"][i].reshape(H, W).astype(np.int64)
            env.room_state = rooms[f"{name}_state"][i].reshape(H, W).astype(np.int64)
            env.num_env_steps, env.reward_last, env.boxes_on_target = 0, 0, 0
            env.player_position = np.argwhere(env.room_state == 5)[0]
            if j % 2 == 0:
                sol = U.get_shortest_action_path(env.room_fixed, env.room_state, MAX_DEPTH=60)
                pre = int(rng.integers(0, 4))
                if sol:
                    seq = list(acts[j, :pre]) + sol
                    acts[j, :min(T, len(seq))] = seq[:T]
            for t in range(T):
                obs, r, d, info = env.step(int(acts[j, t]))
                states[j, t] = env.room_state.reshape(-1)
                rew[j, t] = r; done[j, t] = d
                flags[j, t] = int(info["action_is_effective"]) | (int(info["success"]) << 1)
                if j < 2 and t < 6:
                    texts[f"{name}/{j}/{t}"] = obs
        out[f"{name}_idx"] = idx; out[f"{name}_actions"] = acts; out[f"{name}_states"] = states
        out[f"{name}_reward"] = rew; out[f"{name}_done"] = done; out[f"{name}_flags"] = flags
        out[f"{name}_max_steps"] = np.array(max_steps)
        print(name, "steps ok; wins:", int((rew > 5).sum()), "box-on:", int((np.abs(rew - 0.9) < 1e-9).sum()),
              "box-off:", int((np.abs(rew + 1.1) < 1e-9).sum()), "done:", int(done.sum()))
    np.savez_compressed(os.path.join(OUT, "sokoban_steps.npz"), **out)
    with open(os.path.join(OUT, "sokoban_texts.json"), "w") as fh:
        json.dump(texts, fh, ensure_ascii=False, indent=0)


def frozen_lake():
    fl = R.frozen_lake_env()
    seeds = list(range(0, 64)) + [42, 123, 1000, 2 ** 31 - 1, 2 ** 32 - 1]
    T = 40
    rng = np.random.default_rng(2)
    acts = rng.integers(1, 5, size=(len(seeds), T)).astype(np.uint8)
    lut = {b"S": 0, b"F": 1, b"H": 2, b"G": 3}
    desc = np.zeros((len(seeds), 16), np.uint8); start = np.zeros(len(seeds), np.uint8)
    pos = np.zeros((len(seeds), T), np.uint8); rew = np.zeros((len(seeds), T), np.float32)
    done = np.zeros((len(seeds), T), np.uint8); flags = np.zeros((len(seeds), T), np.uint8)
    texts = {}
    for j, s in enumerate(seeds):
        env = fl.FrozenLakeEnv(fl.FrozenLakeEnvConfig())
        obs0 = env.reset(seed=s)
        desc[j] = [lut[bytes(x)] for x in env.desc.reshape(-1)]
        start[j] = int(env.s)
        if j < 4:
            texts[f"{j}/reset"] = obs0
        for t in range(T):
            obs, r, d, info = env.step(int(acts[j, t]))
            pos[j, t] = int(env.s); rew[j, t] = r; done[j, t] = d
            flags[j, t] = int(info["action_is_effective"]) | (int(bool(info["success"])) << 1)
            if j < 4 and t < 5:
                texts[f"{j}/{t}"] = obs
    # 8x8, non-default p
    big = []
    for s in range(8):
        big.append("".join(R.module("ragen.env.frozen_lake.utils").generate_random_map(size=8, p=0.7, seed=s)))
    np.savez_compressed(os.path.join(OUT, "frozen_lake.npz"), seeds=np.array(seeds, dtype=np.int64), actions=acts, desc=desc,
                        start=start, pos=pos, reward=rew, done=done, flags=flags, maps8=np.array(big))
    with open(os.path.join(OUT, "frozen_lake_texts.json"), "w") as fh:
        json.dump(texts, fh, ensure_ascii=False, indent=0)
    print("frozen_lake ok; goals reached:", int((rew > 0).sum()))


def bandit():
    be = R.bandit_env()
    seeds = np.arange(500, 1012)
    acts = 1 + (np.arange(len(seeds)) % 2)
    lo_first = np.zeros(len(seeds), np.uint8); rew = np.zeros(len(seeds)); succ = np.zeros(len(seeds), np.uint8)
    texts = {}
    for j, s in enumerate(seeds):
        env = be.BanditEnv(be.BanditEnvConfig(lo_arm_name="Phoenix", hi_arm_name="Dragon"))
        obs0 = env.reset(seed=int(s))
        lo_first[j] = env.ACTION_LOOKUP[1] == "Phoenix"
        obs, r, d, info = env.step(int(acts[j]))
        rew[j] = r; succ[j] = info["success"]
        if j < 3:
            texts[str(j)] = [obs0, obs]
    np.savez_compressed(os.path.join(OUT, "bandit.npz"), seeds=seeds, actions=acts.astype(np.uint8), lo_first=lo_first,
                        reward=rew, success=succ)
    with open(os.path.join(OUT, "bandit_texts.json"), "w") as fh:
        json.dump(texts, fh, indent=0)
    print("bandit ok; mean reward", rew.mean())


def advantage():
    ca = R.ragen_core_algos()
    g = torch.Generator().manual_seed(5)
    B, T = 12, 96
    ids_mask = torch.zeros(B, T)
    rewards = torch.zeros(B, T)
    for b in range(B):
        t = int(torch.randint(5, 20, (1,), generator=g))
        n_turn = int(torch.randint(1, 5, (1,), generator=g))
        for k in range(n_turn):
            gap = int(torch.randint(3, 9, (1,), generator=g)); ln = int(torch.randint(3, 12, (1,), generator=g))
            t += gap
            if t + ln >= T:
                break
            ids_mask[b, t:t + ln] = 1
            rewards[b, t + ln - 1] = float(torch.randint(-3, 12, (1,), generator=g)) * 0.1 + 0.05
            t += ln
        if ids_mask[b].sum() == 0:
            ids_mask[b, 10:14] = 1; rewards[b, 13] = 0.7
    values = torch.randn(B, T, generator=g) * ids_mask
    out = {"rewards": rewards.numpy(), "values": values.numpy(), "mask": ids_mask.numpy()}
    for tag, (gamma, lam, hg) in {"a": (1.0, 1.0, 0.95), "b": (0.99, 0.95, 0.9)}.items():
        adv, ret = ca.compute_gae_advantage_return(rewards, values, ids_mask, gamma, lam)
        out[f"gae_{tag}_adv"] = adv.numpy(); out[f"gae_{tag}_ret"] = ret.numpy()
        adv, ret = ca.compute_bi_level_gae_advantage_return(rewards, values, ids_mask, gamma, lam, hg)
        out[f"bi_{tag}_adv"] = adv.numpy(); out[f"bi_{tag}_ret"] = ret.numpy()
        out[f"cfg_{tag}"] = np.array([gamma, lam, hg])
    # outcome rewards for GRPO: one score at the last masked token
    outcome = torch.zeros(B, T)
    for b in range(B):
        last = int(torch.nonzero(ids_mask[b]).flatten()[-1])
        outcome[b, last] = float(torch.randn(1, generator=g))
    uid = np.array(["g%d" % (b // 4) for b in range(B)], dtype=object)
    uid[-1] = "solo"
    out["outcome"] = outcome.numpy(); out["uid"] = np.array([str(u) for u in uid])
    adv, _ = ca.compute_grpo_outcome_advantage(outcome.clone(), ids_mask, uid, norm_adv_by_std_in_grpo=True)
    out["grpo_std_adv"] = adv.numpy()
    adv, _ = ca.compute_grpo_outcome_advantage(outcome.clone(), ids_mask, uid, norm_adv_by_std_in_grpo=False)
    out["grpo_mean_adv"] = adv.numpy()
    # the reference's own print-only example (ragen/trainer/core_algos.py:96-102)
    r = torch.tensor([[0, 0, 0, 0, 1, 0, 0, 0, 0, 1]]); v = torch.tensor([[0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9, 1.0]])
    adv, ret = ca.compute_bi_level_gae_advantage_return(r, v, torch.ones(1, 10), 1, 1, 0.95)
    out["example_bi_adv"] = adv.numpy(); out["example_bi_ret"] = ret.numpy()
    adv, ret = ca.compute_gae_advantage_return(r.float(), v, torch.ones(1, 10), 1, 1)
    out["example_gae_adv"] = adv.numpy(); out["example_gae_ret"] = ret.numpy()
    np.savez_compressed(os.path.join(OUT, "advantage.npz"), **out)
    print("advantage ok")


class FakeQwenTokenizer:
    """Stands in for the HF tokenizer the way the reference's own test does
    (tests/llm_agent/test_context_window.py:6-21): only name_or_path and encode() are reached by
    get_masks_and_scores."""
    name_or_path = "qwen-fake"
    IM_START, IM_END = 151644, 151645

    def encode(self, text):
        return {"<|im_start|>": [self.IM_START], "<|im_end|>": [self.IM_END]}[text]


def ctx_vectors():
    cm = R.ctx_manager()
    cases = [
        "<think>go</think><answer>Up || Down</answer>",
        "<think> a </think>  \n <answer> Left||Right|| Up ||Down||Left||Up||Right </answer> trailing",
        "<think><think></answer> 123. </think><answer> <answer> say | hi </answer></answer>",
        "<think> 456. </think><answer> 789 </answer><think> 10123 </think><answer> 11111 </answer>",
        "<think>no answer close</think><answer>Up",
        "no tags at all",
        "<think>x</think> junk <answer>Up</answer>",
        "<think>x</think><answer></answer>",
        "<think>x</think><answer> || || </answer>",
        "<think>a</think>b</think><answer>Up</answer>",
        "<think>t</think><answer>Up ||Jump|| down || LEFT</answer>",
        "<think>√ unicode  </think> <answer> Up　|| Right </answer>",
        "<think>t<|im_start|></think><answer><|im_end|>Up<think></answer>",
        "<think>t</think><answer>Up | | Down</answer>",
        "<think>first</think><answer>A</answer><answer>B</answer>",
        "<think></think><answer>Up</answer>",
        "<answer>Up</answer>",
        "<think>t</think>\t\r\n\x0b\x0c <answer>Down</answer>",
        "<think>t</think><answer>  </think> Up <|im_end|> </answer>",
        "<think>t</think><answer>Up||||Down</answer>",
        "<think>nested <ans<think>wer> x</think><answer>Right</answer>",
    ]
    out = {"parse": []}
    for think in (True, False):
        for max_actions in (5, 1):
            cfg = C.base_config(enable_think=think, max_actions_per_turn=max_actions)
            ctx = cm.ContextManager.__new__(cm.ContextManager)
            ctx.config = cfg
            ctx.action_sep = cfg.agent_proxy.action_sep
            ctx.special_token_list = ["<think>", "</think>", "<answer>", "</answer>", "<|im_start|>", "<|im_end|>"]
            for text in cases:
                resp, actions = ctx._parse_response(text)
                out["parse"].append({"enable_think": think, "max_actions": max_actions, "in": text, "llm_response": resp,
                                     "actions": actions})
    # masks and scores
    tok = FakeQwenTokenizer()
    rng = np.random.default_rng(5)
    B, L = 6, 64
    ids = np.full((B, L), 151643, dtype=np.int64)
    scores = []
    for b in range(B):
        seq = [tok.IM_START] + list(rng.integers(0, 1000, 3)) + [tok.IM_END, 198]
        nturn = int(rng.integers(1, 4))
        sc = []
        for k in range(nturn):
            seq += [tok.IM_START] + list(rng.integers(0, 1000, int(rng.integers(2, 5)))) + [tok.IM_END, 198]
            seq += [tok.IM_START] + list(rng.integers(0, 1000, int(rng.integers(2, 6)))) + [tok.IM_END, 198]
            sc.append(float(rng.integers(-2, 11)) * 0.1)
        seq = seq[:L]
        ids[b, L - len(seq):] = seq
        scores.append(sc)
    ids_t = torch.tensor(ids)
    ms = {"ids": ids, "scores": scores}
    for uts in (False, True):
        for erm in (True, False):
            s, lm, rm = cm.get_masks_and_scores(ids_t, tok, scores, use_turn_scores=uts, enable_response_mask=erm)
            ms[f"score_{int(uts)}{int(erm)}"] = s.numpy().tolist()
            ms[f"loss_{int(uts)}{int(erm)}"] = lm.numpy().astype(int).tolist()
            ms[f"resp_{int(uts)}{int(erm)}"] = rm.numpy().astype(int).tolist()
    ms["ids"] = ids.tolist()
    out["masks_scores"] = ms
    # group normalisation
    norm = {}
    Bn = 16
    last = torch.tensor(rng.normal(size=Bn).astype(np.float32))
    last[8:12] = 0.5  # a zero-variance group
    pen = torch.tensor(rng.choice([0.0, -0.1, -0.2], size=Bn).astype(np.float32))
    pen[8:12] = 0.0
    env_outputs = [{"group_id": i // 4, "tag": "A" if i < 8 else "B", "penalty": float(pen[i])} for i in range(Bn)]
    norm["last"] = last.tolist(); norm["penalty"] = pen.tolist()
    for grouping in ("state", "inductive", "batch"):
        for method in ("identity", "mean", "mean_std", "asym_clip"):
            cfg = C.base_config(reward_normalization={"grouping": grouping, "method": method})
            ctx = cm.ContextManager.__new__(cm.ContextManager)
            ctx.config = cfg
            st = torch.zeros(Bn, 5); st[:, -1] = last
            res = ctx._normalize_score_tensor(st, env_outputs)
            norm[f"{grouping}/{method}"] = res[:, -1].tolist()
    out["normalize"] = norm
    with open(os.path.join(OUT, "ctx.json"), "w") as fh:
        json.dump(out, fh, ensure_ascii=False, indent=0)
    print("ctx ok:", len(out["parse"]), "parse cases")


def es_manager_script():
    """A scripted 3-turn rollout through the reference EnvStateManager (2 groups x 4 SimpleSokoban,
    seed 123), the recipe of es_manager.py:268-323 extended with invalid / surplus actions."""
    es = R.es_manager()
    cfg = C.base_config(tags=("SimpleSokoban",), n_groups=(2,), group_size=4)
    random.seed(0); np.random.seed(0)
    mgr = es.EnvStateManager(cfg, mode="train")
    cache = mgr.reset(seed=123)
    rooms_fixed = [e["env"].room_fixed.astype(int).reshape(-1).tolist() for e in mgr.envs]
    rooms_state = [e["env"].room_state.astype(int).reshape(-1).tolist() for e in mgr.envs]
    rng = np.random.default_rng(1)
    names = ["Up", "Down", "Left", "Right"]
    script = []
    alive = [c["env_id"] for c in cache]
    for turn in range(4):
        inputs = []
        for env_id in alive:
            k = int(rng.integers(0, 7))
            acts = [names[int(rng.integers(0, 4))] if rng.random() > 0.15 else "Jump" for _ in range(k)]
            if rng.random() < 0.3:
                acts = [a.upper() if rng.random() < 0.5 else a.lower() for a in acts]
            resp = "<think>t</think><answer>" + " || ".join(acts) + "</answer>"
            inputs.append({"env_id": env_id, "llm_raw_response": resp, "llm_response": resp, "actions": acts[:5]})
        script.append(inputs)
        outs = mgr.step(json.loads(json.dumps(inputs)))
        alive = [c["env_id"] for c in outs]
        if not alive:
            break
    final = mgr.get_rollout_states()
    def clean(o):
        if isinstance(o, dict):
            return {k: clean(v) for k, v in o.items()}
        if isinstance(o, (list, tuple)):
            return [clean(v) for v in o]
        if isinstance(o, (np.bool_, bool)):
            return bool(o)
        if isinstance(o, (np.integer,)):
            return int(o)
        if isinstance(o, (np.floating,)):
            return float(o)
        return o
    status = [{"truncated": e["status"].truncated, "terminated": e["status"].terminated, "num_actions": e["status"].num_actions,
               "rewards": e["status"].rewards, "seed": e["status"].seed} for e in mgr.envs]
    with open(os.path.join(OUT, "es_manager.json"), "w") as fh:
        json.dump(clean({"group_size": 4, "seed": 123, "rooms_fixed": rooms_fixed, "rooms_state": rooms_state, "script": script,
                         "final": final, "status": status}), fh, ensure_ascii=False, indent=0)
    print("es_manager ok; turns", len(script))


def _digest(t):
    import hashlib
    a = np.ascontiguousarray(t.numpy() if hasattr(t, "numpy") else t)
    return hashlib.sha256(a.tobytes()).hexdigest()[:24] + ":" + "x".join(map(str, a.shape)) + ":" + str(a.dtype)


LM_VARIANTS = [
    {"name": "update_identity", "update": True, "ap": {}},
    {"name": "update_mean_std", "update": True, "ap": {"reward_normalization": {"grouping": "state", "method": "mean_std"}}},
    {"name": "update_batch_asym", "update": True, "ap": {"reward_normalization": {"grouping": "batch", "method": "asym_clip"}}},
    {"name": "update_turn_scores", "update": True, "ap": {"use_turn_scores": True}},
    {"name": "update_window2", "update": True, "ap": {"max_context_window": 2}},
    {"name": "update_no_think", "update": True, "ap": {"enable_think": False}, "erm": False},
    {"name": "generate", "update": False, "ap": {}},
    {"name": "generate_no_think_window1", "update": False, "ap": {"enable_think": False, "max_context_window": 1}},
]


def lm_inputs():
    """ContextManager.get_lm_inputs / formulate_rollouts / get_env_inputs of the UNMODIFIED reference, driven by the
    byte-level fake tokenizer (tests/fake_tokenizer.py) on the recorded rollout of es_manager.json."""
    import copy
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from fake_tokenizer import FakeQwenTokenizer
    cm = R.ctx_manager()
    with open(os.path.join(OUT, "es_manager.json")) as fh:
        final = json.load(fh)["final"]
    tok = FakeQwenTokenizer()
    out = {"variants": []}
    for v in LM_VARIANTS:
        cfg = C.base_config(tags=("SimpleSokoban",), n_groups=(2,), group_size=4, **v["ap"])
        cfg.enable_response_mask = v.get("erm", True)
        ctx = cm.ContextManager(cfg, tok, mode="train")
        env_outputs = copy.deepcopy(final)
        dp = ctx.get_lm_inputs(env_outputs, prepare_for_update=v["update"])
        rec = {"name": v["name"], "update": v["update"], "ap": v["ap"], "erm": v.get("erm", True), "shape": list(dp.batch["input_ids"].shape),
               "digest": {k: _digest(t) for k, t in dp.batch.items()},
               "env_ids": [int(x) for x in dp.non_tensor_batch["env_ids"]],
               "group_ids": [int(x) for x in dp.non_tensor_batch["group_ids"]],
               "messages_digest": _digest(np.frombuffer(json.dumps([list(m) for m in dp.non_tensor_batch["messages_list"]],
                                                                   ensure_ascii=False).encode(), dtype=np.uint8)),
               "history_len": [len(e["history"]) for e in env_outputs]}
        if "window" in v["name"]:  # the short ones in full, for a readable diff when something moves
            rec["messages_list"] = [list(m) for m in dp.non_tensor_batch["messages_list"]]
        if v["update"]:
            rm = dp.batch["rm_scores"]
            nz = rm.nonzero()
            rec["rm_nonzero"] = [[int(i), int(j), float(rm[i, j])] for i, j in nz.tolist()]
            rec["aliased"] = bool(dp.batch["rm_scores"].data_ptr() == dp.batch["original_rm_scores"].data_ptr())
            rec["loss_mask_sum"] = dp.batch["loss_mask"].sum(-1).tolist()
            rec["metrics"] = {k: float(x) for k, x in dp.meta_info["metrics"].items()}
        out["variants"].append(rec)
    # get_env_inputs: textual responses and token responses (the generation lacks the opening tag, :340)
    texts = ["go</think><answer>Up || Down</answer>", " a </think> <answer> Left||Right|| Up ||Down||Left||Up||Right </answer> x",
             "no close", "t</think><answer>Up ||Jump|| down || LEFT</answer>", "x</think><answer></answer>",
             "\u221a </think><answer> Right </answer>", "t</think><answer>Up<|im_end|></answer>", "</think><answer>up</answer>"]
    env = []
    for think in (True, False):
        cfg = C.base_config(tags=("SimpleSokoban",), n_groups=(2,), group_size=4, enable_think=think)
        ctx = cm.ContextManager(cfg, tok, mode="train")
        body = texts if think else [t.split("<answer>", 1)[-1] for t in texts]
        dp = cm.DataProto(batch=None, non_tensor_batch={"response_texts": body, "env_ids": np.arange(8, dtype=object)})
        a = ctx.get_env_inputs(dp)
        ids = tok([b + "<|im_end|>" for b in body]).input_ids
        dp = cm.DataProto(batch={"responses": ids}, non_tensor_batch={"env_ids": np.arange(8, dtype=object)})
        b = ctx.get_env_inputs(dp)
        for r in a + b:
            r["env_id"] = int(r["env_id"])
        env.append({"enable_think": think, "texts": body, "from_text": a, "from_tokens": b})
    out["env_inputs"] = env
    with open(os.path.join(OUT, "lm_inputs.json"), "w") as fh:
        json.dump(out, fh, ensure_ascii=False, indent=0)
    print("lm_inputs ok:", [(r["name"], r["shape"]) for r in out["variants"]])


ROLLOUT_CASES = [
    {"name": "sokoban_state_meanstd", "tags": ["SimpleSokoban"], "n_groups": [2], "group_size": 4, "seed": 123,
     "ap": {"reward_normalization": {"grouping": "state", "method": "mean_std"}}},
    {"name": "mixed_identity", "tags": ["SimpleSokoban", "FrozenLake", "Bandit", "LargerSokoban"], "n_groups": [2, 2, 1, 1],
     "group_size": 4, "seed": 31, "ap": {}},
    {"name": "mixed_inductive_asym_think_off", "tags": ["FrozenLake", "Bandit", "SimpleSokoban"], "n_groups": [1, 2, 1],
     "group_size": 8, "seed": 7, "ap": {"enable_think": False, "max_turn": 3,
                                         "reward_normalization": {"grouping": "inductive", "method": "asym_clip"}}},
]


def rollouts():
    """The rollout loop of agent_proxy.py:143-159 over the UNMODIFIED reference managers, with the scripted policy of
    tests/fake_tokenizer.py in place of the LLM (agent_proxy.py itself imports vllm / ray and is not loaded)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from fake_tokenizer import FakeQwenTokenizer, ScriptedWg
    cm, es = R.ctx_manager(), R.es_manager()
    tok = FakeQwenTokenizer()
    out = []
    for case in ROLLOUT_CASES:
        cfg = C.base_config(tags=tuple(case["tags"]), n_groups=tuple(case["n_groups"]), group_size=case["group_size"], **case["ap"])
        random.seed(0); np.random.seed(0)
        ctx, mgr = cm.ContextManager(cfg, tok, mode="train"), es.EnvStateManager(cfg, mode="train")
        wg = ScriptedWg(tok, cm.DataProto)
        env_outputs = mgr.reset(seed=case["seed"])
        turns = []
        for _ in range(cfg.agent_proxy.max_turn):
            lm_inputs = ctx.get_lm_inputs(env_outputs, prepare_for_update=False)
            lm_inputs.meta_info = {}
            lm_outputs = wg.generate_sequences(lm_inputs)
            env_inputs = ctx.get_env_inputs(lm_outputs)
            turns.append({"n": len(env_inputs), "input_ids": _digest(lm_inputs.batch["input_ids"]),
                          "responses": lm_outputs.non_tensor_batch["response_texts"][:3]})
            env_outputs = mgr.step(env_inputs)
            if len(env_outputs) == 0:
                break
        dp = ctx.formulate_rollouts(mgr.get_rollout_states())
        rm = dp.batch["rm_scores"]
        out.append({**case, "turns": turns, "shape": list(dp.batch["input_ids"].shape),
                    "digest": {k: _digest(t) for k, t in dp.batch.items()},
                    "rm_last": rm[:, -1].tolist(), "rm_sum": rm.sum(-1).tolist(),
                    "loss_mask_sum": dp.batch["loss_mask"].sum(-1).tolist(),
                    "metrics": {k: float(x) for k, x in dp.meta_info["metrics"].items()}})
    with open(os.path.join(OUT, "rollouts.json"), "w") as fh:
        json.dump(out, fh, ensure_ascii=False, indent=0)
    print("rollouts ok:", [(r["name"], r["shape"], len(r["turns"])) for r in out])


def advantage_more():
    """RLOO / RF++ / RF++-baseline / ReMax / kl_penalty from the UNMODIFIED verl core_algos.py; apply_kl_penalty's arithmetic
    (ray_trainer.py:145-175 imports ray, so its three lines are evaluated here on verl's own kl_penalty and masked_mean)."""
    ca = R.verl_core_algos()
    tf = R.torch_functional()
    g = torch.Generator().manual_seed(21)
    B, T = 12, 80
    mask = torch.zeros(B, T)
    for b in range(B):
        a = int(torch.randint(0, 30, (1,), generator=g)); e = int(torch.randint(a + 5, T + 1, (1,), generator=g))
        mask[b, a:e] = 1
        if b % 3 == 0 and e - a > 12:
            mask[b, a + 4:a + 8] = 0  # a hole, as multi-turn loss masks have
    outcome = torch.zeros(B, T)
    for b in range(B):
        last = int(torch.nonzero(mask[b]).flatten()[-1])
        outcome[b, last] = float(torch.randn(1, generator=g))
    dense = torch.randn(B, T, generator=g) * 0.3 * mask
    uid = np.array(["g%d" % (b // 4) for b in range(B)], dtype=object); uid[-1] = "solo"
    out = {"mask": mask.numpy(), "outcome": outcome.numpy(), "dense": dense.numpy(), "uid": np.array([str(u) for u in uid])}
    out["rloo"] = ca.compute_rloo_outcome_advantage(outcome.clone(), mask, uid)[0].numpy()
    out["rfpp_baseline"] = ca.compute_reinforce_plus_plus_baseline_outcome_advantage(outcome.clone(), mask, uid)[0].numpy()
    for tag, gamma in (("1", 1.0), ("97", 0.97)):
        adv, ret = ca.compute_reinforce_plus_plus_outcome_advantage(dense.clone(), mask, gamma)
        out[f"rfpp_adv_{tag}"] = adv.numpy(); out[f"rfpp_ret_{tag}"] = ret.numpy()
    base = torch.randn(B, generator=g)
    adv, ret = ca.compute_remax_outcome_advantage(dense.clone(), base, mask)
    out["remax_base"] = base.numpy(); out["remax_adv"] = adv.numpy(); out["remax_ret"] = ret.numpy()
    logp = -torch.rand(B, T, generator=g) * 3; ref = logp + torch.randn(B, T, generator=g) * 0.2
    out["logp"] = logp.numpy(); out["ref_logp"] = ref.numpy()
    for kind in ("kl", "abs", "mse", "low_var_kl"):
        kld = ca.kl_penalty(logp, ref, kl_penalty=kind) * mask
        out[f"klrew_{kind}"] = (outcome - 0.05 * kld).numpy()
        out[f"klcur_{kind}"] = np.array(torch.mean(tf.masked_mean(kld, mask=mask, axis=-1), dim=0).item())
    # _filter_rollout statistics: plain torch calls of agent_trainer.py:466-469 on a [G*gs, T] score tensor
    G, gs = 8, 4
    sc = torch.zeros(G * gs, T); sc[:, -1] = torch.randn(G * gs, generator=g); sc[4:8, -1] = 0.5; sc[:, 10] = torch.randn(G * gs, generator=g) * 0.1
    rm = sc.sum(dim=-1).view(G, gs)
    out["filter_scores"] = sc.numpy(); out["filter_std"] = rm.std(dim=-1).numpy(); out["filter_max"] = rm.max(dim=-1).values.numpy()
    out["filter_mean"] = rm.mean(dim=-1).numpy()
    out["filter_top_std"] = rm.std(dim=-1).topk(2).indices.numpy(); out["filter_top_std_rev"] = (-rm.std(dim=-1)).topk(2).indices.numpy()
    np.savez_compressed(os.path.join(OUT, "advantage_more.npz"), **out)
    print("advantage_more ok")


SCORE_PAIRS = [
    ("42", "42"), (" 42 \n", "42"), ("The Answer", "the answer"), ("the answer.", "The Answer"), ("x = 3", "x=3"), ("", "7"), ("7", ""),
    ("   ", "7"), ("3.50", "3.5"), ("003.5", "3.5"), ("about 12 apples", "12"), ("12.0", "12"), ("12.", "12"), ("1,234", "1234"),
    ("no digits", "5"), ("5", "no digits"), ("B", "b"), ("(B) because", "B"), ("the answer is c", "C"), ("E", "A"), ("none", "A"),
    ("none", "none"), ("A or B", "B"), ("\frac{1}{2}", "\\frac{1}{2}"), ("1/2", "0.5"), ("0.5", ".5"), ("-3", "3"), ("3e2", "300"),
    ("10%", "10"), ("Hello, World!", "hello world"), ("tab\there", "tabthere"), ("a\x1cb", "ab"), ("ABCD", "abcd"), ("d", "D."),
    ("123456789012345", "123456789012345.0"), ("x", "X "), ("12 13", "12"), ("13 12", "12"), ("_a_", "a"), ("a-b", "ab"),
]


def text_scores():
    """The UNMODIFIED scorers: MetaMathQAEnv._check_answer (metamathqa/env.py:55-66) and the three compute_score_* functions of
    ragen/env/static/utils.py, on ASCII pairs."""
    su = R.module("ragen.env.static.utils")
    mm = R.module("ragen.env.metamathqa.env")
    rows = []
    for pred, label in SCORE_PAIRS:
        obj = type("X", (), {"correct_answer": label})()
        c, v = mm.MetaMathQAEnv._check_answer(obj, pred)
        row = {"pred": pred, "label": label, "metamathqa": [bool(c), bool(v)]}
        for name, fn in (("exact_match", su.compute_score_exact_match), ("numeric", su.compute_score_numeric),
                         ("multiple_choice", su.compute_score_multiple_choice)):
            r = fn(pred, label)
            row[name] = [bool(r["is_correct"]), bool(r["is_valid"])]
        rows.append(row)
    with open(os.path.join(OUT, "text_scores.json"), "w") as fh:
        json.dump(rows, fh, indent=0)
    print("text_scores ok:", len(rows))


QA_ROWS = [{"query": "What is 6 times 7?", "response": "6 * 7 = 42.\nThe answer is: 42"},
           {"query": "Simplify 1/2 + 1/4.", "response": "Common denominator 4.\nThe answer is: \\frac{3}{4}"},
           {"query": "Name the shape with three sides.", "response": "Three sides.\nThe answer is: Triangle "},
           {"query": "A question whose solution has no marker.", "response": "It is 5."},
           {"query": "What is 10 / 4?", "response": "10/4 = 2.5\nThe answer is: 2.5"}]
QA_ANSWERS = ["42", " 42 ", "forty two", "\\frac{3}{4}", "\\FRAC{3}{4}", "3/4", "triangle", "TRI ANGLE", "", "   ", "5", "2.5", "2.50"]


def metamath_episodes():
    """MetaMathQAEnv.reset / step of the UNMODIFIED reference (ragen/env/metamathqa/env.py:29-53) on an in-memory dataset
    (the constructor's load_dataset download is bypassed with __new__)."""
    import contextlib
    import io
    mm = R.module("ragen.env.metamathqa.env")
    rng = np.random.default_rng(12)
    out = []
    for seed in range(8):
        env = mm.MetaMathQAEnv.__new__(mm.MetaMathQAEnv)
        env.config = mm.MetaMathQAEnvConfig(split="train")
        env.dataset = {"train": QA_ROWS}
        with contextlib.redirect_stdout(io.StringIO()):  # _extract_answer prints the solution
            q = env.reset(seed=seed)
        answers = [QA_ANSWERS[int(i)] for i in rng.integers(0, len(QA_ANSWERS), 4)]
        if seed % 2 == 0 and env.correct_answer:
            answers[2] = env.correct_answer.upper()
        steps = []
        for a in answers:
            obs, reward, done, info = env.step(a)
            steps.append({"answer": a, "obs": obs, "reward": reward, "done": bool(done), "info": {k: bool(v) for k, v in info.items()}})
            if done:
                break
        out.append({"seed": seed, "question": q, "idx": env.current_question_idx, "label": env.correct_answer, "steps": steps})
    with open(os.path.join(OUT, "metamath_episodes.json"), "w") as fh:
        json.dump({"rows": QA_ROWS, "episodes": out}, fh, indent=0)
    print("metamath_episodes ok:", [(e["idx"], len(e["steps"])) for e in out])


def countdown():
    cd = R.countdown_env()
    cases = [("1+2+3", [1, 2, 3], 6), ("1+2", [1, 2, 3], 3), ("3*(2-1)", [1, 2, 3], 3), ("1 + 2 = 3", [1, 2, 3], 3),
             ("6/(3-1)", [1, 3, 6], 3), ("2/0+1", [0, 1, 2], 3), ("", [1], 1), ("012+3", [12, 3], 15), ("-5+ +7 - -2", [5, 7, 2], 4),
             ("(44 - 19) + 5", [44, 19, 5], 30), ("44 - 19 + 5)", [44, 19, 5], 30), ("99*98*97*96", [99, 98, 97, 96], 90345024),
             ("7/2*2", [7, 2, 2], 7), ("10/4", [10, 4], 2), ("  3 + 4", [3, 4], 7), ("3 + 4\n", [3, 4], 7), ("3 4 +", [3, 4], 7),
             ("3 ++ 4", [3, 4], 7), ("3 +* 4", [3, 4], 7), ("((3)) + ((4))", [3, 4], 7), ("3 + 4 answer", [3, 4], 7),
             ("1/3*3", [1, 3, 3], 1), ("50 - 25 - 12", [50, 25, 12], 13), ("5 * 5 / 5", [5, 5, 5], 5), ("0/5 + 2", [0, 5, 2], 2),
             ("8/0", [8, 0], 1), ("00 + 7", [0, 7], 7), ("7 - 00", [0, 7], 7),
             # the rest of what eval() takes from an LLM: floor division, modulo, powers, float literals
             ("8 // 2 + 3", [8, 2, 3], 7), ("2**3", [2, 3], 8), ("7 % 3 + 5", [7, 3, 5], 6), ("-7 // 2", [7, 2], -4), ("-7 % 3", [7, 3], 2),
             ("7 % -3", [7, 3], -2), ("2**-1 * 4", [2, 1, 4], 2), ("-2**2", [2, 2], -4), ("2**3**2", [2, 3, 2], 512), ("(2**3)**2", [2, 3, 2], 64),
             ("1.5*2", [1, 5, 2], 3), (".5*4", [5, 4], 2), ("3e0+4", [3, 0, 4], 7), ("7.0 // 2", [7, 0, 2], 3), ("7.5 % 2", [7, 5, 2], 1),
             ("5 // 0", [5, 0], 1), ("5 % 0", [5, 0], 1), ("0 ** -1", [0, 1], 1), ("2 ** 0.5 * 2 ** 0.5", [2, 0, 5, 2, 0, 5], 2),
             ("9 ** 0.5", [9, 0, 5], 3), ("10 // 3 * 3 + 10 % 3", [10, 3, 3, 10, 3], 10), ("6 / 4 // 1", [6, 4, 1], 1),
             ("012.5 + 1", [12, 5, 1], 13), ("1. + 2", [1, 2], 3), ("4 ** 3 ** 0", [4, 3, 0], 4), ("-3 ** 2 + 18", [3, 2, 18], 9)]
    obj = type("X", (), {"config": type("C", (), {"score": 1, "format_score": 0.1})})()
    rows = [{"answer": a, "nums": n, "target": t, "reward": cd.CountdownEnv.compute_reward(obj, a, {"target": t, "nums": n})}
            for a, n, t in cases]
    with open(os.path.join(OUT, "countdown.json"), "w") as fh:
        json.dump(rows, fh, indent=0)
    print("countdown ok", [r["reward"] for r in rows])


if __name__ == "__main__":
    assert R.available(), "reference not present"
    if os.environ.get("PYTHONHASHSEED") != "0":
        print("warning: PYTHONHASHSEED != 0 (only matters if a reset hits the retry path)")
    rooms = sokoban_rooms()
    sokoban_steps(rooms)
    frozen_lake()
    bandit()
    advantage()
    ctx_vectors()
    es_manager_script()
    lm_inputs()
    rollouts()
    advantage_more()
    text_scores()
    metamath_episodes()
    countdown()
