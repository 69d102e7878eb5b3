"""GPU parity: ragen_parse_actions and ragen_countdown_score (through the C ABI) against the reference-made golden
cases, the oracle's re-based parser and Python's eval, on the C3 corpus (16 384 synthetic LLM-style responses)."""
import numpy as np
import pytest
import torch

from conftest import golden_json
from oracle import managers as M
from oracle import toy_envs as TE

pytestmark = pytest.mark.gpu
SOKOBAN = {1: "Up", 2: "Down", 3: "Left", 4: "Right"}


def test_parse_golden_cases():
    from ragen_b200 import ops

    cases = golden_json("ctx.json")["parse"]
    tables = ops.ActionTables([SOKOBAN], "cuda")
    for think in (True, False):
        for mx in (5, 1):
            sub = [c for c in cases if c["enable_think"] == think and c["max_actions"] == mx]
            res = ops.parse_actions([c["in"] for c in sub], tables, None, think, "||", mx).to_host()
            for c, (resp, actions, codes) in zip(sub, res):
                assert resp == c["llm_response"], c
                assert actions == c["actions"], c
                assert codes == M.ManagerOracle.map_actions(SOKOBAN, actions)


def test_parse_fuzz_against_oracle():
    from ragen_b200 import ops

    rng = np.random.default_rng(12)
    atoms = ["<think>", "</think>", "<answer>", "</answer>", "<|im_start|>", "<|im_end|>", "Up", "down", "LEFT", "Right", "Jump",
             "||", "|", " ", "\n", "\t", "　", "\xa0", "\x85", "√", "<", ">", "</", "<thi", "nk>", "</ans", "wer>", "x", "||||",
             " ", "é", "<answe", "r>", "a" * 40]
    texts = []
    for it in range(6000):
        text = "".join(atoms[int(i)] for i in rng.integers(0, len(atoms), int(rng.integers(1, 16))))
        if it % 3 == 0:
            text = "<think>" + text
        if it % 5 == 0:
            text += "</think><answer>" + "".join(atoms[int(i)] for i in rng.integers(0, len(atoms), 4)) + "</answer>"
        texts.append(text)
    texts += ["", "<", "<think>", "<think></think><answer></answer>"]
    tables = ops.ActionTables([SOKOBAN], "cuda")
    for think in (True, False):
        for mx in (5, 2):
            res = ops.parse_actions(texts, tables, None, think, "||", mx)
            host = res.to_host()
            total = res.n_total.cpu().numpy()
            for text, (resp, actions, codes), tot in zip(texts, host, total):
                want_resp, want_actions = M.parse_response(text, think, "||", mx)
                assert (resp, actions) == (want_resp, want_actions), repr(text)
                assert codes == M.ManagerOracle.map_actions(SOKOBAN, actions), repr(text)


def test_c3_corpus_bandit_countdown():
    """BASELINE config 3: 8192 Bandit + 8192 Countdown responses, per-env lookup tables (arm order differs per env)."""
    from ragen_b200 import ops
    from ragen_b200.synth import synthetic_responses

    c = synthetic_responses(8192, 8192, seed=3)
    n = len(c["responses"])
    rng = np.random.default_rng(0)
    lo_first = rng.integers(0, 2, n)
    lookups = [{1: "Phoenix", 2: "Dragon"}, {1: "Dragon", 2: "Phoenix"}]
    tables = ops.ActionTables(lookups, "cuda")
    tid = np.where(c["kind"] == 0, 1 - lo_first, -1).astype(np.int32)
    res = ops.parse_actions(c["responses"], tables, tid, True, "||", 1)
    host = res.to_host()
    fl = res.flags.cpu().numpy()
    n_match = 0
    for i, (resp, actions, codes) in enumerate(host):
        want_resp, want_actions = M.parse_response(c["responses"][i], True, "||", 1)
        assert (resp, actions) == (want_resp, want_actions), i
        n_match += bool(want_actions)
        if c["kind"][i] == 0:
            assert codes == M.ManagerOracle.map_actions(lookups[tid[i]], actions), i
            assert fl[i] & 16 or not fl[i] & 1
        else:
            assert codes == [] and not fl[i] & 16
    assert n_match > 0.7 * n
    # Countdown scoring of the first parsed action of each Countdown response, spans taken on the device
    cd = np.flatnonzero(c["kind"] == 1)
    idx = torch.from_numpy(cd).cuda()
    assert bool(((res.flags[idx] & 4) != 0).any())  # the nested-tag shape needs the cleaned copy of the answer
    out, uns = ops.countdown_score_parsed(res, idx, torch.from_numpy(c["nums"][cd]).cuda(), torch.from_numpy(c["n_nums"][cd]).cuda(),
                                          torch.from_numpy(c["target"][cd]).cuda())
    out, uns = out.cpu().numpy(), uns.cpu().numpy()
    vals = {0: 0, 1: 0.1, 2: 1}
    hist = {0: 0, 1: 0, 2: 0}
    for j, i in enumerate(cd):
        actions = host[i][1]
        ans = actions[0] if actions else ""
        k = int(c["n_nums"][i])
        want = TE.countdown_reward(ans, c["nums"][i, :k].tolist(), int(c["target"][i]))
        assert vals[int(out[j])] == want, (i, ans)
        hist[int(out[j])] += 1
    assert min(hist.values()) > 100  # all three reward classes occur


def test_countdown_golden_and_grammar_flags():
    from ragen_b200 import ops

    rows = golden_json("countdown.json")
    extra = [("9**99", [9, 99], 8), ("7 & 2", [7, 2], 3), ("1_5*2", [1, 5, 2], 3), ("((((((((((((((((((((((((((((((((((1))))))))))))))))))))))))))))))))))+2", [1, 2], 3)]
    answers = [r["answer"] for r in rows] + [e[0] for e in extra]
    nums_l = [r["nums"] for r in rows] + [e[1] for e in extra]
    targets = [r["target"] for r in rows] + [e[2] for e in extra]
    blob = "".join(answers).encode()
    lens = np.array([len(a.encode()) for a in answers], np.int32)
    start = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
    nums = np.zeros((len(answers), 8), np.int32)
    for i, v in enumerate(nums_l):
        nums[i, :len(v)] = v
    text = torch.frombuffer(bytearray(blob), dtype=torch.uint8).cuda()
    out, uns = ops.countdown_score(text, torch.from_numpy(start).cuda(), torch.from_numpy(lens).cuda(), torch.from_numpy(nums).cuda(),
                                   torch.tensor([len(v) for v in nums_l], dtype=torch.int32).cuda(),
                                   torch.tensor(targets, dtype=torch.int64).cuda())
    out, uns = out.cpu().numpy(), uns.cpu().numpy()
    for i, r in enumerate(rows):
        assert {0: 0, 1: 0.1, 2: 1}[int(out[i])] == r["reward"], r
    flagged = [r["answer"] for r, u in zip(rows, uns) if u]
    assert flagged == ["1 + 2 = 3", "3 + 4 answer"]  # floor division, modulo, powers and float literals are inside the grammar
    assert out[len(rows):].tolist() == [1, 1, 1, 1] and uns[len(rows):].tolist() == [1, 1, 1, 1]


def test_countdown_env_surfaces_the_unsupported_flag():
    """An answer outside the device grammar is scored as an eval() exception AND reported (no host re-evaluation)."""
    import warnings

    from ragen_b200.env.countdown import CountdownEnv, CountdownEnvConfig

    cfg = CountdownEnvConfig(data=[{"nums": [8, 2, 3], "target": 7}, {"nums": [7, 2, 2], "target": 3}])
    env = CountdownEnv(cfg)
    assert len(env.data) == 2
    env.reset(seed=0)
    with warnings.catch_warnings():
        warnings.simplefilter("error")
        assert env.step("8 // 2 + 3")[1] == env.config.score and env.last_unsupported is False
    env.reset(seed=1)
    with pytest.warns(RuntimeWarning, match="outside the device evaluator"):
        _, reward, done, info = env.step("7 ^ 2 ^ 2")  # bit operators: eval() takes them, the device grammar does not
    assert reward == env.config.format_score and done and env.last_unsupported is True and env.n_unsupported == 1
