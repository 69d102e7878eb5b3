"""CPU: the host logic of the device trajectory assembly (ragen_b200/llm_agent/assembly.py) - chat-template glue, the chunk
sequence of a conversation, the init-time decomposability check and the segment view of a prompt - against the reference path
"messages -> chat template -> tokenizer" (ragen/llm_agent/ctx_manager.py:248-280), without a GPU.  The device gather itself
(ragen_traj_assemble / ragen_glyph_tokens) is covered by the -m gpu tests, which compare digests made by the UNMODIFIED
reference ContextManager."""
import copy

import numpy as np
import pytest

from conftest import golden_json
from fake_tokenizer import FakeQwenTokenizer, make_bpe_tokenizer
from oracle import cfg as C

GLYPHS = ("#", "_", "O", "√", "X", "P", "S", "\n")


def _manager(tok, **ap):
    from ragen_b200.llm_agent.ctx_manager import ContextManager

    cfg = C.base_config(tags=("SimpleSokoban",), n_groups=(2,), group_size=4, **ap)
    return ContextManager(cfg, tok, mode="train")


def _assembler(tok, **ap):
    from ragen_b200.llm_agent.assembly import TrajectoryAssembler

    return TrajectoryAssembler(_manager(tok, **ap))


def test_chat_glue_of_the_chatml_template():
    from ragen_b200.llm_agent.assembly import ChatGlue

    g = ChatGlue(FakeQwenTokenizer())
    assert g.start == "<|im_start|>system\n"
    assert g.between[("system", "user")] == "<|im_end|>\n<|im_start|>user\n"
    assert g.between[("user", "assistant")] == "<|im_end|>\n<|im_start|>assistant\n"
    assert g.end_generation == "<|im_end|>\n<|im_start|>assistant\n" and g.end["assistant"] == "<|im_end|>\n"


def test_templates_that_rewrite_contents_are_refused():
    from ragen_b200.llm_agent.assembly import ChatGlue, TokenizerNotDecomposable

    class Trimming(FakeQwenTokenizer):
        def apply_chat_template(self, messages, add_generation_prompt=False, tokenize=False):
            return "".join(f"<{i}|{m['role']}>{m['content']}" for i, m in enumerate(messages))  # glue depends on the position

    with pytest.raises(TokenizerNotDecomposable):
        ChatGlue(Trimming())


@pytest.mark.parametrize("tok_name,want_mode", [("bytes", "cell"), ("bpe", "row")])
def test_decomposability_check_picks_the_mode(tok_name, want_mode):
    """The byte-level tokenizer splits a grid per glyph; the BPE tokenizer merges across glyphs ('##', '#_' ...) but - with
    Qwen's pre-tokenizer - never across a row end, so it gets the row dictionary."""
    tok = FakeQwenTokenizer() if tok_name == "bytes" else make_bpe_tokenizer()
    a = _assembler(tok)
    if tok_name == "bpe":
        assert len(tok.encode("######")) < 6  # merges do cross glyphs
    assert a._validate(GLYPHS, 6, 6) == want_mode


@pytest.mark.parametrize("tok_name", ["bytes", "bpe"])
@pytest.mark.parametrize("variant", golden_json("lm_inputs.json")["variants"], ids=lambda v: v["name"])
def test_segments_equal_the_reference_tokenisation(tok_name, variant):
    """Every row of the reference-recorded rollout (es_manager.json), under all 8 config variants: the concatenation of the
    segments the device gathers is the reference's token row."""
    from ragen_b200.llm_agent.assembly import host_obs_from_text
    from ragen_b200.llm_agent.ctx_manager import ContextManager

    tok = FakeQwenTokenizer() if tok_name == "bytes" else make_bpe_tokenizer()
    cfg = C.base_config(tags=("SimpleSokoban",), n_groups=(2,), group_size=4, **variant["ap"])
    cfg.enable_response_mask = variant["erm"]
    cm = ContextManager(cfg, tok, mode="train")
    from ragen_b200.llm_agent.assembly import TrajectoryAssembler

    a = TrajectoryAssembler(cm)
    a._host_mode = a._validate(GLYPHS, 6, 6)
    env_outputs = copy.deepcopy(golden_json("es_manager.json")["final"])
    pfu = variant["update"]
    for eo in env_outputs:
        cm._trim_history(eo, pfu)
        want = tok([cm._render_text(eo, pfu)], return_tensors="pt", padding=True, padding_side="left", truncation=False).input_ids[0].tolist()
        seg = {"env_id": eo["env_id"], "history": [dict(h, state=host_obs_from_text(h["state"], GLYPHS)) if "state" in h else h
                                                    for h in eo["history"]]}
        assert a.segments_on_host(seg, pfu) == want
        # chunks: a new message starts a new chunk; the chunk holding the response is the only one a turn adds
        items = a.items(seg, pfu, is_obs=lambda st: True)
        n_obs = sum(1 for it in items if type(it).__name__ == "_HostObs")
        assert n_obs == sum(1 for h in eo["history"] if "state" in h)


def test_lazy_messages_list_is_the_reference_value():
    from ragen_b200.llm_agent.ctx_manager import LazyBatch

    calls = []
    lb = LazyBatch({"env_ids": 1}, "messages_list", lambda: calls.append(1) or ["m"])
    assert "messages_list" in lb and not calls
    assert lb["env_ids"] == 1 and not calls
    assert lb["messages_list"] == ["m"] and calls == [1]
    assert dict(lb.items())["messages_list"] == ["m"] and calls == [1]
    lb2 = LazyBatch({}, "messages_list", lambda: ["x"])
    assert list(lb2.values()) == [["x"]]


def test_obs_text_is_a_plain_string_to_everybody_else():
    import json
    import pickle

    from ragen_b200.llm_agent.assembly import ObsText

    class Log:
        epoch = 3

    t = ObsText("#P#\n###", Log(), 3, 0, 5)
    assert t == "#P#\n###" and isinstance(t, str) and json.dumps({"s": t}) == json.dumps({"s": "#P#\n###"})
    assert copy.deepcopy({"state": t})["state"] is t and t.live()
    back = pickle.loads(pickle.dumps(t))
    assert type(back) is str and back == t
    Log.epoch = 4
    assert not t.live()


def test_fast_row_path_equals_the_general_item_walk():
    """row_elements (finished turns resolved once and cached per history entry) must describe the same chunk sequence as
    items() (the reference's message loop) for every stage of a growing trajectory, in generation and in update mode."""
    from ragen_b200.llm_agent.assembly import ObsText

    class Log:  # what ObsText needs from an ObsLog on the host
        epoch, H, W = 1, 6, 6
        glyphs = GLYPHS

    a = _assembler(FakeQwenTokenizer())
    a._modes[(GLYPHS, 6, 6)] = "cell"
    log = Log()
    final = copy.deepcopy(golden_json("es_manager.json")["final"])
    for eo in final:
        full = eo["history"]
        for j, h in enumerate(full):
            if "state" in h:
                h["state"] = ObsText(h["state"], log, 1, j, eo["env_id"])
        for upto in range(1, len(full) + 1):
            for pfu in (False, True):
                hist = [dict(h) for h in full[:upto]]
                if not pfu:  # the open state of a generation call has no response yet
                    hist[-1] = {k: v for k, v in hist[-1].items() if k in ("state", "actions_left")}
                elif "llm_response" not in hist[-1]:
                    continue
                view = {"env_id": eo["env_id"], "history": hist}
                fills = []
                fast = a.row_elements(view, pfu, fills)
                slow = a.items(view, pfu)
                assert fast == slow and [type(x) for x in fast] == [type(x) for x in slow]
                assert len(fills) == len(hist) - 1 and all(f[1] == i for i, f in enumerate(fills))
