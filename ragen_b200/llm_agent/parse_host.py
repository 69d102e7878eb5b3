"""Host-side view of ragen_parse_actions results: builds the (llm_response, actions) pair of
ContextManager._parse_response (ragen/llm_agent/ctx_manager.py:148-173) from the spans the kernel wrote.
Only slicing and joining of byte ranges happens here; every decision was taken on the device."""

F_MATCHED, F_THINK_SCRATCH, F_ANSWER_SCRATCH, F_TRUNCATED, F_CODES = 1, 2, 4, 8, 16


def assemble_response(raw: bytes, scratch: bytes, flags: int, spans, act_spans, kept: int, enable_think: bool, sep: str):
    """raw / scratch: the response's bytes and its region of the scratch blob.  spans = (think_start, think_len,
    answer_start, answer_len), act_spans [max_actions][2]; offsets are relative to the response start."""
    if not flags & F_MATCHED:
        return raw.decode("utf-8", "replace"), []
    abuf = scratch if flags & F_ANSWER_SCRATCH else raw
    tbuf = scratch if flags & F_THINK_SCRATCH else raw
    actions = [abuf[s:s + n].decode("utf-8", "replace") for s, n in act_spans[:kept]]
    if flags & F_TRUNCATED:
        answer = (" " + sep + " ").join(actions)
    else:
        answer = abuf[spans[2]:spans[2] + spans[3]].decode("utf-8", "replace")
    if enable_think:
        think = tbuf[spans[0]:spans[0] + spans[1]].decode("utf-8", "replace")
        return f"<think>{think}</think><answer>{answer}</answer>", actions
    return f"<answer>{answer}</answer>", actions
