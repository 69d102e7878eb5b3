"""Token-trajectory assembly on the device: ContextManager.get_lm_inputs without re-tokenising the history every turn.

The reference (ragen/llm_agent/ctx_manager.py:228-280) rebuilds, for every env and every turn, the whole conversation as
Python strings, renders the chat template and runs the tokenizer over all of it.  Here a prompt is a LIST OF SEGMENTS of a
token pool that lives in HBM:

  * text chunks - chat-template glue, the instruction prefix, "Turn k / State:", "You have N actions left ...", "Reward: r", the
    canonical LLM responses - are tokenised ONCE (memoised by their text) and uploaded; a turn adds one new chunk per env
    (the response), everything else is found again in the pool;
  * grid observations never become text on this path: the glyph codes the env kernels wrote (ObsLog) are turned into token ids
    on the device (ragen_glyph_tokens, "cell" mode) or, for tokenizers whose merges cross glyphs, through a dictionary of row
    segments ("row" mode);
  * one launch (ragen_traj_assemble) gathers the segments of all rows into the left-padded int64 input_ids / attention_mask /
    position_ids, which the mask / score kernels consume in place.

Whether a tokenizer can be decomposed like this is CHECKED, not assumed: the first time a glyph table is seen, random
conversations are assembled both ways on the host (full reference path vs segments) and compared id for id; a tokenizer
that fails is refused with a clear error (ContextManager(assembly="host") is then the explicit choice) - there is no silent
fallback.
"""
from typing import Dict, List, Optional

import numpy as np
import torch

from .. import _lib

SEP = 7  # glyph code of the row separator (include/ragen_b200.h)


class TokenizerNotDecomposable(RuntimeError):
    pass


class _TooManyTokens(Exception):
    """A glyph takes more than the 4 tokens ragen_glyph_tokens holds per code: cell mode is not available."""


class ObsText(str):
    """A text observation that also knows where its glyph codes live (ObsLog slot + env row).  It IS the reference's
    state string for every consumer; the device assembler reads the reference instead of the characters."""

    def __new__(cls, text, log=None, epoch=0, slot=0, index=0):
        self = super().__new__(cls, text)
        self.log, self.epoch, self.slot, self.index = log, epoch, slot, index
        return self

    def __deepcopy__(self, memo):
        return self  # immutable; the log it points to is shared on purpose

    def __reduce__(self):
        return (str, (str(self),))  # leaves the process as a plain string

    def live(self):
        return self.log is not None and self.log.epoch == self.epoch


class ObsLog:
    """Glyph rows of every turn of the current rollout for one block of grid envs, on the device (what the env kernels
    rendered, copied device-to-device) and as unpacked codes on the host (the manager already brought them over to
    build the state strings)."""

    def __init__(self, n, H, W, glyphs, device, capacity=8):
        self.n, self.H, self.W, self.G = int(n), int(H), int(W), int(H) * (int(W) + 1)
        self.glyphs = tuple(glyphs)  # 8 strings: codes 0..6 and the separator
        self.device = device
        self.epoch = 0
        self.dev = torch.empty((capacity, self.n, self.G), dtype=torch.uint8, device=device)
        self.packed: List[bool] = []
        self.host: List[np.ndarray] = []

    def begin(self):
        self.epoch += 1
        self.packed, self.host = [], []

    def record(self, dev_obs, packed, host_codes):
        slot = len(self.host)
        if slot >= self.dev.shape[0]:
            grown = torch.empty((2 * self.dev.shape[0], self.n, self.G), dtype=torch.uint8, device=self.device)
            grown[: self.dev.shape[0]].copy_(self.dev)
            self.dev = grown
        self.dev[slot, :, : dev_obs.shape[1]].copy_(dev_obs)
        self.packed.append(bool(packed))
        self.host.append(host_codes)
        return slot

    def wrap(self, texts, slot):
        return [ObsText(t, self, self.epoch, slot, i) for i, t in enumerate(texts)]


class TokenPool:
    """int32 token ids in HBM + the memo text -> (start, length).  Appends are batched: one host->device copy per call."""

    def __init__(self, device, capacity=1 << 20, limit=1 << 26):
        self.device = device
        self.buf = torch.empty(capacity, dtype=torch.int32, device=device)
        self.used = 0
        self.limit = limit
        self.generation = 0
        self.memo: Dict[str, tuple] = {}

    def reserve(self, n):
        n = int(n)
        if self.used + n > self.buf.numel():
            cap = self.buf.numel()
            while self.used + n > cap:
                cap *= 2
            grown = torch.empty(cap, dtype=torch.int32, device=self.device)
            grown[: self.used].copy_(self.buf[: self.used])
            self.buf = grown
        start = self.used
        self.used += n
        return start

    def put(self, id_lists):
        """Appends the token lists; returns their (start, length) pairs."""
        lens = [len(x) for x in id_lists]
        total = sum(lens)
        start = self.reserve(total)
        if total:
            flat = np.fromiter((t for x in id_lists for t in x), dtype=np.int32, count=total)
            self.buf[start: start + total].copy_(torch.from_numpy(flat), non_blocking=False)
        out, at = [], start
        for n in lens:
            out.append((at, n))
            at += n
        return out

    def full(self):
        return self.used > self.limit

    def reset(self):
        self.used = 0
        self.memo.clear()
        self.generation += 1


class ChatGlue:
    """The constant text a chat template puts around message contents, found by rendering a probe conversation with
    sentinel contents.  Refuses templates that are not a plain concatenation (content-dependent glue)."""

    ROLES = ("system", "user", "assistant", "user", "assistant", "user")

    def __init__(self, tokenizer):
        marks = [f"@@RAGEN{i}MARK@@" for i in range(6)]
        msgs = [{"role": r, "content": m} for r, m in zip(self.ROLES, marks)]

        def cut(text, upto):
            parts, at = [], 0
            for m in marks[:upto]:
                i = text.find(m, at)
                if i < 0:
                    raise TokenizerNotDecomposable("chat template drops or rewrites message contents")
                parts.append(text[at:i])
                at = i + len(m)
            parts.append(text[at:])
            return parts

        gen = cut(tokenizer.apply_chat_template(msgs, add_generation_prompt=True, tokenize=False), 6)
        usr = cut(tokenizer.apply_chat_template(msgs, add_generation_prompt=False, tokenize=False), 6)
        ast = cut(tokenizer.apply_chat_template(msgs[:5], add_generation_prompt=False, tokenize=False), 5)
        short = cut(tokenizer.apply_chat_template(msgs[:2], add_generation_prompt=False, tokenize=False), 2)
        if not (gen[:6] == usr[:6] and gen[:5] == ast[:5] and gen[2] == gen[4] and gen[3] == gen[5] and short[:2] == gen[:2]
                and short[2] == usr[6]):
            raise TokenizerNotDecomposable("chat template glue depends on the position or content of a message")
        self.start = gen[0]
        self.between = {("system", "user"): gen[1], ("user", "assistant"): gen[2], ("assistant", "user"): gen[3]}
        self.end_generation = gen[6]          # after a user message, with the generation prompt
        self.end = {"user": usr[6], "assistant": ast[5]}  # after the last message, no generation prompt


def encode_many(tokenizer, texts):
    """Token ids of each text WITHOUT the specials a tokenizer adds around a whole input."""
    if not texts:
        return []
    try:
        out = tokenizer(list(texts), add_special_tokens=False)["input_ids"]
        return [list(x) for x in out]
    except TypeError:  # a duck-typed tokenizer with the members the reference touches only (tests/fake_tokenizer.py)
        return [list(tokenizer.encode(t)) for t in texts]


class TrajectoryAssembler:
    def __init__(self, ctx_manager, validate_samples=48):
        self.cm = ctx_manager
        self.tokenizer = ctx_manager.tokenizer
        self.device = ctx_manager.device
        self._ctx = self._pool = None  # device objects are made at the first launch: the host logic is testable without a GPU
        self.glue = ChatGlue(self.tokenizer)
        self.pad_id = int(self.tokenizer.pad_token_id)
        self.validate_samples = int(validate_samples)
        self._modes: Dict[tuple, str] = {}       # glyph table -> "cell" | "row"
        self._tables: Dict[tuple, tuple] = {}    # glyph table -> (int32 [8, 4] tokens, int32 [8] lengths), cell mode
        self._rows: Dict[tuple, dict] = {}       # glyph table -> {row bytes: (start, len)}, row mode
        self._rendered: Dict[tuple, tuple] = {}  # (id(log), epoch, slot, pool generation) -> per-env segment arrays
        self._row_cache: Dict[int, list] = {}    # id(rollout_cache dict) -> [dict, finished turns resolved, first entry, last resolved entry, pool generation, elements]
        self._g_ua, self._g_au = self._glue("user", "assistant"), self._glue("assistant", "user")
        self._gen_tail = self.glue.end_generation + ("<think>" if ctx_manager.config.agent_proxy.enable_think else "<answer>")
        self._heads: Dict[int, tuple] = {}       # env_id -> the two chunks every prompt of the env starts with
        self._tails: Dict[tuple, str] = {}       # (env_id, actions_left) -> "You have N actions left ..." text
        self._staging = None
        # the specials the tokenizer adds around a whole input are not part of any chunk: an empty text shows them
        try:
            lone = list(self.tokenizer([""])["input_ids"][0])
        except Exception:  # duck-typed tokenizers (tests/fake_tokenizer.py) only take the reference's exact call
            lone = []
        if len(lone):
            raise TokenizerNotDecomposable("the tokenizer adds special tokens around every input; chunk-wise tokenisation "
                                           "would not reproduce tokenizer(text)")

    @property
    def ctx(self):
        if self._ctx is None:
            self._ctx = _lib.Context.get(self.device.index)  # raises without a B200: there is no CPU assembly path
        return self._ctx

    @property
    def pool(self):
        if self._pool is None:
            self.ctx
            self._pool = TokenPool(self.device)
        return self._pool

    def segments_on_host(self, env_output, prepare_for_update):
        """Token ids of one (already trimmed) trajectory as the concatenation of its segments, computed on the host: what
        assemble() gathers on the device, used by the init-time validation and by the CPU tests."""
        out = []
        for it in self.items(env_output, prepare_for_update, is_obs=lambda st: isinstance(st, _HostObs) and self._host_mode != "text"):
            if isinstance(it, _HostObs):
                out.extend(self._obs_ids_host(it.codes, it.glyphs, self._host_mode))
            else:
                out.extend(encode_many(self.tokenizer, [it])[0])
        return out

    _host_mode = "cell"

    # ---- the conversation as items: text chunks and observation references ---------------------------------------------
    def _is_device_obs(self, st):
        return type(st) is ObsText and st.live() and self.mode(st.log) != "text"

    def items(self, env_output, prepare_for_update, is_obs=None):
        """The reference's message loop (ctx_manager.py:248-274) emitting chunks instead of one string.  A chunk ends where a
        new message starts (the glue begins with a special token there) and at device-rendered observations; the newline
        that follows a state in the prompt belongs to the observation (its last separator code)."""
        cm, glue = self.cm, self.glue
        is_obs = is_obs or self._is_device_obs
        cfg = cm.config.agent_proxy
        think = cfg.enable_think
        fmt = "<think> [Your thoughts] </think> <answer> [your answer] </answer>" if think else "<answer> [your answer] </answer>"
        length = f"Max response length: {cm.env_config_lookup[env_output['env_id']]['max_tokens']} words (tokens)."
        out = [glue.start + "You're a helpful assistant. ", glue.between[("system", "user")] + cm.prefix_lookup[env_output["env_id"]]]
        role = "user"
        history = env_output["history"]
        last = len(history) - 1
        for idx, content in enumerate(history):
            out[-1] += f"\nTurn {idx + 1}:\n"
            if "state" in content:
                st = content["state"]
                tail = (f"You have {content['actions_left']} actions left. Always output: {fmt} with no extra text. "
                        f"Strictly follow this format. {length}\n")
                if is_obs(st):
                    out[-1] += "State:\n"
                    out.append(st)
                    out.append(tail)
                else:
                    out[-1] += f"State:\n{st}\n" + tail
            if "llm_response" in content:
                out.append(self._glue(role, "assistant") + content["llm_response"])
                role = "assistant"
            if "reward" in content and not (prepare_for_update and idx == last):
                out.append(self._glue(role, "user") + f"Reward:\n{content['reward']}\n")
                role = "user"
        if prepare_for_update:
            tailtext = glue.end[role]
        else:
            tailtext = glue.end_generation + ("<think>" if think else "<answer>")
        if tailtext:
            out[-1] += tailtext  # the last item is a text chunk: an observation is always followed by its "You have ..." tail
        return out

    def _glue(self, prev, nxt):
        g = self.glue.between.get((prev, nxt))
        if g is None:  # the reference asserts the same alternation (ctx_manager.py:266)
            raise AssertionError(f"messages do not alternate: a {nxt} message after a {prev} message")
        return g

    # ---- observation tokens ---------------------------------------------------------------------------------------
    def _glyph_ids(self, glyphs):
        return encode_many(self.tokenizer, list(glyphs))

    def mode(self, log):
        """How this tokenizer takes a grid observation: "cell" (one segment rendered on the device from the glyph table), "row"
        (a dictionary of row segments: merges cross glyphs but not row ends) or "text" (merges cross row ends as well: the
        state string is tokenised inside its chunk, once per env and turn)."""
        key = (log.glyphs, log.H, log.W)
        if key not in self._modes:
            self._modes[key] = self._validate(log.glyphs, log.H, log.W)
        return self._modes[key]

    def _obs_ids_host(self, codes_hw, glyphs, mode):
        """Token ids of one observation [H, W + 1] (host, validation only)."""
        if mode == "cell":
            ids = self._glyph_ids(glyphs)
            if any(len(x) > 4 for x in ids):
                raise _TooManyTokens()
            return [t for c in codes_hw.reshape(-1) for t in ids[min(int(c), 7)]]
        rows = ["".join(glyphs[min(int(c), 7)] for c in r) for r in codes_hw]
        return [t for r in encode_many(self.tokenizer, rows) for t in r]

    def _validate(self, glyphs, H, W):
        """Random conversations tokenised the reference's way and the segment way must agree id for id."""
        rng = np.random.default_rng(20260)
        cm = self.cm
        env_ids = list(cm.prefix_lookup.keys())
        if not env_ids:
            return "text"
        rewards = [0.9, -0.1, -0.30000000000000004, 10.9, 0, 1.0, -0.2]
        words = ["Up", "Down", "Left", "Right", "x", "Jump"]
        cases = []
        for _ in range(self.validate_samples):
            turns = int(rng.integers(1, 4))
            hist = []
            for t in range(turns + 1):
                codes = rng.integers(0, 7, (H, W + 1)).astype(np.uint8)
                if rng.random() < 0.5:  # room-like: wall border, sparse interior
                    codes = np.where(rng.random((H, W + 1)) < 0.6, 1, codes).astype(np.uint8)
                    codes[0, :] = codes[-1, :] = 0
                    codes[:, 0] = codes[:, W - 1] = 0
                codes[:, W] = SEP
                text = "".join(glyphs[c] for c in codes.reshape(-1)[:-1])
                entry = {"state": (text, codes), "actions_left": int(rng.integers(0, 11))}
                if t < turns:
                    acts = " || ".join(words[int(i)] for i in rng.integers(0, len(words), int(rng.integers(0, 4))))
                    entry["llm_response"] = (f"<think>plan {t}</think><answer>{acts}</answer>" if rng.random() < 0.8
                                             else f"no tags {acts} ")
                    entry["reward"] = rewards[int(rng.integers(0, len(rewards)))]
                hist.append(entry)
            cases.append({"env_id": env_ids[int(rng.integers(0, len(env_ids)))], "history": hist})
        for mode in ("cell", "row", "text"):
            ok = True
            for case in cases:
                for pfu in (False, True):
                    hist = case["history"][:-1] if pfu else case["history"]
                    if not hist:
                        continue
                    ref_eo = {"env_id": case["env_id"], "history": [dict(h, state=h["state"][0]) for h in hist]}
                    text = self.cm._render_text(ref_eo, pfu)
                    want = list(self.tokenizer([text], return_tensors="pt", padding=True, padding_side="left",
                                               truncation=False).input_ids[0].tolist())
                    seg_eo = {"env_id": case["env_id"],
                              "history": [dict(h, state=_HostObs(h["state"][0], h["state"][1], glyphs)) for h in hist]}
                    self._host_mode = mode
                    try:
                        got = self.segments_on_host(seg_eo, pfu)
                    except _TooManyTokens:
                        got = None
                    finally:
                        self._host_mode = "cell"
                    if not ok or got != want:
                        ok = False
                        break
                if not ok:
                    break
            if ok:
                return mode
        raise TokenizerNotDecomposable(
            f"tokenizer {getattr(self.tokenizer, 'name_or_path', '?')!r}: a grid observation with glyphs {glyphs[:7]} cannot be "
            "assembled from token segments - not per glyph, not per row, and not even message by message (the tokenizer merges "
            "across special tokens or the chat template rewrites contents).  Device trajectory assembly is refused for it; "
            "construct ContextManager(..., assembly='host') to use the reference's host path explicitly.")

    def _cell_table(self, glyphs):
        if glyphs not in self._tables:
            ids = self._glyph_ids(glyphs)
            tab = np.zeros((8, 4), dtype=np.int32)
            ln = np.zeros(8, dtype=np.int32)
            for c, x in enumerate(ids[:8]):
                tab[c, : len(x)] = x
                ln[c] = len(x)
            self._tables[glyphs] = (tab, ln)
        return self._tables[glyphs]

    def _obs_segments(self, log, slot):
        """(starts int64 [n, k], lens int32 [n, k]) of every env's observation of that turn: k = 1 segment (cell mode, rendered
        by ragen_glyph_tokens) or H row segments (row mode)."""
        key = (id(log), log.epoch, slot, self.pool.generation)
        hit = self._rendered.get(key)
        if hit is not None:
            return hit
        mode = self.mode(log)
        codes = log.host[slot]  # [n, G] unpacked codes on the host
        n, G = log.n, log.G
        if mode == "cell":
            tab, ln = self._cell_table(log.glyphs)
            worst = int(ln.max())
            stride = G * max(worst, 1)
            base = self.pool.reserve(n * stride)
            packed = log.packed[slot]
            lib = self.ctx.lib
            _lib.check(lib.ragen_glyph_tokens(self.ctx.handle, n, G, int(packed), _lib._P(log.dev[slot].data_ptr()), G,
                                              tab.ctypes.data_as(_lib._P), ln.ctypes.data_as(_lib._P), stride,
                                              _lib._P(self.pool.buf.data_ptr() + 4 * base), None, self.ctx.stream()))
            starts = (base + stride * np.arange(n, dtype=np.int64))[:, None]
            if worst == 1 and int(ln.min()) == 1:
                lens = np.full((n, 1), G, dtype=np.int32)
            else:
                lens = ln[np.minimum(codes, 7)].sum(axis=1, dtype=np.int64).astype(np.int32)[:, None]
        else:
            RW = log.W + 1
            rows = np.ascontiguousarray(np.minimum(codes, 7).reshape(n * log.H, RW))
            uniq, inv = np.unique(rows, axis=0, return_inverse=True)
            table = self._rows.setdefault((log.glyphs, self.pool.generation), {})
            keys = [u.tobytes() for u in uniq]
            new = [k for k in keys if k not in table]
            if new:
                texts = ["".join(log.glyphs[c] for c in k) for k in new]
                for k, seg in zip(new, self.pool.put(encode_many(self.tokenizer, texts))):
                    table[k] = seg
            seg = np.array([table[k] for k in keys], dtype=np.int64).reshape(-1, 2)
            starts = seg[inv.reshape(-1), 0].reshape(n, log.H)
            lens = seg[inv.reshape(-1), 1].astype(np.int32).reshape(n, log.H)
        self._rendered[key] = (starts, lens)
        if len(self._rendered) > 64:
            for k in list(self._rendered)[:32]:
                del self._rendered[k]
        return starts, lens

    # ---- assembly -------------------------------------------------------------------------------------------------
    def new_rollout(self):
        """Finished-turn segments are cached per rollout_cache dict; a new rollout starts with new histories."""
        self._row_cache.clear()

    def _tail_text(self, env_id, actions_left):
        key = (env_id, actions_left)
        t = self._tails.get(key)
        if t is None:
            think = self.cm.config.agent_proxy.enable_think
            fmt = "<think> [Your thoughts] </think> <answer> [your answer] </answer>" if think else "<answer> [your answer] </answer>"
            length = f"Max response length: {self.cm.env_config_lookup[env_id]['max_tokens']} words (tokens)."
            t = self._tails[key] = (f"You have {actions_left} actions left. Always output: {fmt} with no extra text. "
                                    f"Strictly follow this format. {length}\n")
        return t

    def row_elements(self, env_output, prepare_for_update, fills):
        """items() for the common shape of a grid-env trajectory - every entry carries a device observation, every entry but the
        last is a finished turn (response + reward), the last one is the open state (generation) or a finished turn whose
        reward message is left out (update) - with the per-turn work done ONCE: a finished turn's four chunks [observation |
        "You have N ..." | assistant message | reward message + next turn header] never change, so the resolved segments of
        a row's finished turns are kept (per rollout_cache dict) and a call only builds the chunks of the newest entries.
        Any other shape takes items()."""
        hist = env_output["history"]
        n = len(hist)
        if n == 0:
            return self.items(env_output, prepare_for_update)
        gen = self._pool.generation if self._pool is not None else 0
        k0, out = 0, None
        rc = self._row_cache.get(id(env_output))
        if rc is not None and rc[0] is env_output and rc[4] == gen:
            k = rc[1]
            if 0 < k <= n - 1 and hist[0] is rc[2] and hist[k - 1] is rc[3]:
                k0, out = k, list(rc[5])  # head + the finished turns resolved by earlier calls
        is_obs = self._is_device_obs
        for idx in range(k0, n):
            c = hist[idx]
            st = c.get("state")
            if type(st) is not ObsText or not is_obs(st):
                return self.items(env_output, prepare_for_update)
            if idx < n - 1 and not ("llm_response" in c and "reward" in c):
                return self.items(env_output, prepare_for_update)
        last = hist[-1]
        if prepare_for_update != ("llm_response" in last) or (not prepare_for_update and "reward" in last):
            return self.items(env_output, prepare_for_update)
        glue, env_id = self.glue, env_output["env_id"]
        if out is None:
            head = self._heads.get(env_id)
            if head is None:
                head = self._heads[env_id] = (glue.start + "You're a helpful assistant. ",
                                              glue.between[("system", "user")] + self.cm.prefix_lookup[env_id] + "\nTurn 1:\nState:\n")
            out = [head[0], head[1]]
        g_ua = self._g_ua
        for idx in range(k0, n - 1):
            c = hist[idx]
            a = len(out)
            out.append(c["state"])
            out.append(self._tail_text(env_id, c["actions_left"]))
            out.append(g_ua + c["llm_response"])
            out.append(self._g_au + f"Reward:\n{c['reward']}\n\nTurn {idx + 2}:\nState:\n")
            fills.append((c, idx, out, a, env_output, gen))
        out.append(last["state"])
        tail = self._tail_text(env_id, last["actions_left"])
        if prepare_for_update:
            out.append(tail)
            out.append(g_ua + last["llm_response"] + glue.end["assistant"])
        else:
            out.append(tail + self._gen_tail)
        return out

    def assemble(self, env_outputs, prepare_for_update):
        """-> (input_ids, attention_mask, position_ids) int64 [B, L] CUDA tensors, bit-identical to the reference's
        tokenizer(texts, padding=True, padding_side='left') + attention_mask.cumsum(-1)."""
        pool = self.pool
        if pool.full():
            pool.reset()
            self._rendered.clear()
            self._row_cache.clear()
        fills = []
        rows = [self.row_elements(eo, prepare_for_update, fills) for eo in env_outputs]
        memo = pool.memo
        new = {}
        for items in rows:
            for it in items:
                if type(it) is str and it not in memo and it not in new:
                    new[it] = None
        if new:
            texts = list(new)
            for t, seg in zip(texts, pool.put(encode_many(self.tokenizer, texts))):
                memo[t] = seg
        src, ln, off = [], [], [0]
        obs_lists = {}
        for items in rows:
            for j, it in enumerate(items):
                tp = type(it)
                if tp is tuple:  # segments resolved by an earlier call (entry cache): (starts, lens)
                    src.extend(it[0])
                    ln.extend(it[1])
                elif tp is str:
                    s_, n_ = memo[it]
                    src.append(s_)
                    ln.append(n_)
                else:
                    key = (id(it.log), it.slot)
                    pair = obs_lists.get(key)
                    if pair is None:
                        starts, lens = self._obs_segments(it.log, it.slot)
                        pair = obs_lists[key] = (starts.tolist(), lens.tolist())
                    src.extend(pair[0][it.index])
                    ln.extend(pair[1][it.index])
            off.append(len(src))
        for entry, idx, items, a, eo, gen in fills:  # a finished turn never changes again: keep its resolved segments with the row
            ss, ll = [], []
            for it in items[a:a + 4]:
                if type(it) is str:
                    s_, n_ = memo[it]
                    ss.append(s_)
                    ll.append(n_)
                else:
                    pair = obs_lists[(id(it.log), it.slot)]
                    ss.extend(pair[0][it.index])
                    ll.extend(pair[1][it.index])
            rc = self._row_cache.get(id(eo))
            if rc is None or rc[0] is not eo or rc[4] != gen or rc[1] > idx:
                if idx != 0:
                    continue
                rc = self._row_cache[id(eo)] = [eo, 0, entry, entry, gen, [items[0], items[1]]]
            if rc[1] == idx:
                rc[5].append((tuple(ss), tuple(ll)))
                rc[1], rc[3] = idx + 1, entry
        if len(self._row_cache) > 1 << 20:
            self._row_cache.clear()
        B, E = len(rows), len(src)
        src = np.asarray(src, dtype=np.int64)
        ln = np.asarray(ln, dtype=np.int32)
        off = np.asarray(off, dtype=np.int32)
        csum = np.concatenate([[0], np.cumsum(ln, dtype=np.int64)])
        row_len = (csum[off[1:]] - csum[off[:-1]]).astype(np.int32)
        L = int(row_len.max()) if B else 1
        row_of = np.repeat(np.arange(B), np.diff(off))
        dst = (csum[:-1] - csum[off[:-1]][row_of] + (L - row_len)[row_of]).astype(np.int32)
        # one staging buffer, one host->device copy: [src int64 E | dst int32 E | len int32 E | off int32 B+1 | row_len int32 B]
        words = 2 * E + E + E + (B + 1) + B
        host = np.empty(words + (words & 1), dtype=np.int32)
        host[: 2 * E].view(np.int64)[:] = src
        o = 2 * E
        host[o: o + E] = dst
        host[o + E: o + 2 * E] = ln
        host[o + 2 * E: o + 2 * E + B + 1] = off
        host[o + 2 * E + B + 1: o + 2 * E + 2 * B + 1] = row_len
        dev = torch.from_numpy(host).to(self.device)
        p0 = dev.data_ptr()
        ids = torch.empty((B, L), dtype=torch.int64, device=self.device)
        att = torch.empty((B, L), dtype=torch.int64, device=self.device)
        pos = torch.empty((B, L), dtype=torch.int64, device=self.device)
        P = _lib._P
        _lib.check(self.ctx.lib.ragen_traj_assemble(
            self.ctx.handle, B, L, P(pool.buf.data_ptr()), pool.used, P(p0), P(p0 + 4 * o), P(p0 + 4 * (o + E)),
            P(p0 + 4 * (o + 2 * E)), P(p0 + 4 * (o + 2 * E + B + 1)), self.pad_id, P(ids.data_ptr()), P(att.data_ptr()),
            P(pos.data_ptr()), self.ctx.stream()))
        self._staging = dev  # keep the program alive until the stream has consumed it
        return ids, att, pos


class _HostObs(str):
    """Validation stand-in for ObsText: text + its glyph codes [H, W + 1] on the host."""

    def __new__(cls, text, codes, glyphs=None):
        self = super().__new__(cls, text)
        self.codes, self.glyphs = codes, glyphs
        self.log = None
        return self


def host_obs_from_text(text, glyphs):
    """State string -> _HostObs (tests): the glyph codes are recovered from the characters."""
    rows = text.split("\n")
    W = len(rows[0])
    lut = {g: c for c, g in enumerate(glyphs[:7])}
    codes = np.full((len(rows), W + 1), SEP, dtype=np.uint8)
    for r, row in enumerate(rows):
        codes[r, :W] = [lut[ch] for ch in row]
    return _HostObs(text, codes, tuple(glyphs))
