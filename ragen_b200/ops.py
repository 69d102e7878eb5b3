"""Torch-tensor front ends of the C-ABI entry points that are not environments: advantage pass, ContextManager
tensor passes, response parsing, Countdown scoring.  Every function takes CUDA tensors, launches on the current
stream of their device and returns CUDA tensors; nothing is computed on the host and there is no CPU fallback."""
import numpy as np
import torch

from . import _lib

_METHODS = {"identity": 0, "mean": 1, "mean_std": 2, "asym_clip": 3}


def _ctx(t):
    if not t.is_cuda:
        raise _lib.RagenError("ragen_b200 ops need CUDA tensors (no CPU fallback)")
    return _lib.Context.get(t.device.index or 0)


def _u8(mask):
    return mask if mask.dtype == torch.uint8 else mask.to(torch.uint8)


def _f32(x):
    return x.contiguous() if x.dtype == torch.float32 else x.float().contiguous()


# ---- advantage ---------------------------------------------------------------------------------------------
def gae_raw(rewards, values, mask, gamma, lam, exact=True):
    """Un-whitened GAE + masked moments.  Returns (adv, ret, moments[3] float64).  exact=True: the reference's sequential
    fp32 order, bit for bit (ragen_gae); exact=False: the parallel scan (ragen_gae_scan), equal up to rounding (<< 1e-5)."""
    ctx = _ctx(rewards)
    r, v, m = _f32(rewards), _f32(values), _u8(mask).contiguous()
    B, T = r.shape
    adv, ret = torch.empty_like(r), torch.empty_like(r)
    mom = torch.empty(3, dtype=torch.float64, device=r.device)
    fn = ctx.lib.ragen_gae if exact else ctx.lib.ragen_gae_scan
    _lib.check(fn(ctx.handle, B, T, _lib.ptr(r), _lib.ptr(v), _lib.ptr(m), float(gamma), float(lam),
                                 _lib.ptr(adv), _lib.ptr(ret), _lib.ptr(mom), ctx.stream()))
    return adv, ret, mom


def bilevel_gae_raw(rewards, values, loss_mask, gamma, lam, high_level_gamma, exact=True):
    """exact=True: the reference's fp32 operation order, bit for bit (ragen_bilevel_gae); exact=False: the block-scan kernel
    (ragen_bilevel_gae_scan), equal up to rounding and several times faster from a few thousand rows on."""
    ctx = _ctx(rewards)
    r, v, m = _f32(rewards), _f32(values), _u8(loss_mask).contiguous()
    B, T = r.shape
    adv, ret = torch.empty_like(r), torch.empty_like(r)
    mom = torch.empty(3, dtype=torch.float64, device=r.device)
    fn = ctx.lib.ragen_bilevel_gae if exact else ctx.lib.ragen_bilevel_gae_scan
    _lib.check(fn(ctx.handle, B, T, _lib.ptr(r), _lib.ptr(v), _lib.ptr(m), float(gamma), float(lam),
                  float(high_level_gamma), _lib.ptr(adv), _lib.ptr(ret), _lib.ptr(mom), ctx.stream()))
    return adv, ret, mom


def masked_moments(x, mask):
    ctx = _ctx(x)
    x, m = _f32(x), _u8(mask).contiguous()
    mom = torch.empty(3, dtype=torch.float64, device=x.device)
    _lib.check(ctx.lib.ragen_masked_moments(ctx.handle, x.numel(), _lib.ptr(x), _lib.ptr(m), _lib.ptr(mom), ctx.stream()))
    return mom


def whiten_(x, moments, check=True):
    """In-place masked_whiten from moments (torch_functional.py:139-151).  With check=True the two ValueErrors of
    masked_var are raised (one 24-byte device->host read)."""
    ctx = _ctx(x)
    assert x.dtype == torch.float32 and x.is_contiguous()
    if check:
        n = float(moments[0])
        if n == 0:
            raise ValueError("At least one element in the mask has to be 1.")
        if n == 1:
            raise ValueError("The sum of the mask is one, which can cause a division by zero.")
    _lib.check(ctx.lib.ragen_whiten(ctx.handle, x.numel(), _lib.ptr(x), _lib.ptr(moments), ctx.stream()))
    return x


def all_reduce_moments(moments, group=None):
    """The only exchange of the advantage pass when rows are sharded over ranks: 24 bytes, NCCL all-reduce."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(moments, group=group)
    return moments


def grpo_outcome(rewards, mask, group_index, n_groups, epsilon=1e-6, norm_by_std=True):
    """group_index: int32 CUDA tensor [B] of dense group ids.  Returns (adv, scores)."""
    ctx = _ctx(rewards)
    r, m = _f32(rewards), _u8(mask).contiguous()
    B, T = r.shape
    g = group_index.to(torch.int32).contiguous()
    scores = torch.empty(B, dtype=torch.float32, device=r.device)
    adv = torch.empty_like(r)
    _lib.check(ctx.lib.ragen_grpo_outcome(ctx.handle, B, T, _lib.ptr(r), _lib.ptr(m), _lib.ptr(g), int(n_groups), float(epsilon),
                                          int(bool(norm_by_std)), _lib.ptr(scores), _lib.ptr(adv), ctx.stream()))
    return adv, scores


_OUTCOME_MODES = {"mean": 0, "grpo": 1, "rloo": 2, "rfpp_baseline": 3}
_KL_MODES = {"kl": 0, "abs": 1, "mse": 2, "low_var_kl": 3}


def outcome_advantage(rewards, mask, group_index, n_groups, mode, epsilon=1e-6, want_moments=False):
    """RLOO / REINFORCE++-baseline / GRPO over per-group score statistics.  Returns (adv, scores, moments or None)."""
    ctx = _ctx(rewards)
    r, m = _f32(rewards), _u8(mask).contiguous()
    B, T = r.shape
    g = group_index.to(torch.int32).contiguous()
    scores = torch.empty(B, dtype=torch.float32, device=r.device)
    adv = torch.empty_like(r)
    mom = torch.empty(3, dtype=torch.float64, device=r.device) if want_moments else None
    _lib.check(ctx.lib.ragen_outcome_advantage(ctx.handle, B, T, _lib.ptr(r), _lib.ptr(m), _lib.ptr(g), int(n_groups),
                                               _OUTCOME_MODES[mode], float(epsilon), _lib.ptr(scores), _lib.ptr(adv), _lib.ptr(mom),
                                               ctx.stream()))
    return adv, scores, mom


def discounted_returns(rewards, mask, gamma, exact=True):
    """REINFORCE++ returns (reset after masked-out tokens) + their masked moments.  exact=False: the block-scan kernel (equal up to
    fp32 rounding)."""
    ctx = _ctx(rewards)
    r, m = _f32(rewards), _u8(mask).contiguous()
    B, T = r.shape
    ret = torch.empty_like(r)
    mom = torch.empty(3, dtype=torch.float64, device=r.device)
    fn = ctx.lib.ragen_discounted_returns if exact else ctx.lib.ragen_discounted_returns_scan
    _lib.check(fn(ctx.handle, B, T, _lib.ptr(r), _lib.ptr(m), float(gamma), _lib.ptr(ret), _lib.ptr(mom), ctx.stream()))
    return ret, m, mom


def whiten_masked(x, mask_u8, moments, check=True):
    """masked_whiten(x) * mask into a new tensor."""
    ctx = _ctx(x)
    if check:
        n = float(moments[0])
        if n == 0:
            raise ValueError("At least one element in the mask has to be 1.")
        if n == 1:
            raise ValueError("The sum of the mask is one, which can cause a division by zero.")
    out = torch.empty_like(x)
    _lib.check(ctx.lib.ragen_whiten_masked(ctx.handle, x.numel(), _lib.ptr(x), _lib.ptr(mask_u8), _lib.ptr(moments), _lib.ptr(out),
                                           ctx.stream()))
    return out


def remax(rewards, mask, baselines, exact=True):
    ctx = _ctx(rewards)
    r, m, b = _f32(rewards), _u8(mask).contiguous(), _f32(baselines)
    B, T = r.shape
    adv, ret = torch.empty_like(r), torch.empty_like(r)
    fn = ctx.lib.ragen_remax if exact else ctx.lib.ragen_remax_scan
    _lib.check(fn(ctx.handle, B, T, _lib.ptr(r), _lib.ptr(m), _lib.ptr(b), _lib.ptr(adv), _lib.ptr(ret), ctx.stream()))
    return adv, ret


def kl_rewards(scores, logp, ref_logp, mask, kl_penalty, beta):
    """Returns (token_level_rewards f32 [B, T], row_kl f32 [B])."""
    if kl_penalty not in _KL_MODES:
        raise NotImplementedError(kl_penalty)
    ctx = _ctx(scores)
    s, a, b, m = _f32(scores), _f32(logp), _f32(ref_logp), _u8(mask).contiguous()
    B, T = s.shape
    out = torch.empty_like(s)
    row = torch.empty(B, dtype=torch.float32, device=s.device)
    _lib.check(ctx.lib.ragen_kl_rewards(ctx.handle, B, T, _lib.ptr(s), _lib.ptr(a), _lib.ptr(b), _lib.ptr(m), _KL_MODES[kl_penalty],
                                        float(beta), _lib.ptr(out), _lib.ptr(row), ctx.stream()))
    return out, row


def group_filter_stats(scores, n_groups, group_size):
    """Row sums [B] and per-group (std, max, mean) [n_groups] of a score tensor whose rows are grouped consecutively."""
    ctx = _ctx(scores)
    s = _f32(scores)
    assert s.shape[0] == n_groups * group_size
    dev = s.device
    row = torch.empty(s.shape[0], dtype=torch.float32, device=dev)
    gstd, gmax, gmean = (torch.empty(n_groups, dtype=torch.float32, device=dev) for _ in range(3))
    _lib.check(ctx.lib.ragen_group_filter_stats(ctx.handle, int(n_groups), int(group_size), s.shape[1], _lib.ptr(s), _lib.ptr(row),
                                                _lib.ptr(gstd), _lib.ptr(gmax), _lib.ptr(gmean), ctx.stream()))
    return row, gstd, gmax, gmean


# ---- ContextManager tensor passes ----------------------------------------------------------------------------
def masks_scores(input_ids, turn_scores, n_turns, im_start, im_end, use_turn_scores=False, enable_response_mask=True,
                 roll_scores=True):
    """input_ids int64 [B, L]; turn_scores float64 [B, K] (K may be 0), n_turns int32 [B].
    Returns (score f32 [B, L-1], loss_mask u8 [B, L-1], response_mask u8 [B, L-1])."""
    ctx = _ctx(input_ids)
    ids = input_ids.contiguous()
    B, L = ids.shape
    K = int(turn_scores.shape[1]) if turn_scores is not None and turn_scores.numel() else 0
    ts = turn_scores.to(torch.float64).contiguous() if K else None
    nt = n_turns.to(torch.int32).contiguous()
    dev = ids.device
    lm = torch.empty((B, L - 1), dtype=torch.uint8, device=dev)
    rm = torch.empty((B, L - 1), dtype=torch.uint8, device=dev)
    sc = torch.empty((B, L - 1), dtype=torch.float32, device=dev)
    _lib.check(ctx.lib.ragen_masks_scores(ctx.handle, B, L, _lib.ptr(ids), int(im_start), int(im_end), int(bool(use_turn_scores)),
                                          int(bool(enable_response_mask)), int(bool(roll_scores)), _lib.ptr(ts), _lib.ptr(nt), K,
                                          _lib.ptr(lm), _lib.ptr(rm), _lib.ptr(sc), ctx.stream()))
    return sc, lm, rm


def group_norm(score_last, penalty, group_index, n_groups, method):
    ctx = _ctx(score_last)
    s = _f32(score_last)
    p = _f32(penalty) if penalty is not None else None
    g = group_index.to(torch.int32).contiguous()
    out = torch.empty_like(s)
    _lib.check(ctx.lib.ragen_group_norm(ctx.handle, s.numel(), _lib.ptr(s), _lib.ptr(p), _lib.ptr(g), int(n_groups),
                                        _METHODS[method], _lib.ptr(out), ctx.stream()))
    return out


def group_norm_stats(score_last, penalty, group_index, n_groups, device=None):
    """First half of group_norm for sharded rows: (x = score + penalty f32 [B], per-group {count, sum, sum of squares} f64 [n_groups, 3])."""
    dev = score_last.device if score_last.is_cuda else torch.device("cuda", torch.cuda.current_device() if device is None else int(device))
    ctx = _lib.Context.get(dev.index)
    s = _f32(score_last.to(dev))
    p = _f32(penalty.to(dev)) if penalty is not None else None
    g = group_index.to(dev, torch.int32).contiguous()
    x = torch.empty_like(s)
    gstats = torch.empty((int(n_groups), 3), dtype=torch.float64, device=dev)
    _lib.check(ctx.lib.ragen_group_norm_stats(ctx.handle, s.numel(), _lib.ptr(s), _lib.ptr(p), _lib.ptr(g), int(n_groups), _lib.ptr(x),
                                              _lib.ptr(gstats), ctx.stream()))
    return x, gstats


def group_norm_apply(x, group_index, n_groups, gstats, method):
    """Second half: normalises x in place from (all-reduced) per-group statistics."""
    ctx = _ctx(x)
    g = group_index.to(x.device, torch.int32).contiguous()
    gs = gstats.to(x.device, torch.float64).contiguous()
    _lib.check(ctx.lib.ragen_group_norm_apply(ctx.handle, x.numel(), _lib.ptr(x), _lib.ptr(g), int(n_groups), _lib.ptr(gs), _METHODS[method],
                                              ctx.stream()))
    return x


# ---- response parsing ------------------------------------------------------------------------------------------
class ActionTables:
    """Lookup tables of `config.action_lookup` dicts ({action_id: name}) laid out for ragen_parse_actions."""

    def __init__(self, lookups, device):
        names, off, codes, first = bytearray(), [0], [], [0]
        for lk in lookups:
            for code, name in lk.items():  # dict order: a later duplicate name wins on the device too
                b = name.encode()
                names += b
                off.append(len(names))
                codes.append(int(code))
            first.append(len(codes))
        self.n_tables = len(lookups)
        self.names = torch.tensor(list(names) or [0], dtype=torch.uint8, device=device)
        self.name_off = torch.tensor(off, dtype=torch.int32, device=device)
        self.codes = torch.tensor(codes or [0], dtype=torch.uint8, device=device)
        self.first = torch.tensor(first, dtype=torch.int32, device=device)


class ParseResult:
    """Device tensors written by ragen_parse_actions + host views built lazily."""

    def __init__(self, raw, offsets_np, scratch, flags, spans, act_spans, n_parsed, n_total, codes, n_valid, enable_think, sep):
        self.raw, self.offsets_np, self.scratch = raw, offsets_np, scratch
        self.flags, self.spans, self.act_spans = flags, spans, act_spans
        self.n_parsed, self.n_total, self.codes, self.n_valid = n_parsed, n_total, codes, n_valid
        self.enable_think, self.sep = enable_think, sep

    def to_host(self):
        """[(llm_response, actions, codes)] per response (ctx_manager.py:148-173, es_manager.py:230-240)."""
        from .llm_agent.parse_host import assemble_response

        if getattr(self, "slab", None) is not None:  # one device->host copy for all per-response results
            n, A = self.layout
            h = self.slab.cpu().numpy()
            words = 4 * n + 2 * A * n + n
            i32, b0 = h[: 4 * words].view(np.int32), 4 * words
            sp, asp = i32[: 4 * n].reshape(n, 4), i32[4 * n: 4 * n + 2 * A * n].reshape(n, A, 2)
            fl, kept = h[b0: b0 + n], h[b0 + n: b0 + 2 * n]
            codes, nv = h[b0 + 2 * n: b0 + (2 + A) * n].reshape(n, A), h[b0 + (2 + A) * n:]
            self.flags_host = fl
        else:
            fl = self.flags.cpu().numpy()
            sp = self.spans.cpu().numpy()
            asp = self.act_spans.cpu().numpy()
            kept = self.n_parsed.cpu().numpy()
            codes = self.codes.cpu().numpy()
            nv = self.n_valid.cpu().numpy()
        need_scratch = bool((fl & 6).any())
        scr = self.scratch.cpu().numpy().tobytes() if need_scratch else b""
        # plain Python ints from here on: per-element numpy scalar arithmetic would dominate this loop
        offs, fl_l, sp_l, asp_l = self.offsets_np.tolist(), fl.tolist(), sp.tolist(), asp.tolist()
        kept_l, nv_l, codes_l = kept.tolist(), nv.tolist(), codes.tolist()
        raw, think, sep = self.raw, self.enable_think, self.sep
        out = []
        for i in range(len(fl_l)):
            a, b = offs[i], offs[i + 1]
            resp, actions = assemble_response(raw[a:b], scr[a:b] if need_scratch else b"", fl_l[i], sp_l[i], asp_l[i], kept_l[i], think, sep)
            out.append((resp, actions, codes_l[i][:nv_l[i]]))
        return out


def stage_responses(responses, device):
    """list[str] -> (raw bytes, offsets_np int64 [n+1], text u8 CUDA blob, offsets CUDA): the one host->device copy of a turn."""
    enc = [r.encode() for r in responses]
    raw = b"".join(enc)
    offsets_np = np.zeros(len(enc) + 1, dtype=np.int64)
    np.cumsum([len(e) for e in enc], out=offsets_np[1:])
    text = torch.frombuffer(bytearray(raw or b"\0"), dtype=torch.uint8).to(device)
    return raw, offsets_np, text, torch.from_numpy(offsets_np).to(device)


def parse_actions_device(text, offsets, tables=None, table_id=None, enable_think=True, action_sep="||", max_actions=5, out=None):
    """ragen_parse_actions on a blob already in HBM.  Returns the dict of output tensors (all CUDA); pass a previous
    result as `out` to write into the same buffers again."""
    ctx = _ctx(text)
    dev, n = text.device, offsets.numel() - 1
    if out is None:
        # every per-response result lives in ONE slab [spans i32 4n | act_spans i32 2An | n_total i32 n | flags n | n_parsed n |
        # codes An | n_valid n], so the host gets a turn's parse with a single device->host copy (ParseResult.to_host)
        A = int(max_actions)
        words = 4 * n + 2 * A * n + n
        slab = torch.empty(4 * words + (3 + A) * n, dtype=torch.uint8, device=dev)
        i32 = slab[: 4 * words].view(torch.int32)
        b0 = 4 * words
        out = {"scratch": torch.empty_like(text), "slab": slab, "layout": (n, A),
               "spans": i32[: 4 * n].view(n, 4), "act_spans": i32[4 * n: 4 * n + 2 * A * n].view(n, A, 2),
               "n_total": i32[4 * n + 2 * A * n:], "flags": slab[b0: b0 + n], "n_parsed": slab[b0 + n: b0 + 2 * n],
               "codes": slab[b0 + 2 * n: b0 + (2 + A) * n].view(n, A), "n_valid": slab[b0 + (2 + A) * n:]}
    tid = None
    if table_id is not None:
        tid = torch.as_tensor(table_id, dtype=torch.int32).to(dev).contiguous()
    t = tables
    _lib.check(ctx.lib.ragen_parse_actions(
        ctx.handle, n, _lib.ptr(text), _lib.ptr(offsets), int(bool(enable_think)), action_sep.encode(), int(max_actions),
        _lib.ptr(t.names) if t else None, _lib.ptr(t.name_off) if t else None, _lib.ptr(t.codes) if t else None,
        _lib.ptr(t.first) if t else None, t.n_tables if t else 0, _lib.ptr(tid), _lib.ptr(out["scratch"]), _lib.ptr(out["flags"]),
        _lib.ptr(out["spans"]), _lib.ptr(out["act_spans"]), _lib.ptr(out["n_parsed"]), _lib.ptr(out["n_total"]), _lib.ptr(out["codes"]),
        _lib.ptr(out["n_valid"]), ctx.stream()))
    return out


def parse_actions(responses, tables=None, table_id=None, enable_think=True, action_sep="||", max_actions=5, device=0):
    """responses: list[str].  tables: ActionTables or None; table_id: int32 CUDA tensor [n] / list / None."""
    dev = torch.device("cuda", int(device)) if not isinstance(device, torch.device) else device
    raw, offsets_np, text, offsets = stage_responses(responses, dev)
    o = parse_actions_device(text, offsets, tables, table_id, enable_think, action_sep, max_actions)
    res = ParseResult(raw, offsets_np, o["scratch"], o["flags"], o["spans"], o["act_spans"], o["n_parsed"], o["n_total"], o["codes"],
                      o["n_valid"], enable_think, action_sep)
    res.text, res.offsets = text, offsets
    res.slab, res.layout = o.get("slab"), o.get("layout")
    return res


def countdown_score(text, start, length, nums, n_nums, target):
    """text uint8 CUDA blob; start int64 [n], length int32 [n]; nums int32 [n, 8]; n_nums int32 [n]; target int64 [n].
    Returns (result u8 [n] in {0, 1, 2}, unsupported u8 [n])."""
    ctx = _ctx(text)
    n = start.numel()
    res = torch.empty(n, dtype=torch.uint8, device=text.device)
    uns = torch.empty(n, dtype=torch.uint8, device=text.device)
    _lib.check(ctx.lib.ragen_countdown_score(ctx.handle, n, _lib.ptr(text), _lib.ptr(start.contiguous()), _lib.ptr(length.contiguous()),
                                             _lib.ptr(nums.contiguous()), _lib.ptr(n_nums.contiguous()), _lib.ptr(target.contiguous()),
                                             _lib.ptr(res), _lib.ptr(uns), ctx.stream()))
    return res, uns


def countdown_score_parsed(res, rows, nums, n_nums, target):
    """Scores the first parsed action of rows `rows` (int64 CUDA tensor) of a ParseResult.  The action bytes live in the
    response blob, or in the scratch blob for the rows whose answer had to be cleaned (flag bit 2)."""
    start = res.offsets[rows] + res.act_spans[rows, 0, 0].long()
    length = torch.where(res.n_parsed[rows] > 0, res.act_spans[rows, 0, 1], torch.zeros_like(res.act_spans[rows, 0, 1]))
    out, uns = countdown_score(res.text, start, length, nums, n_nums, target)
    in_scratch = (res.flags[rows] & 4) != 0
    if bool(in_scratch.any()):
        out2, uns2 = countdown_score(res.scratch, start, length, nums, n_nums, target)
        out, uns = torch.where(in_scratch, out2, out), torch.where(in_scratch, uns2, uns)
    return out, uns


_SCORE_MODES = {"metamathqa": 0, "exact_match": 1, "numeric": 2, "multiple_choice": 3}


def text_score(predictions, labels, mode, device=0):
    """Scores n (prediction, label) string pairs on the device.  mode: "metamathqa" (MetaMathQAEnv._check_answer),
    "exact_match", "numeric", "multiple_choice" (ragen/env/static/utils.py).  Returns uint8 CUDA tensor [n]:
    bit0 is_correct, bit1 is_valid, bit2 text_match, bit7 unsupported (see include/ragen_b200.h)."""
    dev = torch.device("cuda", int(device)) if not isinstance(device, torch.device) else device
    ctx = _lib.Context.get(dev.index or 0)
    assert len(predictions) == len(labels)
    _, _, p_text, p_off = stage_responses(list(predictions), dev)
    _, _, l_text, l_off = stage_responses(list(labels), dev)
    out = torch.empty(len(predictions), dtype=torch.uint8, device=dev)
    _lib.check(ctx.lib.ragen_text_score(ctx.handle, len(predictions), _SCORE_MODES[mode], _lib.ptr(p_text), _lib.ptr(p_off),
                                        _lib.ptr(l_text), _lib.ptr(l_off), _lib.ptr(out), ctx.stream()))
    return out
