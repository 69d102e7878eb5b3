"""GPU parity: advantage kernels (GAE, bi-level GAE, GRPO, whitening) and ContextManager tensor passes (masks /
scores, group normalisation) against the reference-made golden fixtures and the torch oracle.

Tolerances (north_star: 1e-5 relative in fp32): raw GAE / bi-level advantages and returns are compared BIT-EXACT (the
kernel keeps the reference's fp32 operation order); whitened / normalised values within 1e-5 of the tensor's scale."""
import numpy as np
import pytest
import torch

from conftest import golden_json, golden_npz
from oracle import advantage as A
from oracle import managers as M

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def close(got, want, scale=None, per_element=True):
    """north_star's bar: 1e-5 RELATIVE error in fp32 - per element where |want| > 1e-3, with two fp32 ulps of the tensor's scale
    as absolute slack (an entry next to the mean is a difference of nearly equal fp32 numbers: the reference's own value
    carries that much rounding); everywhere, 1e-5 of the tensor's scale."""
    got, want = got.detach().cpu().double(), torch.as_tensor(want).double()
    s = max(1.0, float(want.abs().max())) if scale is None else scale
    err = (got - want).abs()
    assert float(err.max()) <= RTOL * s, (float(err.max()), s)
    big = want.abs() > 1e-3
    if per_element and bool(big.any()):
        slack = 2.0 ** -22 * max(s, float(want.abs().max()))
        rel = float(((err[big] - slack).clamp(min=0) / want.abs()[big]).max())
        assert rel <= RTOL, rel


def dev(x, dtype=None):
    t = torch.as_tensor(np.asarray(x))
    return (t.to(dtype) if dtype else t).cuda()


def test_gae_and_bilevel_golden():
    from ragen_b200 import ops

    g = golden_npz("advantage.npz")
    r, v, m = dev(g["rewards"]), dev(g["values"]), dev(g["mask"])
    for tag in ("a", "b"):
        gamma, lam, hg = (float(x) for x in g[f"cfg_{tag}"])
        adv, ret, mom = ops.gae_raw(r, v, m, gamma, lam)
        raw_adv, raw_ret = A.gae_raw(torch.tensor(g["rewards"]), torch.tensor(g["values"]), gamma, lam)
        assert torch.equal(adv.cpu(), raw_adv) and torch.equal(ret.cpu(), raw_ret)  # bit-exact before whitening
        assert torch.equal(ret.cpu(), torch.tensor(g[f"gae_{tag}_ret"]))
        ops.whiten_(adv, mom)
        close(adv, g[f"gae_{tag}_adv"])
        adv, ret, mom = ops.bilevel_gae_raw(r, v, m, gamma, lam, hg)
        raw_adv, raw_ret = A.bilevel_gae_raw(torch.tensor(g["rewards"]), torch.tensor(g["values"]), torch.tensor(g["mask"]), gamma, lam, hg)
        assert torch.equal(adv.cpu(), raw_adv) and torch.equal(ret.cpu(), raw_ret)
        assert torch.equal(ret.cpu(), torch.tensor(g[f"bi_{tag}_ret"]))
        ops.whiten_(adv, mom)
        close(adv, g[f"bi_{tag}_adv"])
    # the reference's own print-only example (ragen/trainer/core_algos.py:96-102)
    r1 = dev([[0, 0, 0, 0, 1, 0, 0, 0, 0, 1]], torch.float32)
    v1 = dev([[0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9, 1.0]], torch.float32)
    m1 = torch.ones(1, 10, device="cuda")
    adv, ret, mom = ops.bilevel_gae_raw(r1, v1, m1, 1, 1, 0.95)
    assert torch.equal(ret.cpu(), torch.tensor(g["example_bi_ret"]))
    close(ops.whiten_(adv, mom), g["example_bi_adv"])
    adv, ret, mom = ops.gae_raw(r1, v1, m1, 1, 1)
    assert torch.equal(ret.cpu(), torch.tensor(g["example_gae_ret"]))
    close(ops.whiten_(adv, mom), g["example_gae_adv"])


@pytest.mark.parametrize("B,T", [(128, 2047), (5, 1), (3, 1025), (33, 3000)])
def test_gae_bilevel_c5_shape_against_oracle(B, T):
    from ragen_b200 import ops

    gen = torch.Generator().manual_seed(B * 7 + T)
    mask = torch.zeros(B, T)
    rewards = torch.zeros(B, T)
    for b in range(B):
        t = int(torch.randint(0, max(1, T // 8), (1,), generator=gen))
        while t < T:
            ln = int(torch.randint(1, 40, (1,), generator=gen))
            e = min(T, t + ln)
            mask[b, t:e] = 1
            rewards[b, e - 1] = float(torch.randint(-3, 12, (1,), generator=gen)) * 0.1 + 0.05
            t = e + int(torch.randint(1, 60, (1,), generator=gen))
    values = torch.randn(B, T, generator=gen) * mask
    for gamma, lam, hg in [(1.0, 1.0, 0.95), (0.99, 0.95, 0.9)]:
        adv, ret, mom = ops.gae_raw(rewards.cuda(), values.cuda(), mask.cuda(), gamma, lam)
        wa, wr = A.gae_raw(rewards, values, gamma, lam)
        assert torch.equal(adv.cpu(), wa) and torch.equal(ret.cpu(), wr)
        if mask.sum() > 1:
            close(ops.whiten_(adv, mom), A.masked_whiten(wa, mask))
        adv, ret, mom = ops.bilevel_gae_raw(rewards.cuda(), values.cuda(), mask.cuda(), gamma, lam, hg)
        if B <= 33:  # the oracle's Python loops are slow
            wa, wr = A.bilevel_gae_raw(rewards, values, mask, gamma, lam, hg)
            assert torch.equal(adv.cpu(), wa) and torch.equal(ret.cpu(), wr)
            if mask.sum() > 1:
                close(ops.whiten_(adv, mom), A.masked_whiten(wa, mask))
    mom = ops.masked_moments(values.cuda(), mask.cuda()).cpu()
    close(mom, torch.stack([mask.sum(), (values * mask).sum(), (values * values * mask).sum()]).double(), scale=float(mask.sum()) + 1)


def test_whiten_raises_like_reference():
    from ragen_b200 import ops

    x = torch.randn(2, 8, device="cuda")
    m = torch.zeros(2, 8, device="cuda")
    with pytest.raises(ValueError):
        ops.whiten_(x.clone(), ops.masked_moments(x, m))
    m[0, 3] = 1
    with pytest.raises(ValueError):
        ops.whiten_(x.clone(), ops.masked_moments(x, m))


def test_grpo_golden_and_random():
    from ragen_b200 import ops

    g = golden_npz("advantage.npz")
    uid = [str(u) for u in g["uid"]]
    ids = {u: i for i, u in enumerate(dict.fromkeys(uid))}
    gi = dev([ids[u] for u in uid], torch.int32)
    for key, by_std in (("grpo_std_adv", True), ("grpo_mean_adv", False)):
        adv, scores = ops.grpo_outcome(dev(g["outcome"]), dev(g["mask"]), gi, len(ids), 1e-6, by_std)
        close(adv, g[key])
    gen = torch.Generator().manual_seed(1)
    B, T = 128, 300
    rewards = torch.zeros(B, T)
    rewards[torch.arange(B), torch.randint(0, T, (B,), generator=gen)] = torch.randn(B, generator=gen)
    mask = (torch.rand(B, T, generator=gen) > 0.4).float()
    index = [i // 16 for i in range(B)]
    adv, _ = ops.grpo_outcome(rewards.cuda(), mask.cuda(), dev(index, torch.int32), 8)
    want, _ = A.grpo_outcome(rewards, mask, index)
    close(adv, want)


def test_masks_scores_golden():
    from ragen_b200 import ops

    ms = golden_json("ctx.json")["masks_scores"]
    ids = dev(ms["ids"], torch.int64)
    scores = ms["scores"]
    K = max(len(s) for s in scores)
    ts = np.zeros((len(scores), K))
    for i, s in enumerate(scores):
        ts[i, :len(s)] = s
    nt = dev([len(s) for s in scores], torch.int32)
    for uts in (0, 1):
        for erm in (1, 0):
            sc, lm, rm = ops.masks_scores(ids, dev(ts), nt, 151644, 151645, bool(uts), bool(erm), True)
            assert np.array_equal(lm.cpu().numpy(), np.array(ms[f"loss_{uts}{erm}"])), (uts, erm)
            assert np.array_equal(rm.cpu().numpy(), np.array(ms[f"resp_{uts}{erm}"])), (uts, erm)
            assert np.array_equal(sc.cpu().numpy(), np.array(ms[f"score_{uts}{erm}"], dtype=np.float32)), (uts, erm)


def test_masks_scores_c5_shape_against_oracle():
    from ragen_b200 import ops

    rng = np.random.default_rng(5)
    B, L, IMS, IME, PAD = 128, 2048, 151644, 151645, 151643
    ids = np.full((B, L), PAD, dtype=np.int64)
    scores = []
    for b in range(B):
        seq = [IMS] + list(rng.integers(0, 1000, 30)) + [IME, 198]
        sc = []
        for k in range(int(rng.integers(1, 6))):
            seq += [IMS] + list(rng.integers(0, 1000, int(rng.integers(20, 200)))) + [IME, 198]
            seq += [IMS] + list(rng.integers(0, 1000, int(rng.integers(5, 150))))
            if rng.random() > 0.1:
                seq += [IME, 198]  # sometimes the assistant turn is cut before <|im_end|>
            sc.append(float(rng.choice([-0.1, -0.2, 0.9, 10.9, 0.0])))
        seq = seq[-L:]
        ids[b, L - len(seq):] = seq
        scores.append(sc)
    K = max(len(s) for s in scores)
    ts = np.zeros((B, K))
    for i, s in enumerate(scores):
        ts[i, :len(s)] = s
    nt = dev([len(s) for s in scores], torch.int32)
    for uts in (False, True):
        for erm in (True, False):
            for roll in (True, False):
                sc, lm, rm = ops.masks_scores(dev(ids), dev(ts), nt, IMS, IME, uts, erm, roll)
                ws, wl, wr = M.masks_and_scores(torch.tensor(ids), scores, IMS, IME, uts, erm, roll)
                assert torch.equal(lm.cpu().bool(), wl) and torch.equal(rm.cpu().bool(), wr)
                assert torch.equal(sc.cpu(), ws), (uts, erm, roll)


def test_group_norm_golden_and_oracle():
    from ragen_b200 import ops

    norm = golden_json("ctx.json")["normalize"]
    last, pen = dev(norm["last"], torch.float32), dev(norm["penalty"], torch.float32)
    B = last.numel()
    keys = {"state": [i // 4 for i in range(B)], "inductive": [0 if i < 8 else 1 for i in range(B)], "batch": [0] * B}
    for grouping, gk in keys.items():
        for method in ("identity", "mean", "mean_std", "asym_clip"):
            out = ops.group_norm(last, pen, dev(gk, torch.int32), len(set(gk)), method)
            close(out, norm[f"{grouping}/{method}"])
    # as many groups as rows: only the penalty is applied; groups of one inside a larger batch
    out = ops.group_norm(last, pen, dev(list(range(B)), torch.int32), B, "mean_std")
    close(out, (last + pen).cpu())
    gk = [0, 0, 0, 1] + [2] * (B - 4)
    for method in ("mean", "mean_std", "asym_clip"):
        out = ops.group_norm(last, pen, dev(gk, torch.int32), 3, method)
        close(out, M.normalize_scores(last.cpu(), pen.cpu(), gk, method))


@pytest.mark.parametrize("B,T,gamma,lam", [(128, 2047, 1.0, 1.0), (5, 1, 1.0, 1.0), (3, 31, 0.99, 0.95), (7, 33, 0.9, 1.0), (64, 3000, 0.97, 0.9),
                                          (1, 4097, 1.0, 0.95), (4099, 257, 0.99, 0.97),
                                          # B * T % 16 == 0: the bulk-copy kernel (tiles cut in aligned flat words) - rows that start at
                                          # every alignment, rows shorter than a 16-byte group, rows of one / two / three tiles
                                          (16, 1, 1.0, 1.0), (16, 5, 0.99, 0.9), (48, 2303, 1.0, 1.0), (16, 2305, 0.99, 0.95),
                                          (32, 4097, 1.0, 0.97), (16, 6911, 0.999, 0.95), (4112, 257, 0.99, 0.97), (16, 2304, 1.0, 1.0)])
def test_gae_as_a_parallel_scan_is_within_the_bar(B, T, gamma, lam):
    """ragen_gae_scan: the GAE recurrence as warp-shuffle suffix sums.  Not bit-identical to the sequential fp32 loop by design;
    un-whitened advantages / returns and the whitened advantages must sit within north_star's 1e-5 of the reference's.
    Shapes with B * T % 16 == 0 take gae_scan_bulk_kernel, the others gae_scan_kernel."""
    from ragen_b200 import ops

    g = torch.Generator().manual_seed(B * 131 + T)
    m = (torch.rand(B, T, generator=g) > 0.3).float()
    m[:, -1] = 1
    r = torch.randn(B, T, generator=g) * 0.1 * m
    v = torch.randn(B, T, generator=g) * m
    adv, ret, mom = ops.gae_raw(r.cuda(), v.cuda(), m.cuda(), gamma, lam, exact=False)
    want_adv, want_ret = A.gae_raw(r, v, gamma, lam)
    # an advantage is a sum of up to T terms: two summation orders in fp32 differ by a few ulps of the LARGEST partial sum, so
    # the bar is 1e-5 of the tensor's scale (an entry near zero next to partial sums of size 5 cannot be had to 1e-5 of itself -
    # the reference's own value is not that close to the exact sum either)
    close(adv, want_adv, per_element=False)
    close(ret, want_ret, per_element=False)
    sel = m.bool()
    want_mom = torch.stack([m.double().sum(), want_adv.double()[sel].sum(), (want_adv.double()[sel] ** 2).sum()])
    assert float((mom.cpu() - want_mom).abs().max()) <= 1e-5 * max(1.0, float(want_mom.abs().max()))
    if int(m.sum()) > 1:
        ops.whiten_(adv, mom)
        close(adv, A.masked_whiten(want_adv, m), per_element=False)
    exact_adv, exact_ret, _ = ops.gae_raw(r.cuda(), v.cuda(), m.cuda(), gamma, lam, exact=True)
    assert torch.equal(exact_adv.cpu(), want_adv) and torch.equal(exact_ret.cpu(), want_ret)  # the exact entry point keeps the bits


def test_many_rows_use_the_lane_per_row_tile_kernel_and_stay_bit_exact():
    """B >= 4096 rows switches GAE / REINFORCE++ / ReMax to tile_scan_kernel (32 rows per warp, one lane per row, cp.async tiles);
    the recurrences must not change.  B is not a multiple of 32 and T not a multiple of 32 (partial tiles), odd T (unaligned mask rows)."""
    from ragen_b200 import ops

    for B, T in [(4100, 1500), (4097, 701), (4099, 31), (4096, 129), (4101, 64)]:
        gen = torch.Generator().manual_seed(B + T)
        mask = (torch.rand(B, T, generator=gen) < 0.6).float()
        mask[:, -1] = 1
        rewards = torch.randn(B, T, generator=gen) * (torch.rand(B, T, generator=gen) < 0.05)
        values = torch.randn(B, T, generator=gen) * mask
        for gamma, lam in [(1.0, 1.0), (0.99, 0.95)]:
            adv, ret, mom = ops.gae_raw(rewards.cuda(), values.cuda(), mask.cuda(), gamma, lam)
            wa, wr = A.gae_raw(rewards, values, gamma, lam)
            assert torch.equal(adv.cpu(), wa) and torch.equal(ret.cpu(), wr)
            close(ops.whiten_(adv, mom), A.masked_whiten(wa, mask))
            ret, _, mom = ops.discounted_returns(rewards.cuda(), mask.cuda(), gamma)
            _, wr = A.rfpp(rewards, mask, gamma)
            assert torch.equal(ret.cpu(), wr)
            close(mom.cpu(), torch.stack([mask.sum(), (wr * mask).sum(), (wr * wr * mask).sum()]).double(), scale=float(mask.sum()) + 1)
        base = torch.randn(B, generator=gen)
        adv, ret = ops.remax(rewards.cuda(), mask.cuda(), base.cuda())
        wa, wr = A.remax(rewards, base, mask)
        assert torch.equal(ret.cpu(), wr) and torch.equal(adv.cpu(), wa)


def test_many_rows_bilevel_equals_the_per_row_kernel_and_the_oracle():
    """Bi-level GAE through tile_scan_kernel (B >= 4096) against the per-row kernel on the same rows (B < 4096 per call) bit for
    bit, and a 12-row sample of it against the oracle's Python loops.  Rewards sit on and off the loss mask."""
    from ragen_b200 import ops

    for B, T in [(4100, 701), (4096, 96)]:
        gen = torch.Generator().manual_seed(B * 3 + T)
        mask = (torch.rand(B, T, generator=gen) < 0.55).float()
        rewards = (torch.randint(-3, 12, (B, T), generator=gen).float() * 0.1 + 0.05) * (torch.rand(B, T, generator=gen) < 0.01)
        last = T - 1 - mask.flip(-1).argmax(-1)  # last loss-mask position of each row carries a reward (the reference indexes past it otherwise)
        rows = torch.arange(B)
        mask[rows, last] = 1
        rewards[rows, last] = 0.7
        values = torch.randn(B, T, generator=gen) * mask
        for gamma, lam, hg in [(1.0, 1.0, 0.95), (0.99, 0.95, 0.9)]:
            adv, ret, mom = ops.bilevel_gae_raw(rewards.cuda(), values.cuda(), mask.cuda(), gamma, lam, hg)
            half = B // 2
            parts = [ops.bilevel_gae_raw(rewards[s].cuda(), values[s].cuda(), mask[s].cuda(), gamma, lam, hg) for s in (slice(0, half), slice(half, B))]
            assert torch.equal(adv.cpu(), torch.cat([p[0].cpu() for p in parts]))
            assert torch.equal(ret.cpu(), torch.cat([p[1].cpu() for p in parts]))
            close(mom.cpu(), parts[0][2].cpu() + parts[1][2].cpu(), scale=float(mask.sum()) + 1)
            pick = torch.tensor([0, 1, 7, 8, 9, 31, 32, half, B - 9, B - 8, B - 2, B - 1])
            wa, wr = A.bilevel_gae_raw(rewards[pick], values[pick], mask[pick], gamma, lam, hg)
            assert torch.equal(adv.cpu()[pick], wa) and torch.equal(ret.cpu()[pick], wr)


@pytest.mark.parametrize("B,T", [(128, 2047), (16, 1), (16, 7), (48, 2303), (16, 2305), (32, 4097), (16, 6911), (4112, 257), (7, 33)])
def test_bilevel_gae_as_block_scans_is_within_the_bar(B, T):
    """ragen_bilevel_gae_scan: the four recurrences of the bi-level estimator as compositions of maps, scanned 2304 words at a time.
    Against the sequential kernel (itself pinned to the oracle bit for bit above) on the same rows, and a sample of rows against the
    oracle's Python loops; rewards on and off the loss mask, rows with one reward and rows with dozens, (7, 33) takes the
    hand-over to ragen_bilevel_gae (B * T % 16 != 0)."""
    from ragen_b200 import ops

    gen = torch.Generator().manual_seed(B * 7 + T)
    mask = (torch.rand(B, T, generator=gen) < 0.55).float()
    rewards = (torch.randint(-3, 12, (B, T), generator=gen).float() * 0.1 + 0.05) * (torch.rand(B, T, generator=gen) < 0.02)
    last = T - 1 - mask.flip(-1).argmax(-1)
    rows = torch.arange(B)
    mask[rows, last] = 1
    rewards[rows, last] = 0.7
    values = torch.randn(B, T, generator=gen) * mask
    for gamma, lam, hg in [(1.0, 1.0, 0.95), (0.99, 0.95, 0.9)]:
        want_adv, want_ret, want_mom = ops.bilevel_gae_raw(rewards.cuda(), values.cuda(), mask.cuda(), gamma, lam, hg, exact=True)
        adv, ret, mom = ops.bilevel_gae_raw(rewards.cuda(), values.cuda(), mask.cuda(), gamma, lam, hg, exact=False)
        close(adv, want_adv.cpu(), per_element=False)  # sums of up to T terms in another order: 1e-5 of the tensor's scale
        close(ret, want_ret.cpu(), per_element=False)
        off = (mask == 0) & (rewards == 0)
        assert not adv.cpu()[off].any() and not ret.cpu()[off].any()  # nothing but zeros off the loss mask and the reward positions
        assert float((mom.cpu() - want_mom.cpu()).abs().max()) <= 1e-5 * max(1.0, float(want_mom.abs().max()))
        pick = torch.tensor(sorted({0, 1, B // 2, B - 2, B - 1} & set(range(B))))
        wa, wr = A.bilevel_gae_raw(rewards[pick], values[pick], mask[pick], gamma, lam, hg)
        close(adv.cpu()[pick], wa, per_element=False)
        close(ret.cpu()[pick], wr, per_element=False)


@pytest.mark.parametrize("B,T", [(128, 2047), (16, 1), (16, 7), (48, 2303), (16, 2305), (32, 4097), (4112, 257), (7, 33)])
def test_reinforce_pp_and_remax_as_block_scans_are_within_the_bar(B, T):
    """ragen_discounted_returns_scan / ragen_remax_scan (returns_scan_bulk_kernel) against the oracle; (7, 33) takes the hand-over
    to the sequential entry points (B * T % 16 != 0).  Masks with holes in the middle of a row (the REINFORCE++ running value is
    reset there) and rows that end in masked-out tokens."""
    from ragen_b200 import ops

    gen = torch.Generator().manual_seed(B * 11 + T)
    mask = (torch.rand(B, T, generator=gen) < 0.7).float()
    mask[: B // 2, T // 2:] = (torch.rand(B // 2, T - T // 2, generator=gen) < 0.2).float()
    rewards = torch.randn(B, T, generator=gen) * 0.1
    base = torch.randn(B, generator=gen)
    for gamma in (1.0, 0.97):
        ret, m_u8, mom = ops.discounted_returns(rewards.cuda(), mask.cuda(), gamma, exact=False)
        want_adv, want_ret = A.rfpp(rewards, mask, gamma)
        close(ret, want_ret, per_element=False)
        sel = mask.bool()
        want_mom = torch.stack([mask.double().sum(), want_ret.double()[sel].sum(), (want_ret.double()[sel] ** 2).sum()])
        assert float((mom.cpu() - want_mom).abs().max()) <= 1e-5 * max(1.0, float(want_mom.abs().max()))
        if int(mask.sum()) > 1:
            close(ops.whiten_masked(ret, m_u8, mom), want_adv, per_element=False)
    adv, ret = ops.remax(rewards.cuda(), mask.cuda(), base.cuda(), exact=False)
    want_adv, want_ret = A.remax(rewards, base, mask)
    close(ret, want_ret, per_element=False)
    close(adv, want_adv, per_element=False)
    exact_adv, exact_ret = ops.remax(rewards.cuda(), mask.cuda(), base.cuda())
    close(exact_ret, want_ret, per_element=False)


@pytest.mark.parametrize("B,T,gsz", [(1030, 301, 10), (2048, 2047, 16)])
def test_outcome_estimators_with_a_warp_per_row(B, T, gsz):
    """From 1024 rows on the row sums and the broadcast run with a warp per row (eight rows per CTA; the last CTA of 1030 rows is
    ragged): GRPO with and without std, RLOO, REINFORCE++-baseline (whitened) and the StarPO-S filter statistics vs the oracle."""
    from ragen_b200 import ops

    gen = torch.Generator().manual_seed(B + T)
    rewards = torch.zeros(B, T)
    rewards[torch.arange(B), torch.randint(0, T, (B,), generator=gen)] = torch.randn(B, generator=gen)
    mask = (torch.rand(B, T, generator=gen) > 0.4).float()
    index = [i // gsz for i in range(B)]
    ng = index[-1] + 1
    gi = dev(index, torch.int32)
    for by_std in (True, False):
        adv, scores = ops.grpo_outcome(rewards.cuda(), mask.cuda(), gi, ng, 1e-6, by_std)
        want, _ = A.grpo_outcome(rewards, mask, index, norm_by_std=by_std)
        close(adv, want)
        close(scores, rewards.sum(-1))
    adv, _, _ = ops.outcome_advantage(rewards.cuda(), mask.cuda(), gi, ng, "rloo")
    close(adv, A.rloo(rewards, mask, index))
    adv, _, mom = ops.outcome_advantage(rewards.cuda(), mask.cuda(), gi, ng, "rfpp_baseline", want_moments=True)
    close(ops.whiten_(adv, mom), A.rfpp_baseline(rewards, mask, index), per_element=False)  # masked_whiten of the tiled scores (:200)
    if B % gsz == 0:
        row, std, mx, mean = ops.group_filter_stats(rewards.cuda(), ng, gsz)
        wstd, wmx, wmean = A.group_filter_stats(rewards, ng, gsz)
        close(std, wstd); close(mx, wmx); close(mean, wmean); close(row, rewards.sum(-1))


@pytest.mark.parametrize("B,T", [(1030, 301), (2048, 2047)])
def test_kl_in_reward_at_many_rows(B, T):
    """apply_kl_penalty's token-level rewards and per-row KL at a thousand and more rows of odd length (128-bit accesses behind the
    alignment head; rows start at every alignment), all four kl_penalty forms, against the oracle.  (A warp per row, as the
    outcome kernels use from 1024 rows on, measured slower for this kernel - 57 vs 53 us at 8192 x 2047 - and was not kept.)"""
    from ragen_b200 import ops

    gen = torch.Generator().manual_seed(3 * B + T)
    scores = torch.zeros(B, T)
    scores[torch.arange(B), torch.randint(0, T, (B,), generator=gen)] = torch.randn(B, generator=gen)
    logp = -torch.rand(B, T, generator=gen)
    ref = logp + 0.1 * torch.randn(B, T, generator=gen)
    mask = (torch.rand(B, T, generator=gen) > 0.4).float()
    for kind in ("kl", "abs", "mse", "low_var_kl"):
        rew, row_kl = ops.kl_rewards(scores.cuda(), logp.cuda(), ref.cuda(), mask.cuda(), kind, 0.05)
        want, cur = A.apply_kl_penalty(scores, logp, ref, mask, kind, 0.05)
        close(rew, want)
        kld = A.kl_penalty(logp, ref, kind) * mask
        close(row_kl, kld.sum(-1) / (mask.sum(-1) + 1e-8))
