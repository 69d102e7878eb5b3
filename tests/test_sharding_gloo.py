"""CPU tests (gloo, world_size 2) of the multi-rank host logic: group ownership, the single fixed-layout all-reduce of
rollout metrics + whitening moments (ragen_b200/sharding.py), against the single-process arithmetic of the reference
(oracle/managers.py:mean_metrics = ctx_manager.py:308-329; oracle/advantage.py:masked_whiten).  The kernels themselves
need a GPU; here the per-rank partial sums are formed with torch on the CPU purely as test input."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import golden_json
from oracle import advantage as A
from oracle import managers as M
from ragen_b200 import sharding as S


def test_group_ownership_partitions_everything():
    for n_groups, gs, world in [(8, 16, 2), (8, 16, 8), (5, 4, 2), (3, 2, 4), (65536, 16, 8)]:
        seen = []
        for r in range(world):
            g = S.owned_groups(n_groups, r, world)
            assert all(x % world == r for x in g) and g == sorted(g)
            seen += g
            if n_groups <= 8:
                ids = S.owned_env_ids(n_groups, gs, r, world)
                assert [i // gs for i in ids[::gs]] == g and len(ids) == len(g) * gs
        assert sorted(seen) == list(range(n_groups))
        sizes = [len(S.owned_groups(n_groups, r, world)) for r in range(world)]
        assert max(sizes) - min(sizes) <= 1  # weak scaling: equal shards


def test_exchange_single_process_equals_reference_arithmetic():
    final = golden_json("es_manager.json")["final"]
    ex = S.RolloutExchange(["SimpleSokoban"], {"SimpleSokoban": 8})
    vec = ex.pack(final, response_tokens=100.0, rows=8, moments=torch.tensor([5.0, 1.0, 2.0], dtype=torch.float64))
    assert vec.numel() == ex.size == 4 * 4 + 2 + 3
    got, mom = ex.unpack(ex.all_reduce(vec))
    want = M.mean_metrics(final, {"SimpleSokoban": 8})
    assert set(got) == set(want) | {"response_length"}
    for k, w in want.items():
        assert got[k] == pytest.approx(float(w), rel=1e-12), k
    assert got["response_length"] == 12.5 and mom.tolist() == [5.0, 1.0, 2.0]
    with pytest.raises(KeyError):
        ex.pack([{"metrics": {"SimpleSokoban/new_key": 1.0}}])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert S.world() == (rank, world)
        # (iii) metrics: this rank's groups of a mixed-tag rollout
        caches = _mixed_caches()
        mine = [c for c in caches if c["group_id"] % world == rank]
        env_nums = {"A": 8, "B": 4}
        ex = S.RolloutExchange(["A", "B"], env_nums)
        gen = torch.Generator().manual_seed(3)
        x = torch.randn(12, 40, generator=gen)
        mask = (torch.rand(12, 40, generator=gen) > 0.4).float()
        rows = [i for i, c in enumerate(caches) if c["group_id"] % world == rank]
        xm, mm = x[rows].double(), mask[rows].double()
        # (ii) the per-rank partial moments the moments kernel would write
        local = torch.stack([mm.sum(), (xm * mm).sum(), (xm * xm * mm).sum()])
        vec = ex.pack(mine, response_tokens=float(mm.sum()), rows=len(rows), moments=local)
        metrics, mom = ex.unpack(ex.all_reduce(vec))
        # ops.all_reduce_moments is the 24-byte variant used by compute_advantage(sharded=True)
        from ragen_b200 import ops
        mom2 = ops.all_reduce_moments(local.clone())
        # (iv) StarPO-S filter statistics: 6 groups of 2 rows, this rank's groups written at their global index
        scores = torch.randn(12, generator=torch.Generator().manual_seed(5)).view(6, 2)
        mine_g = torch.tensor([g for g in range(6) if g % world == rank])
        full = S.all_reduce_group_stats(6, mine_g, (scores[mine_g].std(-1), scores[mine_g].max(-1).values, scores[mine_g].mean(-1)))
        # (iii) tail: tag-wide normalisation statistics of this rank's rows (what ragen_group_norm_stats writes) + row count
        last = torch.randn(12, generator=torch.Generator().manual_seed(6))
        pen = torch.tensor([0.0, -0.1] * 6)
        keys = [0 if c["tag"] == "A" else 1 for c in caches]
        xs = (last + pen)[rows].double()
        tail = torch.zeros(3 * 2 + 1, dtype=torch.float64)
        for v, k in zip(xs.tolist(), [keys[i] for i in rows]):
            tail[3 * k] += 1
            tail[3 * k + 1] += v
            tail[3 * k + 2] += v * v
        tail[-1] = len(rows)
        both = ex.all_reduce(torch.cat([ex.pack(mine, float(mm.sum()), len(rows), moments=local), tail]))
        ret[rank] = (metrics, mom.tolist(), mom2.tolist(), [t.tolist() for t in full], both[ex.size:].tolist())
    finally:
        dist.destroy_process_group()


def _mixed_caches():
    rng = np.random.default_rng(9)
    out = []
    for env_id in range(12):
        tag = "A" if env_id < 8 else "B"
        m = {f"{tag}/success": float(rng.integers(0, 2)), f"{tag}/num_actions": int(rng.integers(0, 11))}
        if rng.random() < 0.8:  # an env that never executed an action reports no info keys
            m[f"{tag}/action_is_effective"] = float(rng.integers(0, 5)) / 4
            m[f"{tag}/action_is_valid"] = 1.0
        out.append({"env_id": env_id, "group_id": env_id // 2, "tag": tag, "metrics": m})
    return out


def test_world_size_2_exchange_matches_single_process():
    world, port = 2, _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        results = [ret[r] for r in range(world)]
    caches = _mixed_caches()
    want = M.mean_metrics(caches, {"A": 8, "B": 4})
    gen = torch.Generator().manual_seed(3)
    x = torch.randn(12, 40, generator=gen)
    mask = (torch.rand(12, 40, generator=gen) > 0.4).float()
    scores = torch.randn(12, generator=torch.Generator().manual_seed(5)).view(6, 2)
    last = torch.randn(12, generator=torch.Generator().manual_seed(6))
    pen = torch.tensor([0.0, -0.1] * 6)
    keys = [0 if c["tag"] == "A" else 1 for c in caches]
    want_norm = M.normalize_scores(last, pen, keys, "mean_std")
    for metrics, mom, mom2, full, tail in results:
        # every rank holds the statistics of ALL groups, exactly the single-process values -> the same top-k everywhere
        assert torch.equal(torch.tensor(full[0]), scores.std(-1)) and torch.equal(torch.tensor(full[1]), scores.max(-1).values)
        assert torch.equal(torch.tensor(full[2]), scores.mean(-1))
        # global per-tag statistics -> the reference's normalisation of the whole batch (norm_apply_kernel's arithmetic)
        assert tail[-1] == 12
        xn = (last + pen).double()
        for k in (0, 1):
            c, s1, s2 = tail[3 * k:3 * k + 3]
            mu = s1 / c
            sd = np.sqrt(max((s2 - c * mu * mu) / (c - 1), 0.0))
            idx = [i for i, kk in enumerate(keys) if kk == k]
            got = (xn[idx] - mu) / (sd + 1e-6)
            assert float((got - want_norm[idx].double()).abs().max()) <= 1e-5
    for metrics, mom, mom2, _, _ in results:
        assert set(metrics) == set(want) | {"response_length"}
        for k, w in want.items():
            assert metrics[k] == pytest.approx(float(w), rel=1e-12), k
        assert metrics["response_length"] == pytest.approx(float(mask.sum(-1).mean()), rel=1e-6)
        assert mom == pytest.approx(mom2, rel=1e-15)
        n, s1, s2 = mom
        assert n == float(mask.sum())
        # whitening from the global moments == reference masked_whiten over the whole batch (1e-5, fp32)
        mean = s1 / (n + 1e-8)
        var = (s2 - 2 * mean * s1 + mean * mean * n) / (n + 1e-8) * (n / (n - 1))
        got = (x.double() - mean) / np.sqrt(var + 1e-8)
        ref = A.masked_whiten(x, mask).double()
        assert float((got - ref).abs().max()) <= 1e-5 * max(1.0, float(ref.abs().max()))
    assert results[0][0] == results[1][0]  # every rank reports the same batch-global numbers
