"""Run under torchrun with N >= 2 GPUs (tests/test_gpu_multi.py launches it): every rank rolls out ITS env groups of a
golden case through LLMAgentProxy(rank=, world_size=), then compute_advantage(sharded=True); rank 0 gathers the per-env
results and checks the union against the single-process reference goldens (tests/golden/rollouts.json) and the oracle."""
import json
import os
import sys

import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
from fake_tokenizer import FakeQwenTokenizer, ScriptedWg  # noqa: E402
from oracle import advantage as A  # noqa: E402
from oracle import cfg as C  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from ragen_b200.llm_agent.agent_proxy import LLMAgentProxy
    from ragen_b200.llm_agent.ctx_manager import DataProto
    from ragen_b200.trainer import AdvantageEstimator, compute_advantage
    from ragen_b200.trainer.agent_trainer import filter_rollout

    with open(os.path.join(HERE, "golden", "rollouts.json")) as fh:
        cases = json.load(fh)
    for case in cases:
        cfg = C.base_config(tags=tuple(case["tags"]), n_groups=tuple(case["n_groups"]), group_size=case["group_size"], **case["ap"])
        tok = FakeQwenTokenizer()
        proxy = LLMAgentProxy(cfg, ScriptedWg(tok, DataProto), tok, device=local, rank=rank, world_size=world)
        dp = proxy.rollout(DataProto(batch=None, non_tensor_batch=None, meta_info={}), val=False, seed=case["seed"])
        batch = dp.batch
        env_ids = [int(e) for e in dp.non_tensor_batch["env_ids"]]
        assert all((e // case["group_size"]) % world == rank for e in env_ids)
        # advantage on this rank's rows with batch-global whitening
        B, T = batch["rm_scores"].shape  # (0, 0) on a rank that owns no group: it still joins both all-reduces
        batch["response_mask"] = batch["loss_mask"].float()
        batch["token_level_rewards"] = batch["rm_scores"].clone()
        gen = torch.Generator().manual_seed(1000 + rank)
        batch["values"] = torch.randn(B, T, generator=gen) * batch["response_mask"]
        compute_advantage(dp, AdvantageEstimator.GAE, gamma=1.0, lam=1.0, device=local, sharded=True)
        # StarPO-S filter on the sharded batch (agent_trainer.py:461-500): top half of ALL groups by in-group std
        import copy as _copy

        n_groups_all = int(sum(case["n_groups"]))
        fdp = DataProto(batch=type(batch)({k: v for k, v in batch.items()}), non_tensor_batch=dict(dp.non_tensor_batch.items()), meta_info={})
        if hasattr(batch, "batch_size"):
            fdp.batch.batch_size = batch.batch_size
        row_sums = batch["original_rm_scores"].sum(-1).tolist()
        fdp, fmetrics = filter_rollout(fdp, n_groups_all, case["group_size"], 0.5, "std", device=local, sharded=True)
        kept_ids = [int(e) for e in fdp.non_tensor_batch["env_ids"]]
        mine = {"env_ids": env_ids, "row_sums": row_sums, "kept_ids": kept_ids,
                "filter_metrics": {k: float(v) for k, v in fmetrics.items()}, "rm_last": batch["rm_scores"][:, -1:].flatten().tolist(), "loss_mask_sum": batch["loss_mask"].sum(-1).tolist(),
                "metrics": {k: float(v) for k, v in dp.meta_info["metrics"].items()},
                "r": batch["token_level_rewards"], "v": batch["values"], "m": batch["response_mask"], "adv": batch["advantages"],
                "ret": batch["returns"]}
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        if rank == 0:
            order = {e: (r, i) for r, g in enumerate(gathered) for i, e in enumerate(g["env_ids"])}
            assert sorted(order) == list(range(len(case["rm_last"])))
            for e in range(len(order)):
                r, i = order[e]
                assert abs(gathered[r]["rm_last"][i] - case["rm_last"][e]) <= 1e-5 * max(1.0, abs(case["rm_last"][e])), (case["name"], e)
                assert gathered[r]["loss_mask_sum"][i] == case["loss_mask_sum"][e], (case["name"], e)
            # the filter: same choice as the reference's arithmetic on the whole batch, same metrics on every rank
            gs_ = case["group_size"]
            sums = torch.tensor([gathered[order[e][0]]["row_sums"][order[e][1]] for e in range(len(order))]).view(-1, gs_)
            std_all = sums.std(dim=-1)
            top = std_all.topk(int(0.5 * sums.shape[0])).indices.tolist()
            want_kept = sorted(g_ * gs_ + i for g_ in top for i in range(gs_))
            assert sorted(e for g in gathered for e in g["kept_ids"]) == want_kept, case["name"]
            for g in gathered:
                assert abs(g["filter_metrics"]["rollout/in_group_std"] - float(std_all.mean())) <= 1e-5 * max(1.0, float(std_all.mean()))
                assert g["filter_metrics"] == gathered[0]["filter_metrics"]
            for g in gathered:
                assert set(g["metrics"]) == set(case["metrics"])
                for k, w in case["metrics"].items():
                    assert abs(g["metrics"][k] - w) <= 1e-6 * max(1.0, abs(w)), (case["name"], k, g["metrics"][k], w)
            # whitening: pad every shard on the left to a common T (GAE runs right-to-left and pads are masked out)
            gathered = [g for g in gathered if len(g["env_ids"])]
            Tm = max(g["r"].shape[1] for g in gathered)
            pad = lambda t: torch.nn.functional.pad(t, (Tm - t.shape[1], 0))
            r_, v_, m_ = (torch.cat([pad(g[k]) for g in gathered]) for k in ("r", "v", "m"))
            wa, wr = A.gae(r_, v_, m_, 1.0, 1.0)
            adv, ret = (torch.cat([pad(g[k]) for g in gathered]) for k in ("adv", "ret"))
            keep = m_.bool()
            assert torch.equal(ret[keep], wr[keep])
            err = float((adv[keep] - wa[keep]).abs().max())
            assert err <= 1e-5 * max(1.0, float(wa[keep].abs().max())), err
            print(f"dist check ok: {case['name']} world={world} rows={[len(g['env_ids']) for g in gathered]}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
