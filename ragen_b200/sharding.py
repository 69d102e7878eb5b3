"""Sharding of the rollout path over ranks (one process per GPU) and its only exchange.

Env groups are independent: rank r owns the groups with `group_id % world_size == r`, keeps their state in
its own HBM for the whole rollout and no env state ever moves.  What couples ranks is a few sets of batch-global scalars:
  (ii)  masked_whiten's moments (sum m, sum x*m, sum x^2*m)            verl/verl/utils/torch_functional.py:122-151
  (iii) the rollout metrics (per-key sums / counts, response length)   ragen/llm_agent/ctx_manager.py:288, 308-329
        and, when the reward normalisation groups by tag or over the whole batch (groups then span ranks), the per-group
        {count, sum, sum of squares} + row count as a tail of the same vector                          ctx_manager.py:175-226
  (iv)  the StarPO-S filter's per-group std / max / mean at their global group index   ragen/trainer/agent_trainer.py:461-500
(ii)+(iii) travel as ONE float64 vector of fixed layout in a single all-reduce per rollout (NCCL on the GPUs, gloo in the CPU
tests; 35 doubles at C5), (iv) as 3 * env_groups doubles when the trainer filters.  Group reward normalisation with
grouping="state" (ctx_manager.py:184-185) and GRPO's per-uid statistics need nothing: whole groups live on one rank.
"""
from typing import Dict, List, Sequence

import numpy as np
import torch

INFO_KEYS = ("success", "num_actions", "action_is_effective", "action_is_valid")  # what the envs of this path report
N_MOMENTS = 3
_PER_KEY = 4  # sum, rows reporting the key, sum of non-zero values, count of non-zero values


def world(group=None):
    """(rank, world_size) of the default / given process group; (0, 1) when torch.distributed is not initialised."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def bind_host_to_gpu(device_index: int) -> Dict:
    """Best effort: run this process on the CPUs of the NUMA node the GPU hangs off, BEFORE any pinned host buffer is
    allocated (first touch then places the staging buffers in that node's memory, so the rank's PCIe traffic does not cross
    the socket interconnect - what limits host-buffer throughput when 8 ranks share one host).  Returns what it found; a host
    that exposes no NUMA topology (single node, or a VM that hides it) is left alone."""
    import os

    info = {"device": int(device_index), "numa_node": None, "cpus": None, "bound": False}
    try:
        bdf = torch.cuda.get_device_properties(device_index)
        bus = f"{getattr(bdf, 'pci_domain_id', 0):04x}:{bdf.pci_bus_id:02x}:{bdf.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as fh:
            node = int(fh.read().strip())
        info["numa_node"] = node
        if node < 0:
            return info
        with open(f"/sys/devices/system/node/node{node}/cpulist") as fh:
            cpus = set()
            for part in fh.read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        info["cpus"] = len(allowed)
        if allowed:
            os.sched_setaffinity(0, allowed)
            info["bound"] = True
    except (OSError, ValueError, AttributeError, RuntimeError):
        pass
    return info


def owned_groups(n_groups: int, rank: int, world_size: int) -> List[int]:
    """Group ids of one rank, ascending.  Round-robin keeps every tag's block spread over all ranks."""
    return list(range(rank, n_groups, world_size))


def owned_env_ids(n_groups: int, group_size: int, rank: int, world_size: int) -> List[int]:
    return [g * group_size + i for g in owned_groups(n_groups, rank, world_size) for i in range(group_size)]


def metric_slots(tags: Sequence[str], extra_keys: Sequence[str] = ()) -> List[str]:
    """The fixed key order of the exchange vector: every rank derives it from the config alone."""
    keys = list(INFO_KEYS) + [k for k in extra_keys if k not in INFO_KEYS]
    return [f"{t}/{k}" for t in dict.fromkeys(tags) for k in keys]


def all_reduce_group_stats(num_groups: int, my_groups: torch.Tensor, stats: Sequence[torch.Tensor], group=None, device=None):
    """(iv) of SURVEY.md 8e: per-group statistics of the groups this rank owns -> the same statistics of ALL groups on every
    rank.  stats: tensors [len(my_groups)]; returns float32 tensors [num_groups].  One all-reduce of len(stats) * num_groups
    doubles (a float32 is exact in a double, and every slot is written by exactly one rank)."""
    import torch.distributed as dist

    k = len(stats)
    on = device if (dist.is_available() and dist.is_initialized() and dist.get_backend(group) == "nccl") else torch.device("cpu")
    vec = torch.zeros(k * num_groups, dtype=torch.float64, device=on)
    idx = my_groups.to(on)
    for j, t in enumerate(stats):
        vec[j * num_groups + idx] = t.to(on, torch.float64)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(vec, group=group)
    return tuple(vec[j * num_groups:(j + 1) * num_groups].to(torch.float32) for j in range(k))


class RolloutExchange:
    """Packs this rank's partial sums, all-reduces them once, unpacks batch-global metrics and moments."""

    def __init__(self, tags: Sequence[str], env_nums: Dict[str, int], extra_keys: Sequence[str] = (), group=None):
        self.slots = metric_slots(tags, extra_keys)
        self.index = {k: i for i, k in enumerate(self.slots)}
        self.env_nums = dict(env_nums)  # GLOBAL number of envs per tag (ctx_manager.py:98-101)
        self.group = group
        self.size = _PER_KEY * len(self.slots) + 2 + N_MOMENTS

    # layout: [per key: sum, present, nz_sum, nz_count] ... [response_mask tokens, rows] [moments x3]
    def pack(self, env_outputs: List[Dict], response_tokens: float = 0.0, rows: int = 0, moments=None, device=None) -> torch.Tensor:
        vec = np.zeros(self.size, dtype=np.float64)
        for env_output in env_outputs:
            for key, value in env_output["metrics"].items():
                if key not in self.index:
                    raise KeyError(f"metric {key!r} has no slot in the exchange vector; pass it in extra_keys")
                o = _PER_KEY * self.index[key]
                vec[o] += value
                vec[o + 1] += 1
                if value != 0:
                    vec[o + 2] += value
                    vec[o + 3] += 1
        o = _PER_KEY * len(self.slots)
        vec[o], vec[o + 1] = response_tokens, rows
        out = torch.from_numpy(vec)
        if device is not None:
            out = out.to(device)
        if moments is not None:
            out[o + 2:] = moments.to(out.device, torch.float64)
        return out

    def all_reduce(self, vec: torch.Tensor) -> torch.Tensor:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(vec, group=self.group)
        return vec

    def unpack(self, vec: torch.Tensor):
        """-> (metrics dict as ctx_manager.py:315-328 would give for the whole batch, moments float64[3])."""
        v = vec.detach().cpu().numpy()
        metrics = {}
        for key, i in self.index.items():
            s, present, nz_sum, nz_count = v[_PER_KEY * i:_PER_KEY * i + 4]
            if present:
                metrics[key] = s / self.env_nums[key.split("/")[0]]
        for key, i in self.index.items():  # the non-zero keys follow the plain ones, as in the reference's dict order
            s, present, nz_sum, nz_count = v[_PER_KEY * i:_PER_KEY * i + 4]
            if present and nz_count:
                prefix, suffix = key.split("/", 1)
                metrics[f"{prefix}/non-zero/{suffix}"] = nz_sum / nz_count
        o = _PER_KEY * len(self.slots)
        metrics["response_length"] = float(v[o] / v[o + 1]) if v[o + 1] else 0.0
        return metrics, vec[o + 2:]
