"""Multi-GPU rollout plumbing: environments sharded by contiguous ranges, one process per GPU.

The step has no collective -- one ODE world per rover in the reference, no cross-environment term anywhere.
`torch.distributed` (NCCL over NVLink on GPUs, gloo in the CPU tests) is used only where the path has a real
exchange: per policy tick the observations / rewards / done flags / finished-episode returns go to the trainer
rank, and after each trainer update the 1.09 MB fp32 actor goes back to every rank
(scripts/rover_training_1.py:101-125 does this through a SavedModel directory on disk).
"""
import numpy as np
import torch
import torch.distributed as dist

from . import abi


def shard_range(n_total, rank, world):
    """Contiguous [first, first+count) of global environment ids owned by `rank`; the first n_total % world ranks hold one more."""
    if n_total < 0 or world < 1 or not 0 <= rank < world:
        raise ValueError("shard_range(%r, %r, %r)" % (n_total, rank, world))
    base, extra = divmod(n_total, world)
    count = base + (1 if rank < extra else 0)
    first = rank * base + min(rank, extra)
    return first, count


def all_shards(n_total, world):
    return [shard_range(n_total, r, world) for r in range(world)]


def _group_info(group):
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def gather_to_trainer(local, n_total, dst=0, group=None):
    """Ragged gather of per-env rows ([n_local, ...] tensor) to rank `dst`, in global environment order.

    Returns the [n_total, ...] tensor on `dst`, None elsewhere.  Shards may differ by one row (shard_range), so the
    exchange is an all_gather over buffers padded to the largest shard (supported by both NCCL and gloo), trimmed on `dst`.
    """
    rank, world = _group_info(group)
    first, count = shard_range(n_total, rank, world)
    if local.shape[0] != count:
        raise ValueError("rank %d holds %d rows, its shard has %d" % (rank, local.shape[0], count))
    if world == 1:
        return local
    cap = shard_range(n_total, 0, world)[1]
    pad = local
    if count < cap:
        pad = torch.cat([local, local.new_zeros((cap - count,) + tuple(local.shape[1:]))], 0)
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad.contiguous(), group=group)
    if rank != dst:
        return None
    return torch.cat([b[: shard_range(n_total, r, world)[1]] for r, b in enumerate(bufs)], 0)


def broadcast_policy(weights, src=0, group=None, device=None):
    """Actor weights (flat fp32, abi.POLICY_NFLOATS) from the trainer rank to every rank; returns a numpy array for load_policy()."""
    rank, world = _group_info(group)
    if world == 1:
        w = np.ascontiguousarray(weights, dtype=np.float32)
        if w.size != abi.POLICY_NFLOATS:
            raise ValueError("actor blob has %d floats, expected %d" % (w.size, abi.POLICY_NFLOATS))
        return w
    if rank == src:
        t = torch.as_tensor(np.ascontiguousarray(weights, dtype=np.float32))
        if t.numel() != abi.POLICY_NFLOATS:
            raise ValueError("actor blob has %d floats, expected %d" % (t.numel(), abi.POLICY_NFLOATS))
    else:
        t = torch.empty(abi.POLICY_NFLOATS, dtype=torch.float32)
    if device is not None:
        t = t.to(device)
    dist.broadcast(t, src=src, group=group)
    return t.cpu().numpy()


def reduce_episode_stats(n_episodes, sum_return, sum_sq_return, n_success, group=None, device=None):
    """Episode-return statistics over all ranks (what scripts/bin/monitor-policies prints per policy): count, mean, std, success rate."""
    t = torch.tensor([float(n_episodes), float(sum_return), float(sum_sq_return), float(n_success)], dtype=torch.float64, device=device)
    rank, world = _group_info(group)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    n, s, ss, ok = t.tolist()
    if n == 0:
        return {"episodes": 0, "mean_return": float("nan"), "std_return": float("nan"), "success_rate": float("nan")}
    mean = s / n
    return {"episodes": int(n), "mean_return": mean, "std_return": max(ss / n - mean * mean, 0.0) ** 0.5, "success_rate": ok / n}


class ShardedRollout:
    """One rank's shard of a rollout job: a RoverBatch over [first, first+count) of `n_total` global environments.

    make_cfg(n_envs, env_id_offset, device) -> TfodeBatchConfig.  `batch_factory` exists for the gloo tests of the host
    logic (no GPU there); the product path always builds a RoverBatch, which raises without a B200.
    """

    def __init__(self, make_cfg, n_total, device=0, group=None, batch_factory=None, trainer_rank=0):
        self.group, self.trainer_rank, self.n_total = group, trainer_rank, n_total
        self.rank, self.world = _group_info(group)
        self.first, self.count = shard_range(n_total, self.rank, self.world)
        cfg = make_cfg(self.count, self.first, device)
        if batch_factory is None:
            from .engine import RoverBatch
            batch_factory = RoverBatch
        self.batch = batch_factory(cfg)
        self.tensor_device = torch.device("cuda", device) if torch.cuda.is_available() and batch_factory.__name__ == "RoverBatch" else torch.device("cpu")
        self._obs = np.zeros((self.count, abi.NOBS), np.float32)
        self._rew = np.zeros(self.count, np.float32)
        self._done = np.zeros(self.count, np.uint8)

    def load_policy(self, weights=None):
        """Trainer rank passes the new actor; every rank installs it."""
        dev = self.tensor_device if self.tensor_device.type == "cuda" else None
        w = broadcast_policy(weights if self.rank == self.trainer_rank else None, self.trainer_rank, self.group, dev)
        self.batch.load_policy(w)
        return w

    def tick(self, n_substeps):
        """n_substeps fused steps on this shard, then gather (obs17, reward, done) of ALL environments on the trainer rank."""
        self.batch.rollout_host(None, n_substeps, self._obs, self._rew, self._done)
        row = torch.from_numpy(np.concatenate([self._obs, self._rew[:, None], self._done[:, None].astype(np.float32)], 1)).to(self.tensor_device)
        full = gather_to_trainer(row, self.n_total, self.trainer_rank, self.group)
        if full is None:
            return None
        full = full.cpu().numpy()
        return full[:, : abi.NOBS], full[:, abi.NOBS], full[:, abi.NOBS + 1] != 0

    def close(self):
        self.batch.close()
