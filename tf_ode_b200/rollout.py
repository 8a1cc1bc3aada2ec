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


class GatherHandle:
    """An in-flight gather of one policy tick.  `wait()` returns the [n_total, abi.ROLLOUT_WIDTH] tensor on the trainer rank
    (device tensor on GPUs; None on the other ranks) once the collective issued on the communication stream has finished."""

    def __init__(self, rows, event, stream, host=None):
        self.rows, self.event, self.stream, self.host = rows, event, stream, host

    def wait(self):
        if self.event is not None:
            self.event.synchronize()
        return self.rows


class ShardedRollout:
    """One rank's shard of a rollout job: a RoverBatch over [first, first+count) of `n_total` global environments.

    make_cfg(n_envs, env_id_offset, device) -> TfodeBatchConfig.  `batch_factory` exists for the gloo tests of the host
    logic (no GPU there); the product path always builds a RoverBatch, which raises without a B200.

    K7: per policy tick the rows of every environment (obs17, reward, done, finished-episode return; the experience
    list of src/rover_training_1.cc:212-224,243-246) travel to the trainer rank.  On GPUs nothing bounces through the host:
    `pack_rollout` writes the shard's rows into a device buffer on the compute stream, and ONE NCCL all_gather moves them on
    a second stream while the compute stream already runs the next tick's sub-steps (two row buffers alternate).
    """

    def __init__(self, make_cfg, n_total, device=0, group=None, batch_factory=None, trainer_rank=0):
        self.group, self.trainer_rank, self.n_total = group, trainer_rank, n_total
        self.rank, self.world = _group_info(group)
        self.first, self.count = shard_range(n_total, self.rank, self.world)
        self.cap = shard_range(n_total, 0, self.world)[1]   # largest shard: every rank contributes a buffer of this many rows
        cfg = make_cfg(self.count, self.first, device)
        on_gpu = batch_factory is None
        if batch_factory is None:
            from .engine import RoverBatch
            batch_factory = RoverBatch
        self.tensor_device = torch.device("cuda", device) if on_gpu else torch.device("cpu")
        if on_gpu:
            torch.cuda.set_device(device)
            self.compute = torch.cuda.current_stream()
            self.comm = torch.cuda.Stream()
            self.batch = batch_factory(cfg, stream=self.compute.cuda_stream)
        else:
            self.compute = self.comm = None
            self.batch = batch_factory(cfg)
        self._rows = [torch.zeros(self.cap, abi.ROLLOUT_WIDTH, dtype=torch.float32, device=self.tensor_device) for _ in range(2)]
        self._full = [torch.zeros(self.world * self.cap, abi.ROLLOUT_WIDTH, dtype=torch.float32, device=self.tensor_device) for _ in range(2)] \
            if self.world > 1 else [None, None]
        self._k = 0
        self._pending = [None, None]

    def load_policy(self, weights=None):
        """Trainer rank passes the new actor; every rank installs it."""
        dev = self.tensor_device if self.tensor_device.type == "cuda" else None
        w = broadcast_policy(weights if self.rank == self.trainer_rank else None, self.trainer_rank, self.group, dev)
        self.batch.load_policy(w)
        return w

    def tick_async(self, n_substeps, actions_dev=None):
        """n_substeps fused steps on this shard, then the gather of every environment's row towards the trainer rank, issued on
        the communication stream.  Returns a GatherHandle; the caller may launch the next tick before waiting on it."""
        slot = self._k & 1
        self._k += 1
        if self._pending[slot] is not None:      # the buffers of this slot are still owned by the tick before last
            self._pending[slot].wait()
        if actions_dev is not None:
            self.batch.set_actions(actions_dev)
        self.batch.step(n_substeps)
        rows = self._rows[slot]
        self.batch.pack_rollout(rows[: self.count] if self.count < self.cap else rows)
        if self.world == 1:
            ev = None
            if self.compute is not None:
                ev = torch.cuda.Event(); ev.record(self.compute)
            h = GatherHandle(rows[: self.count], ev, self.compute)
        elif self.compute is not None:
            ready = torch.cuda.Event(); ready.record(self.compute)
            with torch.cuda.stream(self.comm):
                self.comm.wait_event(ready)
                dist.all_gather_into_tensor(self._full[slot], rows, group=self.group)
                done = torch.cuda.Event(); done.record(self.comm)
            h = GatherHandle(self._trim(self._full[slot]) if self.rank == self.trainer_rank else None, done, self.comm)
        else:   # gloo (CPU tests of the host logic)
            dist.all_gather_into_tensor(self._full[slot], rows, group=self.group)
            h = GatherHandle(self._trim(self._full[slot]) if self.rank == self.trainer_rank else None, None, None)
        self._pending[slot] = h
        return h

    def _trim(self, full):
        """[world*cap, W] in rank order -> [n_total, W] in global environment order (shards may differ by one row)."""
        if self.n_total == self.world * self.cap:
            return full
        parts = [full[r * self.cap: r * self.cap + shard_range(self.n_total, r, self.world)[1]] for r in range(self.world)]
        return torch.cat(parts, 0)

    def tick(self, n_substeps):
        """Blocking form: (obs17, reward, done) of ALL environments as numpy arrays on the trainer rank, None elsewhere."""
        full = self.tick_async(n_substeps).wait()
        if full is None:
            return None
        full = full.cpu().numpy()
        return full[:, : abi.NOBS], full[:, abi.NOBS], full[:, abi.NOBS + 1] != 0

    def gather_experience(self, max_rows_per_rank):
        """The experience tuples appended on every rank since the last call (tfode_batch_drain_transitions), gathered on the trainer
        rank as one [rows, abi.TRANSITION_WIDTH] device tensor (rank order, completion order inside a rank); None elsewhere."""
        buf = torch.zeros(max_rows_per_rank, abi.TRANSITION_WIDTH, dtype=torch.float32, device=self.tensor_device)
        n = self.batch.drain_transitions(buf, max_rows_per_rank)
        if self.world == 1:
            return buf[:n]
        cnt = torch.tensor([n], dtype=torch.int64, device=self.tensor_device)
        cnts = torch.zeros(self.world, dtype=torch.int64, device=self.tensor_device)
        dist.all_gather_into_tensor(cnts, cnt, group=self.group)
        full = torch.zeros(self.world * max_rows_per_rank, abi.TRANSITION_WIDTH, dtype=torch.float32, device=self.tensor_device)
        dist.all_gather_into_tensor(full, buf, group=self.group)
        if self.rank != self.trainer_rank:
            return None
        c = cnts.tolist()
        return torch.cat([full[r * max_rows_per_rank: r * max_rows_per_rank + c[r]] for r in range(self.world)], 0)

    def close(self):
        for h in self._pending:
            if h is not None:
                h.wait()
        self.batch.close()
