"""Host-side multi-process logic (tf_ode_b200/rollout.py) on CPU: world_size 2 over gloo, no GPU, no oracle.

The batch behind ShardedRollout is a stub that fabricates rows from the GLOBAL environment id, so the test can check that
the trainer rank sees every environment exactly once and in global order, that ragged shards work, and that the actor
broadcast reaches every rank.  The device path of the same class is covered by bench.py --gpus N on the GPU box.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tf_ode_b200 import abi, rollout


def test_shard_range_partitions():
    for n in (0, 1, 7, 4096, 65536, 1000003):
        for w in (1, 2, 3, 8):
            sh = rollout.all_shards(n, w)
            assert sh[0][0] == 0 and sum(c for _, c in sh) == n
            for (f0, c0), (f1, _) in zip(sh, sh[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in sh) - min(c for _, c in sh) <= 1
    with pytest.raises(ValueError):
        rollout.shard_range(10, 2, 2)


def test_single_process_paths():
    x = torch.arange(12.0).reshape(4, 3)
    assert rollout.gather_to_trainer(x, 4) is x
    with pytest.raises(ValueError):
        rollout.gather_to_trainer(x, 5)
    w = np.arange(abi.POLICY_NFLOATS, dtype=np.float32)
    assert np.array_equal(rollout.broadcast_policy(w), w)
    with pytest.raises(ValueError):
        rollout.broadcast_policy(w[:10])
    st = rollout.reduce_episode_stats(4, 10.0, 30.0, 1)
    assert st["episodes"] == 4 and abs(st["mean_return"] - 2.5) < 1e-12 and abs(st["success_rate"] - 0.25) < 1e-12


class StubBatch:
    """Stands in for RoverBatch in the gloo test only: rows are a function of the global env id and the step count."""

    def __init__(self, cfg):
        self.first, self.n, self.t, self.policy = cfg["first"], cfg["n"], 0, None

    def load_policy(self, w):
        self.policy = np.array(w, copy=True)

    def step(self, n_substeps):
        self.t += n_substeps

    def pack_rollout(self, rows):
        gid = np.arange(self.first, self.first + self.n)
        out = np.zeros((self.n, abi.ROLLOUT_WIDTH), np.float32)
        out[:, : abi.NOBS] = gid[:, None] * 100 + np.arange(abi.NOBS)[None, :]
        out[:, abi.NOBS] = gid * 0.5 + self.t
        out[:, abi.NOBS + 1] = (gid % 3 == 0)
        out[:, 22] = gid
        rows.copy_(torch.from_numpy(out))

    def drain_transitions(self, buf, max_rows):
        n = 1 + self.first % 3   # ragged: ranks hold different numbers of experience rows
        buf[:n, 0] = torch.arange(n, dtype=torch.float32) + 1000 * self.first
        return n

    def close(self):
        pass


def _worker(rank, world, port, n_total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        r = rollout.ShardedRollout(lambda n, first, dev: {"n": n, "first": first}, n_total, batch_factory=StubBatch)
        assert (r.first, r.count) == rollout.shard_range(n_total, rank, world)
        w0 = np.linspace(-1, 1, abi.POLICY_NFLOATS).astype(np.float32)
        got = r.load_policy(w0 if rank == 0 else None)
        assert np.array_equal(got, w0) and np.array_equal(r.batch.policy, w0)
        out = r.tick(100)
        if rank == 0:
            obs, rew, done = out
            gid = np.arange(n_total)
            assert obs.shape == (n_total, abi.NOBS)
            assert np.array_equal(obs, (gid[:, None] * 100 + np.arange(abi.NOBS)[None, :]).astype(np.float32))
            assert np.array_equal(rew, (gid * 0.5 + 100).astype(np.float32))
            assert np.array_equal(done, gid % 3 == 0)
        else:
            assert out is None
        # pipelined ticks: two gathers in flight (double-buffered rows), waited for out of order
        h1 = r.tick_async(50); h2 = r.tick_async(50)
        f2, f1 = h2.wait(), h1.wait()
        if rank == 0:
            gid = np.arange(n_total)
            assert np.array_equal(f1[:, abi.NOBS].numpy(), (gid * 0.5 + 150).astype(np.float32))
            assert np.array_equal(f2[:, abi.NOBS].numpy(), (gid * 0.5 + 200).astype(np.float32))
            assert np.array_equal(f2[:, 22].numpy(), gid.astype(np.float32))   # global order, every environment once
        else:
            assert f1 is None and f2 is None
        ex = r.gather_experience(8)
        if rank == 0:
            firsts = [rollout.shard_range(n_total, k, world)[0] for k in range(world)]
            want = np.concatenate([np.arange(1 + f % 3) + 1000 * f for f in firsts]).astype(np.float32)
            assert np.array_equal(ex[:, 0].numpy(), want)
        else:
            assert ex is None
        st = rollout.reduce_episode_stats(rank + 1, 2.0 * (rank + 1), 4.0 * (rank + 1), rank)
        assert st["episodes"] == 3 and abs(st["mean_return"] - 2.0) < 1e-12 and st["std_return"] < 1e-6
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [8, 9])
def test_sharded_rollout_world2_gloo(n_total):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, n_total, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in ps]
    for p in ps:
        p.join(30)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
