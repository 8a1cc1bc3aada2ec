"""K7 on real GPUs: world-size-2 NCCL rollout (tf_ode_b200/rollout.py) equals the single-GPU rollout row for row.
Needs 2 visible GPUs (`-m gpu` on a 2-GPU box); skipped otherwise."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402


def _cfg(n, first, dev):
    import bench
    return bench.make_cfg("config3", n, env_id_offset=first, device=dev)


def _worker(rank, world, port, n_total, ticks, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from tf_ode_b200 import abi, glorot_policy
        from tf_ode_b200.rollout import ShardedRollout
        r = ShardedRollout(_cfg, n_total, device=rank)
        r.load_policy(glorot_policy() if rank == 0 else None)     # NCCL broadcast of the actor
        out, hs = [], []
        for _ in range(ticks):                                     # pipelined: the gather of tick k overlaps the sub-steps of tick k+1
            hs.append(r.tick_async(500))
            if len(hs) > 1:
                f = hs.pop(0).wait()
                if rank == 0:
                    out.append(f.cpu().numpy().copy())
        f = hs.pop(0).wait()
        if rank == 0:
            out.append(f.cpu().numpy().copy())
        ex = r.gather_experience(4 * (n_total // world))
        if rank == 0:
            q.put((rank, np.stack(out), ex.cpu().numpy()))
        else:
            assert f is None and ex is None
            q.put((rank, None, None))
        r.close()
    except Exception as e:  # noqa: BLE001
        q.put((rank, repr(e), None))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_world2_nccl_rollout_equals_single_gpu_rows():
    from tf_ode_b200 import RoverBatch, abi, glorot_policy
    n_total, ticks = 1024, 4
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, n_total, ticks, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = {}
    for _ in ps:
        rank, a, b = q.get(timeout=600)
        assert not isinstance(a, str), a
        res[rank] = (a, b)
    for p in ps:
        p.join(60)
    rows2, ex2 = res[0]
    # the same job on one GPU
    g = RoverBatch(_cfg(n_total, 0, 0)); g.load_policy(glorot_policy())
    rows = torch.zeros(n_total, abi.ROLLOUT_WIDTH, device="cuda")
    for k in range(ticks):
        g.step(500); g.pack_rollout(rows)
        assert np.array_equal(rows.cpu().numpy(), rows2[k]), "tick %d" % k
    buf = torch.zeros(4 * n_total, abi.TRANSITION_WIDTH, device="cuda")
    n = g.drain_transitions(buf, buf.shape[0])
    ex1 = buf[:n].cpu().numpy()
    assert n == ex2.shape[0] and n > 100
    for e in range(0, n_total, 61):   # completion order differs between the two runs: compare per environment
        assert np.array_equal(ex1[ex1[:, 0] == e], ex2[ex2[:, 0] == e]), e
    assert np.array_equal(rows2[-1][:, 22], np.arange(n_total, dtype=np.float32))
    g.close()
