"""world_size-2 gloo tests (CPU) of the multi-GPU host plan: view sharding, NCCL-id broadcast plumbing and the
exactness of 'local queries x rank-ordered all-gathered K|V' for global attention (oracle as the checker)."""
import importlib
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

par = importlib.import_module("hunyuanworld-mirror_b200.parallel")
from oracle import aggregator_oracle as O  # noqa: E402


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. id broadcast
        uid = bytes(range(128)) if rank == 0 else None
        got = par.broadcast_unique_id(uid)
        assert got == bytes(range(128))
        # 2. sharded global attention == unsharded
        g = torch.Generator().manual_seed(0)
        n, P = 4, 23
        qf, kf, vf = (torch.randn(1, 2, n * P, 64, generator=g) for _ in range(3))
        full = O.sdpa(qf, kf, vf)
        imgs = torch.arange(n).reshape(1, n, 1, 1, 1).float().expand(1, n, 3, 2, 2)
        loc, _, first = par.shard_inputs(imgs, None, rank, world)
        assert loc.shape[1] == n // world and int(loc[0, 0, 0, 0, 0]) == first
        lo, hi = first * P, (first + n // world) * P
        kv_local = torch.cat([kf[0, :, lo:hi], vf[0, :, lo:hi]], dim=-1).permute(1, 0, 2).reshape(hi - lo, 2 * 2 * 64).contiguous()
        parts = [torch.empty_like(kv_local) for _ in range(world)]
        dist.all_gather(parts, kv_local)
        kv = torch.cat(parts, dim=0)  # rank-major rows, exactly ncclAllGather's in-place layout
        assert par.kv_row_range(hi - lo, rank) == (lo, hi)
        kv = kv.reshape(n * P, 2, 2 * 64).permute(1, 0, 2)
        out = O.sdpa(qf[:, :, lo:hi], kv[None, :, :, :64], kv[None, :, :, 64:])
        torch.testing.assert_close(out, full[:, :, lo:hi], atol=1e-5, rtol=1e-5)
        # 3. output gather along the view axis
        t = torch.full((1, n // world, 2, 4), float(rank))
        cat = par.gather_outputs([t])[0]
        assert cat.shape == (1, n, 2, 4) and cat[0, :, 0, 0].tolist() == [0.0, 0.0, 1.0, 1.0]
        # 4. Gaussian branch under view sharding: splats and the depth confidence gathered in rank order are the scene in
        #    view order, so confidence filter + voxel merge of the gathered rows == the single-process result, bit for bit
        import numpy as np

        from oracle import gs_oracle as G

        syn = G.synthetic_splats(4000)
        conf = (1.0 + np.random.default_rng(3).integers(0, 9, size=4000) * 0.5).astype(np.float32)  # ties on purpose
        full_sp = G.prune_gs(G.apply_confidence_filter(syn, conf, 40.0), voxel_size=0.01)
        per = 4000 // world
        mine = {k: torch.from_numpy(v[rank * per:(rank + 1) * per].copy()) for k, v in dict(syn, conf=conf).items()}
        gathered = {}
        for k, t in mine.items():
            parts = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(parts, t)
            gathered[k] = torch.cat(parts, dim=0).numpy()
        gconf = gathered.pop("conf")
        shard_sp = G.prune_gs(G.apply_confidence_filter(gathered, gconf, 40.0), voxel_size=0.01)
        for k in full_sp:
            assert np.array_equal(full_sp[k], shard_sp[k]), k
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_two_rank_plan_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = dict(q.get(timeout=120) for _ in procs)
    [p.join(30) for p in procs]
    assert res == {0: "ok", 1: "ok"}, res


def test_view_shard_rules():
    assert par.view_shard(32, 3, 8) == (12, 4)
    with pytest.raises(ValueError):
        par.view_shard(10, 0, 4)
    with pytest.raises(ValueError):
        par.view_shard(8, 8, 8)
