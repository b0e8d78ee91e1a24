"""CPU, world_size 2, gloo: the data-parallel path of ControlTraining (batch sharded by rank, one flat all-reduce of the
network gradients per step) and the shard arithmetic bench.py uses.  The physics needs no collective."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from pde_control_b200 import control as C
    rng = np.random.RandomState(0)
    ends = rng.randn(32, 1, 1).astype(np.float32)
    frames = np.stack([ends * t / 2 for t in range(3)], 1)
    torch.manual_seed(100 + rank)                       # different initial weights per rank: prepare() must broadcast rank 0's
    pde = C.IncrementPDE(learnable=True)
    app = C.ControlTraining(2, pde, {"scalar": frames}, None, range(32), obs_loss_frames=(1,), trainable_networks=("OP2",),
                            sequence_class=C.SkipSequence, batch_size=8, device="cpu", learning_rate=0.05)
    assert app.world_size == world and app._train.size == 16          # shard of the simulations
    app.prepare()
    w0 = float(pde.networks["OP2"].weight)
    for _ in range(5):
        app.step()
    np.save(os.path.join(out_dir, "w_%d.npy" % rank), np.array([w0, float(pde.networks["OP2"].weight)]))
    dist.destroy_process_group()


def _single(frames, w0):
    from pde_control_b200 import control as C
    pde = C.IncrementPDE(learnable=True)
    app = C.ControlTraining(2, pde, {"scalar": frames}, None, range(32), obs_loss_frames=(1,), trainable_networks=("OP2",),
                            sequence_class=C.SkipSequence, batch_size=16, device="cpu", learning_rate=0.05)
    app.prepare()
    with torch.no_grad():
        pde.networks["OP2"].weight.fill_(w0)
    return app, pde


def test_two_rank_data_parallel_training_matches_single_process(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a, b = np.load(tmp_path / "w_0.npy"), np.load(tmp_path / "w_1.npy")
    assert a[0] == b[0]                                   # broadcast initial weights
    assert a[1] == b[1]                                   # identical after averaged-gradient steps
    # equals one process training on the union batch (rank 0 sims 0-7 + rank 1 sims 16-23 each step ...):
    rng = np.random.RandomState(0)
    ends = rng.randn(32, 1, 1).astype(np.float32)
    frames = np.stack([ends * t / 2 for t in range(3)], 1)
    order = np.concatenate([np.arange(0, 16).reshape(2, 8), np.arange(16, 32).reshape(2, 8)], 1)   # per step: 8 of each shard
    app, pde = _single(frames[np.concatenate([order[0], order[1]])], float(a[0]))
    # the single process sees [step-0 union batch, step-1 union batch] alternately with batch 16
    for _ in range(5):
        app.step()
    assert abs(float(pde.networks["OP2"].weight) - a[1]) < 1e-5


def _slab_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import slab_cg as S
    rng = np.random.RandomState(5)
    for name, (B, ny, nx, closed) in {"open": (2, 24, 20, False), "closed": (1, 16, 12, True)}.items():
        rhs = rng.randn(B, ny, nx)
        rhs[0] *= 1e-3                                   # elements converge at different iterations (per-element stop rule)
        if closed:
            rhs -= rhs.mean((1, 2), keepdims=True)
        rows = ny // world
        x, iters, exchanges = S.slab_cg(torch.tensor(rhs[:, rank * rows:(rank + 1) * rows]), rank * rows, ny, closed, 1e-9, 500)
        np.save(os.path.join(out_dir, "%s_%d.npy" % (name, rank)), x.numpy())
        if rank == 0:
            np.save(os.path.join(out_dir, "%s_meta.npy" % name), np.array(iters + [exchanges]))
    dist.destroy_process_group()


def test_two_rank_slab_cg_single_exchange_formulation_matches_plain_cg(tmp_path):
    """The algorithm of csrc/cg_tiled.cu (slabs, ONE exchange per iteration, beta from pre-update dot products, neighbour-formed
    halo directions) restated on CPU over gloo == the plain restatement of blas.py:5-64 on the whole domain (fp64)."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from oracle import restate_fluid as F
    port = _free_port()
    mp.spawn(_slab_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    rng = np.random.RandomState(5)
    for name, (B, ny, nx, closed) in {"open": (2, 24, 20, False), "closed": (1, 16, 12, True)}.items():
        rhs = rng.randn(B, ny, nx)
        rhs[0] *= 1e-3
        if closed:
            rhs -= rhs.mean((1, 2), keepdims=True)
        got = np.concatenate([np.load(tmp_path / ("%s_%d.npy" % (name, r))) for r in range(2)], 1)
        meta = np.load(tmp_path / ("%s_meta.npy" % name))
        iters, exchanges = meta[:-1], meta[-1]
        for b in range(B):                               # the reference stops on the whole batch: compare element by element
            ref, k = F.conjugate_gradient(torch.tensor(rhs[b:b + 1]), "closed" if closed else "open", accuracy=1e-9, max_iterations=500)
            assert abs(int(iters[b]) - k) <= 1
            assert np.abs(got[b] - ref[0].numpy()).max() < 1e-7 * max(1.0, np.abs(ref.numpy()).max())
        assert exchanges == max(iters) + 2               # one exchange per iteration (+ the set-up pass and the final test)


def _halo_worker(rank, world, port, out_dir):
    """SlabLayout (pde_control_b200/slab.py: the index arithmetic of the peer-memory halo exchange) driven by a NumPy transport over gloo:
    the haloed arrays must be the global array's rows, and the reverse exchange must be the exact transpose of the forward one."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from pde_control_b200.slab import SlabLayout, CENTRED, YFACE

    def exchange(lay, local, kind, H, reverse=False, g=None):
        """forward: local slab -> haloed array.  reverse: gradient of the haloed array g -> gradient of the slab (accumulated)."""
        hlo, hhi = lay.halo_widths(H)
        plan = lay.plan(kind, H)
        if not reverse:
            out = np.zeros((hlo + local.shape[0] + hhi,) + local.shape[1:], local.dtype)
            out[hlo:hlo + local.shape[0]] = local
            reqs = [dist.isend(torch.tensor(local[e["send_first"]:e["send_first"] + e["n"]].copy()), e["peer"]) for e in plan]
            for e in plan:
                buf = torch.empty((e["n"],) + local.shape[1:], dtype=torch.float64)
                dist.recv(buf, e["peer"])
                out[e["recv_first"]:e["recv_first"] + e["n"]] = buf.numpy()
            for r in reqs:
                r.wait()
            return out
        n_local = lay.local_rows(kind)
        back = g[hlo:hlo + n_local].copy()
        reqs = [dist.isend(torch.tensor(g[e["recv_first"]:e["recv_first"] + e["n"]].copy()), e["peer"]) for e in plan]
        for e in plan:
            buf = torch.empty((e["n"],) + g.shape[1:], dtype=torch.float64)
            dist.recv(buf, e["peer"])
            back[e["send_first"]:e["send_first"] + e["n"]] += buf.numpy()
        for r in reqs:
            r.wait()
        return back

    ny, nx, H = 24, 5, 3
    lay = SlabLayout(ny, world, rank)
    ok = True
    dots = []
    for kind, nrows in ((CENTRED, ny), (YFACE, ny + 1)):
        glob = np.random.RandomState(nrows).randn(nrows, nx)        # the same global array on every rank
        rng = np.random.RandomState(100 + rank)
        local = glob[lay.r0:lay.r0 + lay.local_rows(kind)]
        haloed = exchange(lay, local, kind, H)
        hlo, hhi = lay.halo_widths(H)
        ok = ok and np.array_equal(haloed, glob[lay.r0 - hlo:lay.r0 + lay.local_rows(kind) + hhi])
        # <E x, y> summed over the ranks == <x, E^T y> summed over the ranks
        y = rng.randn(*haloed.shape)
        x = rng.randn(*local.shape)
        lhs = float((exchange(lay, x, kind, H) * y).sum())
        rhs = float((x * exchange(lay, None, kind, H, reverse=True, g=y)).sum())
        t = torch.tensor([lhs, rhs], dtype=torch.float64)
        dist.all_reduce(t)
        dots.append(t.tolist())
    np.save(os.path.join(out_dir, "halo_%d.npy" % rank), np.array([float(ok)] + [v for d in dots for v in d]))
    dist.destroy_process_group()


def test_slab_layout_halo_exchange_and_its_transpose_over_gloo(tmp_path):
    for world in (2, 3):
        port = _free_port()
        mp.spawn(_halo_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
        for rank in range(world):
            r = np.load(tmp_path / ("halo_%d.npy" % rank))
            assert r[0] == 1.0, "haloed array != rows of the global array"
            assert abs(r[1] - r[2]) < 1e-9 * max(1.0, abs(r[1])) and abs(r[3] - r[4]) < 1e-9 * max(1.0, abs(r[3]))
