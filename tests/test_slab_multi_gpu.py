"""GPU, world size 2 (skipped on a one-GPU box; run with `gpurun --gpus 2 -- python -m pytest tests/test_slab_multi_gpu.py -m gpu`):
the slab-decomposed CG (SURVEY 8e / BASELINE config 5) -- two processes, one GPU each, rows split in two slabs, the barrier of
every iteration spanning both GPUs over peer memory.  Oracle: the fp64 direct solve of the whole domain (oracle/restate_fluid.py)
and the single-GPU solve of the same right-hand side."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu


def _smooth(rng, shape, scale, passes=2):
    a = rng.randn(*shape)
    for _ in range(passes):
        a = (a + np.roll(a, 1, -1) + np.roll(a, -1, -1) + np.roll(a, 1, -2) + np.roll(a, -1, -2)) / 5
    return (a / np.abs(a).max() * scale).astype(np.float32)


def _worker(rank, world, port, cases, out_dir):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from pde_control_b200 import ops
    results = {}
    for (B, ny, nx, closed, acc, max_it) in cases:
        rng = np.random.RandomState(ny + nx)
        rhs = _smooth(rng, (B, ny, nx), 1.0)
        if closed:
            rhs -= rhs.mean((1, 2), keepdims=True)
        rows = ny // world
        mine = torch.tensor(rhs[:, rank * rows:(rank + 1) * rows]).cuda().contiguous()
        out = torch.empty_like(mine)
        it = torch.zeros(B, dtype=torch.int32, device="cuda")
        res = torch.zeros(B, device="cuda")
        comm = ops.SlabComm(B, nx)
        for _ in range(2):                       # twice: the second solve exercises the epoch / "done" hand-over
            ops.cg_poisson_slab_raw(mine, out, closed, acc, max_it, ny_global=ny, row0=rank * rows, comm=comm, iterations=it, residual=res)
        torch.cuda.synchronize()
        parts = [None] * world
        dist.all_gather_object(parts, (out.cpu().numpy(), it.cpu().numpy(), res.cpu().numpy()))
        comm.close()
        if rank == 0:
            results[(B, ny, nx, closed)] = (np.concatenate([p[0] for p in parts], 1), [p[1] for p in parts], [p[2] for p in parts], rhs)
    if rank == 0:
        np.save(os.path.join(out_dir, "results.npy"), np.array([results], dtype=object), allow_pickle=True)
    dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_gpu_slab_cg_matches_direct_solve_and_single_gpu(tmp_path):
    import torch.multiprocessing as mp
    from oracle import restate_fluid as F
    cases = [(1, 256, 256, False, 2e-6, 4000), (2, 128, 200, True, 2e-6, 4000), (1, 512, 384, False, 0.0, 50)]
    mp.spawn(_worker, args=(2, 29531, cases, str(tmp_path)), nprocs=2, join=True)
    results = np.load(os.path.join(str(tmp_path), "results.npy"), allow_pickle=True)[0]
    from pde_control_b200 import ops
    for (B, ny, nx, closed, acc, max_it) in cases:
        p, iters, resid, rhs = results[(B, ny, nx, closed)]
        assert all((i >= 0).all() for i in iters), "solve abandoned (time-out)"
        assert np.array_equal(iters[0], iters[1])                        # both ranks decide identically
        single, it1, _ = ops.pressure_solve(torch.tensor(rhs).cuda(), closed, accuracy=acc if acc > 0 else None, max_iterations=max_it)
        if acc > 0:
            bd = "closed" if closed else "open"
            direct = F.pressure_solve(torch.tensor(rhs), bd).numpy()
            got = p - p.mean((1, 2), keepdims=True) if closed else p
            ref = direct - direct.mean((1, 2), keepdims=True) if closed else direct
            assert np.abs(got - ref).max() / np.abs(ref).max() < 2e-4
            assert max(float(r.max()) for r in resid) < acc
            assert abs(int(iters[0].max()) - int(it1.max())) <= max(3, int(0.02 * int(it1.max())))
        else:
            # fixed iteration count: same iterates as the single-GPU tiled kernel up to summation order
            assert int(iters[0].max()) == max_it
            assert np.abs(p - single.cpu().numpy()).max() / np.abs(p).max() < 1e-4


def _step_worker(rank, world, port, out_dir):
    """Two projected steps + adjoint on a 256 x 192 domain split into two row slabs (nothing replicated: halo rows and the pressure halo
    over peer memory, the slab CG) against the single-GPU step of the whole domain computed by the same process."""
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world)
    from pde_control_b200 import ops, slab
    out = {}
    for closed in (False, True):
        B, ny, nx, steps, dt = 2, 256, 192, 2, 0.9
        rng = np.random.RandomState(5 + closed)
        rho = np.abs(_smooth(rng, (B, ny, nx), 1.0, passes=4))
        vy, vx = _smooth(rng, (B, ny + 1, nx), 1.5, passes=4), _smooth(rng, (B, ny, nx + 1), 1.5, passes=4)
        target = np.abs(_smooth(rng, (B, ny, nx), 1.0, passes=4))
        glob = [torch.tensor(a, device="cuda", requires_grad=True) for a in (rho, vy, vx)]
        tgt = torch.tensor(target, device="cuda")
        # single GPU, whole domain
        s = tuple(glob)
        for _ in range(steps):
            s = ops.incompressible_step(*s, dt, closed, accuracy=1e-7, max_iterations=8000)[:3]
        loss_ref = ((s[0] - tgt) ** 2).sum() / (2 * B) + 0.1 * (s[1] ** 2).sum() / (2 * B) + 0.1 * (s[2] ** 2).sum() / (2 * B)
        grads_ref = torch.autograd.grad(loss_ref, glob)
        # slabs
        st = slab.SlabStepper(B, ny, nx, halo=4)
        lay = st.layout
        loc = [t.detach().clone().requires_grad_(True) for t in st.scatter(*[g.detach() for g in glob])]
        z = tuple(loc)
        for _ in range(steps):
            z = st.step(*z, dt, closed, accuracy=1e-7, max_iterations=8000)[:3]
        # the seam row of vy is held twice: count it once in the loss (the lower slab owns it)
        zy = z[1][:, :-1] if lay.has_hi else z[1]
        loss_loc = ((z[0] - tgt[:, lay.r0:lay.r1]) ** 2).sum() / (2 * B) + 0.1 * (zy ** 2).sum() / (2 * B) + 0.1 * (z[2] ** 2).sum() / (2 * B)
        g_loc = torch.autograd.grad(loss_loc, loc)
        loss_all = loss_loc.detach().clone()
        dist.all_reduce(loss_all)
        full = [st.gather(z[0].detach()), st.gather(z[1].detach(), slab.YFACE), st.gather(z[2].detach())]
        gfull = [st.gather_grad(g_loc[0]), st.gather_grad(g_loc[1], slab.YFACE), st.gather_grad(g_loc[2])]
        torch.cuda.synchronize()
        rel = lambda a, b: float((a - b).abs().max() / b.abs().max())
        out[closed] = dict(state=[rel(a, b.detach()) for a, b in zip(full, s)], grad=[rel(a, b) for a, b in zip(gfull, grads_ref)],
                           loss=(float(loss_all), float(loss_ref)), status=int(st.ex.status.item()), violations=int(st.halo_violations.item()))
        # the same with the solver chain on the slabs (whole-domain exact solve as the guess, slab CG on the correction system)
        loc2 = [t.detach().clone().requires_grad_(True) for t in st.scatter(*[g.detach() for g in glob])]
        z, its = tuple(loc2), []
        for _ in range(steps):
            o = st.step(*z, dt, closed, accuracy=1e-7, max_iterations=8000, chained=True)
            z = o[:3]
            its.append(int(o[4].max()))
        zy = z[1][:, :-1] if lay.has_hi else z[1]
        loss2 = ((z[0] - tgt[:, lay.r0:lay.r1]) ** 2).sum() / (2 * B) + 0.1 * (zy ** 2).sum() / (2 * B) + 0.1 * (z[2] ** 2).sum() / (2 * B)
        g2 = torch.autograd.grad(loss2, loc2)
        full2 = [st.gather(z[0].detach()), st.gather(z[1].detach(), slab.YFACE), st.gather(z[2].detach())]
        gfull2 = [st.gather_grad(g2[0]), st.gather_grad(g2[1], slab.YFACE), st.gather_grad(g2[2])]
        torch.cuda.synchronize()
        # its reference: the single-GPU trajectory with the EXACT solver.  (On the closed domain the fp32 CG cannot reach 1e-7 on the rough
        # right-hand sides of the adjoint: it runs to max_iterations and ends 1e-3 off -- in the single-GPU step and in the plain slab CG
        # alike, which is why those two agree with each other; the chain is compared with the exact answer instead.)
        glob2 = [g.detach().clone().requires_grad_(True) for g in glob]
        sd = tuple(glob2)
        for _ in range(steps):
            sd = ops.incompressible_step(*sd, dt, closed, direct=True)[:3]
        loss_d = ((sd[0] - tgt) ** 2).sum() / (2 * B) + 0.1 * (sd[1] ** 2).sum() / (2 * B) + 0.1 * (sd[2] ** 2).sum() / (2 * B)
        grads_d = torch.autograd.grad(loss_d, glob2)
        out[closed].update(chained_state=[rel(a, b.detach()) for a, b in zip(full2, sd)], chained_grad=[rel(a, b) for a, b in zip(gfull2, grads_d)],
                           cg_vs_exact_grad=[rel(a, b) for a, b in zip(grads_ref, grads_d)],
                           chained_iterations=its, status2=int(st.ex.status.item()))
        st.close()
    if rank == 0:
        np.save(os.path.join(out_dir, "step.npy"), np.array([out], dtype=object), allow_pickle=True)
    dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_gpu_slab_step_and_adjoint_match_the_single_gpu_step(tmp_path):
    import torch.multiprocessing as mp
    mp.spawn(_step_worker, args=(2, 29533, str(tmp_path)), nprocs=2, join=True)
    out = np.load(os.path.join(str(tmp_path), "step.npy"), allow_pickle=True)[0]
    for closed in (False, True):
        r = out[closed]
        assert r["status"] == 0 and r["violations"] == 0, r
        assert abs(r["loss"][0] - r["loss"][1]) < 1e-4 * abs(r["loss"][1]), r
        assert max(r["state"]) < 1e-4, r          # sample coordinates of a haloed slab round differently from the global grid's: ~1e-6
        assert max(r["grad"]) < 1e-3, r           # both paths solve to 1e-7: at 1e-6 two CG implementations differ by 1.6e-3 in the gradient
        # solver chain on the slabs: same results, a handful of correction iterations instead of hundreds
        assert r["status2"] == 0 and max(r["chained_state"]) < 1e-4 and max(r["chained_grad"]) < 1e-3, r
        assert max(r["chained_iterations"]) < 200, r
