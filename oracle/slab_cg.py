"""TEST INFRASTRUCTURE (never imported by pde_control_b200): CPU restatement of the slab-decomposed, single-exchange CG that
pde_control_b200/csrc/cg_tiled.cu runs on the GPUs, over a torch.distributed group (gloo in the tests).

It follows conjugate_gradient (PhiFlow/phi/math/blas.py:5-64) on the operator of pressuresolver/sparse.py:114-221, regrouped so
that ONE exchange per iteration carries everything: with q = A m all dot products are taken before the update,
    a = (m.r)/(m.q),  beta = ((r.q) - a (q.q))/(m.q)      [= (r'.q)/(m.q), r' = r - a q]
and the neighbours' new search direction on the halo rows is formed locally from their published rows of (r, q, m).
The tests check it against oracle.restate_fluid.conjugate_gradient (the plain restatement) on the whole domain."""
import torch
import torch.distributed as dist


def _matvec(m, up, down, row0, ny_global, closed):
    """5-point operator on a slab m[B, rows, nx] with halo rows `up` / `down` (zeros at the domain boundary)."""
    B, rows, nx = m.shape
    ext = torch.cat([up[:, None], m, down[:, None]], 1)
    s = ext[:, :-2] + ext[:, 2:]
    s[:, :, 1:] += m[:, :, :-1]
    s[:, :, :-1] += m[:, :, 1:]
    diag = torch.full((rows, nx), -4.0, dtype=m.dtype)
    if closed:
        gi = torch.arange(row0, row0 + rows)
        diag[gi == 0] += 1
        diag[gi == ny_global - 1] += 1
        diag[:, 0] += 1
        diag[:, -1] += 1
    return s + diag * m


def slab_cg(rhs, row0, ny_global, closed, accuracy=1e-5, max_iterations=2000, group=None):
    """rhs: this rank's slab [B, rows, nx].  Returns (x slab, iterations[B] list, exchanges)."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    B, rows, nx = rhs.shape
    r = rhs.clone()
    m = torch.zeros_like(r)
    q = torch.zeros_like(r)
    x = torch.zeros_like(r)
    done = [False] * B
    iters = [0] * B
    k = 0
    zero_row = torch.zeros(B, nx, dtype=rhs.dtype)
    while True:
        # ---- the one exchange: five reductions + edge rows of (r, q, m) ----
        sums = torch.stack([(m * q).sum((1, 2)), (m * r).sum((1, 2)), (r * q).sum((1, 2)), (q * q).sum((1, 2))])
        mx = r.abs().amax((1, 2))
        if world > 1:
            dist.all_reduce(sums, group=group)
            dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
        halo_up = halo_dn = None
        if world > 1:
            top = torch.stack([r[:, 0], q[:, 0], m[:, 0]]).contiguous()
            bot = torch.stack([r[:, -1], q[:, -1], m[:, -1]]).contiguous()
            reqs = []
            if rank > 0:
                halo_up = torch.empty_like(top)
                reqs += [dist.isend(top, rank - 1, group=group), dist.irecv(halo_up, rank - 1, group=group)]
            if rank < world - 1:
                halo_dn = torch.empty_like(bot)
                reqs += [dist.isend(bot, rank + 1, group=group), dist.irecv(halo_dn, rank + 1, group=group)]
            for q_ in reqs:
                q_.wait()
        it = 0 if k == 0 else k - 1
        budget = it < max_iterations
        for b in range(B):
            if not done[b] and not (budget and (accuracy is None or accuracy <= 0 or float(mx[b]) >= accuracy)):
                done[b] = True
                iters[b] = it
        if all(done):
            return x, iters, k + 1
        s1, s2, s4, s5 = sums
        a = torch.where(s1 == 0, torch.zeros_like(s1), s2 / torch.where(s1 == 0, torch.ones_like(s1), s1))
        beta = torch.where(s1 == 0, torch.zeros_like(s1), (s4 - a * s5) / torch.where(s1 == 0, torch.ones_like(s1), s1))
        live = torch.tensor([0.0 if d else 1.0 for d in done], dtype=rhs.dtype)
        a, beta = (a * live)[:, None, None], beta[:, None, None]
        keep = live[:, None, None]
        x = x + a * m
        r_new = r - a * q
        m_new = keep * (r_new - beta * m) + (1 - keep) * m
        up = zero_row if halo_up is None else (halo_up[0] - a[:, 0] * halo_up[1]) - beta[:, 0] * halo_up[2]
        dn = zero_row if halo_dn is None else (halo_dn[0] - a[:, 0] * halo_dn[1]) - beta[:, 0] * halo_dn[2]
        r, m = r_new, m_new
        q = keep * _matvec(m, up, dn, row0, ny_global, closed) + (1 - keep) * q
        k += 1
