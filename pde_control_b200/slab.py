"""Row-slab decomposition of the 2-D incompressible-flow step over the GPUs of one box (BASELINE config 5).

Nothing is replicated: rank k owns rows [r0, r1) of every centred field, the x-face rows [r0, r1) and the y-face rows r0 .. r1 (the
face row ON a seam is held by both neighbours and stays consistent because both compute it from the same inputs).  Per step
  * advection (F1, F2): halo of `halo` rows of rho, vy, vx from both neighbours over NVLink peer memory (pdec_halo_push_rows /
    pdec_halo_pull_rows), then the ordinary single-GPU kernels on the haloed slab with a shifted box, cropped to the slab;
    `halo` must cover max|v| dt / dx + 2 (back-trace + bilinear tap + face interpolation) -- checked on the device, no host sync;
  * mask + divergence (F4): local (the seam face is there);
  * pressure solve (F5): the slab-decomposed on-chip CG, per-iteration exchange inside the kernel (pdec_cg_poisson_2d_slab);
  * gradient subtraction (F6): one halo row of p from both neighbours, local kernel, cropped;
  * density normalisation on CLOSED domains: two all-reduced sums (the only NCCL call of the step, 2 floats per batch element).
The adjoint is autograd over the same operators: the adjoint of a halo exchange is the reverse exchange with accumulation, the adjoint
of the solve is the same slab solve (A symmetric, PF/physics/pressuresolver/sparse.py:93-94).

`SlabLayout` is pure index arithmetic (no CUDA): tests/test_distributed_cpu.py runs it over gloo with a NumPy transport.
"""
import ctypes

import torch

CENTRED, YFACE = "centred", "yface"          # row kinds: cell rows / x-face rows (one per cell row), y-face rows (one more, seam shared)


class SlabLayout(object):
    """Which rows a rank owns and which rows travel in a halo exchange of width H."""

    def __init__(self, ny, world, rank):
        assert ny % world == 0, "equal slabs"
        self.ny, self.world, self.rank = ny, world, rank
        self.rows = ny // world
        self.r0, self.r1 = rank * self.rows, (rank + 1) * self.rows
        self.has_lo, self.has_hi = rank > 0, rank < world - 1

    def local_rows(self, kind):
        return self.rows + (1 if kind == YFACE else 0)

    def halo_widths(self, H):
        assert H <= self.rows, "halo wider than a slab"
        return (H if self.has_lo else 0), (H if self.has_hi else 0)

    def plan(self, kind, H):
        """[(side, neighbour, send_first, recv_first_in_haloed, n)] for the forward exchange; local row indices of the UNhaloed slab for
        the send rows, indices of the haloed array for the received rows.  side 0 = data travelling up (to rank + 1)."""
        hlo, hhi = self.halo_widths(H)
        n_local = self.local_rows(kind)
        seam = 1 if kind == YFACE else 0           # the seam face itself is not sent: the neighbour has its own copy
        out = []
        if self.has_hi:                            # my top rows -> lower halo of rank + 1 ; its bottom rows -> my upper halo
            out.append(dict(side=0, peer=self.rank + 1, send_first=n_local - seam - H, recv_side=1, recv_first=hlo + n_local, n=H))
        if self.has_lo:                            # my bottom rows -> upper halo of rank - 1 ; its top rows -> my lower halo
            out.append(dict(side=1, peer=self.rank - 1, send_first=seam, recv_side=0, recv_first=0, n=H))
        return out


class _Channel(object):
    """One inbox per (field, direction of differentiation): its own sequence counter, so that exchanges of different fields never have to
    agree on an order."""

    def __init__(self, comm, B, row_len, max_rows):
        from . import _lib
        L = _lib.lib()
        self.B, self.row_len, self.max_rows = B, row_len, max_rows
        nbytes = L.pdec_halo_inbox_bytes(B, row_len, max_rows)
        self.local = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        _lib.check(L.pdec_comm_alloc(nbytes, ctypes.byref(self.local), handle), "pdec_comm_alloc")
        _lib.check(L.pdec_halo_inbox_init(self.local, B, row_len, max_rows, _lib.stream_ptr()), "pdec_halo_inbox_init")
        self.handle = bytes(handle)
        self.peers = {}
        self.seq = 0

    def open_peers(self, handles, rank):
        from . import _lib
        L = _lib.lib()
        for r, h in handles.items():
            if r == rank:
                continue
            mapped = ctypes.c_void_p()
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
            _lib.check(L.pdec_comm_open(buf, ctypes.byref(mapped)), "pdec_comm_open")
            self.peers[r] = mapped

    def close(self):
        from . import _lib
        L = _lib.lib()
        for m in self.peers.values():
            L.pdec_comm_close(m)
        self.peers = {}
        if self.local is not None:
            L.pdec_comm_free(self.local)
            self.local = None


class HaloExchanger(object):
    """Peer-memory halo transport for the fields of one decomposed domain.  One process per GPU; `group` defaults to the world group."""

    FIELDS = {"rho": (CENTRED, 0), "vy": (YFACE, 0), "vx": (CENTRED, 1), "p": (CENTRED, 0)}       # kind, extra row length

    def __init__(self, B, ny, nx, halo, group=None):
        import torch.distributed as dist
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.layout = SlabLayout(ny, self.world, self.rank)
        self.B, self.nx, self.halo = B, nx, halo
        self.status = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.channels = {}
        for name, (kind, extra) in self.FIELDS.items():
            for direction in ("fwd", "bwd"):
                self.channels[(name, direction)] = _Channel(self, B, nx + extra, max(halo, 1))
        torch.cuda.synchronize()
        mine = {k: c.handle for k, c in self.channels.items()}
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=group)           # also orders every rank's inbox init before any peer push
        for k, c in self.channels.items():
            neighbours = {r: everyone[r][k] for r in (self.rank - 1, self.rank + 1) if 0 <= r < self.world}
            c.open_peers(neighbours, self.rank)
        dist.barrier(group=group)

    def _run(self, name, direction, src, dst, plan, reverse):
        """forward: push rows of `src` (the slab), pull into `dst` (the haloed array).  reverse (adjoint): push the halo rows of `src`
        (gradient of the haloed array), pull-ACCUMULATE into the rows of `dst` (gradient of the slab) that were sent forward."""
        from . import _lib
        L = _lib.lib()
        ch = self.channels[(name, direction)]
        ch.seq += 1
        st = _lib.stream_ptr()
        B = src.shape[0]
        with torch.cuda.device(src.device):
            for e in plan:          # all pushes first: nobody waits before having sent
                first = e["recv_first"] if reverse else e["send_first"]
                side = e["side"]
                _lib.check(L.pdec_halo_push_rows(_lib.ptr(src), B, src.shape[1], src.shape[2], first, e["n"], ch.peers[e["peer"]], side, ch.seq,
                                                 ch.max_rows, st), "pdec_halo_push_rows")
            for e in plan:
                first = e["send_first"] if reverse else e["recv_first"]
                # forward: the rows that arrive in my upper halo travelled DOWN (side 1) and vice versa; the adjoint reverses the direction
                side = e["recv_side"]
                _lib.check(L.pdec_halo_pull_rows(_lib.ptr(dst), B, dst.shape[1], dst.shape[2], first, e["n"], ch.local, side, ch.seq,
                                                 1 if reverse else 0, ch.max_rows, _lib.ptr(self.status), st), "pdec_halo_pull_rows")

    def exchange(self, field, name, H):
        """field [B, local rows, row_len] -> haloed [B, hlo + local rows + hhi, row_len]; differentiable."""
        return _HaloExchange.apply(field, self, name, int(H))

    def close(self):
        torch.cuda.synchronize()
        for c in self.channels.values():
            c.close()
        self.channels = {}


class _HaloExchange(torch.autograd.Function):
    @staticmethod
    def forward(ctx, field, ex, name, H):
        field = field.contiguous()
        kind = ex.FIELDS[name][0]
        hlo, hhi = ex.layout.halo_widths(H)
        plan = ex.layout.plan(kind, H)
        out = torch.empty((field.shape[0], hlo + field.shape[1] + hhi, field.shape[2]), device=field.device, dtype=field.dtype)
        out[:, hlo:hlo + field.shape[1]] = field
        ex._run(name, "fwd", field, out, plan, reverse=False)
        ctx.ex, ctx.name, ctx.plan, ctx.hlo, ctx.n = ex, name, plan, hlo, field.shape[1]
        return out

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous()
        g_field = g[:, ctx.hlo:ctx.hlo + ctx.n].clone()
        ctx.ex._run(ctx.name, "bwd", g, g_field, ctx.plan, reverse=True)
        return g_field, None, None, None


class _SlabSolve(torch.autograd.Function):
    """The slab-decomposed CG on this rank's rows of the right-hand side; backward = the same solve of the incoming gradient."""

    @staticmethod
    def forward(ctx, div, stepper, closed_code, accuracy, max_iterations):
        from . import ops
        div = div.contiguous()
        p = torch.empty_like(div)
        iters = torch.zeros(div.shape[0], device=div.device, dtype=torch.int32)
        lay = stepper.layout
        ops.cg_poisson_slab_raw(div, p, closed_code, accuracy, max_iterations, ny_global=lay.ny, row0=lay.r0, comm=stepper.comm, iterations=iters)
        ctx.cfg = (stepper, closed_code, accuracy, max_iterations)
        ctx.mark_non_differentiable(iters)
        return p, iters

    @staticmethod
    def backward(ctx, g_p, _gi):
        from . import ops
        stepper, closed_code, accuracy, max_iterations = ctx.cfg
        g_p = g_p.contiguous()
        g_div = torch.empty_like(g_p)
        lay = stepper.layout
        ops.cg_poisson_slab_raw(g_p, g_div, closed_code, accuracy, max_iterations, ny_global=lay.ny, row0=lay.r0, comm=stepper.comm)
        return g_div, None, None, None, None


def laplace_rows(p_full, r0, r1, closed):
    """Rows [r0, r1) of A p for the whole-domain pressure p_full [B, ny, nx]: the 5-point operator of the pressure solve (a zero outside an
    OPEN domain, one neighbour fewer on a CLOSED one: sparse.py:151-221 without obstacles), evaluated with torch slices."""
    q = torch.nn.functional.pad(p_full, (1, 1, 1, 1), mode="replicate" if closed else "constant")
    c = q[:, r0 + 1:r1 + 1, 1:-1]
    return q[:, r0:r1, 1:-1] + q[:, r0 + 2:r1 + 2, 1:-1] + q[:, r0 + 1:r1 + 1, :-2] + q[:, r0 + 1:r1 + 1, 2:] - 4.0 * c


class _SlabSolveChained(torch.autograd.Function):
    """SparseSciPy() & SparseCG(accuracy) on slabs: every rank gathers the right-hand side (one all-gather), runs the exact solve of the
    WHOLE domain (library GEMMs, ops.poisson_direct_raw), and the slab CG then solves the CORRECTION system A d = rhs - A p0 on the
    slabs to the reference's stopping rule (its residual is the residual of the whole system); p = p0 + d.  At 2048 x 2048 the correction
    needs 0-2 iterations instead of ~3000 (scripts/diag/cg_correction_large.py).  Backward = the same chain on the incoming gradient."""

    @staticmethod
    def _solve(rhs, stepper, closed_code, accuracy, max_iterations, iters):
        from . import ops
        lay = stepper.layout
        full = stepper.gather(rhs)
        p0 = torch.empty_like(full)
        ops.poisson_direct_raw(full, p0, closed_code)
        if closed_code:
            # a closed domain's matrix is singular (constants): rounding leaves the correction's right-hand side with a small mean the CG
            # could never remove -- it would run to max_iterations and drift.  Project it out (every rank holds the whole p0 and rhs).
            r_full = full - laplace_rows(p0, 0, lay.ny, True)
            r = (r_full[:, lay.r0:lay.r1] - r_full.mean((1, 2), keepdim=True)).contiguous()
        else:
            r = (rhs - laplace_rows(p0, lay.r0, lay.r1, False)).contiguous()
        d = torch.empty_like(r)
        ops.cg_poisson_slab_raw(r, d, closed_code, accuracy, max_iterations, ny_global=lay.ny, row0=lay.r0, comm=stepper.comm, iterations=iters)
        return p0[:, lay.r0:lay.r1] + d

    @staticmethod
    def forward(ctx, div, stepper, closed_code, accuracy, max_iterations):
        assert closed_code in (0, 1), "the chained slab solve covers uniformly open or closed domains"
        iters = torch.zeros(div.shape[0], device=div.device, dtype=torch.int32)
        p = _SlabSolveChained._solve(div.contiguous(), stepper, closed_code, accuracy, max_iterations, iters)
        ctx.cfg = (stepper, closed_code, accuracy, max_iterations)
        ctx.mark_non_differentiable(iters)
        return p, iters

    @staticmethod
    def backward(ctx, g_p, _gi):
        stepper, closed_code, accuracy, max_iterations = ctx.cfg
        return _SlabSolveChained._solve(g_p.contiguous(), stepper, closed_code, accuracy, max_iterations, None), None, None, None, None


class SlabStepper(object):
    """IncompressibleFlow.step (PF/physics/fluid.py:90-114; no obstacles / effects) on a row-slab decomposed domain."""

    def __init__(self, B, ny, nx, halo=4, box=None, group=None):
        import torch.distributed as dist
        from . import ops
        self.group = group
        self.B, self.ny, self.nx, self.halo = B, ny, nx, halo
        self.box = (0.0, 0.0, float(ny), float(nx)) if box is None else tuple(float(v) for v in box)
        self.ex = HaloExchanger(B, ny, nx, halo, group)
        self.layout = self.ex.layout
        assert self.layout.rows % 8 == 0, "slabs must be a multiple of 8 rows (tiles of the on-chip CG)"
        self.comm = ops.SlabComm(B, nx, group) if self.layout.world > 1 else None
        self.halo_violations = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.dist = dist

    # ---- distribution of global arrays (tests, benchmarks) ------------------------------------------------------------------
    def scatter(self, rho, vy, vx):
        lay = self.layout
        return (rho[:, lay.r0:lay.r1].contiguous(), vy[:, lay.r0:lay.r1 + 1].contiguous(), vx[:, lay.r0:lay.r1].contiguous())

    def gather(self, local, kind=CENTRED):
        """All ranks' slabs -> the global array (every rank gets it; seam rows of y-face fields are taken from the lower slab)."""
        parts = [torch.empty_like(local) for _ in range(self.layout.world)]
        self.dist.all_gather(parts, local.contiguous(), group=self.group)
        if kind == YFACE:
            parts = [p[:, :-1] for p in parts[:-1]] + [parts[-1]]
        return torch.cat(parts, 1)

    def gather_grad(self, g_local, kind=CENTRED):
        """Gradients w.r.t. scattered inputs -> gradient w.r.t. the global array: the adjoint of `scatter` SUMS the two copies of a seam row."""
        parts = [torch.empty_like(g_local) for _ in range(self.layout.world)]
        self.dist.all_gather(parts, g_local.contiguous(), group=self.group)
        if kind != YFACE:
            return torch.cat(parts, 1)
        rows = []
        for k, p in enumerate(parts):
            body = p.clone()
            if k > 0:
                body[:, 0] += parts[k - 1][:, -1]
            rows.append(body if k == len(parts) - 1 else body[:, :-1])
        return torch.cat(rows, 1)

    # ---- the step ---------------------------------------------------------------------------------------------------------------
    def _sides(self, closed, hlo, hhi):
        """Solid sides of a LOCAL array: x sides as the domain's; a y side only where the array really ends at the domain boundary."""
        from . import ops
        c = ops.closed_code(closed)
        bits = (c & 0xF) if (c & 0x10) else (0xF if c else 0)
        ylo = bool(bits & 1) and not self.layout.has_lo and hlo == 0
        yhi = bool(bits & 2) and not self.layout.has_hi and hhi == 0
        return (ylo, yhi, bool(bits & 4), bool(bits & 8))

    def step(self, rho, vy, vx, dt, closed, accuracy=1e-5, max_iterations=20000, conserve_density=True, chained=False):
        """Local slabs in, local slabs out: (rho', vy', vx', p, cg iterations).  chained=True: the solver chain SparseSciPy() & SparseCG
        (_SlabSolveChained: whole-domain exact solve on every rank as the guess, slab CG on the correction)."""
        from . import ops
        from ._lib import MODE_CONSTANT, MODE_REPLICATE
        lay, H = self.layout, self.halo
        ly, lx, sy, sx = self.box
        dy = sy / self.ny
        mode = MODE_REPLICATE if closed else MODE_CONSTANT
        hlo, hhi = lay.halo_widths(H)
        # back-traces must stay inside the halo: |v| dt / dx + 2 rows (tap + face interpolation); counted on the device
        with torch.no_grad():
            reach = torch.maximum(vy.abs().max(), vx.abs().max()) * (abs(float(dt)) / dy) + 2.0
            self.halo_violations += (reach > float(H)).to(torch.int32)
        rho_h = self.ex.exchange(rho, "rho", H)
        vy_h = self.ex.exchange(vy, "vy", H)
        vx_h = self.ex.exchange(vx, "vx", H)
        rows_h = lay.rows + hlo + hhi
        box_h = (ly + (lay.r0 - hlo) * dy, lx, rows_h * dy, sx)
        rho2 = ops.advect_centered(rho_h, vy_h, vx_h, dt, mode, box_h)[:, hlo:hlo + lay.rows]
        ay, ax = ops.advect_staggered(vy_h, vx_h, dt, mode, box_h)
        ay, ax = ay[:, hlo:hlo + lay.rows + 1], ax[:, hlo:hlo + lay.rows]
        if closed and conserve_density:
            # CenteredGrid.normalized (grid.py:205-209): global sums of both densities
            import torch.distributed.nn.functional as dfn
            s_old, s_new = rho.sum((1, 2)), rho2.sum((1, 2))
            if lay.world > 1:
                s_old, s_new = dfn.all_reduce(s_old, group=self.group), dfn.all_reduce(s_new, group=self.group)
            rho2 = rho2 * (s_old / torch.clamp(s_new, min=1e-5))[:, None, None]
        # projection: divergence is local (the seam face is there), the solve is the slab CG, the gradient needs one halo row of p
        div = ops.mask_divergence(ay.contiguous(), ax.contiguous(), self._sides(closed, 0, 0))
        p, iters = (_SlabSolveChained if chained else _SlabSolve).apply(div, self, ops.closed_code(closed), accuracy, int(max_iterations))
        plo, phi = lay.halo_widths(1)
        p_h = self.ex.exchange(p, "p", 1)
        zy = ay.new_zeros((ay.shape[0], 1, ay.shape[2]))
        zx = ax.new_zeros((ax.shape[0], 1, ax.shape[2]))
        ay_h = torch.cat(([zy] if plo else []) + [ay] + ([zy] if phi else []), 1)        # dummy face rows beyond the halo: cropped below
        ax_h = torch.cat(([zx] if plo else []) + [ax] + ([zx] if phi else []), 1)
        rows_p = lay.rows + plo + phi
        oy, ox = ops.subtract_pressure_gradient(ay_h, ax_h, p_h, self._sides(closed, plo, phi), (0.0, 0.0, rows_p * dy, sx))
        return rho2, oy[:, plo:plo + lay.rows + 1], ox[:, plo:plo + lay.rows], p, iters

    def close(self):
        self.ex.close()
        if self.comm is not None:
            self.comm.close()
            self.comm = None
