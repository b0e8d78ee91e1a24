"""CPU: host-side mirror of the reference interface (World / Physics / StateDependency / sequences / executor /
ControlTraining / Scene format) -- no arithmetic kernels involved.  Cases follow the reference's own API tests:
PhiFlow/tests/test_world.py, test_dependencies.py, test_domain.py and scripts/sanity_check.py."""
import os

import numpy as np
import pytest
import torch

from pde_control_b200 import physics as P
from pde_control_b200 import control as C


class _Counter(P.Physics):
    def __init__(self, deps=(), log=None):
        P.Physics.__init__(self, deps)
        self.log = log if log is not None else []

    def step(self, state, dt=1.0, **deps):
        self.log.append((state.name, {k: (v.age if isinstance(v, P.State) else tuple(s.age for s in v)) for k, v in deps.items()}))
        return state.copied_with(age=state.age + dt)


def test_world_add_step_remove_and_proxy():
    world = P.World()
    a = world.add(P.State(name="a", tags=("x",)))
    assert a.age == 0.0 and "a" in world.state
    world.step(dt=0.5)
    assert a.age == 0.5 and world.age == 0.5
    a.age = 3.0                                   # proxy write replaces the state in the world
    assert world.state.a.age == 3.0
    with pytest.raises(AttributeError):
        world.state["a"] = P.State(name="a")      # PhiFlow/phi/physics/collective.py:21-22
    world.remove(a)
    assert "a" not in world.state
    with pytest.raises(AssertionError):
        world.add(P.State(name="b"))
        world.add(P.State(name="b"))


def test_blocking_dependencies_order_and_cycle_detection():
    """PhiFlow/tests/test_dependencies.py: blocking dependencies see the UPDATED state, others the previous one."""
    log = []
    world = P.World()
    world.add(P.State(name="src", tags=("src_tag",)), _Counter(log=log))
    world.add(P.State(name="blocked", tags=("b",)), _Counter([P.StateDependency("d", "src_tag", single_state=True, blocking=True)], log))
    world.add(P.State(name="free", tags=("f",)), _Counter([P.StateDependency("d", "src_tag", blocking=False)], log))
    world.step(dt=1.0)
    seen = dict(log)
    assert seen["blocked"]["d"] == 1.0 and seen["free"]["d"] == (0.0,)
    assert [n for n, _ in log].index("src") < [n for n, _ in log].index("blocked")
    cyc = P.World()
    cyc.add(P.State(name="p", tags=("p",)), _Counter([P.StateDependency("d", "q", blocking=True)]))
    cyc.add(P.State(name="q", tags=("q",)), _Counter([P.StateDependency("d", "p", blocking=True)]))
    with pytest.raises(AssertionError, match="Cyclic"):
        cyc.step()


def test_domain_box_material_modes():
    d = P.Domain([128], box=P.box[0:1])
    assert d.boundaries is P.OPEN and d.rank == 1 and float(d.dx[0]) == 1 / 128
    assert P.OPEN.extrapolation_mode == "constant" and P.CLOSED.extrapolation_mode == "boundary"
    assert P.OPEN.accessible_extrapolation_mode == "boundary" and P.CLOSED.accessible_extrapolation_mode == "constant"
    assert P.CLOSED.vector_extrapolation_mode == "boundary" and P.NO_SLIP.vector_extrapolation_mode == "constant"
    assert P.PERIODIC.extrapolation_mode == "periodic"
    d2 = P.Domain([64, 48])
    assert d2.box == P.AABox([0, 0], [64, 48]) and d2.box_tuple() == (0.0, 0.0, 64.0, 48.0)
    assert P.box[0:4, 0:5].size.tolist() == [4.0, 5.0]


def test_fields_layout_and_staggered_tensor_round_trip():
    dom = P.Domain([6, 5])
    g = P.CenteredGrid.sample(2.0, dom, batch_size=3, device="cpu")
    assert g.data.shape == (3, 6, 5, 1) and g.extrapolation == "constant" and g.tensor.shape == (3, 6, 5)
    pts = g.points
    assert pts.shape == (1, 6, 5, 2) and float(pts[0, 0, 0, 0]) == 0.5 and float(pts[0, 5, 4, 1]) == 4.5
    t = torch.randn(3, 7, 6, 2)
    s = P.StaggeredGrid(t, box=dom.box, extrapolation="constant")
    assert s.data[0].data.shape == (3, 7, 5, 1) and s.data[1].data.shape == (3, 6, 6, 1)
    assert np.allclose(s.data[0].box.lower, [-0.5, 0]) and np.allclose(s.data[1].box.upper, [6, 5.5])
    back = s.staggered_tensor()
    assert torch.equal(back[:, :, :5, 0], t[:, :, :5, 0]) and torch.equal(back[:, :6, :, 1], t[:, :6, :, 1])
    assert float(back[:, :, 5, 0].abs().max()) == 0.0
    summed = (s + s) * 0.5 - s
    assert float(summed.data[0].data.abs().max()) == 0.0
    e = P.FieldEffect(g, ["velocity"])
    assert "velocity_effect" in e.tags and e.mode == P.GROW
    assert torch.equal(P.effect_applied(e, g, 0.5).data, g.data + g.data * 0.5)


class _Recorder(C.PartitioningExecutor):
    def __init__(self):
        self.events = []

    def execute_step(self, a, b, seq):
        C.PartitioningExecutor.execute_step(self, a, b, seq)
        self.events.append(("step", a.index))

    def partition(self, n, a, b, c):
        C.PartitioningExecutor.partition(self, n, a, b, c)
        self.events.append(("op%d" % n, c.index))


def test_sequence_call_orders():
    """Call order of the four schedulers (src/control/sequences.py:77-124), n = 4."""
    r = _Recorder(); C.StaggeredSequence(4, r).execute()
    assert r.events == [("op4", 2), ("op2", 1), ("step", 0), ("step", 1), ("op2", 3), ("step", 2), ("step", 3)]
    r = _Recorder(); C.LinearSequence(4, r).execute()
    assert r.events == [("step", i) for i in range(4)]
    r = _Recorder(); s = C.SkipSequence(4, r); s.execute()
    assert r.events == [("op4", 2)] and [f is not None for f in s] == [True, False, True, False, True]
    r = _Recorder(); C.RefinedSequence(4, r).execute()
    # after a step made frame 1 REAL, the frames ahead are re-predicted from it (update_target)
    assert r.events == [("op4", 2), ("op2", 1), ("step", 0), ("op2", 3), ("op2", 2), ("step", 1), ("op2", 3), ("step", 2), ("step", 3)]
    r = _Recorder(); C.RefinedSequence(8, r).execute()
    assert [e for e in r.events if e[0] == "step"] == [("step", i) for i in range(8)]


@pytest.mark.reference
def test_sequences_match_the_reference_module_event_for_event():
    """src/control/sequences.py has no TensorFlow dependency: load it by path and compare the executor call traces."""
    import contextlib
    import importlib.util
    import io
    from oracle import ref_shim
    path = os.path.join(ref_shim.REFERENCE_ROOT, "src", "control", "sequences.py")
    if not os.path.exists(path):
        pytest.skip("reference scheduler module not staged")
    spec = importlib.util.spec_from_file_location("ref_sequences", path)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)

    class RefRecorder(ref.PartitioningExecutor):
        def __init__(self):
            self.events = []

        def execute_step(self, a, b, seq):
            ref.PartitioningExecutor.execute_step(self, a, b, seq)
            self.events.append(("step", a.index))

        def partition(self, n, a, b, c):
            ref.PartitioningExecutor.partition(self, n, a, b, c)
            self.events.append(("op%d" % n, c.index))

    for name in ("StaggeredSequence", "RefinedSequence", "SkipSequence", "LinearSequence"):
        for n in (1, 2, 4, 8, 16, 32):
            if name == "SkipSequence" and n == 1:
                continue
            mine, theirs = _Recorder(), RefRecorder()
            getattr(C, name)(n, mine).execute()
            with contextlib.redirect_stdout(io.StringIO()):
                getattr(ref, name)(n, theirs).execute()
            assert mine.events == theirs.events, (name, n)


def test_executor_with_increment_pde_reaches_target():
    """scripts/sanity_check.py: the scalar walks linearly from its initial to its target value."""
    for seq_class in (C.StaggeredSequence, C.RefinedSequence, C.LinearSequence):
        pde = C.IncrementPDE()
        app = C.ControlTraining(8, pde, {"scalar": np.stack([np.linspace(0, 8, 9)[:, None]] * 2)}, None, [0, 1], obs_loss_frames=(),
                                trainable_networks=(), sequence_class=seq_class, batch_size=2, device="cpu")
        states = app.infer_all_frames([0, 1])
        values = [float(s.scalar.data[0, 0, 0]) for s in states]
        if seq_class is C.LinearSequence:
            assert values[-1] == 8.0
        else:
            assert np.allclose(values, np.arange(9))
        assert [s.scalar.age for s in states] == [float(i) for i in range(9)]


def test_scene_format_round_trip(tmp_path):
    u = np.random.RandomState(0).randn(3, 5, 16, 1).astype(np.float32)            # 3 sims, 5 frames
    for i in range(3):
        d = C.write_sim(str(tmp_path), i, {"u": u[i][:, None]})                    # Scene arrays carry a batch dim of 1
        assert os.path.basename(d) == "sim_%06d" % i and os.path.exists(os.path.join(d, "u_000004.npz"))
    ds = C.Dataset.load(str(tmp_path), [0, 2])
    assert ds.size == 2 and np.array_equal(ds.channels["u"], u[[0, 2]])
    it = ds.batches(2)
    assert next(it).tolist() == [0, 1] and next(it).tolist() == [0, 1]
    assert C.Dataset.load({"u": u}, range(3)).shard(1, 3).channels["u"].shape[0] == 1


def test_control_training_learns_and_checkpoints(tmp_path):
    torch.manual_seed(0)                     # batch order and initial weight come from the global generators
    np.random.seed(0)
    rng = np.random.RandomState(0)
    ends = rng.randn(16, 1, 1).astype(np.float32)
    frames = np.stack([ends * t / 2 for t in range(3)], 1)                        # [16 sims, 3 frames, 1, 1]: linear walk
    pde = C.IncrementPDE(learnable=True)
    app = C.ControlTraining(2, pde, {"scalar": frames}, [0, 1], range(16), obs_loss_frames=(1,), trainable_networks=("OP2",),
                            sequence_class=C.SkipSequence, batch_size=8, device="cpu", learning_rate=0.05, learning_rate_half_life=50)
    first = app.step()["Loss"]
    for _ in range(60):
        last = app.step()["Loss"]
    assert last < 0.05 * first and abs(float(pde.networks["OP2"].weight) - 0.5) < 0.05
    assert app.float_learning_rate < 0.05 * 0.5
    d = app.save_model(str(tmp_path))
    w = float(pde.networks["OP2"].weight)
    with torch.no_grad():
        pde.networks["OP2"].weight.fill_(0.0)
    app.load_checkpoints({"OP2": os.path.join(d, "OP2.pt")})
    assert float(pde.networks["OP2"].weight) == w
    assert "Loss" in app.infer_scalars([0, 1])


def test_upsample2x_1d_matches_aten_linear_and_the_reference_weights():
    """networks.upsample2x_1d == F.interpolate(linear, align_corners=False) == upsample2x of PF/math/nd.py:353-368."""
    from pde_control_b200.control.networks import upsample2x_1d, OPResNet
    torch.manual_seed(0)
    y = torch.randn(3, 5, 11)
    ref = torch.nn.functional.interpolate(y, scale_factor=2, mode="linear", align_corners=False)
    out = upsample2x_1d(y)
    assert out.shape == ref.shape
    np.testing.assert_allclose(out.numpy(), ref.numpy(), rtol=1e-6, atol=1e-6)
    # reference recipe written out in numpy: replicate pad, (0.25 lower + 0.75 centre, 0.75 centre + 0.25 upper) interleaved
    a = y.numpy(); p = np.pad(a, ((0, 0), (0, 0), (1, 1)), mode="edge")
    exp = np.stack([0.25 * p[..., :-2] + 0.75 * a, 0.75 * a + 0.25 * p[..., 2:]], -1).reshape(3, 5, 22)
    np.testing.assert_allclose(out.numpy(), exp, rtol=1e-6, atol=1e-6)
    net = OPResNet()
    o = net(torch.randn(2, 128, 1), torch.randn(2, 128, 1))
    assert o.shape == (2, 128, 1)
    o.sum().backward()
    assert all(q.grad is not None for q in net.parameters())


def _fluid_target_loss_inputs():
    rng = np.random.RandomState(11)
    a = rng.rand(3, 12, 16, 1).astype(np.float32)
    b = rng.rand(3, 12, 16, 1).astype(np.float32)
    return a, b


def _fluid_target_loss_mine(a, b):
    dom = P.Domain([12, 16])
    pde = C.IncompressibleFluidPDE(dom, 1.0)
    mk = lambda x: P.StateCollection([P.CenteredGrid(torch.tensor(x), box=dom.box, name="density")])
    return float(pde.target_matching_loss(mk(a), mk(b)))


def test_fluid_target_matching_loss_cross_shaped_weights():
    """incompressible_flow.py:23-29 restated in NumPy: fftfreq in 'vector' mode gives one Gaussian per axis, l1_loss sums the two
    channels -> weight G(ky) + G(kx) (the product G(ky) G(kx) is a different objective)."""
    a, b = _fluid_target_loss_inputs()
    diff = (a - b)[..., 0].astype(np.complex64)
    f = np.fft.fft2(diff)
    power = f.imag ** 2 + f.real ** 2
    k = np.stack(np.meshgrid(np.fft.fftfreq(12), np.fft.fftfreq(16), indexing="ij"), -1).astype(np.float32)
    w = np.exp(-0.5 * k ** 2 / 0.01 ** 2)                                  # [ny, nx, 2]
    expected = np.abs(power[..., None] * w[None]).sum() / 3
    mine = _fluid_target_loss_mine(a, b)
    assert abs(mine - expected) <= 1e-5 * expected
    product = np.abs(power * w[..., 0] * w[..., 1]).sum() / 3
    assert abs(product - expected) > 0.1 * expected                         # the two weightings really differ on this input


@pytest.mark.reference
def test_fluid_target_matching_loss_vs_live_reference_math():
    """The same expression evaluated with the reference's own math.fft / math.fftfreq / math.l1_loss (NumPy backend)."""
    from oracle import ref_shim
    flow = ref_shim.flow()
    math = flow.math
    a, b = _fluid_target_loss_inputs()
    diff = a - b
    diff_fft = math.fft(diff)
    psq = math.imag(diff_fft) ** 2 + math.real(diff_fft) ** 2              # psquare, PF/physics/schroedinger.py:34-35
    k = math.fftfreq(diff.shape[1:-1])
    expected = float(math.l1_loss(psq * math.exp(-0.5 * k ** 2 / 0.01 ** 2)))
    mine = _fluid_target_loss_mine(a, b)
    assert abs(mine - expected) <= 1e-5 * abs(expected)
