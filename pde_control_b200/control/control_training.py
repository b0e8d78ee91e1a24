"""ControlTraining: trains observation predictors (OP_n) and the control force estimator (CFE) through the
differentiable physics (src/control/control_training.py:11-142), eager PyTorch edition.

Same constructor arguments and methods (prepare / progress / step / load_checkpoints / save_model / infer_all_frames /
infer_scalars).  What was one static TF graph is rebuilt per optimisation step: world -> pde.create_pde -> the frame
sequence runs the OP/CFE networks and the solver (CUDA kernels behind ops.py) -> losses -> backward -> Adam.
Extras for the B200 build: `device`, in-memory datasets (`datapath` may be a dict), and data-parallel training over
`torch.distributed` (one process per GPU, batch sharded by rank, ONE flat all-reduce of the network gradients per step:
NCCL over NVLink on GPUs, gloo on CPU).
"""
import os

import torch
import torch.distributed as dist

from ..physics import World, CenteredGrid, StaggeredGrid
from .data import Dataset
from .hierarchy import PDEExecutor
from .sequences import StaggeredSequence, SkipSequence, LinearSequence


def property_name(attribute_name):
    return attribute_name


class ControlTraining(object):

    def __init__(self, n, pde, datapath, val_range, train_range, trace_to_channel=None, obs_loss_frames=(-1,),
                 trainable_networks=("CFE", "OP2"), sequence_class=StaggeredSequence, batch_size=16, view_size=16,
                 learning_rate=1e-3, learning_rate_half_life=1000, dt=1.0, new_graph=True, device="cuda", reg_scale=1.0,
                 process_group=None, cuda_graph=False):
        """cuda_graph=True (fixed batch shape, one GPU): the whole optimisation step -- the frame sequence with its network evaluations
        and solver kernels, the losses, backward and the Adam update -- is captured ONCE into a CUDA graph after three eager steps and
        replayed afterwards; per step only the batch is copied into static tensors and the learning rate written into a device scalar.
        The analogue of the reference's one static TF graph per app (src/control/control_training.py:50-84) run by session.run."""
        self.initial_learning_rate = self.float_learning_rate = learning_rate
        self.learning_rate_half_life = learning_rate_half_life
        self.diffphys = sequence_class is not None
        if sequence_class is None:
            assert "CFE" not in trainable_networks, "CFE training requires a sequence_class."
            assert len(obs_loss_frames) > 0, "No loss provided (no obs_loss_frames and no sequence_class)."
            sequence_class = SkipSequence
        self.n, self.dt, self.pde = n, dt, pde
        self.data_path = datapath
        self.sequence_class = sequence_class
        self.trainable_networks = tuple(trainable_networks)
        self.obs_loss_frames = tuple(obs_loss_frames)
        self.trace_to_channel = trace_to_channel or property_name
        self.batch_size = batch_size
        self.device = torch.device(device)
        self.reg_scale = reg_scale
        self.steps = 0
        self.scalars = {}
        self.checkpoint_dict = None
        self.optimizer = None
        self.prepared = False
        self.process_group = process_group
        self.world_size = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if self.world_size > 1 else 0
        self._train = None if train_range is None else Dataset.load(datapath, train_range)
        self._val = None if val_range is None else Dataset.load(datapath, val_range)
        if self.world_size > 1 and self._train is not None:
            self._train = self._train.shard(self.rank, self.world_size)
        self._train_iter = None if self._train is None else self._train.batches(batch_size)
        self.all_states = None
        self.sequence = None
        self.cuda_graph = bool(cuda_graph)
        self._graph = None

    # ---- graph replacement: one forward pass of the sequence --------------------------------------------------
    def _fill(self, worldstate, batch, frame):
        """Write frame `frame` of the batch into every data-holding grid of a world state."""
        states = []
        for s in worldstate.values():
            if isinstance(s, CenteredGrid):
                states.append(s.copied_with(data=self._frame_tensor(batch, s.name, frame, s.data.shape)))
                continue
            new = s
            for key, value in list(vars(s).items()):
                if isinstance(value, CenteredGrid):
                    new = new.copied_with(**{key: value.copied_with(data=self._frame_tensor(batch, key, frame, value.data.shape))})
                elif isinstance(value, StaggeredGrid):
                    raise NotImplementedError("staggered placeholders are not used by the two reference PDEs")
            states.append(new)
        return type(worldstate)(states)

    def _frame_tensor(self, batch, attribute, frame, shape):
        channel = self.trace_to_channel(attribute)
        arr = batch[channel][:, frame]
        t = torch.as_tensor(arr, dtype=torch.float32, device=self.device)
        return t.reshape((t.shape[0],) + tuple(shape[1:]))

    def forward(self, batch):
        """Runs the frame sequence on one batch {channel: array[B, frames, ...]} -> (loss, reg, scalars)."""
        n, pde = self.n, self.pde
        B = next(iter(batch.values())).shape[0]
        world = World(batch_size=B)
        pde.create_pde(world, "CFE" in self.trainable_networks, self.sequence_class != LinearSequence)
        world.state = self._fill(pde.placeholder_state(world, 0), batch, 0)
        target_state = self._fill(pde.placeholder_state(world, n * self.dt), batch, n)
        in_states = [world.state] + [None] * (n - 1) + [target_state]
        for frame in self.obs_loss_frames:
            if in_states[frame] is None:
                in_states[frame] = self._fill(pde.placeholder_state(world, (frame % (n + 1)) * self.dt), batch, frame % (n + 1))
        executor = self.executor = PDEExecutor(world, pde, target_state, self.trainable_networks, self.dt)
        sequence = self.sequence = self.sequence_class(n, executor)
        sequence.execute()
        self.all_states = [f.worldstate for f in sequence if f is not None]
        loss, reg = 0, None
        scalars = {}
        if self.diffphys:
            target_loss = pde.target_matching_loss(target_state, sequence[-1].worldstate)
            if target_loss is not None:
                loss = loss + target_loss
            reg = pde.total_force_loss([s for s in self.all_states if s is not None])
        for frame in self.obs_loss_frames:
            supervised = pde.target_matching_loss(in_states[frame], sequence[frame].worldstate)
            if supervised is not None:
                scalars["GT_obs_%d" % frame] = supervised
                loss = loss + supervised
        scalars.update(pde.scalars)
        return loss, reg, scalars

    # ---- LearningApp surface ---------------------------------------------------------------------------------
    def prepare(self):
        if self.prepared:
            return self
        if self._train is not None:                    # build the networks (variable scopes are created on first use)
            with torch.no_grad():
                self.forward(self._batch(self._train, next(self._train.batches(min(self.batch_size, self._train.size)))))
        self.pde.networks.to(self.device)
        if self.world_size > 1:                        # identical initial weights on every rank
            for p in self.pde.networks.parameters():
                dist.broadcast(p.data, src=0, group=self.process_group)
        self.prepared = True
        return self

    def _trainable_parameters(self):
        params = []
        for scope, net in self.pde.networks.items():
            if scope in self.trainable_networks:
                params += list(net.parameters())
        return params

    @staticmethod
    def _batch(dataset, indices):
        return {k: v[indices] for k, v in dataset.channels.items()}

    def step(self):
        """One optimisation step (control_training.py:122-125 + LearningApp.optimization_step, tf/app.py:248-275)."""
        self.prepare()
        if self.learning_rate_half_life is not None:
            self.float_learning_rate = self.initial_learning_rate * 0.5 ** (self.steps / float(self.learning_rate_half_life))
        if self.cuda_graph:
            return self._graphed_step()
        params = self._trainable_parameters()
        if self.optimizer is None:
            self.optimizer = torch.optim.Adam(params, lr=self.float_learning_rate)
        for group in self.optimizer.param_groups:
            group["lr"] = self.float_learning_rate
        batch = self._batch(self._train, next(self._train_iter))
        self.optimizer.zero_grad(set_to_none=True)
        loss, reg, scalars = self.forward(batch)
        objective = loss if reg is None else loss + reg * self.reg_scale          # add_objective(loss, reg=reg), tf/app.py:219-238
        objective.backward()
        self._allreduce_gradients(params, scalars, objective)
        self.optimizer.step()
        self.steps += 1
        self.scalars = {k: float(v.detach() if isinstance(v, torch.Tensor) else v) for k, v in scalars.items()}
        self.scalars["Loss"] = float(objective.detach())
        return self.scalars

    # ---- the optimisation step as one CUDA graph --------------------------------------------------------------------------------
    _GRAPH_WARMUP = 3

    def _graphed_step(self):
        assert self.world_size == 1, "cuda_graph=True is a single-GPU feature (the gradient all-reduce re-binds .grad every step)"
        assert self.device.type == "cuda"
        host = self._batch(self._train, next(self._train_iter))
        if not hasattr(self, "_static"):
            params = self._trainable_parameters()
            self._lr = torch.tensor(float(self.float_learning_rate), device=self.device)
            self.optimizer = torch.optim.Adam(params, lr=self._lr, capturable=True)
            self._static = {k: torch.empty(v.shape, dtype=torch.float32, device=self.device) for k, v in host.items()}
            self._side = torch.cuda.Stream(self.device)
            self._eager_steps = 0
        self._lr.fill_(float(self.float_learning_rate))
        for k, v in host.items():
            self._static[k].copy_(torch.as_tensor(v, dtype=torch.float32), non_blocking=False)

        def one_step():
            self.optimizer.zero_grad(set_to_none=True)
            loss, reg, scalars = self.forward(self._static)
            objective = loss if reg is None else loss + reg * self.reg_scale
            objective.backward()
            self.optimizer.step()
            out = {k: (v.detach() if isinstance(v, torch.Tensor) else float(v)) for k, v in scalars.items()}
            out["Loss"] = objective.detach()
            return out

        if self._graph is None and self._eager_steps < self._GRAPH_WARMUP:
            # eager steps on the side stream the capture will use (autograd binds gradient sinks to the stream of first use)
            self._side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self._side):
                out = one_step()
            torch.cuda.current_stream(self.device).wait_stream(self._side)
            self._eager_steps += 1
        else:
            if self._graph is None:
                # capture: Adam runs BEFORE it is captured at least once (state initialised above); the capture itself executes
                # nothing, so this step's work happens in the replay right below
                torch.cuda.synchronize(self.device)
                self._graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self._graph, stream=self._side):
                    self._graph_out = one_step()
            self._graph.replay()
            out = self._graph_out
        self.steps += 1
        self.scalars = {k: float(v) for k, v in out.items()}
        return self.scalars

    progress = step

    def _allreduce_gradients(self, params, scalars, objective):
        """Data parallelism: ONE flat all-reduce carries every network gradient (averaged) and the reported scalars."""
        if self.world_size <= 1:
            return
        grads = [p.grad if p.grad is not None else torch.zeros_like(p) for p in params]
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.process_group)
        flat /= self.world_size
        offset = 0
        for p, g in zip(params, grads):
            p.grad = flat[offset:offset + g.numel()].view_as(g).clone()
            offset += g.numel()

    def validation_step(self):
        return self.infer_scalars(None)

    def load_checkpoints(self, checkpoint_dict):
        self.prepare()
        self.checkpoint_dict = checkpoint_dict
        self.executor.load(self.n, checkpoint_dict, preload_n=True, session=None, logf=self.info)

    def save_model(self, directory=None):
        """One state_dict file per variable scope (the reference saves one TF checkpoint per scene directory)."""
        directory = directory or os.path.join(os.path.expanduser("~"), "pde_control_b200_models", "step_%06d" % self.steps)
        os.makedirs(directory, exist_ok=True)
        for scope, net in self.pde.networks.items():
            torch.save(net.state_dict(), os.path.join(directory, "%s.pt" % scope))
        return directory

    def info(self, message):
        if self.rank == 0:
            print(message)

    def infer_all_frames(self, data_range):
        dataset = self._val if data_range is None else Dataset.load(self.data_path, data_range)
        with torch.no_grad():
            self.forward(self._batch(dataset, slice(None)))
        return self.all_states

    def infer_scalars(self, data_range):
        dataset = self._val if data_range is None else Dataset.load(self.data_path, data_range)
        with torch.no_grad():
            loss, reg, scalars = self.forward(self._batch(dataset, slice(None)))
        out = {k: float(v) for k, v in scalars.items()}
        out["Loss"] = float(loss if reg is None else loss + reg * self.reg_scale)
        return out
