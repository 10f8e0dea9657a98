"""Target-sharded direct N^2 across the GPUs of one node: one process per GPU, torch.distributed
for the plumbing, our kernels for the work.  The reference is single-GPU; this
is new capability named by BASELINE.json's north_star.

Rank r owns targets [r*N/R, (r+1)*N/R) and needs all N sources every step.  Two exchanges:

exchange="peer" (default): the all-gather is fused into our own kernels over NVLink peer memory.
  Setup: every rank allocates an exchange block, the CUDA-IPC handles travel once through
  torch.distributed (all_gather_object), every rank maps the others' blocks.  Per step:
  1. pack_push: store-acc + drift of the local slice, each packed float4 {x,y,z,G*m} stored into
     ALL R replicated source buffers (local + R-1 peers), then this rank's step flag raised on
     every peer                                                          (nb_mgpu_peer_step)
  2. force pass over the LOCAL slice (overlaps the peers' pushes)
  3. a one-warp kernel waits for the R-1 peer flags (acquire at system scope)
  4. force pass over the other R-1 slices in ring order
  5. fixed-order reduction of the per-chunk partials + kick

exchange="nccl": the baseline the peer path is measured against.
  1. pack into this rank's slot                                          (nb_mgpu_pack_local)
  2. all_gather_into_tensor of the slots, asynchronous on NCCL's stream  (torch.distributed)
  3. force pass over the LOCAL slot -- overlaps the gather               (nb_mgpu_forces_local)
  4. wait for the gather; force pass over the other R-1 slots            (nb_mgpu_forces_remote)
  5. reduction + kick                                                    (nb_mgpu_finish)
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .ics import FIELDS

# tile quantum of the direct kernel (nbody_b200/csrc/direct.cu: kCfgBig / kCfgSmall)
_SMALL_N = 32768


def shard_layout(n_total: int, rank: int, world: int):
    """Same arithmetic as nb_mgpu_layout: (begin, count, padded_slice)."""
    if world < 1 or not (0 <= rank < world) or n_total <= 0:
        raise ValueError("bad shard layout arguments")
    begin = n_total * rank // world
    end = n_total * (rank + 1) // world
    max_cnt = (n_total + world - 1) // world
    q = 256 if n_total < _SMALL_N else 1024
    padded = (max_cnt + q - 1) // q * q
    return begin, end - begin, padded


def slice_state(state: dict, begin: int, count: int) -> dict:
    return {k: np.ascontiguousarray(state[k][begin:begin + count]) for k in FIELDS}


def pack_sources_host(state_local: dict, padded: int, G: float) -> np.ndarray:
    """Host restatement of the pack kernel (used by the gloo CPU tests of the sharding logic)."""
    n = state_local["pos_x"].size
    out = np.zeros((padded, 4), np.float32)
    out[:n, 0] = state_local["pos_x"]
    out[:n, 1] = state_local["pos_y"]
    out[:n, 2] = state_local["pos_z"]
    out[:n, 3] = np.float32(G) * state_local["mass"]
    return out


class ShardedDirectSystem:
    """ParticleSystem-like driver for one rank of a target-sharded direct-N^2 run."""

    def __init__(self, state_local: dict, n_total: int, rank: int, world: int, G: float = 1.0,
                 softening: float = 0.01, group=None, exchange: str | None = None):
        import torch
        from . import _lib
        from .api import ParticleData, ParticleDataManager, get_context
        self.torch = torch
        self.L = _lib.lib()
        self.check = _lib.check
        self.rank, self.world, self.n_total = rank, world, n_total
        self.G = float(np.float32(G))
        self.eps2 = float(np.float32(softening) * np.float32(softening))
        self.group = group
        self.ctx = get_context()
        b, c, p, tot = C.c_size_t(), C.c_size_t(), C.c_size_t(), C.c_size_t()
        self.check(self.L.nb_mgpu_layout(self.ctx.handle, n_total, rank, world, C.byref(b), C.byref(c),
                                         C.byref(p), C.byref(tot)))
        assert (b.value, c.value, p.value) == shard_layout(n_total, rank, world)
        self.begin, self.count, self.padded = b.value, c.value, p.value
        assert state_local["pos_x"].size == self.count
        self.h = ParticleDataManager.fromState(state_local, pinned=True)
        self.d = ParticleData()
        ParticleDataManager.allocateDevice(self.d, self.count)
        ParticleDataManager.copyToDevice(self.d, self.h)
        self.exchange = exchange or os.environ.get("NB_MGPU_EXCHANGE", "peer")
        if self.exchange not in ("peer", "nccl"):
            raise ValueError("exchange must be 'peer' or 'nccl'")
        self.gathered = None
        if self.exchange == "peer":
            self._setup_peer()
        else:
            self.gathered = torch.zeros(world * self.padded * 4, dtype=torch.float32, device="cuda")
        self.simulation_time = 0.0
        self._evaluated = False

    def _setup_peer(self):
        """Collective: allocate the exchange block, swap CUDA-IPC handles, map the peers' blocks."""
        handle = (C.c_ubyte * 64)()
        self.check(self.L.nb_mgpu_peer_alloc(self.ctx.handle, handle))
        if self.world == 1:
            return
        import torch.distributed as dist
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=self.group)
        for r, h in enumerate(handles):
            if r != self.rank:
                buf = (C.c_ubyte * 64).from_buffer_copy(h)
                self.check(self.L.nb_mgpu_peer_open(self.ctx.handle, r, buf))
        dist.barrier(group=self.group)  # nobody pushes before everybody has mapped and zeroed

    def close(self):
        if self.exchange == "peer":
            self.torch.cuda.synchronize()
            if self.world > 1:
                import torch.distributed as dist
                dist.barrier(group=self.group)  # peers may still be reading our block
            self.check(self.L.nb_mgpu_peer_close(self.ctx.handle))

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    def _evaluate(self, dt: float, step: bool):
        L, ctx = self.L, self.ctx.handle
        s = self._stream()
        self._evaluated = True
        if self.exchange == "peer":
            self.check(L.nb_mgpu_peer_step(ctx, C.byref(self.d.c), dt, self.G, self.eps2, 1 if step else 0, s))
            return
        gp = C.c_void_p(self.gathered.data_ptr())
        self.check(L.nb_mgpu_pack_local(ctx, C.byref(self.d.c), gp, dt, self.G, 1 if step else 0, s))
        work = None
        if self.world > 1:
            import torch.distributed as dist
            mine = self.gathered[self.rank * self.padded * 4:(self.rank + 1) * self.padded * 4]
            work = dist.all_gather_into_tensor(self.gathered, mine, group=self.group, async_op=True)
        self.check(L.nb_mgpu_forces_local(ctx, gp, self.eps2, s))
        if work is not None:
            work.wait()  # current stream waits for the gather; the local pass above overlapped it
            self.check(L.nb_mgpu_forces_remote(ctx, gp, self.eps2, s))
        self.check(L.nb_mgpu_finish(ctx, C.byref(self.d.c), dt, 1 if step else 0, s))

    def compute_forces(self):
        self._evaluate(0.0, False)

    def update(self, dt: float):
        self._evaluate(float(np.float32(dt)), True)
        self.simulation_time += dt

    def host_update(self, dt: float, steps: int = 1):
        """Host-buffer step: H2D of the local state, `steps` steps, D2H (the e2e path at N > 1)."""
        t = self.torch
        for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "acc_x", "acc_y", "acc_z", "mass"):
            self.d.arrays[k].copy_(self.h._pinned[k][:self.count], non_blocking=True)
        for _ in range(steps):
            self.update(dt)
        for k in FIELDS[:12]:
            self.h._pinned[k][:self.count].copy_(self.d.arrays[k], non_blocking=True)
        t.cuda.current_stream().synchronize()

    def local_arrays(self) -> dict:
        return self.d.numpy()

    def compute_energy(self) -> dict:
        """Collective.  Kinetic, potential and total energy of the WHOLE system: every rank
        sums its own bodies' kinetic energy and its targets' share of the pair sum against the replicated sources
        (nb_mgpu_potential_partial), then one all-reduce of two doubles."""
        if not self._evaluated:
            self.compute_forces()  # fills the replicated source array with the current positions
        ke, part = C.c_double(0), C.c_double(0)
        self.check(self.L.nb_kinetic_energy(self.ctx.handle, C.byref(self.d.c), C.byref(ke), self._stream()))
        gp = C.c_void_p(self.gathered.data_ptr()) if self.gathered is not None else None
        eps = float(np.sqrt(np.float32(self.eps2)))
        self.check(self.L.nb_mgpu_potential_partial(self.ctx.handle, C.byref(self.d.c), gp, eps, C.byref(part),
                                                    self._stream()))
        return reduce_energy(ke.value, part.value, self.world, self.group, self.torch)


def reduce_energy(ke_local: float, pair_sum_local: float, world: int, group, torch, device="cuda") -> dict:
    """All-reduce of {local kinetic energy, local share of sum_i m_i sum_{j != i} G m_j / r_ij}: every ordered pair
    appears once in the global sum, so PE = -sum / 2."""
    ke, part = ke_local, pair_sum_local
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ke_local, pair_sum_local], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        ke, part = (float(v) for v in t.cpu())
    pe = -0.5 * part
    return {"kinetic": ke, "potential": pe, "total": ke + pe}


class ShardedBarnesHutSystem:
    """One rank of a target-sharded Barnes-Hut run over REPLICATED state.

    Every rank holds all N bodies.  Per step (nb_mgpu_bh_step): drift all bodies and rebuild the tree
    (identical on every rank), traverse only the groups dealt to this rank with the accelerations
    stored into every rank's exchange buffer from the traversal kernel's epilogue (NVLink peer stores,
    sorted order), wait for the peers' flags, scatter to body order and kick all bodies.  Positions are
    never exchanged: the replicas evolve bit-identically."""

    def __init__(self, state_full: dict, rank: int, world: int, G: float = 1.0, softening: float = 0.01,
                 theta: float = 0.5, group=None):
        import torch
        from . import _lib
        from .api import ParticleData, ParticleDataManager, get_context
        self.torch, self.L, self.check = torch, _lib.lib(), _lib.check
        self.rank, self.world, self.group = rank, world, group
        self.G = float(np.float32(G))
        self.eps = float(np.float32(softening))
        self.theta = float(np.float32(theta))
        self.ctx = get_context()
        self.n = int(state_full["pos_x"].size)
        h = ParticleDataManager.fromState(state_full)
        self.d = ParticleData()
        ParticleDataManager.allocateDevice(self.d, self.n)
        ParticleDataManager.copyToDevice(self.d, h)
        handle = (C.c_ubyte * 64)()
        self.check(self.L.nb_mgpu_bh_alloc(self.ctx.handle, self.n, rank, world, handle))
        if world > 1:
            import torch.distributed as dist
            handles = [None] * world
            dist.all_gather_object(handles, bytes(handle), group=group)
            for r, hb in enumerate(handles):
                if r != rank:
                    self.check(self.L.nb_mgpu_bh_open(self.ctx.handle, r, (C.c_ubyte * 64).from_buffer_copy(hb)))
            dist.barrier(group=group)
        self.simulation_time = 0.0

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    def compute_forces(self):
        self.check(self.L.nb_mgpu_bh_step(self.ctx.handle, C.byref(self.d.c), 0.0, self.theta, self.G, self.eps, 0,
                                          self._stream()))

    def update(self, dt: float):
        self.check(self.L.nb_mgpu_bh_step(self.ctx.handle, C.byref(self.d.c), float(np.float32(dt)), self.theta,
                                          self.G, self.eps, 1, self._stream()))
        self.simulation_time += dt

    def arrays(self) -> dict:
        return self.d.numpy()

    def close(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier(group=self.group)
        self.check(self.L.nb_mgpu_bh_close(self.ctx.handle))
