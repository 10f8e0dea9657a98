"""CPU, world_size = 2, gloo: the host-side logic of the target-sharded direct-N^2 path -- shard
layout, packing into the replicated source buffer, the all-gather, and the local/remote split --
with the oracle standing in for the force kernel."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from nbody_b200 import ics
    from nbody_b200.dist import pack_sources_host, shard_layout, slice_state
    from oracle import bindings as orc

    G, eps2, dt = 1.5, 1e-4, np.float32(1e-3)
    full = ics.plummer(n, seed=13)
    b, c, pad = shard_layout(n, rank, world)
    loc = slice_state(full, b, c)
    acc_hist = []
    for step in range(3):
        if step > 0:   # store-acc + drift of the local slice (the pack kernel's fused part)
            orc.store_acc(loc)
            orc.drift(loc, dt)
        mine = torch.from_numpy(pack_sources_host(loc, pad, G).reshape(-1))
        gathered = torch.zeros(world * pad * 4, dtype=torch.float32)
        dist.all_gather_into_tensor(gathered, mine)
        src = gathered.numpy().reshape(world * pad, 4)
        # every rank must now hold every slice, padding = zero mass
        px, py, pz, gm = (np.ascontiguousarray(src[:, k]) for k in range(4))
        # targets = own slot of the gathered buffer; sources = all slots (local first, then remote)
        t0 = rank * pad
        ax, ay, az = orc.direct_f32(px, py, pz, gm, 1.0, eps2, t0, t0 + c)  # G already folded into the mass
        loc["acc_x"], loc["acc_y"], loc["acc_z"] = (np.ascontiguousarray(a[t0:t0 + c]) for a in (ax, ay, az))
        if step > 0:
            orc.kick(loc, dt)
        acc_hist.append(np.stack([loc["acc_x"], loc["acc_y"], loc["acc_z"]]))
    # energy diagnostic of the sharded run: local KE + local share of the pair sum against the
    # replicated sources, one all-reduce of two doubles (numpy fp64 stands in for nb_mgpu_potential_partial)
    from nbody_b200.dist import reduce_energy
    x = src[:, :3].astype(np.float64)
    gm64 = src[:, 3].astype(np.float64)
    part = 0.0
    for li in range(c):
        r2 = ((x - x[t0 + li]) ** 2).sum(1) + eps2
        inv = gm64 / np.sqrt(r2)
        inv[t0 + li] = 0.0
        part += float(loc["mass"][li]) * inv.sum()
    ke_local = orc.kinetic(loc["vel_x"], loc["vel_y"], loc["vel_z"], loc["mass"])
    energy = reduce_energy(ke_local, part, world, None, torch, device="cpu")
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), begin=b, count=c, acc=np.stack(acc_hist),
             pos=np.stack([loc["pos_x"], loc["pos_y"], loc["pos_z"]]),
             vel=np.stack([loc["vel_x"], loc["vel_y"], loc["vel_z"]]),
             energy=np.array([energy["kinetic"], energy["potential"], energy["total"]]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [1000, 5003])
def test_sharded_direct_matches_single_process(tmp_path, n):
    import torch.multiprocessing as mp
    from nbody_b200 import ics
    from oracle import bindings as orc
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    # single-process reference: same three evaluations / two steps on the full system
    G, eps2, dt = np.float32(1.5), 1e-4, np.float32(1e-3)
    full = ics.plummer(n, seed=13)
    full["acc_x"], full["acc_y"], full["acc_z"] = orc.direct_f32(full["pos_x"], full["pos_y"], full["pos_z"],
                                                                 G * full["mass"], 1.0, eps2)
    a0 = np.stack([full["acc_x"], full["acc_y"], full["acc_z"]]).copy()
    for _ in range(2):
        orc.store_acc(full)
        orc.drift(full, dt)
        full["acc_x"], full["acc_y"], full["acc_z"] = orc.direct_f32(full["pos_x"], full["pos_y"], full["pos_z"],
                                                                     G * full["mass"], 1.0, eps2)
        orc.kick(full, dt)
    covered = 0
    for r in range(world):
        z = np.load(tmp_path / f"rank{r}.npz")
        b, c = int(z["begin"]), int(z["count"])
        covered += c
        sl = slice(b, b + c)
        # padding sources carry zero mass, so the sums agree to rounding of the (different) j order
        np.testing.assert_allclose(z["acc"][0], a0[:, sl], rtol=2e-5, atol=1e-6)
        np.testing.assert_allclose(z["pos"], np.stack([full["pos_x"][sl], full["pos_y"][sl], full["pos_z"][sl]]),
                                   rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(z["vel"], np.stack([full["vel_x"][sl], full["vel_y"][sl], full["vel_z"][sl]]),
                                   rtol=1e-4, atol=1e-6)
    assert covered == n
    # every rank holds the same whole-system energy, equal to the single-process diagnostic of the final state
    ke = orc.kinetic(full["vel_x"], full["vel_y"], full["vel_z"], full["mass"])
    pe = orc.potential(full["pos_x"], full["pos_y"], full["pos_z"], full["mass"], float(G), float(np.sqrt(eps2)))
    for r in range(world):
        e = np.load(tmp_path / f"rank{r}.npz")["energy"]
        assert abs(e[0] - ke) <= 1e-5 * abs(ke) and abs(e[1] - pe) <= 1e-5 * abs(pe)
        assert abs(e[2] - (ke + pe)) <= 1e-5 * abs(pe)
