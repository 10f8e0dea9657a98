"""Run under torchrun on >= 2 GPUs: the peer-memory exchange against the NCCL exchange and the
fp64 oracle on the same inputs.  python -m torch.distributed.run --nproc-per-node N tests/mgpu/peer_check.py [bodies]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from nbody_b200 import ics
    from nbody_b200.dist import ShardedDirectSystem, shard_layout, slice_state
    from oracle import bindings as orc
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 100003
    st = ics.plummer(n, seed=7)
    b, c, _ = shard_layout(n, rank, world)
    out, energy = {}, {}
    for ex in ("peer", "nccl"):
        sysm = ShardedDirectSystem(slice_state(st, b, c), n, rank, world, 1.0, 0.01, exchange=ex)
        sysm.compute_forces()
        torch.cuda.synchronize()
        a0 = {k: v.copy() for k, v in sysm.local_arrays().items()}
        e0 = sysm.compute_energy()
        for _ in range(5):   # odd and even buffer parities, ranks free-running
            sysm.update(1e-3)
        torch.cuda.synchronize()
        energy[ex] = (e0, sysm.compute_energy())   # collective: one all-reduce of two doubles
        out[ex] = (a0, sysm.local_arrays())
        sysm.close()
    idx = np.arange(b, b + c, max(c // 512, 1))
    ref = orc.direct_f64(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, float(np.float32(0.01) ** 2), idx)
    a0 = out["peer"][0]
    err = np.sqrt(sum((a0[k][idx - b].astype(np.float64) - r) ** 2 for k, r in zip(("acc_x", "acc_y", "acc_z"), ref)))
    rel = float((err / np.sqrt(sum(r ** 2 for r in ref))).max())
    assert rel <= 1e-5, rel
    # same kernel, same source order in both exchanges up to the pass split: tiny differences only
    for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z"):
        p, q = out["peer"][1][k], out["nccl"][1][k]
        assert np.allclose(p, q, rtol=0, atol=2e-6 * max(1.0, float(np.abs(q).max()))), k
    # whole-system energy from the sharded diagnostic: same on every rank, equal to the oracle's at t = 0, both
    # exchanges alike, conserved over the five steps
    ke0 = orc.kinetic(st["vel_x"], st["vel_y"], st["vel_z"], st["mass"])
    pe0 = orc.potential(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, 0.01)
    for ex in ("peer", "nccl"):
        e0, e5 = energy[ex]
        assert abs(e0["kinetic"] - ke0) <= 1e-9 * abs(ke0) and abs(e0["potential"] - pe0) <= 2e-6 * abs(pe0), (ex, e0)
        assert abs(e5["total"] - e0["total"]) <= 1e-5 * abs(e0["total"]), (ex, e0, e5)
    assert abs(energy["peer"][1]["total"] - energy["nccl"][1]["total"]) <= 1e-7 * abs(pe0)
    t = torch.tensor([rel], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"peer_check ok: world={world} n={n} max_rel_err_vs_fp64={t.item():.2e} "
              f"E0={energy['peer'][0]['total']:.9f} E5={energy['peer'][1]['total']:.9f}")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
