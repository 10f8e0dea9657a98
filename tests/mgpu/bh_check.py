"""Run under torchrun on >= 2 GPUs: target-sharded Barnes-Hut over replicated state against the
single-GPU Barnes-Hut step on the same inputs (bit-identical), plus a timing line.
python -m torch.distributed.run --nproc-per-node N tests/mgpu/bh_check.py [bodies] [steps]"""
import json
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
    import nbody_b200 as nb
    from nbody_b200 import ics
    from nbody_b200.dist import ShardedBarnesHutSystem
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200003
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    st = ics.plummer(n, seed=7)
    sysm = ShardedBarnesHutSystem(st, rank, world, 1.0, 0.01, 0.5)
    sysm.compute_forces()
    for _ in range(2):
        sysm.update(1e-3)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        sysm.update(1e-3)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    got = sysm.arrays()
    sysm.close()
    # the same steps on this rank alone
    single = nb.ParticleSystem()
    single.config_ = nb.SimulationConfig(particle_count=n, softening=0.01, dt=1e-3, force_method=nb.ForceMethod.BARNES_HUT)
    single.setState(nb.SimulationState(n, 0.0, 1e-3, 1.0, 0.01, nb.ForceMethod.BARNES_HUT,
                                       {k: st[k] for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "mass")}))
    for _ in range(2):
        single.update(1e-3)
    torch.cuda.synchronize()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for _ in range(steps):
        single.update(1e-3)
    s1.record()
    torch.cuda.synchronize()
    want = single.getDeviceData().numpy()
    for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "acc_x", "acc_y", "acc_z"):
        assert np.array_equal(got[k], want[k]), (rank, k)
    if rank == 0:
        print("bh_check ok: " + json.dumps(dict(world=world, n=n, steps=steps, sharded_ms_per_step=t.item(),
                                                single_ms_per_step=s0.elapsed_time(s1) / steps,
                                                speedup=s0.elapsed_time(s1) / steps / t.item())))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
