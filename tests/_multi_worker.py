"""One rank of a multi-GPU run (launched by torch.distributed.run from tests/test_gpu_multi.py):
one process per GPU, faces over NCCL, the 128-byte NCCL id broadcast over gloo."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402


def main():
    L, prec, noise, steps, out_dir = int(sys.argv[1]), sys.argv[2], float(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    sim = entry.load_package().simulation
    s = sim.Settings(L=L, precision=prec, backend="B200", noise=noise, Du=0.2, Dv=0.1, F=0.02, k=0.048, dt=1.0)
    mcd = sim.init_domain(s, rank, world)
    f = sim.init_fields(s, mcd, prec, device=local, seed=9)
    sim.connect_nccl(f, sim.broadcast_unique_id(rank))
    # a few steps piecewise (exchange!, calculate!, swap), the rest through iterate!
    for _ in range(2):
        sim.exchange(f, mcd)
        sim.calculate(f, s, mcd)
        sim.swap(f)
    f.timers_enable(True)
    sim.iterate(f, s, mcd, nsteps=steps - 2)
    t = f.timers()
    u, v = sim.get_fields(f)
    np.save(os.path.join(out_dir, f"u{rank}.npy"), u)
    np.save(os.path.join(out_dir, f"v{rank}.npy"), v)
    np.save(os.path.join(out_dir, f"meta{rank}.npy"),
            np.array(list(mcd.proc_offsets) + list(mcd.proc_sizes) + [f.halo_bytes()], dtype=np.int64))
    print(f"rank {rank}: halo {t['halo_ms']:.3f} ms interior {t['interior_ms']:.3f} ms shell {t['shell_ms']:.3f} ms", flush=True)
    f.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
