#!/usr/bin/env python
"""The reference's driver loop (GrayScott.main, src/GrayScott.jl:68-103) on the B200 backend, for the
Python harness:   python examples/gray_scott.py examples/settings-b200.toml
                  torchrun --nproc-per-node N examples/gray_scott.py cfg.toml      (one rank per GPU)
ADIOS2 is not available here, so write_step! becomes one raw .npy pair per rank and output step
(ghost-free, [z, y, x]); everything else -- settings, decomposition, iterate!, plotgap -- is the reference's."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402


def main(args):
    sim = entry.load_package().simulation
    comm, settings, mcd, fields = sim.initialization(args)          # communication.jl:15-33
    rank = comm["rank"]
    os.makedirs(settings.output, exist_ok=True)
    step = 0
    while step < settings.steps:                                     # GrayScott.jl:81-96
        sim.iterate(fields, settings, mcd)
        step += 1
        if step % settings.plotgap == 0:
            if rank == 0 and settings.verbose:
                print("Simulation at step", step, "writing output step", step // settings.plotgap)
            u, v = sim.get_fields(fields)                            # IO.write_step! -> get_fields
            np.save(os.path.join(settings.output, f"U_step{step:06d}_rank{rank}.npy"), u)
            np.save(os.path.join(settings.output, f"V_step{step:06d}_rank{rank}.npy"), v)
    fields.close()
    sim.finalize()


if __name__ == "__main__":
    main(sys.argv[1:])
