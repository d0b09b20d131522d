#!/usr/bin/env python
"""Generates the golden vectors of tests/golden/ from the CPU oracle (oracle/gs_oracle.c).

The reference itself is Julia and cannot run in this image, and its own tests hold no golden vector
for calculate!/exchange!, so these fixtures pin the ORACLE's output (which is in
turn pinned by the known answers of SURVEY Appendix C and the numpy-scalar restatement in
tests/test_oracle.py).  They let the GPU tests compare against committed bytes without running the
oracle, and they catch any accidental change of the oracle itself.

    python tests/golden/make_golden.py        # rewrites tests/golden/*.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import __graft_entry__ as entry  # noqa: E402

PHYS = dict(Du=0.2, Dv=0.1, F=0.02, k=0.048, dt=1.0)   # examples/settings-files.toml

# name -> (L, precision, steps, noise, seed, dims)
CASES = {
    "L16_f32_n0_s25": (16, "Float32", 25, 0.0, 0, (1, 1, 1)),
    "L16_f64_n0_s25": (16, "Float64", 25, 0.0, 0, (1, 1, 1)),
    "L16_f32_noise_s25": (16, "Float32", 25, 0.1, 0x5EED, (1, 1, 1)),
    "L16_f64_noise_s25": (16, "Float64", 25, 0.1, 0x5EED, (1, 1, 1)),
    "L24_f32_n0_s40_ranks8": (24, "Float32", 40, 0.0, 0, (2, 2, 2)),
}


def main():
    orc = entry.load_oracle()
    orc.build()
    for name, (L, prec, steps, noise, seed, dims) in CASES.items():
        o = orc.Oracle(L, prec, dims=dims, noise=noise, seed=seed, **PHYS)
        o.step(steps)
        u, v = o.gather()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), u=u, v=v, L=L, steps=steps, noise=noise, seed=seed,
                            precision=prec, dims=np.array(dims))
        print(name, u.dtype, float(u.mean()), float(v.mean()))


if __name__ == "__main__":
    main()
