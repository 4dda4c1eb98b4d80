#!/usr/bin/env python
"""Per-iteration {cost, radius, accepted} tables of the CPU oracle on the FULL-SIZE BASELINE.json configs
(VERDICT r1 item 3b): the configurations bench.py times have their LM trajectory pinned in the repository,
not only their shrunken copies.

    python tests/golden/make_fullsize_tables.py [c2 c3 c4 c5]      (rewrites tests/golden/lm_fullsize.npz)

c2-c4: the reference's own choice, DENSE_SCHUR with a dense Cholesky (c3/c4: ~3 min each on 8 cores).
c5: dense S is 54 GB, so (SURVEY.md 8d "C5 deviation") the oracle runs implicit-Schur block-Jacobi PCG for a
FIXED 3 LM iterations; the GPU test runs the same 3 iterations.  Test infrastructure; runs on CPU only."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden", "lm_fullsize.npz")


def main():
    from oracle import pyoracle as po
    from sfm_b200 import synth

    names = sys.argv[1:] or ["c2", "c3", "c4"]
    data = dict(np.load(OUT)) if os.path.exists(OUT) else {}
    threads = os.cpu_count() or 1
    for name in names:
        prob = synth.make_config(name)
        if name == "c5":
            opts = po.default_options(num_threads=threads, linear_solver=2, max_num_iterations=3, pcg_tolerance=1e-10, pcg_max_iterations=1000)
        else:
            opts = po.default_options(num_threads=threads, linear_solver=1)
        t0 = time.time()
        _, _, s, tr = po.solve(prob, opts=opts, trace_capacity=64)
        dt = time.time() - t0
        data[f"{name}_cost"] = np.array([t["cost"] for t in tr])
        data[f"{name}_radius"] = np.array([t["trust_region_radius"] for t in tr])
        data[f"{name}_valid"] = np.array([t["step_is_valid"] for t in tr], dtype=np.int8)
        data[f"{name}_successful"] = np.array([t["step_is_successful"] for t in tr], dtype=np.int8)
        data[f"{name}_model_cost_change"] = np.array([t["model_cost_change"] for t in tr])
        data[f"{name}_linear_iterations"] = np.array([t["linear_iterations"] for t in tr], dtype=np.int32)
        data[f"{name}_summary"] = np.array([s["initial_cost"], s["final_cost"], s["rms_px"], s["num_iterations"], s["num_successful_steps"],
                                            s["termination_type"], prob.n_obs, prob.n_cams, prob.n_pts, dt, threads])
        print(f"{name}: {s['num_iterations']} iterations ({s['num_successful_steps']} successful), final cost {s['final_cost']!r}, "
              f"rms {s['rms_px']!r}, {dt:.1f} s on {threads} threads", flush=True)
        np.savez(OUT, **data)


if __name__ == "__main__":
    main()
