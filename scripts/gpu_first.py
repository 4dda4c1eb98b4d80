"""First GPU bring-up: parity of every stage against the oracle, with verbose diagnostics."""
import sys, os, time, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sfm_b200 import synth, api
from oracle import pyoracle as po

def relerr(a, b):
    return float(np.abs(a - b).max() / max(1e-300, np.abs(b).max()))

def main():
    ctx = api.Context(0)
    for name, kw in [("c1", {}), ("c3", {"scale": 0.05}), ("c2", {"scale": 0.1})]:
        p = synth.make_config(name, **kw)
        print(f"== {name} {kw}: cams {p.n_cams} pts {p.n_pts} obs {p.n_obs}", flush=True)
        try:
            r, J, c = ctx.evaluate(p)
            r0, J0, c0 = po.evaluate(p)
            pred = r0 + p.obs
            er = np.abs(r - r0).max(axis=1) / np.maximum(1, np.abs(pred).max(axis=1))
            ej = np.abs(J - J0).reshape(len(J), -1).max(axis=1) / np.abs(J0).reshape(len(J), -1).max(axis=1)
            print(f"   eval: res err {er.max():.3e} jac err {ej.max():.3e} cost rel {abs(c-c0)/c0:.3e}", flush=True)
        except Exception:
            traceback.print_exc()
        try:
            g = ctx.schur(p, radius=1e4)
            o = po.schur(p, radius=1e4)
            for k in ["scale", "Hpp", "gp", "Hcc", "gc", "rhs", "S"]:
                print(f"   schur {k}: rel err {relerr(g[k], o[k]):.3e}", flush=True)
        except Exception:
            traceback.print_exc()
        for ls, nm in [(api.LINEAR_DENSE_SCHUR, "dense"), (api.LINEAR_PCG, "pcg")]:
            try:
                t0 = time.time()
                cam, pts, s = ctx.solve_packed(p, opts=api.default_options(linear_solver=ls))
                dt = time.time() - t0
                _, _, so, tro = po.solve(p, opts=po.default_options(linear_solver=1, num_threads=8))
                print(f"   solve[{nm}]: its {s['num_iterations']} (oracle {so['num_iterations']}) succ {s['num_successful_steps']} term {s['termination_type']} "
                      f"final {s['final_cost']:.10e} (oracle {so['final_cost']:.10e}, rel {abs(s['final_cost']-so['final_cost'])/so['final_cost']:.2e}) "
                      f"lin {s['total_linear_iterations']} solve_ms {s['solve_ms']:.2f} setup_ms {s['setup_ms']:.2f} h2d_ms {s['h2d_ms']:.2f} wall {dt*1e3:.1f} launches {s['gpu_launches']}", flush=True)
                for a, b in zip(s["trace"], tro):
                    print(f"      cost {a['cost']:.9e} | {b['cost']:.9e}  rho {a['relative_decrease']:.4f} | {b['relative_decrease']:.4f} radius {a['trust_region_radius']:.2e} lin {a['linear_iterations']}")
            except Exception:
                traceback.print_exc()
    # modes 0 / 1 on c1
    p = synth.make_config("c1")
    for mode in (0, 1):
        try:
            cam, pts, s = ctx.solve_packed(p, mode=mode)
            _, _, so, _ = po.solve(p, mode=mode)
            print(f"   mode {mode}: its {s['num_iterations']} (oracle {so['num_iterations']}) final {s['final_cost']:.10e} (oracle {so['final_cost']:.10e})", flush=True)
        except Exception:
            traceback.print_exc()
    # kernel timings on c4
    try:
        t0 = time.time(); p = synth.make_config("c4"); print("gen c4 %.1fs" % (time.time() - t0), flush=True)
        ctx.upload(p)
        s = ctx.solve_resident(api.default_options(linear_solver=api.LINEAR_PCG, max_num_iterations=2, pcg_max_iterations=50, pcg_tolerance=1e-3))
        print("   c4 pcg 2 its:", {k: v for k, v in s.items() if k != "trace"}, flush=True)
        for k in ["rj", "rj_form", "schur_matvec", "backsub", "cam_form", "cost"]:
            ms = ctx.time_kernel(k, 10)
            b = ctx.algorithmic_bytes(k)
            print(f"   kernel {k}: {ms*1e3:.1f} us, {b/1e6:.1f} MB alg, {b/ms/1e6:.1f} GB/s", flush=True)
    except Exception:
        traceback.print_exc()

main()
