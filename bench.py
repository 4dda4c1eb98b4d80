#!/usr/bin/env python
"""bench.py — LM bundle-adjustment throughput of the B200-native solver (BASELINE.json metric:
"LM iters/s & residual+Jacobian obs/s at 1/2/4/8 B200 vs Ceres CPU").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c4] [--impl ours|reference]

A *step* is one complete Levenberg–Marquardt solve of the synthetic problem from its perturbed
initial guess with the reference's options (CTracker.cpp:571-577 = Ceres defaults), restarted
from the same guess every step, so every step does identical work.  ``value`` = trust-region
iterations completed per second of device time (max over ranks, inputs resident in HBM);
``e2e`` = the same through the host-buffer C-ABI call (H2D + index build + solve + D2H inside the
timed region).  Rank 0 prints ONE JSON line.

N > 1: launched by torch.distributed.run, one rank per GPU; observations are sharded by point
(SURVEY.md §8e), cameras replicated, NCCL all-reduces of the camera system inside the library.
Total work is fixed as N grows => "scaling": "strong".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# FP64 tensor (DMMA m8n8k4) peak measured on this pool with scripts/micro/fp64_peaks.cu (profiles/r01_fp64_peaks.txt);
# MEASURED_PEAKS.json carries no FP64 figure (nominal: 128 flop/clk/SM x 148 SMs x 1.965 GHz = 37.2 TFLOP/s).
FP64_DMMA_PEAK_TFLOPS = 37.1
# dram__bytes_read.sum + dram__bytes_write.sum per launch: read from the ncu summary committed with the build
# (profiles/ncu_traffic.json, written by scripts/ncu_traffic.py from an `ncu --set full` capture); null when the
# file has no entry for the kernel / configuration.
TRAFFIC_FILE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_traffic.json")


def load_traffic(config, kernel):
    try:
        with open(TRAFFIC_FILE) as f:
            return json.load(f).get(config, {}).get(kernel)
    except Exception:
        return None


def load_parity(config, summ):
    """Oracle table of the full-size configuration (tests/golden/lm_fullsize.npz, made by
    tests/golden/make_fullsize_tables.py on the CPU) against this run's solve."""
    try:
        g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests", "golden", "lm_fullsize.npz"))
        if f"{config}_summary" not in g:
            return None
        t = g[f"{config}_summary"]
        its = int(t[3])
        if config == "c5":  # the CPU table holds the first 3 iterations only (dense S does not fit the host)
            k = min(its, len(summ["trace"]) - 1)
            ref, got = float(g["c5_cost"][k]), float(summ["trace"][k]["cost"])
            return {"oracle_cost_after_iterations": k, "oracle_cost": ref, "cost": got, "rel_diff": abs(got - ref) / ref,
                    "iterations_equal": None, "source": "tests/golden/lm_fullsize.npz (oracle PCG, 3 iterations)"}
        ref = float(t[1])
        return {"oracle_final_cost": ref, "final_cost": summ["final_cost"], "rel_diff": abs(summ["final_cost"] - ref) / ref,
                "iterations_equal": bool(summ["num_iterations"] == its), "oracle_iterations": its,
                "source": "tests/golden/lm_fullsize.npz (CPU oracle, full size)"}
    except Exception:
        return None

METRIC = "lm_iterations_per_s"
UNIT = "LM iterations/s"


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.proc:
            return out
        try:
            self.proc.terminate()
            self.proc.wait(timeout=5)
            self.f.close()
            rows = [l.strip().split(", ") for l in open(self.path) if l.strip()]
            sm = [float(r[1]) for r in rows if len(r) >= 9]
            if sm:
                out["sm_mhz"] = float(np.median(sm))
                out["sm_max_mhz"] = float(rows[0][2])
                out["samples"] = len(sm)
                names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
                for k, nm in enumerate(names):
                    if any(r[5 + k].strip().lower().startswith("active") for r in rows if len(r) >= 9):
                        out["reasons"].append(nm)
        except Exception:
            pass
        finally:
            try:
                os.unlink(self.path)
            except Exception:
                pass
        return out


def run_reference(args, rank, world):
    """Reference arm: the reference's own CPU implementation of the path.  Real Ceres cannot be
    built here, so this is the oracle port (oracle/), all host threads, on a bounded
    sample of the same workload: ONE trust-region iteration of the named problem per step.  The one-time
    set-up of a solve (index build, Jacobi scaling, initial linearisation) is timed separately and NOT
    charged to the iteration: value = iterations / (solve time - set-up time), i.e. the steady-state cost of
    an iteration (linear solve + candidate evaluation + re-linearisation) — what a 9-iteration solve
    amortises its set-up down to.  ``value_including_setup`` carries the other number."""
    if rank != 0:
        return
    from oracle import pyoracle as po
    from sfm_b200 import synth

    prob = synth.make_config(args.config)
    cores = os.cpu_count() or 1
    dense_ok = prob.n_cams <= 4000
    opts = po.default_options(num_threads=cores, max_num_iterations=1, linear_solver=1 if dense_ok else 2,
                              pcg_tolerance=1e-10, pcg_max_iterations=1000)
    times, its, setup = [], 0, 0.0
    for step in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        _, _, s, _ = po.solve(prob, opts=opts)
        dt = time.perf_counter() - t0
        if step >= args.warmup:
            times.append(dt)
            its += s["num_iterations"]
            setup += s["setup_seconds"]
    total = sum(times)
    value = its / max(total - setup, 1e-9)
    sample = (f"1 LM iteration (dual-number evaluation + {'dense Schur + blocked Cholesky' if dense_ok else 'implicit-Schur block-Jacobi PCG'}"
              f" + candidate evaluation) of {args.config} per step, {cores} OpenMP threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / max(1, args.steps), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": synth.CONFIG_TITLES[args.config], "config_id": args.config, "n_cams": prob.n_cams, "n_pts": prob.n_pts,
                   "n_obs": prob.n_obs},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "value_including_setup": its / total, "setup_seconds_per_step": setup / max(1, args.steps),
        "note": "restated Ceres-1.12 semantics (oracle/), not Ceres itself: the reference cannot be compiled in this image; "
                "one-time set-up of each 1-iteration solve excluded from value (see value_including_setup)",
    }
    _emit(line)


_JSON_FD = None


def _claim_stdout():
    """Everything any library writes to file descriptor 1 (NCCL's version banner, for one) goes to stderr from here
    on; the one JSON line is written to the real stdout by _emit."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def _emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="c4", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--linear", default="auto", choices=["auto", "dense", "pcg"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    from sfm_b200 import api, synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the solver has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    prob = synth.make_config(args.config)
    n_obs_total, n_cams, n_pts_total = prob.n_obs, prob.n_cams, prob.n_pts
    full = prob
    if world > 1:
        prob = synth.shard_by_point(full, world)[rank][0]

    ctx = api.Context(local_rank)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    if world > 1:
        uid = [api.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], rank, world)

    linear = {"auto": api.LINEAR_AUTO, "dense": api.LINEAR_DENSE_SCHUR, "pcg": api.LINEAR_PCG}[args.linear]
    opts = api.default_options(linear_solver=linear)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident arm -------------------------------------------------------------------
    ctx.upload(prob)
    summ = None
    for _ in range(args.warmup):
        ctx.reset()
        summ = ctx.solve_resident(opts)
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    its = launches = 0
    e0.record(stream)
    for _ in range(args.steps):
        ctx.reset()
        summ = ctx.solve_resident(opts)
        its += summ["num_iterations"]
        launches += summ["gpu_launches"]
    e1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    value = its / (ms * 1e-3)

    # ---- kernel timings / roofline (rank 0's shard; live CUDA events inside the library) ---------
    peak, peak_src = load_peaks()
    kernels = {}
    dense = summ["linear_solver_used"] == api.LINEAR_DENSE_SCHUR
    for k in ["rj_form", "rj", "schur_matvec", "schur_rhs", "backsub", "cam_form", "cost"] + (["dense_schur"] if dense else []):
        t = ctx.time_kernel(k, 20 if k != "dense_schur" else 5)
        b = ctx.algorithmic_bytes(k)
        kernels[k] = {"ms": t, "algorithmic_bytes": b, "gbs": b / t / 1e6, "frac_of_hbm_peak": b / t / 1e6 / peak}
    rj = kernels["rj"]
    obs_per_s_local = prob.n_obs / (rj["ms"] * 1e-3)
    rj_obs = torch.tensor([obs_per_s_local], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(rj_obs, op=dist.ReduceOp.SUM)
    chol = None
    if dense and rank == 0 and 6 * n_cams <= 12288:  # dense tap: same kernels on a FULL n x n matrix (not for envelope-only sizes)
        n6 = 6 * n_cams
        f_ms, s_ms, res = ctx.bench_cholesky(n6, 3)
        chol = {"n": n6, "factor_ms": f_ms, "solve_ms": s_ms, "flop": n6 ** 3 / 3.0, "tflops": n6 ** 3 / 3.0 / f_ms * 1e-9,
                "frac_of_fp64_dmma_peak": n6 ** 3 / 3.0 / f_ms * 1e-9 / FP64_DMMA_PEAK_TFLOPS, "solve_residual": res}

    # ---- end-to-end arm: host buffers through the C ABI ----------------------------------------
    e2e = None
    if not args.no_e2e:
        pin = lambda a: torch.from_numpy(a).pin_memory().numpy()  # noqa: E731
        hp = synth.Problem(obs=pin(prob.obs), cam_idx=pin(prob.cam_idx), pt_idx=pin(prob.pt_idx), K=pin(prob.K), cam=pin(prob.cam),
                           pts=pin(prob.pts), cam_true=prob.cam_true, pts_true=prob.pts_true, dist=None, name=prob.name)
        h2d = hp.obs.nbytes + hp.cam_idx.nbytes + hp.pt_idx.nbytes + hp.K.nbytes + hp.cam.nbytes + hp.pts.nbytes
        d2h = hp.cam.nbytes + hp.pts.nbytes
        cam0, pts0 = prob.cam.copy(), prob.pts.copy()
        ctx.solve_packed(hp, opts=opts, inplace=True)
        e2e_its, e2e_s = 0, 0.0
        for _ in range(args.steps):
            np.copyto(hp.cam, cam0); np.copyto(hp.pts, pts0)  # restore the initial guess in the caller's (pinned) storage: not timed
            barrier()
            t0 = time.perf_counter()
            _, _, s2 = ctx.solve_packed(hp, opts=opts, inplace=True)  # H2D + index build + solve + D2H, in place like the reference
            torch.cuda.synchronize()
            e2e_s += time.perf_counter() - t0
            e2e_its += s2["num_iterations"]
        barrier()
        dt = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": e2e_its / float(dt.item()), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "ms_per_step": 1e3 * float(dt.item()) / args.steps}

    # ---- the reference-shaped call: hsfm_ba_solve_ptr from PAGEABLE storage laid out as CSfM::bundleAdjustment builds it
    #      (CSfM.cpp:311-343: vector<Point2d>, vector<int>, vector<Matx33d>, and double* into the keyframes' _rot/_t and
    #      the map's _pts3D — the same pointer for every observation of a point) --------------------------------------
    e2e_ptr = None
    if not args.no_e2e and world == 1:
        frames = np.ascontiguousarray(prob.cam.copy())           # CKeyFrame::_rot.val / _t.val stand-in: [n_cams][6]
        store = np.ascontiguousarray(prob.pts.copy())            # CMap::_pts3D stand-in
        obs_v, cam_v, K_v = prob.obs.copy(), prob.cam_idx.copy(), prob.K.copy()
        R_ptrs = frames.ctypes.data + 48 * np.arange(n_cams, dtype=np.uint64)
        t_ptrs = R_ptrs + np.uint64(24)
        pt_ptrs = store.ctypes.data + np.uint64(24) * prob.pt_idx.astype(np.uint64)
        ctx.solve_ptr(obs_v, cam_v, K_v, R_ptrs, t_ptrs, pt_ptrs, opts=opts)
        p_its, p_s = 0, 0.0
        for _ in range(args.steps):
            np.copyto(frames, prob.cam); np.copyto(store, prob.pts)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            s3 = ctx.solve_ptr(obs_v, cam_v, K_v, R_ptrs, t_ptrs, pt_ptrs, opts=opts)
            torch.cuda.synchronize()
            p_s += time.perf_counter() - t0
            p_its += s3["num_iterations"]
        e2e_ptr = {"value": p_its / p_s, "unit": UNIT, "ms_per_step": 1e3 * p_s / args.steps, "final_cost": s3["final_cost"],
                   "h2d_bytes_per_step": int(obs_v.nbytes + cam_v.nbytes + 8 * prob.n_obs + K_v.nbytes + frames.nbytes + store.nbytes),
                   "d2h_bytes_per_step": int(frames.nbytes + store.nbytes + 8 * n_pts_total),
                   "note": "hsfm_ba_solve_ptr (the drop-in signature of CTracker.h:65) on pageable host storage; pointer identity -> point ids on the device"}

    # ---- CPU baseline (rank 0, N = 1 only): oracle port on a bounded sample ---------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import pyoracle as po

        cores = os.cpu_count() or 1
        dense_ok = n_cams <= 4000
        oo = po.default_options(num_threads=cores, max_num_iterations=1, linear_solver=1 if dense_ok else 2)
        t0 = time.perf_counter()
        _, _, so, _ = po.solve(full, opts=oo)
        dt = time.perf_counter() - t0
        cpu = {"value": so["num_iterations"] / max(dt - so["setup_seconds"], 1e-9), "unit": UNIT, "cores": cores, "kind": "port",
               "value_including_setup": so["num_iterations"] / dt,
               "sample": f"1 LM iteration of {args.config} ({'dense Schur + Cholesky' if dense_ok else 'implicit-Schur PCG'}), {cores} OpenMP threads, "
                         f"{dt:.1f} s of which {so['setup_seconds']:.1f} s one-time set-up (not charged to the iteration)",
               "eval_obs_per_s": 2 * n_obs_total / max(so["eval_seconds"], 1e-9)}

    if rank == 0:
        phases = dict(zip(["linearise", "reduced_system", "factor", "reduced_solve", "step_backsub_cost", "readback"], summ["phase_ms"][:6]))
        # The bandwidth-bound kernels of one LM iteration (north star: residual/Jacobian and Schur kernels).  The
        # headline roofline is the residual+Jacobian kernel; the others are listed in "kernels".
        hbm_dom = "rj_form"
        traffic = load_traffic(args.config, hbm_dom) if world == 1 else None
        roof_hbm = {"kernel": "rj_kernel<write_j,form>", "bound": "hbm", "achieved": kernels[hbm_dom]["gbs"],
                    "peak": peak, "unit": "GB/s", "frac": kernels[hbm_dom]["gbs"] / peak, "traffic": traffic, "peak_source": peak_src}
        step_ms = max(sum(phases.values()), 1e-9)
        its_n = max(summ["num_iterations"], 1)
        roof_hbm["share_of_step"] = (kernels[hbm_dom]["ms"] * its_n) / step_ms
        roofline = dict(roof_hbm)
        roof_tensor = None
        if chol:
            # dense tap: the same factorisation kernels on a FULL n x n matrix (no envelope to exploit): FP64 tensor bound
            roof_tensor = {"kernel": "chol_syrk_kernel (dense Cholesky tap, DMMA m8n8k4)", "bound": "tensor", "achieved": chol["tflops"],
                           "peak": FP64_DMMA_PEAK_TFLOPS, "unit": "TFLOP/s", "frac": chol["frac_of_fp64_dmma_peak"], "traffic": None,
                           "peak_source": "FP64 DMMA peak measured with scripts/micro/fp64_peaks.cu on this pool (MEASURED_PEAKS.json has no FP64 figure)"}
        # where the step goes: with the envelope factorisation the largest phase is the CRITICAL PATH of the
        # per-panel diagonal-block factorisations (latency-bound, one CTA), not a throughput kernel
        dominant = max(phases, key=lambda k: phases[k])
        if dominant == "factor" and summ.get("chol_flops", 0.0) > 0.0:
            # the headline roofline describes what dominates the step: the structured Cholesky of the reduced camera system
            # (a graph of chol_diag / chol_chain_tile / chol_trsm_strip / chol_syrk launches per factorisation, FP64 DMMA)
            f_ms = phases["factor"] / its_n
            tf = summ["chol_flops"] / (f_ms * 1e-3) * 1e-12
            roofline = {"kernel": "structured Cholesky of the reduced camera system (chol_diag / chol_chain_tile / chol_trsm_strip / chol_syrk, DMMA m8n8k4)",
                        "bound": "tensor", "achieved": tf, "peak": FP64_DMMA_PEAK_TFLOPS, "unit": "TFLOP/s", "frac": tf / FP64_DMMA_PEAK_TFLOPS,
                        "traffic": None, "share_of_step": phases["factor"] / step_ms, "ms_per_factorisation": f_ms,
                        "flops_per_factorisation": summ["chol_flops"], "panels": summ["chol_panels"], "critical_panels": summ["chol_critical_panels"], "chains": summ["chol_chains"],
                        "peak_source": "FP64 DMMA peak measured with scripts/micro/fp64_peaks.cu on this pool (MEASURED_PEAKS.json has no FP64 figure)",
                        "note": "concurrent panel chains along the block elimination tree: critical_panels serial panel steps of ~38 us each, the chains' trailing updates share the tensor pipes beside them"}
        dominant_phase = {"phase": dominant, "share_of_step": phases[dominant] / step_ms,
                          "note": "factor = envelope Cholesky of the reduced camera system: a chain of 128-column panel steps "
                                  "(diagonal block on one CTA -> panel solve -> next-column update); latency-bound" if dominant == "factor" else ""}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": synth.CONFIG_TITLES[args.config], "config_id": args.config, "n_cams": n_cams, "n_pts": n_pts_total,
                       "n_obs": n_obs_total, "parallelism": f"point-sharded x{world}", "l2": "inputs larger than L2" if n_obs_total * 184 > 126e6 else "L2-resident working set (latency-bound; no flush)",
                       "linear_solver": {1: "dense_schur", 2: "pcg"}[summ["linear_solver_used"]], "lm_iterations_per_step": summ["num_iterations"],
                       "final_cost": summ["final_cost"], "rms_px": summ["rms_px"], "linear_iterations_per_step": summ["total_linear_iterations"]},
            "parity": load_parity(args.config, summ),
            "rj_obs_per_s": float(rj_obs.item()),
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roofline,
            "roofline_hbm": roof_hbm,
            "roofline_tensor": roof_tensor,
            "dominant_phase": dominant_phase,
            "phase_ms_per_step": phases,
            "kernels": kernels,
            "cholesky": chol,
            "e2e": e2e,
            "e2e_ptr": e2e_ptr,
            "cpu_baseline": cpu,
        }
        _emit(line)
    if world > 1:
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
