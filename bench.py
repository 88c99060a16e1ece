#!/usr/bin/env python
"""bench.py -- the headline benchmark: power iterations per second on the B200 path.

    python bench.py --gpus N --steps K --warmup W [--workload c5|c2|c3|c4] [--impl reference]

A "step" is ONE power iteration (one trip of src/power_method/power_method.hpp:68-96 of the
reference) on the BASELINE.json workload; default workload is the one the metric's target is quoted
on, config 5 (3D 7-point Laplacian 256^3, fp64 CSR), which fits one GPU.  Prints ONE JSON line
(contract in the task statement): `value` = device-resident iterations/s (CUDA events on the launching
stream, max over ranks), `e2e` = the same metric through the reference-facing call
powerMethod(Matrix, SolverOptions) with HOST buffers (matrix + x0 uploaded, eigenvector downloaded,
all inside the timed region), `roofline` for the matvec kernel, `cpu_baseline` = the reference's own
powerMethod (oracle/_ref, compiled from the unmodified reference) timed on this box's host cores.

Under torchrun (N > 1) the matrix is row-sharded, one rank per GPU; total work is fixed ("strong").
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "c2": "2D 5-point Laplacian 1024x1024 (1.05M rows, 5.2M nnz) fp64 CSR [BASELINE configs[1]]",
    "c3": "dense 8192x8192 fp64 random symmetric, GEMV path [BASELINE configs[2]]",
    "c4": "complex Hermitian 2D Laplacian 2048x2048, random phases (4.2M rows) c128 CSR [BASELINE configs[3]]",
    "c5": "3D 7-point Laplacian 256^3 (16.7M rows, 117M nnz) fp64 CSR [BASELINE configs[4]]",
}
DTYPES = {"c2": "f64", "c3": "f64", "c4": "c128", "c5": "f64"}


def workload_rows(name, gen, lo=None, hi=None):
    """CSR row block [lo, hi) (global columns) of a sparse workload; returns (n, rowptr, colidx, vals)."""
    if name == "c2":
        g = int(os.environ.get("BENCH_GRID", 1024))
        n = g * g
        return (n,) + gen.laplacian_2d(g, lo or 0, n if hi is None else hi)
    if name == "c4":
        g = int(os.environ.get("BENCH_GRID", 2048))
        n = g * g
        return (n,) + gen.hermitian_laplacian_2d(g, 1234, False, lo or 0, n if hi is None else hi)
    if name == "c5":
        g = int(os.environ.get("BENCH_GRID", 256))
        n = g ** 3
        return (n,) + gen.laplacian_3d(g, lo or 0, n if hi is None else hi)
    raise ValueError(name)


def workload_order(name):
    g = os.environ.get("BENCH_GRID")
    return {"c2": int(g or 1024) ** 2, "c3": int(g or 8192), "c4": int(g or 2048) ** 2, "c5": int(g or 256) ** 3}[name]


def workload_nnz(name, n):
    g = int(os.environ.get("BENCH_GRID", 0))
    if name == "c2":
        g = g or 1024
        return 5 * n - 4 * g
    if name == "c4":
        g = g or 2048
        return 5 * n - 4 * g
    if name == "c5":
        g = g or 256
        return 7 * n - 6 * g * g
    return n * n


def workload_config(name):
    """`config` of the JSON line: identical in both arms (the launch geometry of the GPU arm is under `launch`)."""
    n = workload_order(name)
    return {"workload": WORKLOADS[name], "n": n, "nnz": workload_nnz(name, n),
            "l2": "inputs larger than L2 (1.74 GB per matvec)" if name == "c5" else
                  ("matrix + vectors larger than L2" if name in ("c3", "c4") else "working set ~ L2 size: effective GB/s"),
            "x0": "Eigen Random after srand(1)", "tolerance": -1.0}


def golden_lambda(name, trips):
    """Eigenvalue of the compiled reference after `trips` fixed trips (tests/golden/bench_lambda.json), or None."""
    try:
        t = json.load(open(os.path.join(ROOT, "tests", "golden", "bench_lambda.json")))["lambda"].get(f"{name}_{trips}")
        return complex(t[0], t[1]) if t else None
    except Exception:
        return None


def parity_of(lam, ref, trips, source):
    lam, ref = complex(lam), complex(ref)
    return {"lambda_rel_err": abs(lam - ref) / abs(ref), "trips": trips, "lambda": repr(lam), "reference_lambda": repr(ref),
            "reference": source, "tolerance": 1e-10, "ok": bool(abs(lam - ref) <= 1e-10 * abs(ref))}


def alg_bytes_global(name, n, nnz):
    ss = 16 if DTYPES[name] == "c128" else 8
    if name == "c3":
        return ss * n * n + 2 * ss * n
    return nnz * (ss + 4) + 4 * (n + 1) + 2 * ss * n


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "50", "-i", str(index)], stdout=subprocess.PIPE,
                                      stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
            out, _ = self.p.communicate()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [t.strip() for t in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
                pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def recorded_traffic(name, nranks):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu capture, if any."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return t.get(f"{name}_n{nranks}")
    except Exception:
        return None


# --------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's own powerMethod on host cores (single-threaded by
# construction: eigen/Eigen/src/SparseCore/SparseDenseProduct.h:94-111 has no OpenMP branch).
# --------------------------------------------------------------------------------------------------
def reference_rate(name, budget_s, max_trips, warmup=1):
    """iterations/s of EigSol::powerMethod (oracle/_ref) on the full workload matrix, same x0 seed,
    tolerance = -1 (fixed trips).  The sample is `trips` reference iterations of the FULL matrix;
    x0 generation is subtracted with a 0-trip call.  `warmup` untimed trips run first."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _oracle import REF_SO, Reference
    from pcsc_eigenvalue_solver_project_b200 import generators as gen
    if not os.path.exists(REF_SO):
        raise RuntimeError("oracle/_ref/libeigsol_ref.so missing (built by __graft_entry__.build() where /root/reference exists)")
    ref = Reference()
    if name == "c3":
        n = workload_order(name)
        a = gen.dense_random_symmetric(n)
        h = ref.create_dense(a)
        dt = np.float64
        del a
    else:
        n, rp, ci, v = workload_rows(name, gen)
        outer, inner, vals = gen.csr_to_csc_hermitian(rp, ci, v)
        h = ref.create_csc((n, n), outer, inner, np.ascontiguousarray(vals))
        dt = v.dtype
        del rp, ci, v, outer, inner, vals
    t0 = ref.power(h, dt, n, 1, 0, -1.0, want_x=False)[5]
    t1 = ref.power(h, dt, n, 1, 1, -1.0, want_x=False)[5]
    per_trip = max(t1 - t0, 1e-9)
    trips = int(max(1, min(max_trips, budget_s / per_trip)))
    if warmup > 1:  # the 1-trip probe above already is one warm-up trip
        ref.power(h, dt, n, 1, int(min(warmup - 1, max(1, budget_s / 4 / per_trip))), -1.0, want_x=False)
    st, lam, _, iters, _, sec = ref.power(h, dt, n, 1, trips, -1.0, want_x=False)
    ref.destroy(h)
    assert st == 0 and iters == trips
    elapsed = max(sec - t0, 1e-9)
    flags = open(os.path.join(os.path.dirname(REF_SO), "FLAGS.txt")).read().strip() if os.path.exists(
        os.path.join(os.path.dirname(REF_SO), "FLAGS.txt")) else "?"
    return {"value": trips / elapsed, "unit": "iter/s", "cores": 1, "kind": "reference",
            "sample": f"{trips} trips of EigSol::powerMethod on the full matrix (2 matvecs per reference trip), "
                      f"x0 generation ({t0:.2f}s) subtracted; host has {os.cpu_count()} cores, the reference path is "
                      f"single-threaded; built with: {flags}",
            "lambda": repr(lam), "trips": trips, "seconds": elapsed}


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    cfg = workload_config(args.workload)
    try:
        cb = reference_rate(args.workload, budget_s=float(os.environ.get("BENCH_REF_BUDGET_S", 60)),
                            max_trips=max(1, args.steps), warmup=args.warmup)
    except Exception as e:  # the oracle always exists; this is a packaging error worth seeing
        emit({"impl": "reference", "unavailable": f"{type(e).__name__}: {e}"})
        return
    line = {"impl": "reference", "metric": "power_iterations_per_s", "value": cb["value"], "unit": "iter/s",
            "n_gpus": args.gpus, "steps": cb["trips"], "warmup": args.warmup, "ms_per_step": 1e3 / cb["value"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": DTYPES[args.workload],
            "data": "synthetic", "config": cfg,
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": "iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "lambda": cb["lambda"]}
    emit(line)


# --------------------------------------------------------------------------------------------------
_REAL_STDOUT = None


def quiet_stdout():
    """Library chatter (NCCL prints its version banner on stdout) must not reach the ONE JSON line the
    contract asks for: fd 1 is pointed at stderr for the run and emit() writes to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        os.write(1, data)
    else:
        os.write(_REAL_STDOUT, data)


def e2e_leg(name, steps, ngpu, host=None):
    """The reference-facing call with HOST buffers, timed end to end: powerMethod(Matrix, SolverOptions{steps, -1}) in
    ONE process -- matrix + x0 uploaded, CSC -> CSR, `steps` trips, eigenvector downloaded.  With ngpu > 1 the library
    shards the call over the first ngpu devices itself (EIGSOL_B200_NGPU).  One untimed 1-trip call of the same API runs
    first (CUDA contexts of the group's devices, pinned staging buffers and the device memory pool are warm, as they are
    for any caller's second solve); each call gets its own Matrix (a Matrix keeps its device copy, and the timed call
    must upload)."""
    import pcsc_eigenvalue_solver_project_b200 as es
    from pcsc_eigenvalue_solver_project_b200 import generators as gen
    n = workload_order(name)
    dt = np.complex128 if DTYPES[name] == "c128" else np.float64
    x0 = es.random_unit(n, dt, seed=1)
    if name == "c3":
        a = host["a"] if host else gen.dense_random_symmetric(n)
        mk = lambda: es.Matrix(a)
        h2d = a.nbytes + x0.nbytes
    else:
        if host and "rp" in host and host.get("whole"):
            rp, ci, v = host["rp"], host["ci"], host["v"]
        else:
            _, rp, ci, v = workload_rows(name, gen)
        outer, inner, vals = gen.csr_to_csc_hermitian(rp, ci, v)
        del rp, ci, v
        mk = lambda: es.Matrix(((n, n), outer, inner, vals))
        h2d = outer.nbytes + inner.nbytes + vals.nbytes + x0.nbytes
    os.environ["EIGSOL_B200_NGPU"] = str(ngpu)
    try:
        for _ in range(2 if ngpu > 1 else 1):  # a group's first call creates the devices' contexts, its second warms the pools
            es.powerMethod(mk(), es.SolverOptions(1, -1.0), x0=x0)
        M = mk()
        t0 = time.perf_counter()
        r = es.powerMethod(M, es.SolverOptions(steps, -1.0), x0=x0)
        t_e2e = time.perf_counter() - t0
    finally:
        os.environ["EIGSOL_B200_NGPU"] = "1"
    assert r.iterations == steps
    del M
    return {"value": steps / t_e2e, "unit": "iter/s", "h2d_bytes_per_step": h2d / steps,
            "d2h_bytes_per_step": x0.nbytes / steps, "seconds": t_e2e, "lambda": repr(r.eigenvalue),
            "call": "powerMethod(Matrix, SolverOptions{steps, -1}) one-shot in one process: matrix + x0 uploaded, "
                    "CSC->CSR, steps trips, eigenvector downloaded (after untimed 1-trip calls of the same API: one on one GPU, two on a group)" +
                    (f"; EIGSOL_B200_NGPU={ngpu}: the library shards the call over {ngpu} GPUs" if ngpu > 1 else "")}


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("BENCH_WORKLOAD", "c5"), choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-child", type=int, default=0, help=argparse.SUPPRESS)  # internal: run only the e2e leg on N GPUs
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.e2e_child:
        emit(e2e_leg(args.workload, args.steps, args.e2e_child))
        return
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
        args.gpus = world

    import pcsc_eigenvalue_solver_project_b200 as es
    from pcsc_eigenvalue_solver_project_b200 import generators as gen
    if es.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback")

    comm = None
    dist = None
    cpu_group = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))
        cpu_group = dist.new_group(backend="gloo")  # host-side barrier: waiting ranks leave their GPUs idle
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(es.Comm.unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        comm = es.Comm(rank, world, bytes(idt.cpu().numpy().tobytes()), local_rank)

    def barrier():
        if dist is not None:
            dist.barrier()

    def cpu_barrier():
        if dist is not None:
            dist.barrier(group=cpu_group)

    name = args.workload
    n = workload_order(name)
    dt = np.complex128 if DTYPES[name] == "c128" else np.float64
    x0 = es.random_unit(n, dt, seed=1)  # the reference's own start vector for std::srand(1)

    # ---- build this rank's part and make it resident ----
    t_build = time.time()
    if name == "c3":
        a = gen.dense_random_symmetric(n)
        host = {"a": a}
        dm = es.DeviceMatrix.from_matrix(es.Matrix(a), comm)
        nnz = n * n
    else:
        lo, hi = 0, n
        if world > 1:
            # contiguous row blocks; the stencils have (nearly) constant row length so rows == nnz balance
            lo, hi = n * rank // world, n * (rank + 1) // world
        _, rp, ci, v = workload_rows(name, gen, lo, hi)
        host = {"rp": rp, "ci": ci, "v": v, "lo": lo}
        dm = es.DeviceMatrix.from_csr_rows(n, lo, rp, ci, v, comm)
    info = dm.info()
    nnz = info["nnz"]
    t_build = time.time() - t_build

    # ---- device-resident timing: W warm-up trips, then exactly K trips ----
    # the clock sampler (a subprocess) starts BEFORE the warm-up: anything slow on one rank between the barrier
    # and the timed launches would be waited for, in-kernel, by every other rank and land in their timers
    sampler = ClockSampler(local_rank) if rank == 0 else None
    dm.begin(x0, 2**31 - 2, -1.0)
    dm.step(args.warmup)
    barrier()
    launches0 = dm.launch_count()
    ms = dm.step(args.steps)  # CUDA events on the launching stream, stream-synchronised on both sides
    launches = dm.launch_count() - launches0
    barrier()
    clocks = sampler.stop() if sampler else None
    if comm is not None:
        ms = comm.max_float(ms)
    res = dm.end(want_vector=False)
    ms_per_step = ms / args.steps
    value = 1e3 / ms_per_step

    # ---- roofline of the matvec kernel ----
    peak, peak_src = measured_peak()
    alg = alg_bytes_global(name, n, nnz)
    # per-launch duration of the dominant kernel: on one GPU a step IS one launch of it
    # (launches back to back in a CUDA graph); on N GPUs the rank's share of the bytes over the step.
    achieved = (info["alg_bytes"] if world > 1 else alg) / (ms_per_step * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": recorded_traffic(name, world), "peak_source": peak_src,
                "alg_bytes_per_launch": info["alg_bytes"] if world > 1 else alg,
                "frac_of_8TBps": achieved / 8000.0, "kernel": info["kernel_name"],
                "note": "per-rank bytes / step time" if world > 1 else "one launch per step"}

    # ---- e2e: the reference-facing call with host buffers ----
    # powerMethod(Matrix, SolverOptions) in ONE process, exactly what a caller of the reference writes.  On N > 1 GPUs
    # rank 0 makes that call with EIGSOL_B200_NGPU=N (the library shards the solve over the N devices itself,
    # single-process group).  The device-resident part is over by now: the other ranks close their matrices and NCCL
    # communicators and leave, so that the ONE process of this leg has the box to itself (with the other ranks parked on
    # a barrier the same call measured 55 ms instead of 44 ms on 2 GPUs).
    if world > 1:
        dm.close()
        dm = None
        comm.close()
        cpu_barrier()
        dist.destroy_process_group()
        dist = None
        if rank != 0:
            return
        try:
            import torch
            torch.cuda.synchronize()
            torch.cuda.empty_cache()
        except Exception:
            pass
        time.sleep(3.0)  # the other ranks' processes tear down NCCL / CUDA and exit; outside every timed region
        # NCCL / the launcher may have narrowed this process's CPU affinity to the cores next to GPU 0; the library's
        # uploader threads (one set per GPU) inherit it.  The one-process leg gets the whole host back.
        try:
            affinity_before = len(os.sched_getaffinity(0))
            os.sched_setaffinity(0, range(os.cpu_count()))
            sys.stderr.write(f"bench.py: CPU affinity {affinity_before} -> {len(os.sched_getaffinity(0))} cores\n")
        except Exception as exc:
            sys.stderr.write(f"bench.py: sched_setaffinity: {exc}\n")
    e2e = None
    if not args.no_e2e:
        try:
            if world == 1:
                e2e = e2e_leg(name, args.steps, 1, dict(host, whole=True))
            else:
                # (The same call made from a fresh child process of rank 0 was tried: no faster than in this process.
                # Under torchrun the leg measures 55-85 ms on 2 GPUs against 43-49 ms for the identical call from a
                # stand-alone script on the same box -- one GPU's upload runs at half speed; CPU affinity, torch, NCCL
                # and OMP_NUM_THREADS were ruled out, the cause was not found: profiles/r02_notes.md.)
                del host
                e2e = e2e_leg(name, args.steps, world)
        except Exception as exc:  # the device-resident figure must survive a failure of the end-to-end leg
            sys.stderr.write(f"bench.py: e2e leg failed: {type(exc).__name__}: {exc}\n")
            e2e = None

    cpu_baseline = None
    parity = None
    if world == 1 and not args.no_cpu_baseline:
        del host
        try:
            cb = reference_rate(name, budget_s=float(os.environ.get("BENCH_CPU_BUDGET_S", 20)), max_trips=20)
            cpu_baseline = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
            # parity (outside every timed region): the GPU path on the resident matrix for exactly the trips the
            # reference just ran, same x0 -- eigenvalue within the north star's 1e-10
            rp_ = dm.power(x0, es.SolverOptions(cb["trips"], -1.0))
            parity = parity_of(rp_.eigenvalue, complex(cb["lambda"]), cb["trips"],
                               "EigSol::powerMethod (oracle/_ref) run beside the bench")
        except Exception as e:
            cpu_baseline = {"value": None, "unit": "iter/s", "cores": 1, "kind": "reference",
                            "sample": f"unavailable: {type(e).__name__}: {e}"}
    if parity is None and e2e is not None:
        gl = golden_lambda(name, args.steps)
        if gl is not None:
            parity = parity_of(complex(e2e["lambda"]), gl, args.steps,
                               "tests/golden/bench_lambda.json (compiled reference, generated in the authoring container)")
    # a read-dominated HBM ceiling measured in the same run (12 read planes per written plane = the matvec's mix)
    try:
        roofline["peak_read_mix"] = es.hbm_probe(128 << 20, 12, 20)[0]
        roofline["frac_of_read_mix"] = achieved / roofline["peak_read_mix"]
    except Exception as e:  # measurement extra: never fails the bench
        roofline["peak_read_mix"] = None
        sys.stderr.write(f"hbm_probe: {e}\n")

    line = {"metric": "power_iterations_per_s", "value": value, "unit": "iter/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": DTYPES[name], "data": "synthetic",
            "config": workload_config(name),
            "launch": {"partition": f"rows/{world}", "kernel": info["kernel_name"], "grid": info["grid"],
                       "block": info["block"], "halo_per_rank": info["halo"], "exchange": info["exchange"]},
            "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "parity": parity, "gpu_launches": launches,
            "clocks": clocks, "lambda": repr(res.eigenvalue), "trips_done": res.iterations, "build_s": round(t_build, 2)}
    emit(line)
    if dm is not None:
        dm.close()


if __name__ == "__main__":
    main()
