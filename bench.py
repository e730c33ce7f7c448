#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE.json's config.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun, one rank per GPU)
    python bench.py --impl reference ...                      (the reference's own CPU path, oracle/_ref)

Workload ("step") = one CholeskyQR2 factorization (qr::cacqr::factor, variant 2, 1D grid) of the
2^22 x 1024 FP64 matrix of BASELINE config 2, rows dealt over the N GPUs (strong scaling: the global
matrix is fixed, m_loc = 2^22 / N), synthetic input from the reference's own drand48 generator.
Metric = algorithmic FP64 TFLOP/s, flops = 4 m n^2 + (5/3) n^3, time = CUDA events,
max over ranks.  One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

M_GLOBAL = 1 << 22
N_COLS = 1024
METRIC = "FP64 TFLOP/s, CholeskyQR2 2^22x1024 (algorithmic flops 4mn^2 + 5n^3/3)"
FP64_PEAKS = os.path.join(ROOT, "profiles", "fp64_peak.json")


def flops_cacqr2(m, n):
    return 4.0 * m * n * n + (5.0 / 3.0) * n ** 3


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference_sample(m, n, ranks, threads, timeout=1500):
    """One timed qr::cacqr::factor of the UNMODIFIED reference (oracle/_ref/ref_driver: warm-up call,
    barrier, timed call, max over ranks -- bench/qr/cacqr.cpp:42-52) on the host cores."""
    drv = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
    if not os.path.exists(drv):
        raise RuntimeError("oracle/_ref/ref_driver missing: run `make -C oracle` in the build container")
    stage_mb = max(64, (m // ranks) * n * 8 // (1 << 20) + 64)
    env = dict(os.environ, MINIMPI_NP=str(ranks), OPENBLAS_NUM_THREADS=str(threads), MINIMPI_STAGE_MB=str(stage_mb), REF_NO_DUMP="1")
    prefix = "/tmp/capital_ref_%d" % os.getpid()
    out = subprocess.run([drv, "cacqr", prefix, "2", str(m), str(n), "1", "1", "1", "0", "0", "0", "0", "0"],
                         env=env, capture_output=True, text=True, timeout=timeout, check=True).stdout
    fields = dict(kv.split("=") for kv in out.split() if "=" in kv)
    return float(fields["seconds"])


def cpu_reference_rate(m_sample, n, cores):
    """Reference CPU throughput (TFLOP/s) on a bounded sample; ranks x BLAS threads = all host cores."""
    ranks = 4 if cores >= 4 else 1
    threads = max(1, cores // ranks)
    sec = run_reference_sample(m_sample, n, ranks, threads)
    return flops_cacqr2(m_sample, n) / sec * 1e-12, ranks, threads, sec


def reference_arm(args):
    """--impl reference: the reference's own CPU implementation of the path, all host cores, each
    step one bounded sample (2^19 x 1024 rows = one eighth of the 2^22 x 1024 workload)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = host_cores()
    m_sample = 1 << 19
    rates, secs = [], []
    ranks = threads = 1
    for _ in range(max(1, args.steps)):
        rate, ranks, threads, sec = cpu_reference_rate(m_sample, N_COLS, cores)
        rates.append(rate); secs.append(sec)
    value = sum(rates) / len(rates)
    sample = "CholeskyQR2 %dx%d (1/%d of the rows of the 2^22x1024 workload), %d minimpi ranks x %d OpenBLAS threads" % (
        m_sample, N_COLS, M_GLOBAL // m_sample, ranks, threads)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "TFLOP/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sum(secs) / len(secs), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic (reference drand48 generator)",
        "config": {"workload": "CholeskyQR2 2^22x1024 FP64 (BASELINE config 2), bounded sample per step", "m": M_GLOBAL, "n": N_COLS,
                   "sample_m": m_sample, "grid": "1D, %d CPU ranks" % ranks},
        "cpu_baseline": {"value": value, "unit": "TFLOP/s", "cores": ranks * threads, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc = gpu_index, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200", "-i", str(self.gpu)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill(); out = ""
        sm, mx, reasons, watts = [], 0.0, set(), 0.0
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for row in out.strip().splitlines():
            f = [x.strip() for x in row.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx = max(mx, float(f[2])); watts = max(watts, float(f[3]))
            except ValueError:
                continue
            for name, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm), "power_w_max": watts}


def measure_fp64_peak(device_index):
    """DMMA.8x8x4 issue-rate peak measured NOW on this GPU by tools/dmma_peak (built by build()); MEASURED_PEAKS.json
    carries no FP64 entry.  Falls back to the round-1 measurement file, and says so."""
    exe = os.path.join(ROOT, "tools", "dmma_peak")
    try:
        env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(device_index))
        out = subprocess.run([exe], env=env, capture_output=True, text=True, timeout=120, check=True).stdout
        burst, sustained, clock = 0.0, None, None
        for row in out.splitlines():
            if row.startswith("warps/SM"):
                burst = max(burst, float(row.split("DMMA")[1].split()[0]))
            elif row.startswith("sustained DMMA"):
                sustained = float(row.split(":")[1].split()[0])
            elif row.startswith("device="):
                clock = row.strip()
        if sustained:
            return {"dmma_tflops_sustained": sustained, "dmma_tflops_burst": burst, "device": clock,
                    "source": "tools/dmma_peak run inside this bench (register-only DMMA.8x8x4 loop, 16 warps/SM, ~3 s back to back)"}
    except Exception as exc:  # noqa: BLE001
        err = str(exc)[:120]
    else:
        err = "no sustained line"
    if os.path.exists(FP64_PEAKS):
        pk = json.load(open(FP64_PEAKS))
        return {"dmma_tflops_sustained": float(pk["dmma_tflops_sustained"]), "dmma_tflops_burst": float(pk["dmma_tflops_burst"]),
                "source": "profiles/fp64_peak.json (round-1 measurement; live tools/dmma_peak failed: %s)" % err}
    raise SystemExit("bench.py: no FP64 peak available (tools/dmma_peak failed: %s)" % err)


def bind_to_gpu_numa_node(device_index):
    """Pin this process to the CPUs of the NUMA node the GPU hangs off (sysfs), BEFORE the pinned host buffers of the
    end-to-end leg are allocated: round 1 had all 8 ranks' buffers on node 0 and 107 GB/s aggregate over 8 PCIe links."""
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(device_index)],
                             capture_output=True, text=True, timeout=20).stdout.strip()
        bdf = out[-12:]                                          # 00000000:1B:00.0 -> 0000:1B:00.0
        base = "/sys/bus/pci/devices/" + bdf.lower()
        node = int(open(base + "/numa_node").read().strip())
        cpus = open(base + "/local_cpulist").read().strip()
        ids = set()
        for part in cpus.split(","):
            lo, _, hi = part.partition("-")
            ids.update(range(int(lo), int(hi or lo) + 1))
        allowed = ids & set(os.sched_getaffinity(0))
        if node < 0 or not allowed:
            return {"node": node, "bound": False}
        os.sched_setaffinity(0, allowed)
        return {"node": node, "cpus": len(allowed), "bound": True}
    except Exception as exc:  # noqa: BLE001 -- binding is an optimisation, never a reason to lose the number
        return {"bound": False, "error": str(exc)[:120]}


def cpu_extra_baselines(cores):
    """cfg1 exactly through the reference's own bench main, and scaled cholinv / SUMMA runs of the
    unmodified reference (oracle/_ref).  Seconds are the reference's own MPI_Wtime bracket, max over its ranks."""
    ref = os.path.join(ROOT, "oracle", "_ref")
    out = {}

    def run(tag, cmd, ranks, threads, flops, parse):
        try:
            env = dict(os.environ, MINIMPI_NP=str(ranks), OPENBLAS_NUM_THREADS=str(threads), REF_NO_DUMP="1", MINIMPI_STAGE_MB="512")
            txt = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300, check=True).stdout
            sec = parse(txt)
            out[tag] = {"seconds": sec, "tflops": flops / sec * 1e-12, "ranks": ranks, "blas_threads": threads, "cores": ranks * threads,
                        "cmd": " ".join(os.path.basename(c) if i == 0 else c for i, c in enumerate(cmd))}
        except Exception as exc:  # noqa: BLE001
            out[tag] = {"error": str(exc)[:160]}

    field = lambda txt: float(dict(kv.split("=") for kv in txt.split() if "=" in kv)["seconds"])
    last_col = lambda txt: float(txt.strip().splitlines()[-1].split()[-1])
    m1, n1 = 65536, 256
    for threads in sorted({1, max(1, cores // 4)}):        # bench/qr/cacqr.cpp as `cacqr 2 65536 256 1 1 1 1 0 0 0 0 5` (BASELINE config 1)
        run("cfg1_cacqr_65536x256_4ranks_x%d" % threads, [os.path.join(ref, "b_cacqr"), "2", str(m1), str(n1), "1", "1", "1", "1", "0", "0", "0", "0", "5"],
            4, threads, flops_cacqr2(m1, n1), last_col)
    nc = 8192
    run("cholinv_8192_p1", [os.path.join(ref, "ref_driver"), "cholinv", "/tmp/capital_ref_x", str(nc), "1", "0", "1", "0", "0", "0", "2", "0"],
        1, cores, nc ** 3 / 3.0, field)
    if cores >= 8:
        run("cholinv_8192_2x2x2", [os.path.join(ref, "ref_driver"), "cholinv", "/tmp/capital_ref_x", str(nc), "1", "0", "1", "0", "0", "0", "2", "0"],
            8, max(1, cores // 8), nc ** 3 / 3.0, field)
    run("summa_8192_p1", [os.path.join(ref, "ref_driver"), "summa", "/tmp/capital_ref_x", str(nc), str(nc), str(nc), "1", "0", "0"],
        1, cores, 2.0 * nc ** 3, field)
    return out


def ours_arm(args):
    import torch
    import torch.distributed as dist
    from capital_b200 import checks
    from capital_b200 import dist as cdist
    from capital_b200.engine import Engine, colmajor

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("bench.py --gpus %d needs %d ranks (launch with torch.distributed.run); WORLD_SIZE=%d" % (args.gpus, args.gpus, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    eng = Engine(local_rank)
    comm = cdist.Comm(eng) if world > 1 else None
    comm_ptr = comm.ptr if comm else None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # FP64 tensor peak of THIS box, measured before anything else (rank 0's GPU; the others wait at the barrier)
    peaks = measure_fp64_peak(local_rank) if rank == 0 else None
    barrier()
    if world > 1:
        box = [peaks]
        dist.broadcast_object_list(box, src=0)
        peaks = box[0]
    peak = float(peaks["dmma_tflops_sustained"])
    nominal = 40.0                                     # BASELINE.json: "~40 TFLOP/s FP64 tensor" per B200

    m, n = M_GLOBAL, N_COLS
    m_loc, _ = cdist.local_rows(m, world, rank)
    A = colmajor(m_loc, n); Q = colmajor(m_loc, n); R = colmajor(n, n)
    eng.fill_random(A, rank)                       # bench/qr/cacqr.cpp:34: distribute_random(key = rank / c), c = 1

    def step():
        eng.cacqr_factor_1d(A, Q, R, 2, comm_ptr)

    for _ in range(max(3, args.warmup)):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = eng.launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms_total = cdist.max_over_ranks(e0.elapsed_time(e1))
    launches = (eng.launches() - launches0) * world
    clocks = sampler.stop() if rank == 0 else None
    info = eng.info()
    ms_step = ms_total / args.steps
    value = flops_cacqr2(m, n) / (ms_step * 1e-3) * 1e-12

    # ---- parity of the timed result, at every N: the reference's own validators (test/qr/validate.hpp:7-52) on the device
    orth, resid = eng.validate_cacqr_1d(A, Q, R, comm_ptr)
    tol = 1e-13 * n
    parity = {"orthogonality": orth, "residual": resid, "tol": tol, "lapack_info": info,
              "metric": "||Q^T Q - I||_F / n and ||Q R - A||_F / ||A||_F of the timed factorization (cap_validate_cacqr_1d, all ranks)",
              "ok": bool(orth <= tol and resid <= tol and info == 0)}

    # ---- roofline of the dominant kernels, timed live on the launching stream ------------------
    X = colmajor(n, n); G = colmajor(n, n)
    eng.cqr_gram(Q, G); eng.cholinv(G, X, n, True)
    def time_kernel(fn, reps=5):
        fn(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / reps
    ms_apply = time_kernel(lambda: eng.cqr_apply(X, A, Q))
    ms_gram = time_kernel(lambda: eng.cqr_gram(A, G))
    ms_small = time_kernel(lambda: eng.cholinv(G.clone(), X, n, True))
    kflops = float(m_loc) * n * n                  # algorithmic flops of one launch: triangular multiply / syrk of m_loc x n
    traffic, traffic_src = None, None
    for tname in ("r02_traffic.json", "r01_traffic.json"):
        tpath = os.path.join(ROOT, "profiles", tname)
        if os.path.exists(tpath):
            tj = json.load(open(tpath))
            traffic = float(tj["dram_bytes_per_element"]) * m_loc * n     # measured DRAM bytes per panel element x elements of this launch
            traffic_src = tj["source"]
            break
    roofline = {"bound": "tensor", "kernel": "dmma_tma_gemm_kernel<A_MC,B_KC> as Q=A*Rinv, row-panel schedule (cap_cqr_apply)", "achieved": kflops / ms_apply * 1e-9, "peak": peak,
                "unit": "TFLOP/s", "frac": kflops / ms_apply * 1e-9 / peak, "frac_of_nominal_peak": kflops / ms_apply * 1e-9 / nominal,
                "traffic": traffic, "traffic_unit": "bytes per launch (dram read + write)",
                "traffic_source": traffic_src, "algorithmic_bytes_per_launch": 16.0 * m_loc * n,
                "peak_source": peaks["source"], "peak_burst": peaks.get("dmma_tflops_burst"), "peak_device": peaks.get("device"),
                "ms_per_launch": ms_apply, "algorithmic_flops_per_launch": kflops,
                "gram": {"kernel": "dmma_tma_gemm_kernel<A_KC,B_KC> split-K + splitk_reduce (cap_cqr_gram)", "achieved": kflops / ms_gram * 1e-9, "frac": kflops / ms_gram * 1e-9 / peak, "ms_per_launch": ms_gram},
                "potrf_trtri_ms": ms_small}
    del X, G

    # ---- end to end through the C ABI with HOST buffers (H2D + factor + D2H inside the timed region)
    e2e = None
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else None      # pinned buffers are first-touched on the GPU's own node
    try:
        del A, Q
        torch.cuda.empty_cache()
        hA = torch.empty((n, m_loc), dtype=torch.float64, pin_memory=True)
        hQ = torch.empty((n, m_loc), dtype=torch.float64, pin_memory=True)
        hR = torch.empty((n, n), dtype=torch.float64, pin_memory=True)
        stage = colmajor(m_loc, n); eng.fill_random(stage, rank); hA.copy_(stage.t()); torch.cuda.synchronize()
        del stage
        torch.cuda.empty_cache()
        e2e_steps = max(1, min(args.steps, 3))
        eng.cacqr_factor_1d_host(hA, hQ, hR, m_loc, n, 2, comm_ptr)       # warm-up (allocates the staging workspace)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            eng.cacqr_factor_1d_host(hA, hQ, hR, m_loc, n, 2, comm_ptr)   # synchronises before returning
        barrier()
        sec = cdist.max_over_ranks(time.perf_counter() - t0) / e2e_steps
        e2e = {"value": flops_cacqr2(m, n) / sec * 1e-12, "unit": "TFLOP/s", "h2d_bytes_per_step": m * n * 8,
               "d2h_bytes_per_step": m * n * 8 + world * n * n * 8, "ms_per_step": sec * 1e3, "steps": e2e_steps,
               "h2d_plus_d2h_gbs_per_gpu": 2.0 * m_loc * n * 8 / sec * 1e-9, "numa_binding": numa,
               "api": "cap_cacqr_factor_1d_host (pinned host A -> Q, R)"}
        del hA, hQ, hR
    except Exception as exc:  # noqa: BLE001 -- report, never fake, a missing end-to-end number
        e2e = {"value": None, "unit": "TFLOP/s", "error": str(exc)[:200]}
    eng.release_workspace()
    torch.cuda.empty_cache()

    # ---- the other algorithms BASELINE.json's metric and configs name, at this N, each with its own parity check ---------
    extras = {}
    def extra(name, fn):
        if args.no_extras:
            return
        t0 = time.perf_counter()
        try:
            extras[name] = fn()
        except Exception as exc:  # noqa: BLE001 -- an extra never breaks the headline line; a failed one is reported as failed
            extras[name] = {"ok": False, "error": "%s: %s" % (type(exc).__name__, str(exc)[:300])}
        extras[name]["wall_s"] = round(time.perf_counter() - t0, 2)
        try:
            eng.release_workspace()
        except Exception:  # noqa: BLE001
            pass
        torch.cuda.empty_cache()

    n_chol = 32768
    extra("parity_small", lambda: {"cacqr2_1d": checks.cqr2_small_check(eng, comm_ptr, world, rank),
                                   "potrf_1d": checks.potrf1d_small_check(eng, comm_ptr, world, rank),
                                   "grid": checks.grid_small_check(eng, comm, world, rank, 1024),
                                   "note": "product path vs numpy on assembled global matrices, run inside this bench process group at %d rank(s)" % world})
    if "parity_small" in extras and "cacqr2_1d" in extras["parity_small"]:
        ps = extras["parity_small"]
        ps["ok"] = bool(ps["cacqr2_1d"]["ok"] and ps["potrf_1d"]["ok"] and ps["grid"]["ok"])
    def cholesky():
        # block width: the bulk update prefers wide panels, the diagonal chain narrow ones; both candidates are run and named
        cands = [2048, 1024] if world <= 2 else [1024, 512]
        runs = [checks.cholesky_block(eng, comm_ptr, world, rank, n_chol, nb, 3, peak) for nb in cands]
        best = min(runs, key=lambda r: r["ms"])
        best = dict(best, tried=[{"nb": r["nb"], "ms": r["ms"], "resid_sampled": r["resid_sampled"], "ok": r["ok"]} for r in runs])
        best["ok"] = bool(all(r["ok"] for r in runs))
        return best
    extra("cholesky", cholesky)
    extra("cholinv_grid", lambda: checks.cholinv_grid_block(eng, comm, world, rank, n_chol, 2, peak))
    extra("summa", lambda: checks.summa_block(eng, comm, world, rank, 65536 if world == 8 else 32768, 2, peak))

    # ---- BASELINE config 5 (one GPU only): Cholesky and triangular inverse of N = 16384 through the LAPACK-style entry points
    if world == 1 and not args.no_extras:
        def cfg5():
            nf = 16384
            S = colmajor(nf, nf); eng.fill_symmetric(S, 0, 0, 1, 1, nf, True); W = colmajor(nf, nf)
            ms_copy = time_kernel(lambda: W.copy_(S), 3)
            ms_potrf = time_kernel(lambda: (W.copy_(S), eng._potrf(W, nf, "U")), 3) - ms_copy
            ms_trtri = time_kernel(lambda: (S.copy_(W), eng._trtri(S, nf, "U", "N")), 3) - ms_copy
            fl = nf ** 3 / 3.0
            return {"workload": "potrf + trtri N=16384 FP64 (BASELINE config 5), one GPU, algorithmic flops N^3/3 each",
                    "potrf_ms": ms_potrf, "potrf_tflops": fl / ms_potrf * 1e-9, "potrf_frac_of_peak": fl / ms_potrf * 1e-9 / peak,
                    "trtri_ms": ms_trtri, "trtri_tflops": fl / ms_trtri * 1e-9, "trtri_frac_of_peak": fl / ms_trtri * 1e-9 / peak,
                    "lapack_info": eng.info(), "ok": eng.info() == 0}
        extra("config5", cfg5)

    # ---- HBM-bound helpers of the path against the measured copy bandwidth of this pool's B200s
    if world == 1 and not args.no_extras:
        def helpers():
            hbm = None
            mp = os.path.join(ROOT, "MEASURED_PEAKS.json")
            if os.path.exists(mp):
                hbm = float(json.load(open(mp)).get("hbm_gbs", 0.0)) or None
            nh = 16384
            src = colmajor(nh, nh); dst = colmajor(nh, nh); src.uniform_()
            packed = torch.empty(nh * (nh + 1) // 2, dtype=torch.float64, device="cuda")
            full = 16.0 * nh * nh                                 # read + write of every element
            rows = {}
            def put(name, fn, nbytes):
                ms = time_kernel(fn, 5)
                rows[name] = {"ms": ms, "gbs": nbytes / ms * 1e-6, "frac_of_hbm_peak": (nbytes / ms * 1e-6 / hbm) if hbm else None}
            put("copy2d mode 0 (all)", lambda: eng.copy2d(dst, src, nh, nh, 0), full)
            put("copy2d mode 1 (upper only)", lambda: eng.copy2d(dst, src, nh, nh, 1), full / 2)
            put("copy2d mode 3 (upper + zero lower)", lambda: eng.copy2d(dst, src, nh, nh, 3), 12.0 * nh * nh)
            put("copy2d mode 0, strided view (ld 16384 -> 8192 rows)", lambda: eng.copy2d(dst[:8192], src[8192:], 8192, nh, 0), full / 2)
            put("pack_upper", lambda: eng.pack_upper(packed, src, nh), full / 2)
            put("unpack_upper + zero lower", lambda: eng.unpack_upper(dst, packed, nh, True), 12.0 * nh * nh)
            put("axpby (y = a x + b y)", lambda: eng.axpby(dst, src, 0.5, 0.5), 24.0 * nh * nh)
            return {"workload": "16384 x 16384 FP64 operands; algorithmic bytes / CUDA-event time, 5 launches each", "hbm_peak_gbs": hbm,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs (driver-written)" if hbm else None, "kernels": rows, "ok": True}
        extra("hbm_helpers", helpers)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cores = host_cores()
            m_sample = 1 << 19
            rate, ranks_c, threads_c, sec = cpu_reference_rate(m_sample, n, cores)
            cpu = {"value": rate, "unit": "TFLOP/s", "cores": ranks_c * threads_c, "kind": "reference", "seconds": sec,
                   "sample": "CholeskyQR2 %dx%d (1/%d of the rows), unmodified reference via oracle/_ref, %d minimpi ranks x %d OpenBLAS threads" % (
                       m_sample, n, m // m_sample, ranks_c, threads_c)}
            if not args.no_extras:
                cpu["extra"] = cpu_extra_baselines(cores)
        except Exception as exc:  # noqa: BLE001
            cpu = {"value": None, "unit": "TFLOP/s", "error": str(exc)[:200]}

    # a wrong number fails the run; an extra that could not run (exception: reported in its block) does not
    all_ok = parity["ok"] and all(v.get("ok", True) or "error" in v for v in extras.values())
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "TFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (reference drand48 generator, bit-exact on device)",
            "config": {"workload": "CholeskyQR2 2^22x1024 FP64 (BASELINE config 2), rows dealt over %d GPU(s), NCCL allreduce of the packed Gram" % world,
                       "m": m, "n": n, "m_loc": m_loc, "grid": "1D c=1 d=%d" % world, "l2": "inputs (%.0f GiB per GPU) exceed the 126 MB L2; no flush needed" % (m_loc * n * 8 / 2 ** 30)},
            "frac_of_fp64_peak": value / (peak * world), "frac_of_measured_peak": value / (peak * world), "frac_of_nominal_peak": value / (nominal * world),
            "fp64_peak": peaks, "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "parity": parity, "parity_ok": bool(all_ok),
            "roofline": roofline, "cpu_baseline": cpu, "vs_cpu_baseline": (value / cpu["value"]) if cpu and cpu.get("value") else None,
            "lapack_info": info, "cholesky": extras.get("cholesky"), "cholinv_grid": extras.get("cholinv_grid"), "summa": extras.get("summa"),
            "parity_small": extras.get("parity_small"), "config5": extras.get("config5"), "hbm_helpers": extras.get("hbm_helpers"),
        }
        print(json.dumps(line), flush=True)
    eng.close()                 # the handle may own a duplicate of comm: it goes first
    if comm:
        comm.close()
    if world > 1:
        dist.destroy_process_group()
    return 0 if all_ok else 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline + parity + roofline + e2e only (skip the Cholesky / cholinv / SUMMA blocks)")
    args = ap.parse_args()
    return reference_arm(args) if args.impl == "reference" else ours_arm(args)


if __name__ == "__main__":
    sys.exit(main())
