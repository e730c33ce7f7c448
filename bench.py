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


def ours_arm(args):
    import torch
    import torch.distributed as dist
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

    m, n = M_GLOBAL, N_COLS
    m_loc, _ = cdist.local_rows(m, world, rank)
    A = colmajor(m_loc, n); Q = colmajor(m_loc, n); R = colmajor(n, n)
    eng.fill_random(A, rank)                       # bench/qr/cacqr.cpp:34: distribute_random(key = rank / c), c = 1

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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
    peaks = json.load(open(FP64_PEAKS)) if os.path.exists(FP64_PEAKS) else {"dmma_tflops_sustained": 36.6, "source": "fallback constant"}
    peak = float(peaks["dmma_tflops_sustained"])
    kflops = float(m_loc) * n * n                  # algorithmic flops of one launch: triangular multiply / syrk of m_loc x n
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "r01_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        traffic = float(tj["dram_bytes_per_element"]) * m_loc * n     # measured DRAM bytes per panel element x elements of this launch
        traffic_src = tj["source"]
    roofline = {"bound": "tensor", "kernel": "dmma_tma_gemm_kernel<A_MC,B_KC> as Q=A*Rinv, row-panel schedule (cap_cqr_apply)", "achieved": kflops / ms_apply * 1e-9, "peak": peak,
                "unit": "TFLOP/s", "frac": kflops / ms_apply * 1e-9 / peak, "traffic": traffic, "traffic_unit": "bytes per launch (dram read + write)",
                "traffic_source": traffic_src, "algorithmic_bytes_per_launch": 16.0 * m_loc * n,
                "peak_source": "measured DMMA.8x8x4 issue-rate microkernel, sustained (tools/dmma_peak.cu -> profiles/fp64_peak.json); MEASURED_PEAKS.json has no FP64 entry",
                "ms_per_launch": ms_apply, "algorithmic_flops_per_launch": kflops,
                "gram": {"kernel": "dmma_tma_gemm_kernel<A_KC,B_KC> split-K + splitk_reduce (cap_cqr_gram)", "achieved": kflops / ms_gram * 1e-9, "frac": kflops / ms_gram * 1e-9 / peak, "ms_per_launch": ms_gram},
                "potrf_trtri_ms": ms_small}
    del X, G

    # ---- BASELINE config 5 beside the headline (rank 0 of a 1-GPU run only): Cholesky and triangular inverse of N = 16384
    factor = None
    if world == 1:
        try:
            nf = 16384
            S = colmajor(nf, nf); eng.fill_symmetric(S, 0, 0, 1, 1, nf, True); W = colmajor(nf, nf)
            ms_copy = time_kernel(lambda: W.copy_(S), 3)
            ms_potrf = time_kernel(lambda: (W.copy_(S), eng._potrf(W, nf, "U")), 3) - ms_copy
            ms_trtri = time_kernel(lambda: (S.copy_(W), eng._trtri(S, nf, "U", "N")), 3) - ms_copy
            fl = nf ** 3 / 3.0
            factor = {"workload": "potrf + trtri N=16384 FP64 (BASELINE config 5), one GPU, algorithmic flops N^3/3 each",
                      "potrf_ms": ms_potrf, "potrf_tflops": fl / ms_potrf * 1e-9, "potrf_frac_of_peak": fl / ms_potrf * 1e-9 / peak,
                      "trtri_ms": ms_trtri, "trtri_tflops": fl / ms_trtri * 1e-9, "trtri_frac_of_peak": fl / ms_trtri * 1e-9 / peak,
                      "lapack_info": eng.info()}
            del S, W
            torch.cuda.empty_cache()
            # the matrix size of BASELINE config 3 on ONE GPU (the 8-GPU numbers of that config: profiles/r01_potrf1d_summary.txt)
            nf = 32768
            S = colmajor(nf, nf); eng.fill_symmetric(S, 0, 0, 1, 1, nf, True); W = colmajor(nf, nf)
            ms_copy = time_kernel(lambda: W.copy_(S), 2)
            ms_big = time_kernel(lambda: (W.copy_(S), eng._potrf(W, nf, "U")), 2) - ms_copy
            fl = nf ** 3 / 3.0
            factor.update({"potrf_n32768_ms": ms_big, "potrf_n32768_tflops": fl / ms_big * 1e-9, "potrf_n32768_frac_of_peak": fl / ms_big * 1e-9 / peak,
                           "lapack_info_n32768": eng.info()})
            del S, W
            torch.cuda.empty_cache()
        except Exception as exc:  # noqa: BLE001 -- an extra, never allowed to break the headline line
            factor = dict(factor or {}, error=str(exc)[:200])

    # ---- end to end through the C ABI with HOST buffers (H2D + factor + D2H inside the timed region)
    e2e = None
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
               "api": "cap_cacqr_factor_1d_host (pinned host A -> Q, R)"}
    except Exception as exc:  # noqa: BLE001 -- report, never fake, a missing end-to-end number
        e2e = {"value": None, "unit": "TFLOP/s", "error": str(exc)[:200]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cores = host_cores()
            m_sample = 1 << 19
            rate, ranks_c, threads_c, sec = cpu_reference_rate(m_sample, n, cores)
            cpu = {"value": rate, "unit": "TFLOP/s", "cores": ranks_c * threads_c, "kind": "reference", "seconds": sec,
                   "sample": "CholeskyQR2 %dx%d (1/%d of the rows), unmodified reference via oracle/_ref, %d minimpi ranks x %d OpenBLAS threads" % (
                       m_sample, n, m // m_sample, ranks_c, threads_c)}
        except Exception as exc:  # noqa: BLE001
            cpu = {"value": None, "unit": "TFLOP/s", "error": str(exc)[:200]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "TFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (reference drand48 generator, bit-exact on device)",
            "config": {"workload": "CholeskyQR2 2^22x1024 FP64 (BASELINE config 2), rows dealt over %d GPU(s), NCCL allreduce of the packed Gram" % world,
                       "m": m, "n": n, "m_loc": m_loc, "grid": "1D c=1 d=%d" % world, "l2": "inputs (%.0f GiB per GPU) exceed the 126 MB L2; no flush needed" % (m_loc * n * 8 / 2 ** 30)},
            "frac_of_fp64_peak": value / (peak * world), "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "roofline": roofline, "cpu_baseline": cpu, "lapack_info": info, "config5": factor,
        }
        print(json.dumps(line), flush=True)
    if comm:
        comm.close()
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    return reference_arm(args) if args.impl == "reference" else ours_arm(args)


if __name__ == "__main__":
    sys.exit(main())
