"""cfg5 probe (BASELINE config 5: Cholesky + triangular inverse of N = 16384 on one GPU): times cap_dpotrf (N^3/3
schedule) over its block size, the reference's schedule for the same R (cap_dcholinv, complete_inv = 0), cap_dtrtri
and cap_dtrsm, through the C ABI with CUDA events.  The restoring copy of the input is timed alone and subtracted.

    python tools/factor_probe.py [--N 8192 16384] [--nb 512 1024 2048] [--reps 3]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from capital_b200.engine import Engine, colmajor  # noqa: E402

PEAK = 36.6   # measured sustained DMMA issue rate, TFLOP/s (profiles/fp64_peak.json)


def best_ms(fn, reps):
    fn(); torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--N", type=int, nargs="+", default=[8192, 16384])
    ap.add_argument("--nb", type=int, nargs="+", default=[512, 1024, 2048])
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    E = Engine(0)
    rows = []

    def report(what, n, ms, flops, extra=""):
        tf = flops / ms * 1e-9
        rows.append({"case": what, "n": n, "ms": ms, "tflops": tf, "frac_of_peak": tf / PEAK})
        print("%-34s n=%-6d %9.3f ms  %6.2f TFLOP/s  %5.1f %% of DMMA peak %s" % (what, n, ms, tf, 100 * tf / PEAK, extra), flush=True)

    for n in args.N:
        S = colmajor(n, n); E.fill_symmetric(S, 0, 0, 1, 1, n, True)
        W = colmajor(n, n); X = colmajor(n, n)
        t_copy = best_ms(lambda: W.copy_(S), args.reps)
        fl = n ** 3 / 3
        E.reset_info()
        t = best_ms(lambda: (W.copy_(S), E.cholinv(W, X, n, False)), args.reps) - t_copy
        report("cholinv ci=0 (reference schedule)", n, t, fl, "(executes 5N^3/12)")
        ref = W.clone()
        for sched, name in ((1, "recursive"), (0, "look-ahead")):
            E.set_potrf_schedule(sched)
            for nb in args.nb:
                if nb >= n:
                    continue
                E.set_block_sizes(nb, 512)
                t = best_ms(lambda: (W.copy_(S), E._potrf(W, n, "U")), args.reps) - t_copy
                diff = float((torch.triu(W) - torch.triu(ref)).norm() / torch.triu(ref).norm())
                report("potrf %s nb=%d" % (name, nb), n, t, fl, "| R vs reference schedule %.1e" % diff)
        E.set_block_sizes(0, 512)
        t = best_ms(lambda: (W.copy_(S), E._potrf(W, n, "U")), args.reps) - t_copy
        report("potrf as shipped (defaults)", n, t, fl)
        t = best_ms(lambda: (W.copy_(S), E.cholinv(W, X, n, True)), args.reps) - t_copy
        report("cholinv ci=1 (R and R^-1)", n, t, 2 * fl)
        R = torch.triu(ref); del ref
        Rm = colmajor(n, n); Rm.copy_(R); del R
        t = best_ms(lambda: (W.copy_(Rm), E._trtri(W, n, "U", "N")), args.reps) - t_copy
        report("trtri (batched leaves)", n, t, fl)
        for nb in (256, 512, 1024):
            E.set_block_sizes(0, nb)
            t = best_ms(lambda: (W.copy_(S), E.trsm(Rm, W, n, n, "L", "U", "T", "N", 1.0)), args.reps) - t_copy
            report("trsm L/U/T nb=%d, %d rhs" % (nb, n), n, t, float(n) ** 3)
        E.set_block_sizes(0, 512)
        print("info", E.info(), flush=True)
        del S, W, X, Rm
        torch.cuda.empty_cache()
    if args.json:
        with open(args.json, "w") as f:
            json.dump({"peak_tflops": PEAK, "rows": rows}, f, indent=1)
    E.close()


if __name__ == "__main__":
    main()
