"""Kernel probe: time selected hot kernels through the C ABI (CUDA events) -- the command ncu wraps.

    python tools/kernel_probe.py [case ...] [--reps R] [--m M]
cases: gemm_nn gemm_tn gemm_nt apply gram cholinv cqr2 potrf trtri   (default: all small ones)
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from capital_b200.engine import Engine, colmajor  # noqa: E402


PROFILE = False


def timeit(fn, reps):
    fn(); torch.cuda.synchronize()
    best, tot = 1e30, 0.0
    if PROFILE:   # ncu --profile-from-start off: only the timed launches are captured
        torch.cuda.profiler.start()
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        t = a.elapsed_time(b); best = min(best, t); tot += t
    if PROFILE:
        torch.cuda.profiler.stop()
    return best, tot / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("cases", nargs="*", default=["gemm_nn", "gemm_tn", "gemm_nt", "apply", "gram", "cholinv"])
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--m", type=int, default=1 << 20)
    ap.add_argument("--n", type=int, default=1024)
    ap.add_argument("--N", type=int, default=8192)
    ap.add_argument("--profile", action="store_true")
    ap.add_argument("--K", type=int, nargs="+", default=[1024], help="contraction lengths of the `update` case")
    args = ap.parse_args()
    global PROFILE
    PROFILE = args.profile
    E = Engine(0)
    N, m, n = args.N, args.m, args.n
    for case in args.cases:
        if case.startswith("gemm_"):
            ta, tb = case[5].upper(), case[6].upper()
            A = colmajor(N, N); B = colmajor(N, N); Cm = colmajor(N, N); A.uniform_(); B.uniform_()
            best, avg = timeit(lambda: E._gemm(A, B, Cm, N, N, N, ta, tb), args.reps)
            print("%s N=%d: best %.3f ms avg %.3f ms  %.2f TFLOP/s" % (case, N, best, avg, 2 * N ** 3 / best * 1e-9), flush=True)
            del A, B, Cm
        elif case == "update":
            # the shape of the Cholesky trailing update: C(M x Nc) -= A^T B, A: K x M, B: K x Nc, both k-contiguous, beta = 1
            M, Nc = 2 * N, N // 2
            for K in args.K:
                A = colmajor(K, M); B = colmajor(K, Nc); Cm = colmajor(M, Nc); A.uniform_(); B.uniform_(); Cm.zero_()
                best, avg = timeit(lambda: E._gemm(A, B, Cm, M, Nc, K, "T", "N", -1.0, 1.0), args.reps)
                print("update M=%d N=%d K=%d: best %.3f ms avg %.3f ms  %.2f TFLOP/s" % (M, Nc, K, best, avg, 2.0 * M * Nc * K / best * 1e-9), flush=True)
                del A, B, Cm
        elif case in ("apply", "gram", "cqr2", "apply_inplace"):
            A = colmajor(m, n); E.fill_random(A, 1); Q = colmajor(m, n); R = colmajor(n, n); G = colmajor(n, n); X = colmajor(n, n)
            E.cqr_gram(A, G); E.cholinv(G, X, n, True)
            if case == "apply":
                best, avg = timeit(lambda: E.cqr_apply(X, A, Q), args.reps); fl = m * n * n
            elif case == "apply_inplace":
                Q.copy_(A)
                best, avg = timeit(lambda: E.cqr_apply(X, Q, Q), args.reps); fl = m * n * n
            elif case == "gram":
                best, avg = timeit(lambda: E.cqr_gram(A, G), args.reps); fl = m * n * n
            else:
                best, avg = timeit(lambda: E.cacqr_factor_1d(A, Q, R, 2), args.reps); fl = 4 * m * n * n + 5 * n ** 3 / 3
            print("%s %dx%d: best %.3f ms avg %.3f ms  %.2f TFLOP/s" % (case, m, n, best, avg, fl / best * 1e-9), flush=True)
            del A, Q
        elif case in ("cholinv", "potrf", "trtri"):
            for nn in ([n] if case == "cholinv" else [args.N]):
                S = colmajor(nn, nn); E.fill_symmetric(S, 0, 0, 1, 1, nn, True); X = colmajor(nn, nn); W = colmajor(nn, nn)
                if case == "cholinv":
                    best, avg = timeit(lambda: (W.copy_(S), E.cholinv(W, X, nn, True)), args.reps); fl = 2 * nn ** 3 / 3
                elif case == "potrf":
                    best, avg = timeit(lambda: (W.copy_(S), E._potrf(W, nn, "U")), args.reps); fl = nn ** 3 / 3
                else:
                    E._potrf(S, nn, "U")
                    best, avg = timeit(lambda: (W.copy_(S), E._trtri(W, nn, "U", "N")), args.reps); fl = nn ** 3 / 3
                print("%s n=%d: best %.3f ms avg %.3f ms  %.2f TFLOP/s (incl. copy)" % (case, nn, best, avg, fl / best * 1e-9), flush=True)
        else:
            raise SystemExit("unknown case " + case)
    print("launches", E.launches(), "info", E.info())
    E.close()


if __name__ == "__main__":
    main()
