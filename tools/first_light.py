"""First GPU contact: engine parity vs torch fp64 (cuBLAS) + numpy oracle, and first timings."""
import sys, os, time, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import numpy as np, torch
from capital_b200.engine import Engine, colmajor, to_colmajor
import capital_oracle as orc

torch.manual_seed(0)
E = Engine(0)
rel = lambda a, b: (torch.linalg.norm(a - b) / torch.linalg.norm(b)).item()

def tgemm(m, n, k, ta, tb, alpha=1.0, beta=0.0):
    A = to_colmajor(torch.rand((k, m) if ta == "T" else (m, k), dtype=torch.float64))
    B = to_colmajor(torch.rand((n, k) if tb == "T" else (k, n), dtype=torch.float64))
    Cm = to_colmajor(torch.rand((m, n), dtype=torch.float64)); C0 = Cm.clone()
    E._gemm(A, B, Cm, m, n, k, ta, tb, alpha, beta); torch.cuda.synchronize()
    ref = alpha * ((A.t() if ta == "T" else A) @ (B.t() if tb == "T" else B)) + beta * C0
    return rel(Cm, ref)

for (m, n, k) in [(128, 128, 64), (256, 384, 512), (100, 77, 130), (513, 255, 1025), (64, 64, 8192)]:
    for ta in "NT":
        for tb in "NT":
            print("gemm", m, n, k, ta, tb, "%.2e" % tgemm(m, n, k, ta, tb, 1.5, 0.5), flush=True)

# syrk / trmm
for n, k in [(256, 4096), (300, 1000), (1024, 65536)]:
    A = to_colmajor(torch.rand((k, n), dtype=torch.float64)); G = colmajor(n, n, fill=7.0)
    E._syrk(A, G, n, k, "U", "T"); torch.cuda.synchronize()
    ref = torch.triu(A.t() @ A) + torch.tril(torch.full((n, n), 7.0, dtype=torch.float64, device="cuda"), -1)
    print("syrk UT", n, k, "%.2e" % rel(G, ref), flush=True)
for side in "LR":
    for uplo in "UL":
        for tr in "NT":
            m, n = 384, 200
            nt = m if side == "L" else n
            T = to_colmajor(torch.rand((nt, nt), dtype=torch.float64)); B = to_colmajor(torch.rand((m, n), dtype=torch.float64)); B0 = B.clone()
            Tt = torch.triu(T) if uplo == "U" else torch.tril(T); op = Tt.t() if tr == "T" else Tt
            ref = 2.0 * (op @ B0 if side == "L" else B0 @ op)
            E._trmm(T, B, m, n, side, uplo, tr, "N", 2.0); torch.cuda.synchronize()
            print("trmm", side, uplo, tr, "%.2e" % rel(B, ref), flush=True)

# cholinv / potrf / trtri
for n in [64, 128, 200, 512, 1024, 2048]:
    A_np = orc.symmetric_global(n); A = to_colmajor(A_np); X = colmajor(n, n)
    E.cholinv(A, X, n, True); torch.cuda.synchronize()
    R = torch.triu(A); Rref = torch.linalg.cholesky(torch.as_tensor(A_np, device="cuda")).t()
    print("cholinv", n, "R %.2e" % rel(R, Rref), "Rinv %.2e" % rel(X, torch.linalg.inv(Rref)), "info", E.info(), flush=True)
n = 700
Tn = np.triu(np.random.rand(n, n)) + n * np.eye(n); T = to_colmajor(Tn); E._trtri(T, n, "U", "N"); torch.cuda.synchronize()
print("trtri U", "%.2e" % rel(torch.triu(T), torch.linalg.inv(torch.as_tensor(Tn, device="cuda"))))
T = to_colmajor(Tn.T.copy()); E._trtri(T, n, "L", "N"); torch.cuda.synchronize()
print("trtri L", "%.2e" % rel(torch.tril(T), torch.linalg.inv(torch.as_tensor(Tn.T.copy(), device="cuda"))))

# generators + CQR2 vs numpy oracle
m, n = 8192, 128
A = colmajor(m, n); E.fill_random(A, 0); A_np = orc.distribute_random(n, m, 0, 0, 1, 1, 0)
print("fill_random bit-exact:", bool((A.cpu().numpy() == A_np).all()))
S = colmajor(300, 300); E.fill_symmetric(S, 0, 0, 1, 1, 300, True)
print("fill_symmetric bit-exact:", bool((S.cpu().numpy() == orc.symmetric_global(300)).all()))
Q = colmajor(m, n); R = colmajor(n, n); E.cacqr_factor_1d(A, Q, R, 2); torch.cuda.synchronize()
qo, ro = orc.cacqr_factor_1d([A_np], 2)
print("cacqr2 vs oracle: Q %.2e R %.2e orth %.2e" % (rel(Q, torch.as_tensor(qo[0], device="cuda")), rel(R, torch.as_tensor(ro, device="cuda")),
      orc.qr_orthogonality(Q.cpu().numpy())), flush=True)

def timeit(fn, reps=3):
    fn(); torch.cuda.synchronize(); best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
for N in [4096, 8192, 16384]:
    A = colmajor(N, N); B = colmajor(N, N); Cm = colmajor(N, N); A.uniform_(); B.uniform_()
    for ta, tb in [("N", "N"), ("T", "N"), ("N", "T")]:
        ms = timeit(lambda: E._gemm(A, B, Cm, N, N, N, ta, tb))
        print("our dgemm %s%s N=%d: %.2f ms  %.2f TFLOP/s" % (ta, tb, N, ms, 2 * N ** 3 / ms * 1e-9), flush=True)
    ms = timeit(lambda: torch.matmul(A, B, out=Cm))
    print("cuBLAS dgemm N=%d: %.2f ms  %.2f TFLOP/s" % (N, ms, 2 * N ** 3 / ms * 1e-9), flush=True)
    del A, B, Cm
for (m, n) in [(1 << 19, 1024), (1 << 22, 256)]:
    A = colmajor(m, n); E.fill_random(A, 1); Q = colmajor(m, n); R = colmajor(n, n); G = colmajor(n, n); X = colmajor(n, n)
    ms = timeit(lambda: E.cqr_gram(A, G)); print("gram %dx%d: %.2f ms %.2f TF (mn^2)" % (m, n, ms, m * n * n / ms * 1e-9))
    E.cholinv(G, X, n, True)
    ms = timeit(lambda: E.cholinv(G.clone(), X, n, True)); print("cholinv n=%d: %.3f ms" % (n, ms))
    ms = timeit(lambda: E.cqr_apply(X, A, Q)); print("apply oop: %.2f ms %.2f TF (mn^2)" % (ms, m * n * n / ms * 1e-9))
    ms = timeit(lambda: E.cqr_apply(X, Q, Q)); print("apply inplace: %.2f ms %.2f TF" % (ms, m * n * n / ms * 1e-9))
    ms = timeit(lambda: E.cacqr_factor_1d(A, Q, R, 2)); print("CQR2 %dx%d: %.2f ms %.2f TFLOP/s; orth %.2e" % (m, n, ms, orc.flops_cacqr2(m, n) / ms * 1e-9,
          (torch.linalg.norm(Q.t() @ Q - torch.eye(n, device="cuda", dtype=torch.float64)) / n).item()), flush=True)
    del A, Q
print("launches", E.launches())
