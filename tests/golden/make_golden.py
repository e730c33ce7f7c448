"""Generates tests/golden/*.npz from the UNMODIFIED reference compiled in oracle/_ref (run in the build
container only: needs /root/reference).  Usage: make -C oracle && python tests/golden/make_golden.py

Each fixture stores the reference's own inputs and outputs (per rank) for a small configuration; the
tests compare the numpy oracle (CPU) and the CUDA path (GPU) against them."""
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")


def run(np_ranks, args, tmp):
    env = dict(os.environ, MINIMPI_NP=str(np_ranks), OPENBLAS_NUM_THREADS="1", MINIMPI_STAGE_MB="64")
    out = subprocess.run([DRIVER] + [str(a) for a in args], env=env, check=True, capture_output=True, text=True, timeout=300)
    return out.stdout.strip()


def rd(prefix, rank, name, shape):
    return np.fromfile("%s.r%d.%s.f64" % (prefix, rank, name)).reshape(shape, order="F")


def ceil_div(a, b):
    return -(-a // b)


def gen_cacqr3d(tmp):
    """3D CholeskyQR2 of the reference on 8 mini-MPI ranks, topo::rect(c = 2) = the 2x2x2 cube (cacqr.hpp:197-216 -> sweep_3d
    :77-122), cholinv base case ReplicateComp, complete_inv = 1 and 0 (the latter goes through cacqr::solve, :46-73)."""
    q = {}
    m, n, c, d = 256, 32, 2, 2
    for tag, ci in (("ci1", 1), ("ci0", 0)):
        p = os.path.join(tmp, "q3" + tag)
        line = run(8, ["cacqr", p, 2, m, n, c, ci, 1, 0, 0, 0, 2, 1], tmp)
        q[tag + "_meta"] = np.array([8, 2, m, n, c, d, ci])
        q[tag + "_line"] = np.array(line)
        for r in range(8):
            q["%s_A_r%d" % (tag, r)] = rd(p, r, "A", (m // d, n // c))
            q["%s_Q_r%d" % (tag, r)] = rd(p, r, "Q", (m // d, n // c))
            q["%s_R_r%d" % (tag, r)] = rd(p, r, "R", (n // c, n // c))
    np.savez_compressed(os.path.join(HERE, "cacqr3d.npz"), **q)


def main():
    import sys
    if "--only-cacqr3d" in sys.argv:
        with tempfile.TemporaryDirectory() as tmp:
            gen_cacqr3d(tmp)
        return
    rng = np.random.default_rng(20261104)
    with tempfile.TemporaryDirectory() as tmp:
        # ---- generators -------------------------------------------------------------------------
        g = {}
        for ranks, x, y in ((1, 7, 20), (3, 5, 10), (4, 6, 32)):
            p = os.path.join(tmp, "gr%d" % ranks); run(ranks, ["gen", p, "random", x, y], tmp)
            for r in range(ranks):
                g["random_p%d_x%d_y%d_r%d" % (ranks, x, y, r)] = rd(p, r, "A", (ceil_div(y, ranks), x))
        for ranks, n in ((1, 9), (4, 10), (4, 7)):
            d = int(round(ranks ** 0.5))
            p = os.path.join(tmp, "gs%d_%d" % (ranks, n)); run(ranks, ["gen", p, "symmetric", n, n], tmp)
            for r in range(ranks):
                g["symmetric_p%d_n%d_r%d" % (ranks, n, r)] = rd(p, r, "A", (ceil_div(n, d), ceil_div(n, d)))
        np.savez_compressed(os.path.join(HERE, "generators.npz"), **g)

        # ---- CholeskyQR / CholeskyQR2, 1D -------------------------------------------------------
        q = {}
        for tag, ranks, variant, m, n in (("cqr2_p1", 1, 2, 384, 24), ("cqr2_p4", 4, 2, 512, 32), ("cqr1_p2", 2, 1, 256, 16)):
            p = os.path.join(tmp, tag)
            line = run(ranks, ["cacqr", p, variant, m, n, 1, 1, 1, 0, 0, 0, 0, 1], tmp)
            q[tag + "_meta"] = np.array([ranks, variant, m, n])
            q[tag + "_line"] = np.array(line)
            q[tag + "_R"] = np.triu(rd(p, 0, "R", (n, n)))
            for r in range(ranks):
                q["%s_A_r%d" % (tag, r)] = rd(p, r, "A", (m // ranks, n))
                q["%s_Q_r%d" % (tag, r)] = rd(p, r, "Q", (m // ranks, n))
        np.savez_compressed(os.path.join(HERE, "cacqr.npz"), **q)
        gen_cacqr3d(tmp)

        # ---- cholinv ----------------------------------------------------------------------------
        c = {}
        for tag, ranks, n, ci, bc, policy in (("p1_ci1", 1, 96, 1, -3, 2), ("p1_ci0", 1, 96, 0, -3, 2), ("p8_ci1", 8, 64, 1, -2, 2), ("p8_ci0", 8, 64, 0, -2, 2)):
            p = os.path.join(tmp, "c" + tag)
            line = run(ranks, ["cholinv", p, n, 1, ci, 1, bc, 0, 0, policy, 1], tmp)
            d = 1 if ranks == 1 else 2
            c[tag + "_meta"] = np.array([ranks, n, ci, bc, d])
            c[tag + "_line"] = np.array(line)
            for r in range(ranks):
                for name in ("A", "R", "Rinv"):
                    c["%s_%s_r%d" % (tag, name, r)] = rd(p, r, name, (n // d, n // d))
        np.savez_compressed(os.path.join(HERE, "cholinv.npz"), **c)

        # ---- SUMMA gemm -------------------------------------------------------------------------
        s = {}
        for tag, ranks, (M, N, K) in (("p1", 1, (24, 20, 28)), ("p8", 8, (32, 24, 40))):
            p = os.path.join(tmp, "s" + tag); run(ranks, ["summa", p, M, N, K, 1, 0, 0], tmp)
            d = 1 if ranks == 1 else 2
            s[tag + "_meta"] = np.array([ranks, M, N, K, d])
            for r in range(ranks):
                s["%s_A_r%d" % (tag, r)] = rd(p, r, "A", (M // d, K // d))
                s["%s_B_r%d" % (tag, r)] = rd(p, r, "B", (K // d, N // d))
                s["%s_C_r%d" % (tag, r)] = rd(p, r, "C", (M // d, N // d))
        np.savez_compressed(os.path.join(HERE, "summa.npz"), **s)

        # ---- engines (blas::engine / lapack::engine on OpenBLAS) --------------------------------
        e = {}

        def put(name, arr):
            arr = np.asfortranarray(arr)
            arr.T.tofile(os.path.join(tmp, "eng.in.%s.f64" % name))
            return arr

        pre = os.path.join(tmp, "eng")
        for ta, tb in (("N", "N"), ("T", "N"), ("N", "T"), ("T", "T")):
            m, n, k = 37, 29, 45
            A = put("A", rng.random((k, m) if ta == "T" else (m, k))); B = put("B", rng.random((n, k) if tb == "T" else (k, n))); Cm = put("C", rng.random((m, n)))
            run(1, ["engine", pre, "gemm", ta, tb, m, n, k, A.shape[0], B.shape[0], m, 1.25, -0.5], tmp)
            e["gemm_%s%s_A" % (ta, tb)], e["gemm_%s%s_B" % (ta, tb)], e["gemm_%s%s_C0" % (ta, tb)] = A, B, Cm
            e["gemm_%s%s_C" % (ta, tb)] = rd(pre, 0, "C", (m, n))
        n, k = 33, 70
        A = put("A", rng.random((k, n))); Cm = put("C", rng.random((n, n)))
        run(1, ["engine", pre, "syrk", "U", "T", n, k, k, n, 1.0, 0.0], tmp)
        e["syrk_A"], e["syrk_C0"], e["syrk_C"] = A, Cm, rd(pre, 0, "C", (n, n))
        for side, uplo, tr in (("R", "U", "N"), ("L", "U", "T"), ("L", "U", "N"), ("R", "L", "T")):
            m, n = 41, 26
            nt = m if side == "L" else n
            T = put("A", rng.random((nt, nt)) + np.eye(nt)); B = put("B", rng.random((m, n)))
            run(1, ["engine", pre, "trmm", side, uplo, tr, "N", m, n, nt, m, -1.0], tmp)
            key = "trmm_%s%s%s" % (side, uplo, tr)
            e[key + "_T"], e[key + "_B0"], e[key + "_B"] = T, B, rd(pre, 0, "B", (m, n))
        n = 50
        X = rng.random((n + 20, n)); S = put("A", X.T @ X + n * np.eye(n))
        run(1, ["engine", pre, "potrf", "U", n, n], tmp)
        e["potrf_A"], e["potrf_R"] = S, rd(pre, 0, "A", (n, n))
        Tm = put("A", np.triu(rng.random((n, n))) + n * np.eye(n) + np.tril(rng.random((n, n)), -1))
        run(1, ["engine", pre, "trtri", "U", "N", n, n], tmp)
        e["trtri_T"], e["trtri_X"] = Tm, rd(pre, 0, "A", (n, n))
        np.savez_compressed(os.path.join(HERE, "engines.npz"), **e)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
