"""C++ host mirror (capital_b200/host/capital.hpp), the parts that are pure host arithmetic, compiled with g++ and run on the
CPU: the generators' edge-padding rule (structure.hpp:42-43,72-73,110-111; ADVICE round 1: the drivers used to fill the padded
row / column too) against the oracle's restatement, over sizes the grid does and does not divide."""
import os
import subprocess

from conftest import ROOT
import capital_oracle as orc

SRC = r'''
#include "capital_b200/host/capital.hpp"
#include <cstdio>
int main() {
  for (long g = 1; g <= 40; ++g)
    for (long p = 1; p <= 6; ++p)
      for (long c = 0; c < p; ++c) {
        const long d = (long)capital::local_dim(g, p);
        std::printf("%ld %ld %ld %ld %ld\n", g, p, c, d, (long)capital::pad_len(d, g, c, p));
      }
  return 0;
}
'''


def test_pad_len_and_local_dim_match_the_oracle(tmp_path):
    src = tmp_path / "padlen.cpp"
    src.write_text(SRC)
    exe = str(tmp_path / "padlen")
    subprocess.run(["g++", "-std=c++17", "-O1", "-I" + ROOT, str(src), "-o", exe, "-L" + os.path.join(ROOT, "capital_b200"),
                    "-l:libcapital_b200.so", "-Wl,--allow-shlib-undefined", "-Wl,-rpath," + os.path.join(ROOT, "capital_b200")], check=True)
    env = dict(os.environ, LD_LIBRARY_PATH=os.pathsep.join(filter(None, [os.environ.get("LD_LIBRARY_PATH"), "/usr/local/cuda/lib64"])))
    import torch  # noqa: F401  (its bundled libnccl is what the library links against)
    import sysconfig
    env["LD_LIBRARY_PATH"] += os.pathsep + os.path.join(sysconfig.get_paths()["purelib"], "nvidia", "nccl", "lib")
    out = subprocess.run([exe], capture_output=True, text=True, check=True, env=env).stdout
    rows = [tuple(int(v) for v in line.split()) for line in out.strip().splitlines()]
    assert len(rows) == 40 * (1 + 2 + 3 + 4 + 5 + 6)
    for g, p, c, d, pad in rows:
        assert d == orc.local_dim(g, p)
        assert pad == orc._pad_len(g, p, c, d), (g, p, c)
    # the cases ADVICE named: m = 10001 rows over 2 ranks -> rank 1 holds 5001 rows of which the last is padding
    assert orc._pad_len(10001, 2, 1, 5001) == 5000 and orc._pad_len(10001, 2, 0, 5001) == 5001
