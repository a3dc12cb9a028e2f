"""CPU test of the store-thread choice (csrc/qcb_host_tuner.h, plain C++): a g++ harness feeds it call times."""
import os
import subprocess

from conftest import ROOT

HARNESS = r'''
#include <cstdio>
#include <cstdlib>
#include "qcb_host_tuner.h"
// argv: pool, then one time per thread count "t:ms" (the call time when `t` threads are used)
int main(int argc, char **argv) {
    const int pool = atoi(argv[1]);
    HostTuner t;
    for (int call = 0; call < 12; ++call) {
        const int n = t.next(pool);
        double ms = 100.0;
        for (int a = 2; a < argc; ++a) { int th; double v; if (sscanf(argv[a], "%d:%lf", &th, &v) == 2 && th == n) ms = v; }
        if (call == 0) ms *= 3.0;                 // the settling call (allocations) must not count
        printf("%d ", n);
        t.record(pool, ms);
    }
    printf("-> %d\n", t.chosen);
    return 0;
}
'''


def _run(tmp_path, *args):
    src = tmp_path / "tuner.cpp"
    exe = tmp_path / "tuner"
    if not exe.exists():
        src.write_text(HARNESS)
        subprocess.run(["g++", "-O1", "-std=c++17", "-I", os.path.join(ROOT, "quantumcollocation.jl_b200", "csrc"),
                        str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe), *map(str, args)], check=True, capture_output=True, text=True).stdout.split()
    return [int(x) for x in out[:-2]], int(out[-1])


def test_exploration_order_and_choice(tmp_path):
    seq, chosen = _run(tmp_path, 16)                       # every candidate equally fast: all threads stay
    assert seq[:9] == [16, 16, 16, 12, 12, 8, 8, 6, 6] and chosen == 16 and seq[9:] == [16, 16, 16]
    seq, chosen = _run(tmp_path, 16, "8:90")               # half the threads 10 % faster
    assert chosen == 8 and seq[9:] == [8, 8, 8]
    _, chosen = _run(tmp_path, 16, "12:98", "6:99")        # gains below 3 % do not move it
    assert chosen == 16
    _, chosen = _run(tmp_path, 16, "12:95", "8:93", "6:120")
    assert chosen == 8
    _, chosen = _run(tmp_path, 16, "16:80")                # the slow settling call alone must not favour fewer threads
    assert chosen == 16
    seq, chosen = _run(tmp_path, 4, "2:10")                # small pools are not tuned
    assert chosen == 4 and set(seq) == {4}
    seq, chosen = _run(tmp_path, 9)                        # candidates of an odd pool stay distinct and >= 2
    assert seq[:9] == [9, 9, 9, 6, 6, 4, 4, 3, 3] and chosen == 9
