// qcb_host_tuner.h -- plain C++ (no CUDA): included by qcb200.cu and by tests/test_host_tuner.py's g++ harness.
#pragma once
#include <algorithm>

// Chooses the number of host threads that expand the arenas for one kind of host-buffer call on the large-d
// path by timing the calls themselves: one settling call, then two calls with each candidate (all pool threads,
// 3/4, 1/2 and 3/8 of them), then the fastest stays -- fewer threads only if that is at least 3 % faster.
// Why it can pay: the cores' streaming stores and the PCIe writes share the host's DRAM write bandwidth
// (scratch/dma_under_stores.py: D2H 57 GB/s alone, 32 GB/s beside 8 or 16 store threads, 44 GB/s beside 4), and
// a fused call needs both to finish.
struct HostTuner {
    static constexpr int NC = 4;
    int calls = 0, chosen = 0;
    double best[NC] = {1e30, 1e30, 1e30, 1e30};
    static int candidate(int pool, int i) {
        const int num[NC] = {8, 6, 4, 3};
        return i == 0 ? pool : std::max(2, pool * num[i] / 8);
    }
    int next(int pool) const {
        if (chosen) return chosen;
        if (pool < 8 || calls == 0) return pool;
        return candidate(pool, ((calls - 1) / 2) % NC);
    }
    void record(int pool, double ms) {
        if (chosen) return;
        if (pool < 8) { chosen = pool; return; }
        if (calls > 0) { const int i = ((calls - 1) / 2) % NC; best[i] = std::min(best[i], ms); }
        if (++calls == 1 + 2 * NC) {
            int b = 0;
            for (int i = 1; i < NC; ++i) if (best[i] < 0.97 * best[0] && best[i] < best[b]) b = i;
            chosen = candidate(pool, b);
        }
    }
};
