// Host-layout timing probe (diagnostic): rtl::build_layout / build_top / build_refit_graph on given scene bytes.  See tools/layout_probe.py.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <cmath>
#include <vector>
#include <thread>
#include <atomic>
#include <algorithm>
#include "rt_flat.h"
#include "rt_layout.h"
extern "C" void time_layout(const rt_flat_node* nd, int nNodes, const int32_t* idx, int nIdx, const rt_flat_shape* sh, int nShapes, double omax, double* out) {
    using clk = std::chrono::steady_clock;
    const int reps = 10;
    double t[4] = {0,0,0,0};
    for (int r = 0; r < reps; ++r) {
        rtl::Layout L;
        auto a = clk::now();
        rtl::build_layout(nd, nNodes, idx, (size_t)nIdx, sh, nShapes, true, omax, L);
        auto b = clk::now();
        rtl::build_top(L, nd, nNodes);
        auto c = clk::now();
        rtl::build_refit_graph(L);
        auto d = clk::now();
        t[0] += std::chrono::duration<double, std::milli>(b - a).count();
        t[1] += std::chrono::duration<double, std::milli>(c - b).count();
        t[2] += std::chrono::duration<double, std::milli>(d - c).count();
        out[4] = (double)L.src.size(); out[5] = (double)L.srcB.size(); out[6] = (double)L.nSubA; out[7] = L.costA; out[3] = L.costB;
    }
    for (int i = 0; i < 3; ++i) out[i] = t[i] / reps;
}
