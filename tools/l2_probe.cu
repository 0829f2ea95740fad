// l2_probe.cu — read-bandwidth microbenchmark for the roofline denominators SURVEY.md §8(d) asks for beside the HBM
// copy peak of MEASURED_PEAKS.json: the L2 -> SM read bandwidth (working set well inside the 126 MB L2, re-read many
// times) and, for orientation, L1-resident reads with the same kernel.  The render kernel's node
// fetches hit in L1 / L2, so these are the ceilings its achieved fetch bandwidth is to be compared with.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/l2_probe tools/l2_probe.cu && tools/l2_probe > profiles/l2_peak.json
// Prints one JSON object: {"sm_count": .., "detail": [...], "l1_gbs": .., "l2_gbs": ..}.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

// Every block sweeps the whole working set `reps` times with 16-byte loads (ld.global.cg when CG: L1 bypassed, so the
// traffic is served by L2; plain ld.global otherwise), starting at a block-dependent offset so that the blocks do not
// walk in lockstep.  words = number of float4 in the set (a multiple of the block size).
template <bool CG>
__global__ void __launch_bounds__(256) sweep(const float4* __restrict__ buf, size_t words, int reps, float* sink) {
    float acc = 0.f;
    const size_t start = ((size_t)blockIdx.x * 9973u * 256u) % words;
    for (int r = 0; r < reps; ++r) {
        for (size_t i = threadIdx.x; i < words; i += 256) {
            size_t j = start + i; if (j >= words) j -= words;
            const float4 v = CG ? __ldcg(buf + j) : __ldg(buf + j);
            acc += v.x + v.y + v.z + v.w;
        }
    }
    if (acc == 123456.789f) *sink = acc;          // never true: keeps the loads alive
}

int main() {
    int dev = 0, sms = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const size_t maxBytes = (size_t)64 << 20;
    float4* buf = nullptr; float* sink = nullptr;
    CK(cudaMalloc(&buf, maxBytes)); CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(buf, 0, maxBytes));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    struct Case { const char* name; size_t bytes; bool cg; int blocksPerSm; };
    // l1: a 32 KB set per block with plain loads (L1-resident after the first sweep);
    // l2_*: one set shared by all blocks, L1 bypassed (ld.global.cg), well inside the 126 MB L2
    const Case cases[] = {{"l1_32KB", (size_t)32 << 10, false, 4}, {"l2_8MB", (size_t)8 << 20, true, 8},
                          {"l2_32MB", (size_t)32 << 20, true, 8}, {"l2_64MB", (size_t)64 << 20, true, 8}};
    const int nCases = 4;
    double best[4] = {0, 0, 0, 0};
    printf("{\"sm_count\": %d, \"detail\": [", sms);
    for (int c = 0; c < nCases; ++c) {
        const Case& k = cases[c];
        const size_t words = k.bytes / 16;
        const int blocks = sms * k.blocksPerSm;
        int reps = (int)(40e9 / ((double)blocks * (double)k.bytes));      // about 40 GB of loads per launch
        if (reps < 1) reps = 1;
        for (int t = 0; t < 5; ++t) {
            CK(cudaEventRecord(e0));
            if (k.cg) sweep<true><<<blocks, 256>>>(buf, words, reps, sink);
            else sweep<false><<<blocks, 256>>>(buf, words, reps, sink);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            CK(cudaGetLastError());
            float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
            const double gbs = (double)blocks * (double)k.bytes * reps / (ms * 1e-3) / 1e9;
            if (t > 0 && gbs > best[c]) best[c] = gbs;                     // the first launch warms the set
        }
        printf("%s{\"case\": \"%s\", \"bytes\": %zu, \"blocks\": %d, \"reps\": %d, \"gbs\": %.1f}", c ? ", " : "", k.name, k.bytes, blocks, reps, best[c]);
    }
    double l2 = best[1]; for (int c = 2; c < nCases; ++c) if (best[c] > l2) l2 = best[c];
    printf("], \"l1_gbs\": %.1f, \"l2_gbs\": %.1f}\n", best[0], l2);
    return 0;
}
