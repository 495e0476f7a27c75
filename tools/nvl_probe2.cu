// NVLink probe 2 (measurement tool): does a CTA barrier drain outstanding REMOTE stores?
// The x-edge CTAs of the two-step push sweep ran 3.5x longer with remote neighbours (CTA timelines,
// profiles/r02q_cta_trace_n4.jsonl) although plain remote stores cost nothing extra (tools/nvl_probe.cu).
// That sweep has a __syncthreads() per 4-row chunk; the one-step sweep (mbarrier ring only) showed no
// slow-down.  CTAs of 128 threads; lanes 0..W-1 of warp 0 store into F arrays at stride pitch every row;
// every 4 rows: nothing / __syncthreads() / fence.acq_rel.cta / mbarrier arrive+wait.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)
struct Arrs { double *a[6]; };

__global__ void __launch_bounds__(128) probe(Arrs dst, long long pitch, int rows, int W, int F, int sync, int work, double *sink)
{
    __shared__ uint64_t bar;
    __shared__ double xb[128];
    const int tid = threadIdx.x;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(&bar)), "r"(128));
    }
    __syncthreads();
    const long long row0 = (static_cast<long long>(blockIdx.x) * rows) % 16000;
    double v = 1.0 + tid;
    unsigned phase = 0;
    for (int r = 0; r < rows; ++r) {
        for (int k = 0; k < work; ++k) v = v * 1.0000001 + 1e-9;
        if (tid < W)
            for (int f = 0; f < F; ++f) dst.a[f][(row0 + r) * pitch + tid] = v;
        if ((r & 3) == 3) {
            xb[tid] = v;
            if (sync == 1) {
                __syncthreads();
            } else if (sync == 2) {
                asm volatile("fence.acq_rel.cta;" ::: "memory");
            } else if (sync == 3) {
                uint64_t st;
                asm volatile("mbarrier.arrive.shared::cta.b64 %0, [%1];" : "=l"(st) : "r"((unsigned)__cvta_generic_to_shared(&bar)) : "memory");
                unsigned done = 0;
                while (!done)
                    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                                 : "=r"(done) : "r"((unsigned)__cvta_generic_to_shared(&bar)), "r"(phase) : "memory");
                phase ^= 1;
            } else if (sync == 4) {
                // named barrier over the three warps that do NOT store remotely + warp 0 only arrives
                if (tid < 32) asm volatile("bar.arrive 1, 128;" ::: "memory");
                else asm volatile("bar.sync 1, 128;" ::: "memory");
            }
            v += xb[(tid + 1) & 127];
        }
    }
    if (v == 123.456) *sink = v;
}

int main()
{
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (ndev < 2) { printf("needs 2 GPUs\n"); return 0; }
    const long long pitch = 16448, ny = 16400;
    Arrs loc{}, rem{};
    double *sink;
    CK(cudaSetDevice(1));
    for (int f = 0; f < 6; ++f) CK(cudaMalloc(&rem.a[f], pitch * ny * 8));
    CK(cudaSetDevice(0));
    for (int f = 0; f < 6; ++f) CK(cudaMalloc(&loc.a[f], pitch * ny * 8));
    CK(cudaMalloc(&sink, 8));
    CK(cudaDeviceEnablePeerAccess(1, 0));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int rows = 124;
    const char *names[] = {"none", "syncthreads", "fence.cta", "mbarrier", "bar.arrive(storing warp)+bar.sync(others)"};
    printf("where W F work ctas sync us_per_launch\n");
    for (int where = 0; where < 2; ++where)
        for (int W : {1, 2})
            for (int F : {3, 6})
                for (int work : {0, 100})
                    for (int ctas : {1, 64, 264})
                        for (int sync = 0; sync < 5; ++sync) {
                            const Arrs &d = where ? rem : loc;
                            for (int it = 0; it < 3; ++it) probe<<<ctas, 128>>>(d, pitch, rows, W, F, sync, work, sink);
                            CK(cudaDeviceSynchronize());
                            CK(cudaEventRecord(e0));
                            for (int it = 0; it < 10; ++it) probe<<<ctas, 128>>>(d, pitch, rows, W, F, sync, work, sink);
                            CK(cudaEventRecord(e1));
                            CK(cudaEventSynchronize(e1));
                            float ms;
                            CK(cudaEventElapsedTime(&ms, e0, e1));
                            printf("%s %d %d %d %d \"%s\" %.1f\n", where ? "remote" : "local", W, F, work, ctas, names[sync], ms * 100.0);
                        }
    return 0;
}
