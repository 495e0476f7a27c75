// NVLink small-store probe (measurement tool, not product code).
// How expensive are the x-edge halo stores of a push sweep when the neighbour is another GPU?  One
// process, two devices, peer access; CTAs of one warp store W consecutive doubles per row into each of
// F arrays at stride `pitch` (the shape of the x-edge push: W = 1 or 2, F = 3 or 6), or one packed strip
// row of F*W doubles (the alternative).  Reports time per launch, ns per store instruction per CTA and
// the aggregate store rate, local vs remote.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/nvl_probe.cu -o tools/_build/nvl_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

struct Arrs { double *a[6]; };

// mode 0: F arrays, W lanes store one double each per row at stride pitch
// mode 1: one packed strip: row r of the CTA's segment is F*W consecutive doubles
// work: dependent FP64 operations per row between the stores (the sweep's own arithmetic stands in)
__global__ void probe(Arrs dst, long long pitch, int rows, int W, int F, int mode, int work, int fence, double *sink)
{
    const int lane = threadIdx.x;
    const long long row0 = (static_cast<long long>(blockIdx.x) * rows) % 16000;
    double v = 1.0 + lane;
    for (int r = 0; r < rows; ++r) {
        for (int k = 0; k < work; ++k) v = v * 1.0000001 + 1e-9;
        if (mode == 0) {
            if (lane < W)
                for (int f = 0; f < F; ++f) dst.a[f][(row0 + r) * pitch + lane] = v;
        } else {
            if (lane < W * F) dst.a[0][(row0 + r) * (W * F) + lane] = v;
        }
    }
    if (fence) __threadfence_system();
    if (v == 123.456) *sink = v;
}

__global__ void busy_copy(double2 *__restrict__ d, const double2 *__restrict__ s, size_t n)
{
    for (size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < n; t += static_cast<size_t>(gridDim.x) * blockDim.x)
        d[t] = s[t];
}
static double2 *busy_src, *busy_dst;

int main(int argc, char **argv)
{
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    const long long pitch = 16448, ny = 16400;
    Arrs loc{}, rem{};
    double *sink;
    CK(cudaSetDevice(0));
    for (int f = 0; f < 6; ++f) CK(cudaMalloc(&loc.a[f], pitch * ny * 8));
    CK(cudaMalloc(&sink, 8));
    const bool have_peer = ndev > 1;
    if (have_peer) {
        int can = 0;
        CK(cudaDeviceCanAccessPeer(&can, 0, 1));
        printf("peer access 0->1: %d\n", can);
        CK(cudaSetDevice(1));
        for (int f = 0; f < 6; ++f) CK(cudaMalloc(&rem.a[f], pitch * ny * 8));
        CK(cudaSetDevice(0));
        CK(cudaDeviceEnablePeerAccess(1, 0));
    }
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int rows = 124;
    if (argc > 1) goto both_busy;
    printf("where mode W F work ctas fence us_per_launch ns_per_store_instr_per_cta Mstores_per_s\n");
    for (int where = 0; where < (have_peer ? 2 : 1); ++where)
        for (int mode = 0; mode < 2; ++mode)
            for (int W : {1, 2, 16})
                for (int F : {3, 6})
                    for (int work : {0, 100})
                        for (int ctas : {1, 8, 64, 132, 264, 1056})
                            for (int fence : {0, 1}) {
                                if (mode == 1 && W == 16) continue;
                                if (fence == 0 && work != 0) continue;
                                const Arrs &d = where ? rem : loc;
                                for (int it = 0; it < 3; ++it) probe<<<ctas, 32>>>(d, pitch, rows, W, F, mode, work, fence, sink);
                                CK(cudaDeviceSynchronize());
                                const int reps = 10;
                                CK(cudaEventRecord(e0));
                                for (int it = 0; it < reps; ++it) probe<<<ctas, 32>>>(d, pitch, rows, W, F, mode, work, fence, sink);
                                CK(cudaEventRecord(e1));
                                CK(cudaEventSynchronize(e1));
                                float ms;
                                CK(cudaEventElapsedTime(&ms, e0, e1));
                                const double us = ms * 1000.0 / reps;
                                const double instr = static_cast<double>(rows) * (mode == 0 ? F : 1);
                                printf("%s %d %d %d %d %d %d %.2f %.1f %.1f\n", where ? "remote" : "local", mode, W, F, work, ctas,
                                       fence, us, us * 1000.0 / instr, instr * ctas / us);
                            }
    // ---- destination busy: device 1 streams a copy over its own arrays (HBM saturated) while device 0
    // stores into device 1's memory.  Source time and the slow-down of the destination's copy.
    if (have_peer) {
        cudaStream_t s1;
        cudaEvent_t c0, c1;
        CK(cudaSetDevice(1));
        CK(cudaStreamCreate(&s1));
        CK(cudaEventCreate(&c0));
        CK(cudaEventCreate(&c1));
        const size_t nbytes = static_cast<size_t>(pitch) * ny * 8;
        auto dest_copy = [&](int reps) {   // arrays 3,4,5 of device 1 <- arrays 0,1,2 shifted: untouched by the probe rows? no: plain traffic
            CK(cudaSetDevice(1));
            CK(cudaEventRecord(c0, s1));
            for (int it = 0; it < reps; ++it) busy_copy<<<148 * 16, 256, 0, s1>>>(busy_dst, busy_src, nbytes / 16);
            CK(cudaEventRecord(c1, s1));
            CK(cudaSetDevice(0));
        };
        CK(cudaSetDevice(1));
        CK(cudaMalloc(&busy_src, nbytes));
        CK(cudaMalloc(&busy_dst, nbytes));
        CK(cudaSetDevice(0));
        printf("busy-destination: mode W F ctas | src_us_idle_dest src_us_busy_dest | dest_copy_ms_alone dest_copy_ms_with_stores\n");
        for (int mode = 0; mode < 2; ++mode)
            for (int W : {1, 2, 16})
                for (int F : {3, 6})
                    for (int ctas : {132, 264, 1056, 2112}) {
                        if (mode == 1 && W == 16) continue;
                        const int creps = 6;
                        // destination alone
                        dest_copy(creps);
                        CK(cudaSetDevice(1)); CK(cudaEventSynchronize(c1));
                        float alone; CK(cudaEventElapsedTime(&alone, c0, c1));
                        CK(cudaSetDevice(0));
                        // source alone
                        probe<<<ctas, 32>>>(rem, pitch, rows, W, F, mode, 100, 1, sink);
                        CK(cudaDeviceSynchronize());
                        CK(cudaEventRecord(e0));
                        for (int it = 0; it < 20; ++it) probe<<<ctas, 32>>>(rem, pitch, rows, W, F, mode, 100, 1, sink);
                        CK(cudaEventRecord(e1));
                        CK(cudaEventSynchronize(e1));
                        float idle; CK(cudaEventElapsedTime(&idle, e0, e1));
                        // both
                        dest_copy(creps);
                        CK(cudaEventRecord(e0));
                        for (int it = 0; it < 20; ++it) probe<<<ctas, 32>>>(rem, pitch, rows, W, F, mode, 100, 1, sink);
                        CK(cudaEventRecord(e1));
                        CK(cudaEventSynchronize(e1));
                        float busy; CK(cudaEventElapsedTime(&busy, e0, e1));
                        CK(cudaSetDevice(1)); CK(cudaEventSynchronize(c1));
                        float with; CK(cudaEventElapsedTime(&with, c0, c1));
                        CK(cudaSetDevice(0));
                        printf("busy %d %d %d %d | %.1f %.1f | %.3f %.3f\n", mode, W, F, ctas, idle * 1000 / 20, busy * 1000 / 20,
                               alone / creps, with / creps);
                    }
    }
both_busy:
    // ---- source AND destination busy: both devices stream copies; the probe CTAs run beside the source's copy
    if (have_peer && argc > 1) {
        cudaStream_t s0, s1b, sp;
        double2 *src0, *dst0;
        const size_t nbytes = static_cast<size_t>(pitch) * ny * 8;
        CK(cudaSetDevice(0));
        CK(cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&sp, cudaStreamNonBlocking));
        CK(cudaMalloc(&src0, nbytes));
        CK(cudaMalloc(&dst0, nbytes));
        CK(cudaSetDevice(1));
        CK(cudaStreamCreate(&s1b));
        if (!busy_src) { CK(cudaMalloc(&busy_src, nbytes)); CK(cudaMalloc(&busy_dst, nbytes)); }
        CK(cudaSetDevice(0));
        printf("both-busy: where mode W F ctas burst | probe_us (20 launches back to back beside a streaming copy on the source; dest also copying)\n");
        for (int where = 0; where < 2; ++where)
            for (int mode = 0; mode < 2; ++mode)
                for (int W : {1, 2})
                    for (int F : {3, 6})
                        for (int ctas : {66, 132}) {
                            const Arrs &d = where ? rem : loc;
                            CK(cudaSetDevice(1));
                            for (int it = 0; it < 8; ++it) busy_copy<<<148 * 16, 256, 0, s1b>>>(busy_dst, busy_src, nbytes / 16);
                            CK(cudaSetDevice(0));
                            // the source's copy leaves room for the probe CTAs: 148 x 6 CTAs of 256 threads
                            for (int it = 0; it < 8; ++it) busy_copy<<<148 * 6, 256, 0, s0>>>(dst0, src0, nbytes / 16);
                            CK(cudaEventRecord(e0, sp));
                            for (int it = 0; it < 20; ++it) probe<<<ctas, 32, 0, sp>>>(d, pitch, rows, W, F, mode, 100, 1, sink);
                            CK(cudaEventRecord(e1, sp));
                            CK(cudaEventSynchronize(e1));
                            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
                            const bool overlapped = cudaStreamQuery(s0) == cudaErrorNotReady;
                            CK(cudaDeviceSynchronize());
                            CK(cudaSetDevice(1)); CK(cudaDeviceSynchronize()); CK(cudaSetDevice(0));
                            printf("bothbusy %s %d %d %d %d | %.1f overlapped=%d\n", where ? "remote" : "local", mode, W, F, ctas, ms * 1000 / 20, overlapped ? 1 : 0);
                        }
    }
    return 0;
}
