// dlesm_static_scratch.cu -- dlesm_dev_copy_patch_static: the in-field patch copy of
// copy_field(field, src_region, dest_region) for callers that must not allocate device memory.
//
// Who this is for.  dlesm_dev_copy_patch stages an overlapping copy through a scratch buffer that
// the library cudaMalloc's (and grows) on demand.  Some hosts of this library cannot afford that
// call: a model whose own pool (an OpenACC / CUDA Fortran runtime, a caching allocator) has already
// taken all of the GPU's memory, so that a further cudaMalloc fails; a time-step loop that must not
// take the device-wide synchronisation a cudaMalloc / cudaFree implies; a code that runs under a
// fixed memory budget it has to account for up front.  For them the staging area below is part of
// the library image: it exists as soon as the library is loaded on a GPU, costs
// DLESM_STATIC_SCRATCH_DOUBLES * 8 bytes on every GPU that loads the library until the process
// exits, and is never resized.  Nothing in this file calls cudaMalloc, cudaFree, the library's
// context (which allocates on first use) or any other allocating routine.
//
// The price: ONE staging area per GPU, shared by every call.  It keeps whatever the last call left
// in it; that is harmless, because the stage-in kernel overwrites every cell the stage-out kernel
// reads (nothing here has to start at zero, so nothing is cleared).  Two calls whose kernels can be
// in flight at the same time on one GPU (different streams, different host threads) race on it:
// keep calls on one stream, or order them yourself.
//
// No counterpart in the reference (it has no device code); the result is the one of
// src/field_mod.f90:1179-1187, array-assignment semantics (right-hand side read first).
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "dlesm_host.hpp"
#include "dlesm_kernels.cuh"

namespace {

// The staging area.  In HOST code this name is only a handle for the symbol -- not an address a
// kernel can use -- so the launches below get the pointer cudaGetSymbolAddress returns for the
// current device (looked up once per device, see staging_area()).
static __device__ __align__(256) double g_static_scratch[DLESM_STATIC_SCRATCH_DOUBLES];

constexpr int kMaxDevices = 64;
constexpr int kSMs = 148;
std::mutex g_mu;
double *g_area[kMaxDevices] = {};   // per device: device address of g_static_scratch

// Device address of the staging area on the CURRENT device.  The first call on a device asks the
// runtime for it (this also makes the runtime load the library's device code on that device).
cudaError_t staging_area(double **area, int *device)
{
    cudaError_t e = cudaGetDevice(device);
    if (e != cudaSuccess) return e;
    if (*device < 0 || *device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_area[*device]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_area[*device] = static_cast<double *>(p);
    }
    *area = g_area[*device];
    return cudaSuccess;
}

__device__ inline size_t cell(int pitch, int i, int j)   // 1-based (i, j), column-major
{
    return (static_cast<size_t>(j) - 1) * static_cast<size_t>(pitch) + (static_cast<size_t>(i) - 1);
}

// w x h cells from (si0, sj0) of `src` to (di0, dj0) of `dst`; one of the two is the packed
// staging area (pitch w, origin (1,1)), or neither when the patches are disjoint.  Coalesced along i.
__global__ void __launch_bounds__(256) static_patch_kernel(const double *src, int spitch, int si0,
                                                           int sj0, double *dst, int dpitch, int di0,
                                                           int dj0, int w, int h)
{
    const long long n = static_cast<long long>(w) * h;
    for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < n;
         t += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int jj = static_cast<int>(t / w), ii = static_cast<int>(t - static_cast<long long>(jj) * w);
        dst[cell(dpitch, di0 + ii, dj0 + jj)] = src[cell(spitch, si0 + ii, sj0 + jj)];
    }
}

int fail(int code, const std::string &msg)
{
    dlesm::set_error(code, "dlesm_dev_copy_patch_static: " + msg);
    return code;
}

int fail_cuda(cudaError_t e, const char *what)
{
    return fail(DLESM_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e) + " (no CPU fallback)");
}

}  // namespace

// field(dx.., dy..) = field(sx.., sy..) inside ONE device array of nx x ny doubles (column-major,
// 1-based inclusive boxes), like dlesm_dev_copy_patch -- but `field` may be any device memory of
// the caller (it is not looked up in the library's registry, hence nx and ny), the work goes to the
// caller's `stream` (a cudaStream_t; NULL = the default stream) on the CURRENT device, and an
// overlapping copy is staged through the library's built-in staging area instead of an allocation.
// `blocking` != 0 waits for the stream before returning.
//
// Limit: overlapping patches of more than DLESM_STATIC_SCRATCH_DOUBLES cells are refused with
// DLESM_ERR_ARGS (disjoint patches go directly and have no limit).
// First call on a device: looks up the staging area's device address (cudaGetSymbolAddress).  Make
// that first call during set-up, before any stream capture is begun on the device.
extern "C" int dlesm_dev_copy_patch_static(void *field, int nx, int ny, int sxs, int sxe, int sys, int sye, int dxs,
                                           int dxe, int dys, int dye, void *stream, int blocking)
{
    if (!field || nx < 1 || ny < 1) return fail(DLESM_ERR_ARGS, "null field or empty allocation");
    const int w = sxe - sxs + 1, h = sye - sys + 1;
    if (w != dxe - dxs + 1 || h != dye - dys + 1)
        return fail(DLESM_ERR_ARGS, "source and destination patches differ in shape");
    if (w <= 0 || h <= 0) return DLESM_OK;
    if (sxs < 1 || sys < 1 || dxs < 1 || dys < 1 || sxe > nx || dxe > nx || sye > ny || dye > ny)
        return fail(DLESM_ERR_ARGS, "patch outside the field allocation");
    const bool overlap = !(sxe < dxs || dxe < sxs || sye < dys || dye < sys);
    const long long n = static_cast<long long>(w) * h;
    if (overlap && n > static_cast<long long>(DLESM_STATIC_SCRATCH_DOUBLES))
        return fail(DLESM_ERR_ARGS, "overlapping patches of " + std::to_string(n) + " cells exceed the built-in staging area of " +
                                        std::to_string(DLESM_STATIC_SCRATCH_DOUBLES) +
                                        " cells (DLESM_STATIC_SCRATCH_DOUBLES); split the copy, or use dlesm_dev_copy_patch");
    double *area = nullptr;
    int device = -1;
    cudaError_t e = staging_area(&area, &device);
    if (e != cudaSuccess) return fail_cuda(e, "staging area set-up");
    // a clean refusal instead of a faulting kernel when handed host memory
    cudaPointerAttributes at{};
    e = cudaPointerGetAttributes(&at, field);
    if (e != cudaSuccess) return fail_cuda(e, "cudaPointerGetAttributes");
    if (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged)
        return fail(DLESM_ERR_UNKNOWN_PTR, "field is not device memory");
    if (at.type == cudaMemoryTypeDevice && at.device != device)
        return fail(DLESM_ERR_STATE, "field lives on device " + std::to_string(at.device) + ", the current device is " +
                                         std::to_string(device));

    cudaStream_t s = static_cast<cudaStream_t>(stream);
    double *f = static_cast<double *>(field);
    const unsigned blocks = n / 256 + 1 < kSMs * 8 ? static_cast<unsigned>(n / 256 + 1) : kSMs * 8;
    if (overlap) {
        static_patch_kernel<<<blocks, 256, 0, s>>>(f, nx, sxs, sys, area, w, 1, 1, w, h);
        static_patch_kernel<<<blocks, 256, 0, s>>>(area, w, 1, 1, f, nx, dxs, dys, w, h);
        dlesm::k::note_launch(2);
    } else {
        static_patch_kernel<<<blocks, 256, 0, s>>>(f, nx, sxs, sys, f, nx, dxs, dys, w, h);
        dlesm::k::note_launch(1);
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "kernel launch");
    if (blocking) {
        e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) return fail_cuda(e, "cudaStreamSynchronize");
    }
    return DLESM_OK;
}
