// dlesm_kernels.cu -- hand-written sm_100a FP64 kernels for the field_mod operations, the
// halo pack/unpack, and the un-fused shallow-water sweeps.  Everything here is HBM-bound
// byte shuffling or low-intensity FP64 arithmetic: coalesced along the contiguous i
// dimension, 16-byte vectors where alignment allows, grids sized in multiples of 148 SMs,
// no tensor cores.  Compiled with -fmad=false so that each arithmetic operation rounds once,
// like the reference's Fortran loops built without FMA (compiler_setup/gnu.sh:7-8) and like
// the CPU oracle: stencil results are bit-identical, not merely within 1e-12.
#include <algorithm>
#include "dlesm_kernels.cuh"

#include <atomic>

namespace dlesm {
namespace k {

namespace {
std::atomic<long long> g_launches{0};
inline void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
constexpr int kSMs = 148;

__host__ __device__ inline size_t idx(int pitch, int i, int j)
{
    return (static_cast<size_t>(j) - 1) * static_cast<size_t>(pitch) + (static_cast<size_t>(i) - 1);
}
inline unsigned cdiv(long long a, long long b) { return static_cast<unsigned>((a + b - 1) / b); }
}  // namespace

// Bound of every in-kernel wait for another rank, in SM clock cycles; 0 = wait for ever (like an MPI wait).
// Set once per device by set_wait_limit (DLESM_WAIT_TIMEOUT_S, dlesm_runtime.cu); read on the slow path only.
__device__ long long g_wait_limit = 0;
__device__ __forceinline__ bool wait_expired(long long t0)
{
    const long long lim = *reinterpret_cast<volatile long long *>(&g_wait_limit);
    return lim > 0 && clock64() - t0 > lim;
}
cudaError_t set_wait_limit(long long cycles) { return cudaMemcpyToSymbol(g_wait_limit, &cycles, sizeof(cycles)); }

long long launch_count() { return g_launches.load(); }
void note_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---------------------------------------------------------------------------------------
// copy_field (src/field_mod.f90:1152-1159): out(:,:) = in(:,:), padding included.
// 16 B/point.  Every CTA copies one contiguous 16 KB chunk (4 x 16-byte loads in flight per
// thread) and CTAs are laid out in memory order, so the CTAs resident at any moment sweep one
// contiguous window of DRAM -- the same lockstep effect that lifted the fused sweep (DESIGN.md
// section 4); the earlier grid-stride form (accesses ~10 MB apart) reached 0.95 of the measured
// copy bandwidth, the generated PSy copy sweep with this layout 1.05.
// ---------------------------------------------------------------------------------------
constexpr int kCopyUnroll = 4;
__global__ void __launch_bounds__(256) copy_kernel(const double2 *__restrict__ in,
                                                   double2 *__restrict__ out, size_t n2,
                                                   const double *__restrict__ in_tail,
                                                   double *__restrict__ out_tail, int tail)
{
    const size_t base = static_cast<size_t>(blockIdx.x) * (256 * kCopyUnroll) + threadIdx.x;
    double2 v[kCopyUnroll];
#pragma unroll
    for (int q = 0; q < kCopyUnroll; ++q)
        if (base + q * 256 < n2) v[q] = __ldcs(in + base + q * 256);
#pragma unroll
    for (int q = 0; q < kCopyUnroll; ++q)
        if (base + q * 256 < n2) __stcs(out + base + q * 256, v[q]);
    if (blockIdx.x == 0 && static_cast<int>(threadIdx.x) < tail) out_tail[threadIdx.x] = in_tail[threadIdx.x];
}

cudaError_t copy_field(const double *in, double *out, size_t n, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    const size_t n2 = n / 2;
    const int tail = static_cast<int>(n - 2 * n2);
    const size_t per_cta = 256 * kCopyUnroll;
    const unsigned blocks = static_cast<unsigned>((n2 + per_cta - 1) / per_cta);
    copy_kernel<<<blocks ? blocks : 1, 256, 0, s>>>(reinterpret_cast<const double2 *>(in),
                                                   reinterpret_cast<double2 *>(out), n2, in + 2 * n2,
                                                   out + 2 * n2, tail);
    count_launch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// set_field (src/field_mod.f90:1191-1202): fld%data = val.  8 B/point.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) set_kernel(double2 *__restrict__ out, size_t n2, double val,
                                                  double *__restrict__ out_tail, int tail)
{
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    const double2 v2 = make_double2(val, val);
    for (size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < n2; t += stride)
        __stcs(out + t, v2);
    if (blockIdx.x == 0 && static_cast<int>(threadIdx.x) < tail) out_tail[threadIdx.x] = val;
}

cudaError_t set_field(double *f, size_t n, double val, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    const size_t n2 = n / 2;
    const unsigned blocks = static_cast<unsigned>(
        n2 / 256 + 1 < static_cast<size_t>(kSMs) * 16 ? n2 / 256 + 1 : static_cast<size_t>(kSMs) * 16);
    set_kernel<<<blocks, 256, 0, s>>>(reinterpret_cast<double2 *>(f), n2, val, f + 2 * n2,
                                      static_cast<int>(n - 2 * n2));
    count_launch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// copy_field patch (src/field_mod.f90:1179-1187): f(dest box) = f(src box) inside one array.
// Fortran array assignment evaluates the right-hand side first, so overlapping boxes go
// through a scratch buffer; halo patches never overlap and are copied directly.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) patch_kernel(const double *__restrict__ src, int spitch,
                                                    int si0, int sj0, double *__restrict__ dst,
                                                    int dpitch, int di0, int dj0, int w, int h)
{
    const long long n = static_cast<long long>(w) * h;
    for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < n;
         t += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int jj = static_cast<int>(t / w), ii = static_cast<int>(t - static_cast<long long>(jj) * w);
        dst[idx(dpitch, di0 + ii, dj0 + jj)] = src[idx(spitch, si0 + ii, sj0 + jj)];
    }
}

cudaError_t copy_patch(double *f, int pitch, Box src, Box dst, double *scratch, cudaStream_t s)
{
    const int w = src.xe - src.xs + 1, h = src.ye - src.ys + 1;
    if (w <= 0 || h <= 0) return cudaSuccess;
    const long long n = static_cast<long long>(w) * h;
    const unsigned blocks = n / 256 + 1 < kSMs * 8 ? static_cast<unsigned>(n / 256 + 1) : kSMs * 8;
    const bool overlap = !(src.xe < dst.xs || dst.xs + w - 1 < src.xs || src.ye < dst.ys ||
                           dst.ys + h - 1 < src.ys);
    if (overlap) {
        if (!scratch) return cudaErrorInvalidValue;
        patch_kernel<<<blocks, 256, 0, s>>>(f, pitch, src.xs, src.ys, scratch, w, 1, 1, w, h);
        patch_kernel<<<blocks, 256, 0, s>>>(scratch, w, 1, 1, f, pitch, dst.xs, dst.ys, w, h);
        count_launch();
    } else {
        patch_kernel<<<blocks, 256, 0, s>>>(f, pitch, src.xs, src.ys, f, pitch, dst.xs, dst.ys, w, h);
    }
    count_launch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// field_checksum (src/field_mod.f90:1289-1302): SUM(ABS(f(box))).  8 B/point.
// Deterministic two-pass tree: pass 1 gives every block a fixed set of rows, four
// accumulators per thread, warp-shuffle then shared-memory tree; pass 2 folds the
// kChecksumBlocks partials in a fixed order.  No floating-point atomics, so the result does
// not depend on scheduling.
// ---------------------------------------------------------------------------------------
__device__ inline double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ inline double block_sum_256(double v)
{
    __shared__ double sh[8];
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) sh[wid] = v;
    __syncthreads();
    double r = 0.0;
    if (wid == 0) {
        r = lane < 8 ? sh[lane] : 0.0;
        r = warp_sum(r);
    }
    return r;  // valid in warp 0
}

// One launch: every block leaves its partial, takes a ticket, and the block that draws the last
// ticket folds all partials in index order (so the sum does not depend on which block that is) and
// stores the result on the device and, when asked, straight into mapped pinned host memory -- the
// caller then needs a stream synchronise, no copy.  Blocks beyond the row count leave 0.0: the
// fold tree is the one of a kChecksumBlocks-wide launch whatever the field's height.
__global__ void __launch_bounds__(256) abs_sum_fold_kernel(const double *__restrict__ f, int pitch, Box b,
                                                           double *__restrict__ partials, int npartials,
                                                           unsigned *ticket, double *__restrict__ result,
                                                           double *host_result, unsigned long long seq,
                                                           const __grid_constant__ GlobalSumPlan gs)
{
    const int w = b.xe - b.xs + 1;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    for (int j = b.ys + blockIdx.x; j <= b.ye; j += npartials) {
        const double *row = f + idx(pitch, b.xs, j);
        int i = threadIdx.x;
        for (; i + 768 < w; i += 1024) {
            const double v0 = __ldcs(row + i), v1 = __ldcs(row + i + 256);
            const double v2 = __ldcs(row + i + 512), v3 = __ldcs(row + i + 768);
            a0 += fabs(v0); a1 += fabs(v1); a2 += fabs(v2); a3 += fabs(v3);
        }
        for (; i < w; i += 256) a0 += fabs(__ldcs(row + i));
    }
    const double r = block_sum_256((a0 + a1) + (a2 + a3));
    __shared__ bool last;
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = r;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    double a = 0.0;
    for (int t = threadIdx.x; t < npartials; t += 256) a += t < static_cast<int>(gridDim.x) ? __ldcg(partials + t) : 0.0;
    double tot = block_sum_256(a);
    if (gs.nranks > 1) {
        // global sum over the ranks, through peer memory
        __shared__ double vals[kGsumMaxRanks];
        if (threadIdx.x == 0) vals[0] = tot;
        __syncthreads();
        const double mine = vals[0];
        __syncthreads();
        const int t = threadIdx.x;
        if (t < gs.nranks) {
            gs.peer_cell[t][0] = mine;
            __threadfence_system();
            asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(gs.peer_cell[t] + 1), "l"(gs.seq) : "memory");
            const double *cell = gs.own_cells + 2 * t;
            const long long t0 = clock64();
            for (;;) {
                unsigned long long v;
                asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(cell + 1) : "memory");
                if (v >= gs.seq) break;
                if (wait_expired(t0)) {   // DLESM_WAIT_TIMEOUT_S: a rank never reached this checksum -- fatal
                    atomicExch(gs.err, 1);
                    __threadfence_system();
                    __trap();
                }
                __nanosleep(40);
            }
            vals[t] = __ldcg(cell);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            tot = 0.0;
            for (int r = 0; r < gs.nranks; ++r) tot += vals[r];
        }
    }
    if (threadIdx.x == 0) {
        *result = tot;
        if (host_result) {
            // value, then the sequence number the host polls for (it never has to enter the driver)
            *host_result = tot;
            __threadfence_system();
            *reinterpret_cast<volatile unsigned long long *>(host_result + 1) = seq;
        }
        *ticket = 0u;
    }
}

cudaError_t checksum(const double *f, int pitch, Box b, double *partials, unsigned *ticket, double *result,
                     double *host_result, unsigned long long seq, cudaStream_t s, const GlobalSumPlan *gs)
{
    const int rows = b.ye - b.ys + 1;
    const int blocks = rows < 1 || b.xe < b.xs ? 1 : (rows < kChecksumBlocks ? rows : kChecksumBlocks);
    GlobalSumPlan none{};
    abs_sum_fold_kernel<<<blocks, 256, 0, s>>>(f, pitch, b, partials, kChecksumBlocks, ticket, result, host_result, seq,
                                               gs ? *gs : none);
    count_launch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Periodic halo fill: the copy_field(fld, halo%source, halo%dest) sequence of SURVEY CS-6
// (E, W columns, then N, S rows over xstart-1:xstop+1) for up to 8 fields in one launch.
// Every destination reads its final source directly (the N/S rows take the wrapped
// column for their two end cells), so there is no ordering between the four strips and
// the bits equal those of the sequential list.
// ---------------------------------------------------------------------------------------
struct HaloArgs {
    double *f[kMaxExchangeFields];
    int nfields, pitch;
    Box in;
    int px, py;
};

__global__ void __launch_bounds__(256) periodic_halo_kernel(HaloArgs a)
{
    const int w = a.in.xe - a.in.xs + 1, h = a.in.ye - a.in.ys + 1;
    const int ncol = a.px ? 2 * h : 0;          // E then W column cells
    const int nrow = a.py ? 2 * (w + 2) : 0;    // N then S row cells
    const int per_field = ncol + nrow;
    const long long total = static_cast<long long>(per_field) * a.nfields;
    for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < total;
         t += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int kf = static_cast<int>(t / per_field);
        int c = static_cast<int>(t - static_cast<long long>(kf) * per_field);
        int di, dj, si, sj;
        if (c < ncol) {
            const bool east = c < h;
            const int jj = east ? c : c - h;
            dj = sj = a.in.ys + jj;
            di = east ? a.in.xe + 1 : a.in.xs - 1;
            si = east ? a.in.xs : a.in.xe;
        } else {
            c -= ncol;
            const bool north = c < w + 2;
            const int ii = north ? c : c - (w + 2);
            di = a.in.xs - 1 + ii;
            dj = north ? a.in.ye + 1 : a.in.ys - 1;
            sj = north ? a.in.ys : a.in.ye;
            si = di;
            if (a.px) {  // the end cells were themselves filled by the E/W copies
                if (di == a.in.xs - 1) si = a.in.xe;
                if (di == a.in.xe + 1) si = a.in.xs;
            }
        }
        double *f = a.f[kf];
        f[idx(a.pitch, di, dj)] = f[idx(a.pitch, si, sj)];
    }
}

cudaError_t periodic_halos(double *const *fields, int nfields, int pitch, Box in, bool px, bool py,
                           cudaStream_t s)
{
    if (nfields <= 0 || (!px && !py)) return cudaSuccess;
    if (nfields > kMaxExchangeFields) return cudaErrorInvalidValue;
    HaloArgs a;
    for (int q = 0; q < nfields; ++q) a.f[q] = fields[q];
    a.nfields = nfields; a.pitch = pitch; a.in = in; a.px = px; a.py = py;
    const long long total = static_cast<long long>(nfields) *
                            ((px ? 2LL * (in.ye - in.ys + 1) : 0) + (py ? 2LL * (in.xe - in.xs + 3) : 0));
    const unsigned blocks = total / 256 + 1 < kSMs * 4 ? static_cast<unsigned>(total / 256 + 1) : kSMs * 4;
    periodic_halo_kernel<<<blocks, 256, 0, s>>>(a);
    count_launch();
    return cudaGetLastError();
}

// The same periodic fill to `depth` cells (extension: the reference's halo lists are one cell
// deep, src/field_mod.f90:1394-1464).  Every frame cell reads its wrapped internal source, so
// corners need no ordering.
struct DeepHaloArgs {
    double *f[kMaxExchangeFields];
    int nfields, pitch, depth;
    Box in;
};

__global__ void __launch_bounds__(256) periodic_halo_deep_kernel(DeepHaloArgs a)
{
    const int w = a.in.xe - a.in.xs + 1, h = a.in.ye - a.in.ys + 1, d = a.depth;
    const int wf = w + 2 * d;
    const long long nband = 2LL * d * wf;        // S then N bands, full frame width
    const long long nside = 2LL * d * h;         // W then E side bands
    const long long per_field = nband + nside;
    const long long total = per_field * a.nfields;
    for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < total;
         t += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int kf = static_cast<int>(t / per_field);
        long long c = t - kf * per_field;
        int di, dj;
        if (c < nband) {
            const int rr = static_cast<int>(c / wf), ii = static_cast<int>(c - static_cast<long long>(rr) * wf);
            di = a.in.xs - d + ii;
            dj = rr < d ? a.in.ys - d + rr : a.in.ye + 1 + (rr - d);
        } else {
            c -= nband;
            const int jj = static_cast<int>(c / (2 * d)), ii = static_cast<int>(c - static_cast<long long>(jj) * 2 * d);
            dj = a.in.ys + jj;
            di = ii < d ? a.in.xs - d + ii : a.in.xe + 1 + (ii - d);
        }
        int si = di, sj = dj;
        if (si < a.in.xs) si += w; else if (si > a.in.xe) si -= w;
        if (sj < a.in.ys) sj += h; else if (sj > a.in.ye) sj -= h;
        double *f = a.f[kf];
        f[idx(a.pitch, di, dj)] = f[idx(a.pitch, si, sj)];
    }
}

cudaError_t periodic_halos_deep(double *const *fields, int nfields, int pitch, Box in, int depth, cudaStream_t s)
{
    if (nfields <= 0 || depth <= 0) return cudaSuccess;
    if (nfields > kMaxExchangeFields) return cudaErrorInvalidValue;
    const int w = in.xe - in.xs + 1, h = in.ye - in.ys + 1;
    if (depth > w || depth > h) return cudaErrorInvalidValue;
    DeepHaloArgs a;
    for (int q = 0; q < nfields; ++q) a.f[q] = fields[q];
    a.nfields = nfields; a.pitch = pitch; a.depth = depth; a.in = in;
    const long long total = static_cast<long long>(nfields) * (2LL * depth * (w + 2 * depth) + 2LL * depth * h);
    const unsigned blocks = total / 256 + 1 < kSMs * 4 ? static_cast<unsigned>(total / 256 + 1) : kSMs * 4;
    periodic_halo_deep_kernel<<<blocks, 256, 0, s>>>(a);
    count_launch();
    return cudaGetLastError();
}

// east / west halo columns, `depth` deep, of rows j0..j1 only (periodic in x): what a row band of the
// multi-pass host step needs before the next pass reads it, while other bands are still in flight
__global__ void __launch_bounds__(256) periodic_x_rows_kernel(DeepHaloArgs a)
{
    const int w = a.in.xe - a.in.xs + 1, h = a.in.ye - a.in.ys + 1, d = a.depth;
    const long long per_field = 2LL * d * h;
    const long long total = per_field * a.nfields;
    for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < total;
         t += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int kf = static_cast<int>(t / per_field);
        const long long c = t - kf * per_field;
        const int jj = static_cast<int>(c / (2 * d)), ii = static_cast<int>(c - static_cast<long long>(jj) * 2 * d);
        const int dj = a.in.ys + jj;
        const int di = ii < d ? a.in.xs - d + ii : a.in.xe + 1 + (ii - d);
        const int si = di < a.in.xs ? di + w : di - w;
        double *f = a.f[kf];
        f[idx(a.pitch, di, dj)] = f[idx(a.pitch, si, dj)];
    }
}
cudaError_t periodic_x_rows(double *const *fields, int nfields, int pitch, Box rows, int depth, cudaStream_t s)
{
    if (nfields <= 0 || depth <= 0 || rows.ye < rows.ys) return cudaSuccess;
    if (nfields > kMaxExchangeFields || depth > rows.xe - rows.xs + 1) return cudaErrorInvalidValue;
    DeepHaloArgs a;
    for (int q = 0; q < nfields; ++q) a.f[q] = fields[q];
    a.nfields = nfields; a.pitch = pitch; a.depth = depth; a.in = rows;
    const long long total = static_cast<long long>(nfields) * 2LL * depth * (rows.ye - rows.ys + 1);
    const unsigned blocks = total / 256 + 1 < kSMs ? static_cast<unsigned>(total / 256 + 1) : kSMs;
    periodic_x_rows_kernel<<<blocks, 256, 0, s>>>(a);
    count_launch();
    return cudaGetLastError();
}

// tmask <-> FP64 image of it, so that the mask can ride the FP64 halo machinery once (the
// two-step sweep needs the periodic mask two cells deep; grid_init replicates it outward instead)
__global__ void __launch_bounds__(256) mask_convert_kernel(int *__restrict__ m, double *__restrict__ d, size_t n,
                                                           int to_double)
{
    for (size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < n;
         t += static_cast<size_t>(gridDim.x) * blockDim.x) {
        if (to_double) d[t] = static_cast<double>(m[t]);
        else m[t] = static_cast<int>(d[t]);
    }
}
cudaError_t mask_convert(int *m, double *d, size_t n, bool to_double, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    const unsigned blocks = n / 256 + 1 < kSMs * 8 ? static_cast<unsigned>(n / 256 + 1) : kSMs * 8;
    mask_convert_kernel<<<blocks, 256, 0, s>>>(m, d, n, to_double ? 1 : 0);
    count_launch();
    return cudaGetLastError();
}

// byte image of the tmask for the TMA ring of the fused sweeps: 1 = wet (tmask == 1), 0 otherwise;
// its own row pitch (a multiple of 16 bytes: TMA rows must start 16-byte aligned whatever
// DL_ESM_ALIGNMENT is); columns beyond the array are dry
__global__ void __launch_bounds__(256) mask_to_bytes_kernel(const int *__restrict__ m, int pitch, int ny,
                                                            unsigned char *__restrict__ m8, int pitch8)
{
    const size_t n = static_cast<size_t>(pitch8) * ny;
    for (size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < n;
         t += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const size_t j = t / pitch8, i = t - j * pitch8;
        m8[t] = (i < static_cast<size_t>(pitch) && m[j * pitch + i] == 1) ? 1 : 0;
    }
}
cudaError_t mask_to_bytes(const int *m, int pitch, int ny, unsigned char *m8, int pitch8, cudaStream_t s)
{
    const size_t n = static_cast<size_t>(pitch8) * ny;
    if (n == 0) return cudaSuccess;
    const unsigned blocks = n / 256 + 1 < kSMs * 8 ? static_cast<unsigned>(n / 256 + 1) : kSMs * 8;
    mask_to_bytes_kernel<<<blocks, 256, 0, s>>>(m, pitch, ny, m8, pitch8);
    count_launch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Halo pack / unpack (src/parallel_comms_mod.f90:1678-1683, 1788-1793): every patch of every
// field in one launch; within a patch j is the outer and i the inner index, as the
// reference packs.  blockIdx.y = segment * nfields + field.
// ---------------------------------------------------------------------------------------
template <bool kPack>
__global__ void __launch_bounds__(256) pack_kernel(PackPlan plan, double *__restrict__ buffer)
{
    const int sidx = blockIdx.y / plan.nfields, kf = blockIdx.y - sidx * plan.nfields;
    const Segment sg = plan.seg[sidx];
    const long long n = static_cast<long long>(sg.nx) * sg.ny;
    double *buf = buffer + sg.offset + static_cast<long long>(kf) * n;
    double *f = plan.fields[kf];
    for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < n;
         t += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int jj = static_cast<int>(t / sg.nx), ii = static_cast<int>(t - static_cast<long long>(jj) * sg.nx);
        const size_t q = idx(plan.pitch, sg.i0 + ii, sg.j0 + jj);
        if (kPack) buf[t] = f[q];
        else f[q] = buf[t];
    }
}

static cudaError_t pack_launch(const PackPlan &plan, double *buffer, bool do_pack, cudaStream_t s)
{
    if (plan.nseg <= 0 || plan.nfields <= 0) return cudaSuccess;
    long long nmax = 1;
    for (int q = 0; q < plan.nseg; ++q) {
        const long long n = static_cast<long long>(plan.seg[q].nx) * plan.seg[q].ny;
        if (n > nmax) nmax = n;
    }
    dim3 grid(nmax / 256 + 1 < 64 ? static_cast<unsigned>(nmax / 256 + 1) : 64u,
              static_cast<unsigned>(plan.nseg * plan.nfields));
    if (do_pack) pack_kernel<true><<<grid, 256, 0, s>>>(plan, buffer);
    else pack_kernel<false><<<grid, 256, 0, s>>>(plan, buffer);
    count_launch();
    return cudaGetLastError();
}
cudaError_t pack(const PackPlan &plan, double *buffer, cudaStream_t s) { return pack_launch(plan, buffer, true, s); }
cudaError_t unpack(const PackPlan &plan, const double *buffer, cudaStream_t s)
{
    return pack_launch(plan, const_cast<double *>(buffer), false, s);
}

// ---------------------------------------------------------------------------------------
// Direct halo exchange over peer memory (see XPlan in the header).  blockIdx.y = segment * nfields
// + field as in pack_kernel; the packed layout inside a slot is the same (field-major inside a
// segment, j outer / i inner inside a field), so what lands in the neighbour's slot is byte for
// byte what NCCL would have delivered.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) xdirect_send_kernel(const __grid_constant__ XPlan plan, unsigned *ticket)
{
    const int sidx = blockIdx.y / plan.nfields, kf = blockIdx.y - sidx * plan.nfields;
    const XSegment &sg = plan.seg[sidx];
    const long long n = static_cast<long long>(sg.nx) * sg.ny;
    double *buf = sg.buf + static_cast<long long>(kf) * n;
    const double *f = plan.fields[kf];
    for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < n;
         t += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int jj = static_cast<int>(t / sg.nx), ii = static_cast<int>(t - static_cast<long long>(jj) * sg.nx);
        buf[t] = f[idx(plan.pitch, sg.i0 + ii, sg.j0 + jj)];
    }
    // every thread's peer stores ordered before the ticket; the last block publishes
    __threadfence_system();
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x * gridDim.y - 1;
    __syncthreads();
    if (!last) return;
    if (threadIdx.x < plan.npeers) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(plan.flag[threadIdx.x]), "l"(plan.count[threadIdx.x])
                     : "memory");
    }
    if (threadIdx.x == 0) *ticket = 0u;
}

__global__ void __launch_bounds__(256) xdirect_recv_kernel(const __grid_constant__ XPlan plan, int *err)
{
    if (threadIdx.x < plan.npeers) {
        const long long t0 = clock64();
        for (;;) {
            unsigned long long v;
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(plan.flag[threadIdx.x]) : "memory");
            if (v >= plan.count[threadIdx.x]) break;
            if (wait_expired(t0)) {   // DLESM_WAIT_TIMEOUT_S: the ranks are out of step -- fatal, as in the halo push
                atomicExch(err, 1);
                __threadfence_system();
                __trap();
            }
            __nanosleep(100);
        }
    }
    __syncthreads();
    const int sidx = blockIdx.y / plan.nfields, kf = blockIdx.y - sidx * plan.nfields;
    const XSegment &sg = plan.seg[sidx];
    const long long n = static_cast<long long>(sg.nx) * sg.ny;
    const double *buf = sg.buf + static_cast<long long>(kf) * n;
    double *f = plan.fields[kf];
    for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < n;
         t += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int jj = static_cast<int>(t / sg.nx), ii = static_cast<int>(t - static_cast<long long>(jj) * sg.nx);
        f[idx(plan.pitch, sg.i0 + ii, sg.j0 + jj)] = __ldcg(buf + t);   // written by a peer: not through L1
    }
}

// Both phases in ONE launch.  The grid is small enough to be co-resident (<= 2 CTAs per SM), so CTAs that
// have finished their share of the sends may spin on the neighbours' flags while others still send: the
// flags this rank publishes depend only on its own sends, never on a neighbour's.
// An east / west patch is one element per ROW of the field: every load of the pack is its own DRAM page, ~1.5 us
// of latency.  Work items are therefore small (1024 elements: a 16384-row edge spreads over 16 CTAs, not 4) and a
// thread keeps four loads in flight before it stores (with 4096-element items a thread walked 16 dependent
// load -> peer-store pairs: 16-25 us per exchange whatever its size, profiles/r02w_bench_n8.json config5).
constexpr int kXChunk = 1024;   // elements per work item
__device__ __forceinline__ void xdirect_items(const XPlan &plan, bool send)
{
    int base = 0;
    for (int sidx = 0; sidx < plan.nseg; ++sidx) {
        const XSegment &sg = plan.seg[sidx];
        const int n = sg.nx * sg.ny;
        const int chunks = (n + kXChunk - 1) / kXChunk;
        const int items = chunks * plan.nfields;
        // first work item of this segment that belongs to this CTA
        int w = (static_cast<int>(blockIdx.x) - base % static_cast<int>(gridDim.x) + static_cast<int>(gridDim.x)) %
                static_cast<int>(gridDim.x);
        for (; w < items; w += gridDim.x) {
            const int kf = w / chunks, ch = w - kf * chunks;
            double *buf = sg.buf + static_cast<long long>(kf) * n;
            double *f = plan.fields[kf];
            const int hi = min(n, (ch + 1) * kXChunk);
            for (int t0 = ch * kXChunk + threadIdx.x; t0 < hi; t0 += 4 * blockDim.x) {
                double v[4];
                size_t q[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int t = t0 + k * blockDim.x;
                    if (t < hi) {
                        const int jj = t / sg.nx, ii = t - jj * sg.nx;
                        q[k] = idx(plan.pitch, sg.i0 + ii, sg.j0 + jj);
                        v[k] = send ? f[q[k]] : __ldcg(buf + t);
                    }
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int t = t0 + k * blockDim.x;
                    if (t < hi) {
                        if (send) buf[t] = v[k];
                        else f[q[k]] = v[k];
                    }
                }
            }
        }
        base += items;
    }
}

__global__ void __launch_bounds__(256) xdirect_fused_kernel(const __grid_constant__ XPlan sp,
                                                            const __grid_constant__ XPlan rp, unsigned *ticket, int *err)
{
    xdirect_items(sp, true);
    // one system fence per CTA, by the thread that takes the ticket: the barrier orders the other threads'
    // peer stores before it and the fence is cumulative (the grid-sync pattern); a fence in every thread
    // made every warp wait for its own acknowledgements first
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) {
        __threadfence_system();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last) {
        if (threadIdx.x < sp.npeers) {
            __threadfence_system();
            asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(sp.flag[threadIdx.x]), "l"(sp.count[threadIdx.x])
                         : "memory");
        }
        if (threadIdx.x == 0) *ticket = 0u;
    }
    if (threadIdx.x < rp.npeers) {
        const long long t0 = clock64();
        for (;;) {
            unsigned long long v;
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(rp.flag[threadIdx.x]) : "memory");
            if (v >= rp.count[threadIdx.x]) break;
            if (wait_expired(t0)) {
                atomicExch(err, 1);
                __threadfence_system();
                __trap();
            }
            __nanosleep(40);
        }
    }
    __syncthreads();
    xdirect_items(rp, false);
}

cudaError_t xdirect_exchange(const XPlan &sp, const XPlan &rp, unsigned *ticket, int *err, cudaStream_t s)
{
    if (sp.nfields <= 0 || (sp.nseg <= 0 && rp.nseg <= 0)) return cudaSuccess;
    long long items = 0;
    for (int q = 0; q < sp.nseg; ++q)
        items += static_cast<long long>((sp.seg[q].nx * sp.seg[q].ny + kXChunk - 1) / kXChunk) * sp.nfields;
    long long ritems = 0;
    for (int q = 0; q < rp.nseg; ++q)
        ritems += static_cast<long long>((rp.seg[q].nx * rp.seg[q].ny + kXChunk - 1) / kXChunk) * rp.nfields;
    items = std::max(items, ritems);
    const unsigned blocks = static_cast<unsigned>(std::max(1LL, std::min<long long>(items, kSMs * 2)));
    xdirect_fused_kernel<<<blocks, 256, 0, s>>>(sp, rp, ticket, err);
    count_launch();
    return cudaGetLastError();
}

static dim3 xdirect_grid(const XPlan &plan)
{
    long long nmax = 1;
    for (int q = 0; q < plan.nseg; ++q)
        nmax = std::max(nmax, static_cast<long long>(plan.seg[q].nx) * plan.seg[q].ny);
    return dim3(nmax / 256 + 1 < 32 ? static_cast<unsigned>(nmax / 256 + 1) : 32u,
                static_cast<unsigned>(plan.nseg * plan.nfields));
}
cudaError_t xdirect_send(const XPlan &plan, unsigned *ticket, cudaStream_t s)
{
    if (plan.nseg <= 0 || plan.nfields <= 0) return cudaSuccess;
    xdirect_send_kernel<<<xdirect_grid(plan), 256, 0, s>>>(plan, ticket);
    count_launch();
    return cudaGetLastError();
}
cudaError_t xdirect_recv(const XPlan &plan, int *err, cudaStream_t s)
{
    if (plan.nseg <= 0 || plan.nfields <= 0) return cudaSuccess;
    xdirect_recv_kernel<<<xdirect_grid(plan), 256, 0, s>>>(plan, err);
    count_launch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// grid metrics: broadcast a 1-D coordinate vector over the 2-D array (xt(ji,:) / yt(:,jj))
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) broadcast_kernel(double *__restrict__ out, int nx, int ny,
                                                        const double *__restrict__ vec, int along_y)
{
    const size_t n = static_cast<size_t>(nx) * ny;
    for (size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < n;
         t += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const size_t j = t / nx, i = t - j * nx;
        out[t] = vec[along_y ? j : i];
    }
}
cudaError_t broadcast_x(double *out, int nx, int ny, const double *xvec, cudaStream_t s)
{
    broadcast_kernel<<<kSMs * 8, 256, 0, s>>>(out, nx, ny, xvec, 0);
    count_launch();
    return cudaGetLastError();
}
cudaError_t broadcast_y(double *out, int nx, int ny, const double *yvec, cudaStream_t s)
{
    broadcast_kernel<<<kSMs * 8, 256, 0, s>>>(out, nx, ny, yvec, 1);
    count_launch();
    return cudaGetLastError();
}

// r2d_field(init_global_data=...) scatter (src/field_mod.f90:378-389) from a device copy of
// the replicated global array
__global__ void __launch_bounds__(256) scatter_kernel(double *__restrict__ f, int pitch, Box b,
                                                      const double *__restrict__ g, int gnx, int dx, int dy)
{
    const int w = b.xe - b.xs + 1, h = b.ye - b.ys + 1;
    const long long n = static_cast<long long>(w) * h;
    for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < n;
         t += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int jj = static_cast<int>(t / w), ii = static_cast<int>(t - static_cast<long long>(jj) * w);
        const int i = b.xs + ii, j = b.ys + jj;
        f[idx(pitch, i, j)] = g[idx(gnx, i + dx, j + dy)];
    }
}
cudaError_t scatter_global(double *f, int pitch, Box b, const double *g, int gnx, int dx, int dy,
                           cudaStream_t s)
{
    if (b.xe < b.xs || b.ye < b.ys) return cudaSuccess;
    scatter_kernel<<<kSMs * 4, 256, 0, s>>>(f, pitch, b, g, gnx, dx, dy);
    count_launch();
    return cudaGetLastError();
}
// gather_inner_data pack (src/field_mod.f90:1361-1368): box -> contiguous, i fastest
cudaError_t gather_box(const double *f, int pitch, Box b, double *packed, cudaStream_t s)
{
    const int w = b.xe - b.xs + 1, h = b.ye - b.ys + 1;
    if (w <= 0 || h <= 0) return cudaSuccess;
    patch_kernel<<<kSMs * 4, 256, 0, s>>>(f, pitch, b.xs, b.ys, packed, w, 1, 1, w, h);
    count_launch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Tile-set exchange: every same-rank halo copy, pack or unpack of one phase in one launch.
// Patches are halo strips (a few KB to a few hundred KB): blockIdx.y picks the copy, blockIdx.x
// strides over its elements; rows of the wide (north / south) strips are read and written coalesced.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) tile_ops_kernel(const __grid_constant__ TileOpBatch b)
{
    const TileOp &o = b.op[blockIdx.y];
    const long long n = static_cast<long long>(o.nx) * o.ny;
    for (long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; t < n;
         t += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int jj = static_cast<int>(t / o.nx), ii = static_cast<int>(t - static_cast<long long>(jj) * o.nx);
        o.dst[static_cast<size_t>(jj) * o.dpitch + ii] = o.src[static_cast<size_t>(jj) * o.spitch + ii];
    }
}

cudaError_t tile_ops(const TileOpBatch &b, cudaStream_t s)
{
    if (b.n <= 0) return cudaSuccess;
    long long nmax = 1;
    for (int q = 0; q < b.n; ++q) nmax = std::max(nmax, static_cast<long long>(b.op[q].nx) * b.op[q].ny);
    const unsigned bx = static_cast<unsigned>(std::min<long long>((nmax + 255) / 256, 32));
    tile_ops_kernel<<<dim3(bx, b.n), 256, 0, s>>>(b);
    count_launch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Neighbour hand-shake of the halo push.  Launched after the fused sweep on the same stream:
// the sweep's peer stores are complete when this kernel starts, so a system-scope release of
// the epoch into each neighbour's flag array publishes them; then wait (acquire) until every
// neighbour has published the same epoch into ours.  One thread per neighbour.  The wait is
// bounded (DLESM_WAIT_TIMEOUT_S, default 300 s; 0 = unbounded): on time-out *err is set and the kernel TRAPS -- the run ends with
// a CUDA error on the next call instead of hanging the GPU or sweeping over stale halos.
// ---------------------------------------------------------------------------------------
__global__ void peer_signal_wait_kernel(PeerFlags pf, unsigned long long epoch, int *err, int publish)
{
    const int q = threadIdx.x;
    if (q >= pf.n) return;
    if (publish) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(pf.remote[q]), "l"(epoch) : "memory");
    }
    const long long t0 = clock64();
    for (;;) {
        unsigned long long v;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(pf.local[q]) : "memory");
        if (v >= epoch) break;
        if (wait_expired(t0)) {
            // DLESM_WAIT_TIMEOUT_S without the neighbour's epoch: the ranks are out of step.  Carrying
            // on would sweep over halos that were never pushed, so the time-out is FATAL: the flag is
            // set for dlesm_dev_sync's message and the kernel traps, which fails every later CUDA
            // call of this process (checksum, get_data, gather ... all return DLESM_ERR_CUDA).
            atomicExch(err, 1);
            __threadfence_system();
            __trap();
        }
        __nanosleep(200);
    }
}
__global__ void __launch_bounds__(256) peer_wait_scatter_kernel(const __grid_constant__ PeerFlags pf, unsigned long long epoch,
                                                                int *err, int publish, const __grid_constant__ StripScatter sp)
{
    const int q = threadIdx.x;
    if (q < pf.n) {
        if (publish && blockIdx.x == 0) {
            __threadfence_system();
            asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(pf.remote[q]), "l"(epoch) : "memory");
        }
        const long long t0 = clock64();
        for (;;) {
            unsigned long long v;
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(pf.local[q]) : "memory");
            if (v >= epoch) break;
            if (wait_expired(t0)) {   // fatal, like peer_signal_wait_kernel
                atomicExch(err, 1);
                __threadfence_system();
                __trap();
            }
            __nanosleep(200);
        }
    }
    __syncthreads();
    // the slots were written by other GPUs (or by this one's sweep): read them past L1
    for (int b = 0; b < sp.n; ++b) {
        const StripScatter::Item &it = sp.item[b];
        const int total = it.nx * it.ny * 6;
        for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
            const int c = e % it.nx, f = (e / it.nx) % 6, r = e / (6 * it.nx);
            sp.dst[f][idx(sp.pitch, it.ides + c, it.jdes + r)] = __ldcg(it.slot + e);
        }
    }
}
cudaError_t peer_wait_scatter(const PeerFlags &pf, unsigned long long epoch, int *err, bool publish,
                              const StripScatter &sp, cudaStream_t s)
{
    if (sp.n <= 0) return pf.n > 0 ? peer_signal_wait(pf, epoch, err, s, publish) : cudaSuccess;
    long long total = 0;
    for (int b = 0; b < sp.n; ++b) total += static_cast<long long>(sp.item[b].nx) * sp.item[b].ny * 6;
    const unsigned blocks = static_cast<unsigned>(std::max(1LL, std::min<long long>((total + 2047) / 2048, kSMs)));
    peer_wait_scatter_kernel<<<blocks, 256, 0, s>>>(pf, epoch, err, publish ? 1 : 0, sp);
    count_launch();
    return cudaGetLastError();
}

cudaError_t peer_signal_wait(const PeerFlags &pf, unsigned long long epoch, int *err, cudaStream_t s, bool publish)
{
    if (pf.n <= 0) return cudaSuccess;
    peer_signal_wait_kernel<<<1, 32, 0, s>>>(pf, epoch, err, publish ? 1 : 0);
    count_launch();
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Un-fused shallow-water sweeps.  One launch per reference-style kernel; the loop nest of
// SURVEY CS-3 (jj outer, ji inner over the output field's internal region) becomes a 2-D
// grid with i along threadIdx.x; the per-point functor is the "kernel code".
// Formulas: SURVEY.md a-S, identical expression trees to oracle/dlesm_oracle_sw.c.
// ---------------------------------------------------------------------------------------
constexpr int kSweepRows = 4;

template <class Op>
__global__ void __launch_bounds__(256) sweep_kernel(Op op, Box b)
{
    const int i = b.xs + blockIdx.x * 256 + threadIdx.x;
    if (i > b.xe) return;
    const int j0 = b.ys + blockIdx.y * kSweepRows;
#pragma unroll
    for (int r = 0; r < kSweepRows; ++r) {
        const int j = j0 + r;
        if (j <= b.ye) op(i, j);
    }
}

template <class Op>
static cudaError_t sweep(const Op &op, Box b, cudaStream_t s)
{
    if (b.xe < b.xs || b.ye < b.ys) return cudaSuccess;
    dim3 grid(cdiv(b.xe - b.xs + 1, 256), cdiv(b.ye - b.ys + 1, kSweepRows));
    sweep_kernel<Op><<<grid, 256, 0, s>>>(op, b);
    count_launch();
    return cudaGetLastError();
}

struct Mask {
    const int *t;
    int pitch;
    __device__ bool at(int i, int j) const { return t[idx(pitch, i, j)] == 1; }
    __device__ bool wet_t(int i, int j) const { return !t || at(i, j); }
    __device__ bool wet_u(int i, int j) const { return !t || (at(i, j) && at(i - 1, j)); }
    __device__ bool wet_v(int i, int j) const { return !t || (at(i, j) && at(i, j - 1)); }
    __device__ bool wet_f(int i, int j) const
    {
        return !t || (at(i, j) && at(i - 1, j) && at(i, j - 1) && at(i - 1, j - 1));
    }
};

struct CuOp {
    double *cu; const double *p, *u; int n; Mask m;
    __device__ void operator()(int i, int j) const
    {
        const size_t q = idx(n, i, j);
        cu[q] = m.wet_u(i, j) ? 0.5 * (p[q] + p[q - 1]) * u[q] : 0.0;
    }
};
struct CvOp {
    double *cv; const double *p, *v; int n; Mask m;
    __device__ void operator()(int i, int j) const
    {
        const size_t q = idx(n, i, j);
        cv[q] = m.wet_v(i, j) ? 0.5 * (p[q] + p[q - n]) * v[q] : 0.0;
    }
};
struct ZOp {
    double *z; const double *p, *u, *v; int n; double fsdx, fsdy; Mask m;
    __device__ void operator()(int i, int j) const
    {
        const size_t q = idx(n, i, j);
        z[q] = m.wet_f(i, j) ? (fsdx * (v[q] - v[q - 1]) - fsdy * (u[q] - u[q - n])) /
                                   (p[q - n - 1] + p[q - n] + p[q] + p[q - 1])
                             : 0.0;
    }
};
struct HOp {
    double *h; const double *p, *u, *v; int n; Mask m;
    __device__ void operator()(int i, int j) const
    {
        const size_t q = idx(n, i, j);
        h[q] = m.wet_t(i, j) ? p[q] + 0.25 * (u[q + 1] * u[q + 1] + u[q] * u[q] + v[q + n] * v[q + n] + v[q] * v[q])
                             : 0.0;
    }
};
struct UnewOp {
    double *unew; const double *uold, *z, *cv, *h, *ucur; int n; double tdts8, tdtsdx; Mask m;
    __device__ void operator()(int i, int j) const
    {
        const size_t q = idx(n, i, j);
        unew[q] = m.wet_u(i, j)
                      ? uold[q] + tdts8 * (z[q + n] + z[q]) * (cv[q + n] + cv[q + n - 1] + cv[q - 1] + cv[q]) -
                            tdtsdx * (h[q] - h[q - 1])
                      : ucur[q];
    }
};
struct VnewOp {
    double *vnew; const double *vold, *z, *cu, *h, *vcur; int n; double tdts8, tdtsdy; Mask m;
    __device__ void operator()(int i, int j) const
    {
        const size_t q = idx(n, i, j);
        vnew[q] = m.wet_v(i, j)
                      ? vold[q] - tdts8 * (z[q + 1] + z[q]) * (cu[q + 1] + cu[q] + cu[q - n] + cu[q - n + 1]) -
                            tdtsdy * (h[q] - h[q - n])
                      : vcur[q];
    }
};
struct PnewOp {
    double *pnew; const double *pold, *cu, *cv, *pcur; int n; double tdtsdx, tdtsdy; Mask m;
    __device__ void operator()(int i, int j) const
    {
        const size_t q = idx(n, i, j);
        pnew[q] = m.wet_t(i, j) ? pold[q] - tdtsdx * (cu[q + 1] - cu[q]) - tdtsdy * (cv[q + n] - cv[q])
                                : pcur[q];
    }
};
struct SmoothOp {
    const double *f, *fnew; double *fold; int n; double alpha; Mask m; int points;
    __device__ void operator()(int i, int j) const
    {
        const bool wet = points == 0 ? m.wet_u(i, j) : points == 1 ? m.wet_v(i, j) : m.wet_t(i, j);
        if (!wet) return;
        const size_t q = idx(n, i, j);
        fold[q] = f[q] + alpha * (fnew[q] - 2.0 * f[q] + fold[q]);
    }
};

cudaError_t compute_cu(double *cu, const double *p, const double *u, int n, Box b, const int *t, cudaStream_t s)
{
    return sweep(CuOp{cu, p, u, n, Mask{t, n}}, b, s);
}
cudaError_t compute_cv(double *cv, const double *p, const double *v, int n, Box b, const int *t, cudaStream_t s)
{
    return sweep(CvOp{cv, p, v, n, Mask{t, n}}, b, s);
}
cudaError_t compute_z(double *z, const double *p, const double *u, const double *v, int n, Box b,
                      double fsdx, double fsdy, const int *t, cudaStream_t s)
{
    return sweep(ZOp{z, p, u, v, n, fsdx, fsdy, Mask{t, n}}, b, s);
}
cudaError_t compute_h(double *h, const double *p, const double *u, const double *v, int n, Box b,
                      const int *t, cudaStream_t s)
{
    return sweep(HOp{h, p, u, v, n, Mask{t, n}}, b, s);
}
cudaError_t compute_unew(double *unew, const double *uold, const double *z, const double *cv,
                         const double *h, int n, Box b, double tdts8, double tdtsdx, const int *t,
                         const double *ucur, cudaStream_t s)
{
    return sweep(UnewOp{unew, uold, z, cv, h, ucur, n, tdts8, tdtsdx, Mask{t, n}}, b, s);
}
cudaError_t compute_vnew(double *vnew, const double *vold, const double *z, const double *cu,
                         const double *h, int n, Box b, double tdts8, double tdtsdy, const int *t,
                         const double *vcur, cudaStream_t s)
{
    return sweep(VnewOp{vnew, vold, z, cu, h, vcur, n, tdts8, tdtsdy, Mask{t, n}}, b, s);
}
cudaError_t compute_pnew(double *pnew, const double *pold, const double *cu, const double *cv, int n,
                         Box b, double tdtsdx, double tdtsdy, const int *t, const double *pcur,
                         cudaStream_t s)
{
    return sweep(PnewOp{pnew, pold, cu, cv, pcur, n, tdtsdx, tdtsdy, Mask{t, n}}, b, s);
}
cudaError_t time_smooth(const double *f, const double *fnew, double *fold, int n, Box b, double alpha,
                        const int *t, int points, cudaStream_t s)
{
    return sweep(SmoothOp{f, fnew, fold, n, alpha, Mask{t, n}, points}, b, s);
}

// ---------------------------------------------------------------------------------------
// synthetic initial conditions (classic "shallow"), device sin/cos
// ---------------------------------------------------------------------------------------
struct InitOp {
    double *u, *v, *p; int n; Box b; int gx0, gy0, M, N; double dx, dy;
    __device__ void operator()(int i, int j) const
    {
        const double a = 1.0e6, pi = 3.14159265358979323846;
        const double di = 2.0 * pi / static_cast<double>(M), dj = 2.0 * pi / static_cast<double>(N);
        const double el = static_cast<double>(N) * dx;
        const double pcf = pi * pi * a * a / (el * el);
        const int gi = gx0 + (i - b.xs), gj = gy0 + (j - b.ys);
        const double s_i = sin((gi - 0.5) * di), s_i1 = sin((gi + 0.5) * di);
        const double s_j = sin((gj - 0.5) * dj), s_j1 = sin((gj + 0.5) * dj);
        const double psi_ij = a * s_i * s_j, psi_ij1 = a * s_i * s_j1, psi_i1j = a * s_i1 * s_j;
        const size_t q = idx(n, i, j);
        p[q] = pcf * (cos(2.0 * (gi - 1.0) * di) + cos(2.0 * (gj - 1.0) * dj)) + 50000.0;
        u[q] = -(psi_ij1 - psi_ij) / dy;
        v[q] = (psi_i1j - psi_ij) / dx;
    }
};
cudaError_t sw_init(double *u, double *v, double *p, int n, Box b, int gx0, int gy0, int M, int N,
                    double dx, double dy, cudaStream_t s)
{
    return sweep(InitOp{u, v, p, n, b, gx0, gy0, M, N, dx, dy}, b, s);
}

}  // namespace k
}  // namespace dlesm
