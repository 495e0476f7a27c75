// dlesm_fused.cu -- the whole shallow-water step as ONE sweep (96 B / point).
//
// Every thread owns one column i and marches along j, carrying in registers everything the
// next row reuses: the previous raw row of p,u,v (3 columns wide) and the intermediates
// cu, cv, z, h of the previous one/two rows.  Each arriving raw row N (= row j) yields the
// intermediates of row j and the outputs of row j-1:
//
//   unew,vnew,pnew(i,j-1) and, unless it is the first step, the Robert-Asselin filtered
//   uold,vold,pold(i,j-1) in place.
//
// Intermediates on the halo ring are recomputed from the (periodic / exchanged) halo values
// of p,u,v instead of being wrapped after the fact; with identical operands and -fmad=false
// this gives the same bits as the reference-style sequence "sweep, halo-copy, sweep".
//
// Variants (dlesm_sw_step_fused's `variant`):
//   1  direct : raw rows come straight from global memory through L1 (any pitch, tmask ok)
//   2+ tma    : a producer warp streams row chunks of all six inputs into a shared-memory
//               ring with cp.async.bulk.tensor (TMA) + mbarrier full/empty pairs; consumer
//               warps march out of shared memory.  Needs a 16-byte-multiple pitch
//               (DL_ESM_ALIGNMENT even), no tmask.
//   0  auto   : tma when eligible, else direct.
#include <cuda.h>

#include <cstdint>
#include <map>
#include <mutex>
#include <tuple>

#include "dlesm_kernels.cuh"

namespace dlesm {
namespace k {

namespace {
__host__ __device__ inline size_t idx(int pitch, int i, int j)
{
    return (static_cast<size_t>(j) - 1) * static_cast<size_t>(pitch) + (static_cast<size_t>(i) - 1);
}

template <bool M>
struct Raw {
    double pW, pC, pE, uW, uC, uE, vW, vC, vE;
    bool tW, tC, tE;  // only meaningful when M
};

template <bool M>
struct Carry {
    Raw<M> P;                                  // raw row j-1
    double cuC_P, cuE_P, cuC_PP, cuE_PP;       // cu(i,j-1), cu(i+1,j-1), cu(i,j-2), cu(i+1,j-2)
    double cvC_P, cvW_P;                       // cv(i,j-1), cv(i-1,j-1)
    double zC_P, zE_P;                         // z(i,j-1), z(i+1,j-1)
    double hC_PP;                              // h(i,j-2)
    bool tC_PP;
};

struct Outs {
    double unew, vnew, pnew, uold, vold, pold;
    bool wu, wv, wt;  // which filtered values are to be stored (masked runs)
};

// One marching step: raw row N = row j arrives; produces outputs of row j-1.
// Expression trees are those of oracle/dlesm_oracle_sw.c (SURVEY.md a-S), term by term.
template <bool M>
__device__ __forceinline__ void ingest(Carry<M> &c, const Raw<M> &N, const SwScalars &sc, double uo,
                                       double vo, double po, Outs &o)
{
    const Raw<M> &P = c.P;
    double cuC_N = 0.5 * (N.pC + N.pW) * N.uC;
    double cuE_N = 0.5 * (N.pE + N.pC) * N.uE;
    double cvC_N = 0.5 * (N.pC + P.pC) * N.vC;
    double cvW_N = 0.5 * (N.pW + P.pW) * N.vW;
    double zC_N = (sc.fsdx * (N.vC - N.vW) - sc.fsdy * (N.uC - P.uC)) / (P.pW + P.pC + N.pC + N.pW);
    double zE_N = (sc.fsdx * (N.vE - N.vC) - sc.fsdy * (N.uE - P.uE)) / (P.pC + P.pE + N.pE + N.pC);
    double hC_P = P.pC + 0.25 * (P.uE * P.uE + P.uC * P.uC + N.vC * N.vC + P.vC * P.vC);
    double hW_P = P.pW + 0.25 * (P.uC * P.uC + P.uW * P.uW + N.vW * N.vW + P.vW * P.vW);
    if (M) {
        if (!(N.tC && N.tW)) cuC_N = 0.0;
        if (!(N.tE && N.tC)) cuE_N = 0.0;
        if (!(N.tC && P.tC)) cvC_N = 0.0;
        if (!(N.tW && P.tW)) cvW_N = 0.0;
        if (!(N.tC && N.tW && P.tC && P.tW)) zC_N = 0.0;
        if (!(N.tE && N.tC && P.tE && P.tC)) zE_N = 0.0;
        if (!P.tC) hC_P = 0.0;
        if (!P.tW) hW_P = 0.0;
    }
    o.unew = uo + sc.tdts8 * (zC_N + c.zC_P) * (cvC_N + cvW_N + c.cvW_P + c.cvC_P) -
             sc.tdtsdx * (hC_P - hW_P);
    o.vnew = vo - sc.tdts8 * (c.zE_P + c.zC_P) * (c.cuE_P + c.cuC_P + c.cuC_PP + c.cuE_PP) -
             sc.tdtsdy * (hC_P - c.hC_PP);
    o.pnew = po - sc.tdtsdx * (c.cuE_P - c.cuC_P) - sc.tdtsdy * (cvC_N - c.cvC_P);
    o.wu = o.wv = o.wt = true;
    if (M) {
        o.wu = P.tC && P.tW;
        o.wv = P.tC && c.tC_PP;
        o.wt = P.tC;
        if (!o.wu) o.unew = P.uC;
        if (!o.wv) o.vnew = P.vC;
        if (!o.wt) o.pnew = P.pC;
    }
    o.uold = P.uC + sc.alpha * (o.unew - 2.0 * P.uC + uo);
    o.vold = P.vC + sc.alpha * (o.vnew - 2.0 * P.vC + vo);
    o.pold = P.pC + sc.alpha * (o.pnew - 2.0 * P.pC + po);
    // rotate
    c.cuC_PP = c.cuC_P; c.cuE_PP = c.cuE_P;
    c.cuC_P = cuC_N; c.cuE_P = cuE_N;
    c.cvC_P = cvC_N; c.cvW_P = cvW_N;
    c.zC_P = zC_N; c.zE_P = zE_N;
    c.hC_PP = hC_P;
    c.tC_PP = P.tC;
    c.P = N;
}

template <bool M>
__device__ __forceinline__ void carry_init(Carry<M> &c)
{
    // values before the first real row only feed outputs that are never stored; keep the
    // denominators non-zero so no inf/nan is manufactured on the way
    c.P.pW = c.P.pC = c.P.pE = 1.0;
    c.P.uW = c.P.uC = c.P.uE = c.P.vW = c.P.vC = c.P.vE = 0.0;
    c.P.tW = c.P.tC = c.P.tE = true;
    c.cuC_P = c.cuE_P = c.cuC_PP = c.cuE_PP = c.cvC_P = c.cvW_P = c.zC_P = c.zE_P = c.hC_PP = 0.0;
    c.tC_PP = true;
}

// ---------------------------------------------------------------------------------------
// variant 1: direct
// ---------------------------------------------------------------------------------------
template <bool M, int ROWS>
__global__ void __launch_bounds__(128) fused_direct_kernel(FusedArgs a)
{
    const int i = a.b.xs + blockIdx.x * 128 + threadIdx.x;
    if (i > a.b.xe) return;
    const int j0 = a.b.ys + blockIdx.y * ROWS;
    const int j1 = min(j0 + ROWS - 1, a.b.ye);
    const int n = a.pitch;
    const SwScalars sc = a.sc;

    Carry<M> c;
    carry_init(c);
    size_t q = idx(n, i, j0 - 1);  // raw row pointer
    // rows j0-1 .. j1+1 arrive; outputs appear from the third arrival on
    for (int j = j0 - 1; j <= j1 + 1; ++j, q += n) {
        Raw<M> N;
        N.pW = a.p[q - 1]; N.pC = a.p[q]; N.pE = a.p[q + 1];
        N.uW = a.u[q - 1]; N.uC = a.u[q]; N.uE = a.u[q + 1];
        N.vW = a.v[q - 1]; N.vC = a.v[q]; N.vE = a.v[q + 1];
        if (M) {
            N.tW = a.tmask[q - 1] == 1; N.tC = a.tmask[q] == 1; N.tE = a.tmask[q + 1] == 1;
        }
        const bool emit = j >= j0 + 1;  // outputs of row j-1 >= j0
        const size_t qo = q - n;
        double uo = 0.0, vo = 0.0, po = 0.0;
        if (emit) {
            uo = a.uold[qo]; vo = a.vold[qo]; po = a.pold[qo];
        }
        Outs o;
        ingest<M>(c, N, sc, uo, vo, po, o);
        if (emit) {
            a.unew[qo] = o.unew; a.vnew[qo] = o.vnew; a.pnew[qo] = o.pnew;
            if (!sc.first) {
                if (o.wu) a.uold[qo] = o.uold;
                if (o.wv) a.vold[qo] = o.vold;
                if (o.wt) a.pold[qo] = o.pold;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// variants 2+: TMA ring
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int x, int y, uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(x), "r"(y), "r"(smem_u32(bar))
        : "memory");
}

struct TmaStepParams {
    double *unew, *vnew, *pnew, *uold, *vold, *pold;
    int pitch;
    Box b;
    int seg_rows;
    int i_origin;   // first output column (1-based) of strip 0: b.xs, or lower to align strips to 128 B
    SwScalars sc;
};

constexpr int kDefaultTmaVariant = 25;
constexpr int align128(int v) { return (v + 127) & ~127; }

template <int TX, int RY, int STAGES>
struct TmaCfg {
    // TMA needs the first element of a box 16-byte aligned (even FP64 column; measured with
    // tools/tma_probe.cu: an odd start column raises "illegal instruction").  Every tile
    // therefore starts at the even column at or below the strip's west halo column and is
    // TX + 4 wide, which covers west halo + TX + east halo for either parity.
    static constexpr int BW = TX + 4;
    static constexpr int A = align128(BW * RY * 8);
    static constexpr int STAGE_BYTES = 6 * A;
    static constexpr int TX_BYTES = 6 * BW * RY * 8;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 2 * STAGES * 8 + 128;
    static constexpr int THREADS = TX + 32;
};

template <int TX, int RY, int STAGES, int MINB>
__global__ void __launch_bounds__(TX + 32, MINB)
fused_tma_kernel(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_v,
                 const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_uo,
                 const __grid_constant__ CUtensorMap tm_vo, const __grid_constant__ CUtensorMap tm_po,
                 const TmaStepParams prm)
{
    using Cfg = TmaCfg<TX, RY, STAGES>;
    constexpr int BW = Cfg::BW;
    extern __shared__ unsigned char smem_raw[];
    // 128-byte alignment for the TMA destinations
    unsigned char *smem = reinterpret_cast<unsigned char *>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~static_cast<uintptr_t>(127));
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + STAGES * Cfg::STAGE_BYTES);
    uint64_t *empty = full + STAGES;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int j0 = prm.b.ys + blockIdx.y * prm.seg_rows;
    const int j1 = min(j0 + prm.seg_rows - 1, prm.b.ye);
    const int nsteps = (j1 - j0 + 1) + 2;            // raw rows j0-1 .. j1+1
    const int nchunks = (nsteps + RY - 1) / RY;
    const int i_first = prm.i_origin + blockIdx.x * TX;   // first output column of this strip

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], TX / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (warp == TX / 32) {
        // ---- producer warp: one lane feeds the ring ----
        if (lane == 0) {
            const int x_box = (i_first - 2) & ~1;   // even 0-based column at/below the west halo column
            for (int c = 0; c < nchunks; ++c) {
                const int s = c % STAGES;
                if (c >= STAGES) mbar_wait(&empty[s], ((c / STAGES) - 1) & 1);
                unsigned char *st = smem + s * Cfg::STAGE_BYTES;
                const int y_raw = (j0 - 1) + c * RY - 1;   // 0-based row of the chunk's first raw row
                const int y_old = y_raw - 1;               // outputs trail the raw rows by one
                mbar_expect_tx(&full[s], Cfg::TX_BYTES);
                tma_load_2d(st + 0 * Cfg::A, &tm_p, x_box, y_raw, &full[s]);
                tma_load_2d(st + 1 * Cfg::A, &tm_u, x_box, y_raw, &full[s]);
                tma_load_2d(st + 2 * Cfg::A, &tm_v, x_box, y_raw, &full[s]);
                tma_load_2d(st + 3 * Cfg::A, &tm_uo, x_box, y_old, &full[s]);
                tma_load_2d(st + 4 * Cfg::A, &tm_vo, x_box, y_old, &full[s]);
                tma_load_2d(st + 5 * Cfg::A, &tm_po, x_box, y_old, &full[s]);
            }
        }
        return;
    }

    // ---- consumer warps ----
    const int i = i_first + tid;
    const bool active = i <= prm.b.xe && i >= prm.b.xs;
    const int xoff = (i_first - 2) & 1;
    const SwScalars sc = prm.sc;
    Carry<false> cr;
    carry_init(cr);
    size_t qo = idx(prm.pitch, i, j0 - 2);  // output row trails: step t writes row j0-2+t
    int t = 0;
    for (int c = 0; c < nchunks; ++c) {
        const int s = c % STAGES;
        mbar_wait(&full[s], (c / STAGES) & 1);
        const unsigned char *st = smem + s * Cfg::STAGE_BYTES;
        const double *sp = reinterpret_cast<const double *>(st + 0 * Cfg::A);
        const double *su = reinterpret_cast<const double *>(st + 1 * Cfg::A);
        const double *sv = reinterpret_cast<const double *>(st + 2 * Cfg::A);
        const double *suo = reinterpret_cast<const double *>(st + 3 * Cfg::A);
        const double *svo = reinterpret_cast<const double *>(st + 4 * Cfg::A);
        const double *spo = reinterpret_cast<const double *>(st + 5 * Cfg::A);
#pragma unroll
        for (int r = 0; r < RY; ++r) {
            if (t < nsteps) {
                Raw<false> N;
                const int o = r * BW + tid + xoff;   // west neighbour's column inside the tile
                N.pW = sp[o]; N.pC = sp[o + 1]; N.pE = sp[o + 2];
                N.uW = su[o]; N.uC = su[o + 1]; N.uE = su[o + 2];
                N.vW = sv[o]; N.vC = sv[o + 1]; N.vE = sv[o + 2];
                const double uo = suo[o + 1], vo = svo[o + 1], po = spo[o + 1];
                Outs out;
                ingest<false>(cr, N, sc, uo, vo, po, out);
                if (t >= 2 && active) {
                    __stcs(prm.unew + qo, out.unew);
                    __stcs(prm.vnew + qo, out.vnew);
                    __stcs(prm.pnew + qo, out.pnew);
                    if (!sc.first) {
                        __stcs(prm.uold + qo, out.uold);
                        __stcs(prm.vold + qo, out.vold);
                        __stcs(prm.pold + qo, out.pold);
                    }
                }
            }
            ++t;
            qo += prm.pitch;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
}

// ---- tensor-map cache ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn()
{
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

std::mutex g_map_mutex;
std::map<std::tuple<const void *, int, int, int, int>, CUtensorMap> g_maps;

bool get_map(const double *base, int nx, int ny, int boxw, int boxh, CUtensorMap &out)
{
    std::lock_guard<std::mutex> lock(g_map_mutex);
    auto key = std::make_tuple(static_cast<const void *>(base), nx, ny, boxw, boxh);
    auto it = g_maps.find(key);
    if (it != g_maps.end()) {
        out = it->second;
        return true;
    }
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    cuuint64_t gdim[2] = {static_cast<cuuint64_t>(nx), static_cast<cuuint64_t>(ny)};
    cuuint64_t gstride[1] = {static_cast<cuuint64_t>(nx) * sizeof(double)};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(boxw), static_cast<cuuint32_t>(boxh)};
    cuuint32_t estr[2] = {1, 1};
    CUtensorMap m;
    CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double *>(base), gdim, gstride, box,
                    estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    if (g_maps.size() > 256) g_maps.clear();
    g_maps[key] = m;
    out = m;
    return true;
}

// pick the segment height so that the grid is a whole number of waves of `slots` resident CTAs
int pick_seg_rows(int w, int h, int tx, int slots)
{
    const int strips = (w + tx - 1) / tx;
    // aim for segments of ~256 rows (prologue overhead 2/256), rounded to fill whole waves
    int nseg = (h + 255) / 256;
    if (nseg < 1) nseg = 1;
    long long ctas = static_cast<long long>(strips) * nseg;
    const long long waves = (ctas + slots - 1) / slots;
    nseg = static_cast<int>((waves * slots) / strips);
    if (nseg < 1) nseg = 1;
    if (nseg > h) nseg = h;
    return (h + nseg - 1) / nseg;
}

template <int TX, int RY, int STAGES, int MINB>
cudaError_t launch_tma(const FusedArgs &a, bool aligned_origin, cudaStream_t s)
{
    using Cfg = TmaCfg<TX, RY, STAGES>;
    CUtensorMap mu, mv, mp, muo, mvo, mpo;
    if (!get_map(a.u, a.pitch, a.ny, Cfg::BW, RY, mu) || !get_map(a.v, a.pitch, a.ny, Cfg::BW, RY, mv) ||
        !get_map(a.p, a.pitch, a.ny, Cfg::BW, RY, mp) || !get_map(a.uold, a.pitch, a.ny, Cfg::BW, RY, muo) ||
        !get_map(a.vold, a.pitch, a.ny, Cfg::BW, RY, mvo) || !get_map(a.pold, a.pitch, a.ny, Cfg::BW, RY, mpo))
        return cudaErrorNotSupported;
    TmaStepParams prm;
    prm.unew = a.unew; prm.vnew = a.vnew; prm.pnew = a.pnew;
    prm.uold = a.uold; prm.vold = a.vold; prm.pold = a.pold;
    // aligned origin: strips start on 128-byte lines of the row (1-based column 1 + 16k), the
    // columns below xs are masked; otherwise strips start at xs
    const int i_origin = aligned_origin ? ((a.b.xs - 1) / 16) * 16 + 1 : a.b.xs;
    const int w = a.b.xe - i_origin + 1, h = a.b.ye - a.b.ys + 1;
    const int seg_rows = pick_seg_rows(w, h, TX, 148 * MINB);
    prm.pitch = a.pitch; prm.b = a.b; prm.seg_rows = seg_rows; prm.i_origin = i_origin; prm.sc = a.sc;
    auto kern = fused_tma_kernel<TX, RY, STAGES, MINB>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    dim3 grid((w + TX - 1) / TX, (h + seg_rows - 1) / seg_rows);
    kern<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, s>>>(mu, mv, mp, muo, mvo, mpo, prm);
    return cudaGetLastError();
}

}  // namespace

namespace {
struct TmaVariant {
    int id;
    const char *name;
    cudaError_t (*launch)(const FusedArgs &, bool, cudaStream_t);
};
// {rows per chunk, stages, resident CTAs per SM}; ids + 20 use the 128-byte aligned strip origin
const TmaVariant kTmaVariants[] = {
    {2, "tma-128x4x2-occ3", launch_tma<128, 4, 2, 3>},  {3, "tma-128x8x2-occ2", launch_tma<128, 8, 2, 2>},
    {4, "tma-128x4x2-occ4", launch_tma<128, 4, 2, 4>},  {5, "tma-128x4x3-occ2", launch_tma<128, 4, 3, 2>},
    {6, "tma-128x4x4-occ2", launch_tma<128, 4, 4, 2>},  {7, "tma-128x2x6-occ2", launch_tma<128, 2, 6, 2>},
    {8, "tma-128x8x3-occ1", launch_tma<128, 8, 3, 1>},  {9, "tma-128x4x6-occ1", launch_tma<128, 4, 6, 1>},
    {10, "tma-128x2x4-occ3", launch_tma<128, 2, 4, 3>}, {11, "tma-128x16x2-occ1", launch_tma<128, 16, 2, 1>},
    {12, "tma-128x2x8-occ2", launch_tma<128, 2, 8, 2>}, {13, "tma-128x1x12-occ2", launch_tma<128, 1, 12, 2>},
};
const TmaVariant *find_tma(int id)
{
    for (const auto &v : kTmaVariants)
        if (v.id == id) return &v;
    return nullptr;
}
}  // namespace

const char *fused_variant_name(int variant)
{
    if (variant == 0) return "auto";
    if (variant == 1) return "direct-128x64";
    const TmaVariant *v = find_tma(variant >= 20 ? variant - 20 : variant);
    return v ? v->name : "unknown";
}

cudaError_t sw_step_fused(const FusedArgs &a, int variant, cudaStream_t s)
{
    if (a.b.xe < a.b.xs || a.b.ye < a.b.ys) return cudaSuccess;
    const bool tma_ok = a.tmask == nullptr && (a.pitch % 2 == 0) && encode_fn() != nullptr;
    if (variant == 0) variant = tma_ok ? kDefaultTmaVariant : 1;
    if (variant >= 2 && !tma_ok) return cudaErrorNotSupported;
    cudaError_t e;
    if (variant == 1) {
        constexpr int ROWS = 64;
        const int w = a.b.xe - a.b.xs + 1, h = a.b.ye - a.b.ys + 1;
        dim3 grid((w + 127) / 128, (h + ROWS - 1) / ROWS);
        if (a.tmask) fused_direct_kernel<true, ROWS><<<grid, 128, 0, s>>>(a);
        else fused_direct_kernel<false, ROWS><<<grid, 128, 0, s>>>(a);
        e = cudaGetLastError();
    } else {
        const bool aligned = variant >= 20;
        const TmaVariant *v = find_tma(aligned ? variant - 20 : variant);
        if (!v) return cudaErrorInvalidValue;
        e = v->launch(a, aligned, s);
    }
    note_launch();
    return e;
}

}  // namespace k
}  // namespace dlesm
