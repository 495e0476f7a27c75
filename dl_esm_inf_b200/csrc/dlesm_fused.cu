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
#include <cstdlib>
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
// APPLY = false on a masked run skips the mask selects (the caller has established that every
// mask value this step looks at is wet); the masks are still carried along.
template <bool M, bool APPLY = M>
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
    if (APPLY) {
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
    if (APPLY) {
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
    PushPlan push;
};

// edge outputs of row j, column i -> the neighbours' halo cells (rare: only threads whose
// column/segment meets a push box get here)
__device__ __noinline__ void push_row(const PushPlan &pp, unsigned mask, int i, int j, double un, double vn,
                                      double pn)
{
    for (unsigned m = mask; m; m &= m - 1) {
        const PushSeg &g = pp.seg[__ffs(m) - 1];
        if (j >= g.j0 && j <= g.j1) {
            const size_t d = idx(g.dpitch, g.ides + (i - g.i0), g.jdes + (j - g.j0));
            g.dst[0][d] = un;
            g.dst[1][d] = vn;
            g.dst[2][d] = pn;
        }
    }
}

constexpr int kDefaultTmaVariant = 25;
constexpr int align128(int v) { return (v + 127) & ~127; }

template <int TX, int RY, int STAGES, bool M, bool ALN>
struct TmaCfg {
    // TMA needs the first element of a box 16-byte aligned (even FP64 column; measured with
    // tools/tma_probe.cu: an odd start column raises "illegal instruction").  Every tile
    // therefore starts at the even column at or below the strip's west halo column and is
    // TX + 4 wide, which covers west halo + TX + east halo for either parity.  The int32
    // tmask tile starts at a multiple of four columns and is TX + 8 wide for the same reason.
    // uold/vold/pold need no halo columns: with the 128-byte aligned strip origin (ALN) their
    // tiles are exactly the TX output columns (whole 32-byte sectors, nothing shared with the
    // neighbouring strip); otherwise they start at the even column at or below the first one.
    static constexpr int BW = TX + 4;
    static constexpr int BWO = ALN ? TX : TX + 2;
    static constexpr int BWM = TX + 8;
    static constexpr int A = align128(BW * RY * 8);
    static constexpr int AO = align128(BWO * RY * 8);
    static constexpr int AM = M ? align128(BWM * RY * 4) : 0;
    static constexpr int OFF_OLD = 3 * A;
    static constexpr int OFF_MASK = 3 * A + 3 * AO;
    static constexpr int STAGE_BYTES = 3 * A + 3 * AO + AM;
    static constexpr int TX_BYTES = 3 * BW * RY * 8 + 3 * BWO * RY * 8 + (M ? BWM * RY * 4 : 0);
    static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
    static constexpr int SMEM_BYTES = BAR_OFFSET + 2 * STAGES * 8;
    static constexpr int THREADS = TX + 32;
};

// Dynamic shared memory, addressed straight off this symbol so that the compiler keeps the
// shared state space (LDS, not generic LD).  No static shared memory exists in this kernel,
// so the window starts at offset 0 of the CTA's allocation, which satisfies the 128-byte
// alignment of TMA destinations; the kernel traps if that ever stops being true.
template <int TX, int RY, int STAGES, int MINB, bool M, bool FIRST, bool ALN>
__global__ void __launch_bounds__(TX + 32, MINB)
fused_tma_kernel(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_v,
                 const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_uo,
                 const __grid_constant__ CUtensorMap tm_vo, const __grid_constant__ CUtensorMap tm_po,
                 const __grid_constant__ CUtensorMap tm_mask, const __grid_constant__ TmaStepParams prm)
{
    using Cfg = TmaCfg<TX, RY, STAGES, M, ALN>;
    constexpr int BW = Cfg::BW;
    constexpr int BWO = Cfg::BWO;
    constexpr int BWM = Cfg::BWM;
    extern __shared__ __align__(1024) unsigned char fused_smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(fused_smem + Cfg::BAR_OFFSET);
    uint64_t *empty = full + STAGES;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int j0 = prm.b.ys + blockIdx.y * prm.seg_rows;
    const int j1 = min(j0 + prm.seg_rows - 1, prm.b.ye);
    const int nsteps = (j1 - j0 + 1) + 2;            // raw rows j0-1 .. j1+1
    const int nchunks = (nsteps + RY - 1) / RY;
    const int i_first = prm.i_origin + blockIdx.x * TX;   // first output column of this strip

    if (tid == 0) {
        if (smem_u32(fused_smem) & 127u) __trap();
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
            const int x_box_m = (i_first - 2) & ~3;
            const int x_old = (i_first - 1) & ~1;   // == i_first - 1 when ALN
            for (int c = 0; c < nchunks; ++c) {
                const int s = c % STAGES;
                if (c >= STAGES) mbar_wait(&empty[s], ((c / STAGES) - 1) & 1);
                unsigned char *st = fused_smem + s * Cfg::STAGE_BYTES;
                const int y_raw = (j0 - 1) + c * RY - 1;   // 0-based row of the chunk's first raw row
                const int y_old = y_raw - 1;               // outputs trail the raw rows by one
                mbar_expect_tx(&full[s], Cfg::TX_BYTES);
                tma_load_2d(st + 0 * Cfg::A, &tm_p, x_box, y_raw, &full[s]);
                tma_load_2d(st + 1 * Cfg::A, &tm_u, x_box, y_raw, &full[s]);
                tma_load_2d(st + 2 * Cfg::A, &tm_v, x_box, y_raw, &full[s]);
                tma_load_2d(st + Cfg::OFF_OLD + 0 * Cfg::AO, &tm_uo, x_old, y_old, &full[s]);
                tma_load_2d(st + Cfg::OFF_OLD + 1 * Cfg::AO, &tm_vo, x_old, y_old, &full[s]);
                tma_load_2d(st + Cfg::OFF_OLD + 2 * Cfg::AO, &tm_po, x_old, y_old, &full[s]);
                if (M) tma_load_2d(st + Cfg::OFF_MASK, &tm_mask, x_box_m, y_raw, &full[s]);
            }
        }
        return;
    }

    // ---- consumer warps ----
    // Every chunk is processed in full (rows past the segment's last raw row hold the next
    // rows of the array, or TMA's zero fill outside it); only the stores are predicated, so
    // the unrolled row loop is straight-line code.
    const int i = i_first + tid;
    const bool active = i <= prm.b.xe && i >= prm.b.xs;
    const int col = tid + ((i_first - 2) & 1);        // west neighbour's column inside the tile
    const int colm = tid + ((i_first - 2) & 3);
    const int colo = tid + ((i_first - 1) & 1);       // own column inside the old-field tile
    const SwScalars sc = prm.sc;
    Carry<M> cr;
    carry_init(cr);
    // output row trails the raw row by one: raw step t produces row j0-2+t
    size_t qo = idx(prm.pitch, i, j0 - 2);
    const size_t pitch = static_cast<size_t>(prm.pitch);
    // steps whose outputs are stored: 2 <= t < nsteps
    int t_lo = active ? 2 : (1 << 30), t_hi = nsteps;
    // push boxes this thread's column and this CTA's rows can meet (0 for almost every thread)
    unsigned pmask = 0;
    if (active)
        for (int q = 0; q < prm.push.n; ++q) {
            const PushSeg &g = prm.push.seg[q];
            if (i >= g.i0 && i <= g.i1 && g.j1 >= j0 && g.j0 <= j1) pmask |= 1u << q;
        }
    int jrow = j0 - 2;
    bool wet0 = true, wet1 = true, wet2 = true;   // warp-wide "row all wet", newest first (masked runs)
    for (int c = 0; c < nchunks; ++c) {
        const int s = c % STAGES;
        mbar_wait(&full[s], (c / STAGES) & 1);
        const unsigned char *st = fused_smem + s * Cfg::STAGE_BYTES;
        const double *sp = reinterpret_cast<const double *>(st + 0 * Cfg::A) + col;
        const double *su = reinterpret_cast<const double *>(st + 1 * Cfg::A) + col;
        const double *sv = reinterpret_cast<const double *>(st + 2 * Cfg::A) + col;
        const double *suo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 0 * Cfg::AO) + colo;
        const double *svo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 1 * Cfg::AO) + colo;
        const double *spo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 2 * Cfg::AO) + colo;
        const int *sm = reinterpret_cast<const int *>(st + Cfg::OFF_MASK) + colm;
#pragma unroll
        for (int r = 0; r < RY; ++r) {
            Raw<M> N;
            N.pW = sp[r * BW]; N.pC = sp[r * BW + 1]; N.pE = sp[r * BW + 2];
            N.uW = su[r * BW]; N.uC = su[r * BW + 1]; N.uE = su[r * BW + 2];
            N.vW = sv[r * BW]; N.vC = sv[r * BW + 1]; N.vE = sv[r * BW + 2];
            if (M) {
                N.tW = sm[r * BWM] == 1; N.tC = sm[r * BWM + 1] == 1; N.tE = sm[r * BWM + 2] == 1;
            }
            const double uo = suo[r * BWO], vo = svo[r * BWO], po = spo[r * BWO];
            Outs out;
            if (M) {
                // rows j, j-1, j-2 all wet across the warp's columns and their neighbours
                // (warp-uniform): the unmasked march gives the same bits without the selects
                wet2 = wet1; wet1 = wet0;
                wet0 = __all_sync(0xffffffffu, N.tW && N.tC && N.tE);
                if (wet0 && wet1 && wet2) ingest<M, false>(cr, N, sc, uo, vo, po, out);
                else ingest<M, true>(cr, N, sc, uo, vo, po, out);
            } else {
                ingest<M>(cr, N, sc, uo, vo, po, out);
            }
            if (r >= t_lo && r < t_hi) {
                __stcs(prm.unew + qo, out.unew);
                __stcs(prm.vnew + qo, out.vnew);
                __stcs(prm.pnew + qo, out.pnew);
                if (!FIRST) {
                    if (!M || out.wu) __stcs(prm.uold + qo, out.uold);
                    if (!M || out.wv) __stcs(prm.vold + qo, out.vold);
                    if (!M || out.wt) __stcs(prm.pold + qo, out.pold);
                }
                if (pmask) push_row(prm.push, pmask, i, jrow, out.unew, out.vnew, out.pnew);
            }
            qo += pitch;
            ++jrow;
        }
        t_lo -= RY;
        t_hi -= RY;
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
std::map<std::tuple<const void *, int, int, int, int, int>, CUtensorMap> g_maps;

// 2-D tiled tensor map over a column-major (nx, ny) array of FP64 (elem_bytes 8) or int32 (4)
bool get_map(const void *base, int elem_bytes, int nx, int ny, int boxw, int boxh, CUtensorMap &out)
{
    std::lock_guard<std::mutex> lock(g_map_mutex);
    auto key = std::make_tuple(base, elem_bytes, nx, ny, boxw, boxh);
    auto it = g_maps.find(key);
    if (it != g_maps.end()) {
        out = it->second;
        return true;
    }
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    cuuint64_t gdim[2] = {static_cast<cuuint64_t>(nx), static_cast<cuuint64_t>(ny)};
    cuuint64_t gstride[1] = {static_cast<cuuint64_t>(nx) * elem_bytes};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(boxw), static_cast<cuuint32_t>(boxh)};
    cuuint32_t estr[2] = {1, 1};
    // L2 promotion of the TMA reads; DLESM_TMA_L2PROMO=0..3 (none, 64 B, 128 B, 256 B) overrides
    // the default for the experiments recorded in DESIGN.md
    static const int promo = [] {
        const char *e = std::getenv("DLESM_TMA_L2PROMO");
        const int v = e ? std::atoi(e) : 3;
        return v < 0 || v > 3 ? 3 : v;
    }();
    const CUtensorMapL2promotion promos[4] = {CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_64B,
                                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B};
    CUtensorMap m;
    CUresult r = fn(&m, elem_bytes == 8 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_INT32, 2,
                    const_cast<void *>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, promos[promo], CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    if (g_maps.size() > 256) g_maps.clear();
    g_maps[key] = m;
    out = m;
    return true;
}

// Pick the segment height.  Every segment costs two extra raw rows (its prologue) and its
// row count is rounded up to whole chunks of `ry` rows, so: segments of about `target` rows,
// the CTA count rounded up to whole waves of `slots` resident CTAs, and seg_rows + 2 a
// multiple of `ry` whenever that does not add a segment.
int pick_seg_rows(int w, int h, int tx, int slots, int ry, int target, int extra = 2)
{
    const int strips = (w + tx - 1) / tx;
    int nseg = (h + target - 1) / target;
    if (nseg < 1) nseg = 1;
    long long ctas = static_cast<long long>(strips) * nseg;
    const long long waves = (ctas + slots - 1) / slots;
    nseg = static_cast<int>((waves * slots) / strips);
    if (nseg < 1) nseg = 1;
    if (nseg > h) nseg = h;
    int seg = (h + nseg - 1) / nseg;
    // seg + extra a multiple of ry: round up if the segment count stays the same, else down
    const int up = ((seg + extra + ry - 1) / ry) * ry - extra;
    const int down = ((seg + extra) / ry) * ry - extra;
    if (up >= 1 && (h + up - 1) / up == (h + seg - 1) / seg) seg = up;
    else if (down >= ry) seg = down;
    return seg;
}

template <int TX, int RY, int STAGES, int MINB, bool M, bool FIRST, bool ALN>
cudaError_t launch_tma_mf(const FusedArgs &a, cudaStream_t s)
{
    using Cfg = TmaCfg<TX, RY, STAGES, M, ALN>;
    constexpr bool aligned_origin = ALN;
    CUtensorMap mu, mv, mp, muo, mvo, mpo, mm;
    if (!get_map(a.u, 8, a.pitch, a.ny, Cfg::BW, RY, mu) || !get_map(a.v, 8, a.pitch, a.ny, Cfg::BW, RY, mv) ||
        !get_map(a.p, 8, a.pitch, a.ny, Cfg::BW, RY, mp) || !get_map(a.uold, 8, a.pitch, a.ny, Cfg::BWO, RY, muo) ||
        !get_map(a.vold, 8, a.pitch, a.ny, Cfg::BWO, RY, mvo) || !get_map(a.pold, 8, a.pitch, a.ny, Cfg::BWO, RY, mpo))
        return cudaErrorNotSupported;
    if (M) {
        if (!get_map(a.tmask, 4, a.pitch, a.ny, Cfg::BWM, RY, mm)) return cudaErrorNotSupported;
    } else {
        mm = mp;  // unused by the unmasked kernel
    }
    TmaStepParams prm;
    prm.unew = a.unew; prm.vnew = a.vnew; prm.pnew = a.pnew;
    prm.uold = a.uold; prm.vold = a.vold; prm.pold = a.pold;
    // aligned origin: strips start on 128-byte lines of the row (1-based column 1 + 16k), the
    // columns below xs are masked; otherwise strips start at xs
    const int i_origin = aligned_origin ? ((a.b.xs - 1) / 16) * 16 + 1 : a.b.xs;
    const int w = a.b.xe - i_origin + 1, h = a.b.ye - a.b.ys + 1;
    static const int target = [] {
        // Short segments keep the strips of a band in lockstep, so that together they sweep whole
        // contiguous rows of DRAM and re-read each other's halo rows out of L2 (measured at
        // 16384^2: 256 rows 0.93, 128 rows 0.95, 48 rows 0.995, 24-32 rows 1.01-1.02 of the
        // measured copy bandwidth).  DLESM_SEG_ROWS overrides for experiments.
        const char *e = std::getenv("DLESM_SEG_ROWS");
        const int v = e ? std::atoi(e) : 32;
        return v < 8 ? 32 : v;
    }();
    const int seg_rows = pick_seg_rows(w, h, TX, 148 * MINB, RY, target);
    prm.pitch = a.pitch; prm.b = a.b; prm.seg_rows = seg_rows; prm.i_origin = i_origin; prm.sc = a.sc;
    prm.push = a.push;
    auto kern = fused_tma_kernel<TX, RY, STAGES, MINB, M, FIRST, ALN>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    dim3 grid((w + TX - 1) / TX, (h + seg_rows - 1) / seg_rows);
    kern<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, s>>>(mu, mv, mp, muo, mvo, mpo, mm, prm);
    return cudaGetLastError();
}

template <int TX, int RY, int STAGES, int MINB, bool ALN>
cudaError_t launch_tma(const FusedArgs &a, cudaStream_t s)
{
    const bool first = a.sc.first != 0;
    if (a.tmask)
        return first ? launch_tma_mf<TX, RY, STAGES, MINB, true, true, ALN>(a, s)
                     : launch_tma_mf<TX, RY, STAGES, MINB, true, false, ALN>(a, s);
    return first ? launch_tma_mf<TX, RY, STAGES, MINB, false, true, ALN>(a, s)
                 : launch_tma_mf<TX, RY, STAGES, MINB, false, false, ALN>(a, s);
}

// ---------------------------------------------------------------------------------------
// Two model steps per sweep (temporal blocking).  Stage A is the march above, run on a strip
// one column wider on each side and one row longer at each end, producing the intermediate
// state (u1,v1,p1 and the once-filtered old fields) of step n+1; its rows go through a small
// shared-memory exchange buffer so that stage B -- the same march again, fed by u1,v1,p1 of the
// west / own / east column -- produces step n+2 two rows behind.  The intermediate state never
// reaches memory: 6 reads + 6 writes per point for TWO steps.  Thread t of TX owns column
// i_first - 1 + t in both stages; threads 1 .. TX-2 store.
// ---------------------------------------------------------------------------------------
struct Tma2Params {
    double *u2, *v2, *p2, *uold2, *vold2, *pold2;
    int pitch;
    Box b;
    int seg_rows;
    SwScalars sc;
    PushPlan2 push;
};

// edge outputs of row j, column i -> the neighbours' halo cells, two cells deep (rare path)
__device__ __noinline__ void push_row2(const PushPlan2 &pp, unsigned mask, int i, int j, double un, double vn,
                                       double pn, double uo, double vo, double po)
{
    for (unsigned m = mask; m; m &= m - 1) {
        const PushSeg2 &g = pp.seg[__ffs(m) - 1];
        if (j >= g.j0 && j <= g.j1) {
            const size_t d = idx(g.dpitch, g.ides + (i - g.i0), g.jdes + (j - g.j0));
            g.dst[0][d] = un;
            g.dst[1][d] = vn;
            g.dst[2][d] = pn;
            g.dst[3][d] = uo;
            g.dst[4][d] = vo;
            g.dst[5][d] = po;
        }
    }
}

template <int TX, int RY, int STAGES>
struct Tma2Cfg {
    static constexpr int BW = TX + 4;    // raw tiles: west halo + TX + east halo, even start column
    static constexpr int BWO = TX + 2;   // old tiles: TX columns, even start column
    static constexpr int A = align128(BW * RY * 8);
    static constexpr int AO = align128(BWO * RY * 8);
    static constexpr int OFF_OLD = 3 * A;
    static constexpr int STAGE_BYTES = 3 * A + 3 * AO;
    static constexpr int TX_BYTES = 3 * BW * RY * 8 + 3 * BWO * RY * 8;
    static constexpr int XW = TX + 2;    // exchange row: one pad column on each side
    static constexpr int XBUF_BYTES = 3 * RY * XW * 8;   // u1, v1, p1 rows of one chunk
    static constexpr int OFF_X = STAGES * STAGE_BYTES;
    static constexpr int BAR_OFFSET = OFF_X + 2 * XBUF_BYTES;
    static constexpr int SMEM_BYTES = BAR_OFFSET + STAGES * 8;
    static constexpr int THREADS = TX;
    static constexpr int OUT = TX - 2;   // output columns per strip
};

template <int TX, int RY, int STAGES, int MINB>
__global__ void __launch_bounds__(TX, MINB)
fused2_tma_kernel(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_v,
                  const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_uo,
                  const __grid_constant__ CUtensorMap tm_vo, const __grid_constant__ CUtensorMap tm_po,
                  const __grid_constant__ Tma2Params prm)
{
    using Cfg = Tma2Cfg<TX, RY, STAGES>;
    constexpr int BW = Cfg::BW, BWO = Cfg::BWO, XW = Cfg::XW;
    extern __shared__ __align__(1024) unsigned char fused_smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(fused_smem + Cfg::BAR_OFFSET);

    const int tid = threadIdx.x;
    const int j0 = prm.b.ys + blockIdx.y * prm.seg_rows;
    const int j1 = min(j0 + prm.seg_rows - 1, prm.b.ye);
    const int nsteps = (j1 - j0 + 1) + 4;            // raw rows j0-2 .. j1+2
    const int nchunks = (nsteps + RY - 1) / RY;
    const int i_first = prm.b.xs + blockIdx.x * Cfg::OUT;   // first output column of this strip
    const int x_box = (i_first - 3) & ~1;   // even 0-based column at/below thread 0's west neighbour
    const int x_old = (i_first - 2) & ~1;   // even 0-based column at/below thread 0's own column

    // No producer warp here: the per-chunk barrier between the two stages already tells thread 0
    // that every warp is done with a ring stage, so it refills the stage right after it.  (A
    // fifth warp would also cap the kernel at 168 registers: 10 warps on 4 schedulers.)
    auto issue = [&](int c) {
        const int s = c % STAGES;
        unsigned char *st = fused_smem + s * Cfg::STAGE_BYTES;
        const int y_raw = (j0 - 2) + c * RY - 1;   // 0-based row of the chunk's first raw row
        const int y_old = y_raw - 1;               // stage A's outputs trail the raw rows by one
        mbar_expect_tx(&full[s], Cfg::TX_BYTES);
        tma_load_2d(st + 0 * Cfg::A, &tm_p, x_box, y_raw, &full[s]);
        tma_load_2d(st + 1 * Cfg::A, &tm_u, x_box, y_raw, &full[s]);
        tma_load_2d(st + 2 * Cfg::A, &tm_v, x_box, y_raw, &full[s]);
        tma_load_2d(st + Cfg::OFF_OLD + 0 * Cfg::AO, &tm_uo, x_old, y_old, &full[s]);
        tma_load_2d(st + Cfg::OFF_OLD + 1 * Cfg::AO, &tm_vo, x_old, y_old, &full[s]);
        tma_load_2d(st + Cfg::OFF_OLD + 2 * Cfg::AO, &tm_po, x_old, y_old, &full[s]);
    };
    if (tid == 0) {
        if (smem_u32(fused_smem) & 127u) __trap();
        for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int c = 0; c < STAGES && c < nchunks; ++c) issue(c);
    }
    // pad columns of the exchange rows (read by threads 0 and TX-1 only, whose stage-B results
    // are never stored): finite values so that no inf/nan is manufactured
    for (int q = tid; q < 2 * 3 * RY * 2; q += TX) {
        const int row = q >> 1, side = q & 1;        // row over (buffer, field, r)
        double *xr = reinterpret_cast<double *>(fused_smem + Cfg::OFF_X) + row * XW;
        xr[side ? XW - 1 : 0] = 1.0;
    }
    __syncthreads();

    const int i = i_first - 1 + tid;
    const bool active = tid >= 1 && tid <= TX - 2 && i <= prm.b.xe;
    const int col = tid + ((i_first - 3) & 1);        // west neighbour's column inside the raw tile
    const int colo = tid + ((i_first - 2) & 1);       // own column inside the old-field tile
    const SwScalars sc = prm.sc;
    Carry<false> ca, cb;
    carry_init(ca);
    carry_init(cb);
    double uo1 = 0.0, vo1 = 0.0, po1 = 0.0;   // once-filtered old fields of the row stage B emits next
    // raw step t ingests raw row j0-2+t; stage A emits row j0-3+t; stage B emits row j0-4+t
    long long qo = (static_cast<long long>(j0) - 5) * prm.pitch + (i - 1);
    const long long pitch = prm.pitch;
    int t_lo = active ? 4 : (1 << 30), t_hi = nsteps;
    for (int c = 0; c < nchunks; ++c) {
        const int s = c % STAGES;
        mbar_wait(&full[s], (c / STAGES) & 1);
        const unsigned char *st = fused_smem + s * Cfg::STAGE_BYTES;
        const double *sp = reinterpret_cast<const double *>(st + 0 * Cfg::A) + col;
        const double *su = reinterpret_cast<const double *>(st + 1 * Cfg::A) + col;
        const double *sv = reinterpret_cast<const double *>(st + 2 * Cfg::A) + col;
        const double *suo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 0 * Cfg::AO) + colo;
        const double *svo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 1 * Cfg::AO) + colo;
        const double *spo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 2 * Cfg::AO) + colo;
        double *xb = reinterpret_cast<double *>(fused_smem + Cfg::OFF_X + (c & 1) * Cfg::XBUF_BYTES) + tid;
        double uo_r[RY], vo_r[RY], po_r[RY];
        // ---- stage A: RY raw rows -> RY rows of the intermediate state ----
#pragma unroll
        for (int r = 0; r < RY; ++r) {
            Raw<false> N;
            N.pW = sp[r * BW]; N.pC = sp[r * BW + 1]; N.pE = sp[r * BW + 2];
            N.uW = su[r * BW]; N.uC = su[r * BW + 1]; N.uE = su[r * BW + 2];
            N.vW = sv[r * BW]; N.vC = sv[r * BW + 1]; N.vE = sv[r * BW + 2];
            Outs o;
            ingest<false>(ca, N, sc, suo[r * BWO], svo[r * BWO], spo[r * BWO], o);
            xb[(0 * RY + r) * XW + 1] = o.unew;
            xb[(1 * RY + r) * XW + 1] = o.vnew;
            xb[(2 * RY + r) * XW + 1] = o.pnew;
            uo_r[r] = o.uold; vo_r[r] = o.vold; po_r[r] = o.pold;
        }
        __syncthreads();
        if (tid == 0 && c + STAGES < nchunks) issue(c + STAGES);
        // ---- stage B: those RY rows (west / own / east column) -> RY rows of step n+2 ----
#pragma unroll
        for (int r = 0; r < RY; ++r) {
            Raw<false> N;
            N.uW = xb[(0 * RY + r) * XW]; N.uC = xb[(0 * RY + r) * XW + 1]; N.uE = xb[(0 * RY + r) * XW + 2];
            N.vW = xb[(1 * RY + r) * XW]; N.vC = xb[(1 * RY + r) * XW + 1]; N.vE = xb[(1 * RY + r) * XW + 2];
            N.pW = xb[(2 * RY + r) * XW]; N.pC = xb[(2 * RY + r) * XW + 1]; N.pE = xb[(2 * RY + r) * XW + 2];
            Outs o;
            ingest<false>(cb, N, sc, uo1, vo1, po1, o);
            uo1 = uo_r[r]; vo1 = vo_r[r]; po1 = po_r[r];
            if (r >= t_lo && r < t_hi) {
                __stcs(prm.u2 + qo, o.unew);
                __stcs(prm.v2 + qo, o.vnew);
                __stcs(prm.p2 + qo, o.pnew);
                __stcs(prm.uold2 + qo, o.uold);
                __stcs(prm.vold2 + qo, o.vold);
                __stcs(prm.pold2 + qo, o.pold);
            }
            qo += pitch;
        }
        t_lo -= RY;
        t_hi -= RY;
    }
}

// ---------------------------------------------------------------------------------------
// Two steps per sweep, second form: the two-step sweep is bound by the FP64 pipe, not by HBM
// (ncu: 75 % FP64 pipe, 27 GB of DRAM traffic in 4.5 ms), and a third of its FP64 work is the
// east / west copies of cu, z / cv, h every thread recomputes for itself.  Here every lane
// computes the four intermediates of its own column only and takes the neighbours' through
// warp shuffles.  Warps overlap by two columns (lane 0 and lane 31 only feed their
// neighbours), so a warp yields 30 columns: 4 warps = 120 columns of the intermediate state,
// 118 output columns per CTA.  The expression trees are unchanged -- lane l+1's cu(own) IS
// lane l's cu(east), operand for operand -- so the bits are the same.
// ---------------------------------------------------------------------------------------
struct RawS {
    double pW, pC, uC, uE, vW, vC;
};
struct CarryS {
    RawS P;
    double cuC_P, cuE_P, cuC_PP, cuE_PP, cvC_P, cvW_P, zC_P, zE_P, hC_PP;
    bool tC_P, tW_P, tC_PP;   // tmask of row j-1 (own, west) and j-2 (own); masked runs only
};
__device__ __forceinline__ void carry_init(CarryS &c)
{
    c.tC_P = c.tW_P = c.tC_PP = true;
    c.P.pW = c.P.pC = 1.0;
    c.P.uC = c.P.uE = c.P.vW = c.P.vC = 0.0;
    c.cuC_P = c.cuE_P = c.cuC_PP = c.cuE_PP = c.cvC_P = c.cvW_P = c.zC_P = c.zE_P = c.hC_PP = 0.0;
}
// tC, tW: tmask == 1 at row N for the own and the west column (M only; the caller rotates the
// carried masks with rotate_masks afterwards, so that an all-wet warp can take ingest_sh<false>).  The masked rule is the
// one of ingest<true>: intermediates are 0 unless every T cell they straddle is wet, a dry
// prognostic point keeps its value and is not filtered -- the neighbours' copies arrive masked.
template <bool M>
__device__ __forceinline__ void ingest_sh(CarryS &c, const RawS &N, bool tC, bool tW, const SwScalars &sc,
                                          double uo, double vo, double po, Outs &o)
{
    const RawS &P = c.P;
    double cuC_N = 0.5 * (N.pC + N.pW) * N.uC;
    double cvC_N = 0.5 * (N.pC + P.pC) * N.vC;
    double zC_N = (sc.fsdx * (N.vC - N.vW) - sc.fsdy * (N.uC - P.uC)) / (P.pW + P.pC + N.pC + N.pW);
    double hC_P = P.pC + 0.25 * (P.uE * P.uE + P.uC * P.uC + N.vC * N.vC + P.vC * P.vC);
    if (M) {
        if (!(tC && tW)) cuC_N = 0.0;
        if (!(tC && c.tC_P)) cvC_N = 0.0;
        if (!(tC && tW && c.tC_P && c.tW_P)) zC_N = 0.0;
        if (!c.tC_P) hC_P = 0.0;
    }
    const double cuE_N = __shfl_down_sync(0xffffffffu, cuC_N, 1);
    const double zE_N = __shfl_down_sync(0xffffffffu, zC_N, 1);
    const double cvW_N = __shfl_up_sync(0xffffffffu, cvC_N, 1);
    const double hW_P = __shfl_up_sync(0xffffffffu, hC_P, 1);
    o.unew = uo + sc.tdts8 * (zC_N + c.zC_P) * (cvC_N + cvW_N + c.cvW_P + c.cvC_P) -
             sc.tdtsdx * (hC_P - hW_P);
    o.vnew = vo - sc.tdts8 * (c.zE_P + c.zC_P) * (c.cuE_P + c.cuC_P + c.cuC_PP + c.cuE_PP) -
             sc.tdtsdy * (hC_P - c.hC_PP);
    o.pnew = po - sc.tdtsdx * (c.cuE_P - c.cuC_P) - sc.tdtsdy * (cvC_N - c.cvC_P);
    bool wu = true, wv = true, wt = true;
    if (M) {
        wu = c.tC_P && c.tW_P;
        wv = c.tC_P && c.tC_PP;
        wt = c.tC_P;
        if (!wu) o.unew = P.uC;
        if (!wv) o.vnew = P.vC;
        if (!wt) o.pnew = P.pC;
    }
    o.uold = P.uC + sc.alpha * (o.unew - 2.0 * P.uC + uo);
    o.vold = P.vC + sc.alpha * (o.vnew - 2.0 * P.vC + vo);
    o.pold = P.pC + sc.alpha * (o.pnew - 2.0 * P.pC + po);
    if (M) {
        if (!wu) o.uold = uo;
        if (!wv) o.vold = vo;
        if (!wt) o.pold = po;
    }
    c.cuC_PP = c.cuC_P; c.cuE_PP = c.cuE_P;
    c.cuC_P = cuC_N; c.cuE_P = cuE_N;
    c.cvC_P = cvC_N; c.cvW_P = cvW_N;
    c.zC_P = zC_N; c.zE_P = zE_N;
    c.hC_PP = hC_P;
    c.P = N;
}

__device__ __forceinline__ void rotate_masks(CarryS &c, bool tC, bool tW)
{
    c.tC_PP = c.tC_P;
    c.tC_P = tC;
    c.tW_P = tW;
}

template <int TX, int RY, int STAGES, bool M = false>
struct Tma2sCfg {
    static constexpr int NW = TX / 32;
    static constexpr int VA = 30 * NW;   // columns of the intermediate state a CTA produces
    static constexpr int OUT = VA - 2;   // output columns per CTA
    static constexpr int BW = VA + 6;    // raw tiles: lanes' columns (VA + 2) + west + east, even start
    static constexpr int BWO = VA + 4;   // old tiles: lanes' columns, even start
    static constexpr int BWM = VA + 8;   // int32 tmask tile: lanes' columns, start column a multiple of four
    static constexpr int A = align128(BW * RY * 8);
    static constexpr int AO = align128(BWO * RY * 8);
    static constexpr int AM = M ? align128(BWM * RY * 4) : 0;
    static constexpr int OFF_OLD = 3 * A;
    static constexpr int OFF_MASK = 3 * A + 3 * AO;
    static constexpr int STAGE_BYTES = 3 * A + 3 * AO + AM;
    static constexpr int TX_BYTES = 3 * BW * RY * 8 + 3 * BWO * RY * 8 + (M ? BWM * RY * 4 : 0);
    static constexpr int XW = VA + 4;    // exchange row: index of column x = x - i_first + 3
    static constexpr int XBUF_BYTES = 3 * RY * XW * 8;
    static constexpr int OFF_X = STAGES * STAGE_BYTES;
    static constexpr int BAR_OFFSET = OFF_X + 2 * XBUF_BYTES;
    static constexpr int SMEM_BYTES = BAR_OFFSET + STAGES * 8;
    static constexpr int THREADS = TX;
};

template <int TX, int RY, int STAGES, int MINB, bool M>
__global__ void __launch_bounds__(TX, MINB)
fused2s_tma_kernel(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_v,
                   const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_uo,
                   const __grid_constant__ CUtensorMap tm_vo, const __grid_constant__ CUtensorMap tm_po,
                   const __grid_constant__ CUtensorMap tm_mask, const __grid_constant__ Tma2Params prm)
{
    using Cfg = Tma2sCfg<TX, RY, STAGES, M>;
    constexpr int BW = Cfg::BW, BWO = Cfg::BWO, BWM = Cfg::BWM, XW = Cfg::XW;
    extern __shared__ __align__(1024) unsigned char fused_smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(fused_smem + Cfg::BAR_OFFSET);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // (Mapping the x-edge strips, which push on every row, to the lowest block ids removed the tail of
    // short sweeps when those neighbours are the rank itself, but with REMOTE x neighbours it bunches
    // ~100k single-cell NVLink stores into the first waves: 2x2 tiles, 4.8 -> 7.0 ms per launch.  Natural
    // order spreads them over the sweep.)
    const int bx = blockIdx.x, by = blockIdx.y;
    const int j0 = prm.b.ys + by * prm.seg_rows;
    const int j1 = min(j0 + prm.seg_rows - 1, prm.b.ye);
    const int nsteps = (j1 - j0 + 1) + 4;            // raw rows j0-2 .. j1+2
    const int nchunks = (nsteps + RY - 1) / RY;
    const int i_first = prm.b.xs + bx * Cfg::OUT;   // first output column of this CTA
    // lane column (both stages): i_first + 30*warp + lane - 2
    const int x_box = (i_first - 4) & ~1;   // even 0-based column at/below stage A lane 0's west neighbour
    const int x_old = (i_first - 3) & ~1;   // even 0-based column at/below stage A lane 0's own column
    const int x_msk = (i_first - 3) & ~3;   // the int32 tmask tile starts on a multiple of four columns

    auto issue = [&](int c) {
        const int s = c % STAGES;
        unsigned char *st = fused_smem + s * Cfg::STAGE_BYTES;
        const int y_raw = (j0 - 2) + c * RY - 1;
        const int y_old = y_raw - 1;
        mbar_expect_tx(&full[s], Cfg::TX_BYTES);
        tma_load_2d(st + 0 * Cfg::A, &tm_p, x_box, y_raw, &full[s]);
        tma_load_2d(st + 1 * Cfg::A, &tm_u, x_box, y_raw, &full[s]);
        tma_load_2d(st + 2 * Cfg::A, &tm_v, x_box, y_raw, &full[s]);
        tma_load_2d(st + Cfg::OFF_OLD + 0 * Cfg::AO, &tm_uo, x_old, y_old, &full[s]);
        tma_load_2d(st + Cfg::OFF_OLD + 1 * Cfg::AO, &tm_vo, x_old, y_old, &full[s]);
        tma_load_2d(st + Cfg::OFF_OLD + 2 * Cfg::AO, &tm_po, x_old, y_old, &full[s]);
        if (M) tma_load_2d(st + Cfg::OFF_MASK, &tm_mask, x_msk, y_raw, &full[s]);
    };
    if (tid == 0) {
        if (smem_u32(fused_smem) & 127u) __trap();
        for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int c = 0; c < STAGES && c < nchunks; ++c) issue(c);
    }
    // exchange columns no stage-A lane writes (index 0, 1, VA+2, VA+3): finite values
    for (int q = tid; q < 2 * 3 * RY * 4; q += TX) {
        const int row = q >> 2, k4 = q & 3;
        double *xr = reinterpret_cast<double *>(fused_smem + Cfg::OFF_X) + row * XW;
        xr[k4 < 2 ? k4 : Cfg::VA + k4] = 1.0;
    }
    __syncthreads();

    const int base = 30 * warp + lane;
    const int i = i_first + base - 2;                 // this lane's column
    // lanes 0 and 31 only feed their neighbours; the CTA's first and last intermediate columns
    // (i_first - 1, i_first + OUT) have no valid neighbour on one side in stage B
    const bool a_writes = lane >= 1 && lane <= 30;
    const bool active = a_writes && i >= i_first && i < i_first + Cfg::OUT && i <= prm.b.xe;
    const int xo = base + 1;                          // exchange index of this lane's column
    const int col = base + ((i_first - 4) & 1);       // west neighbour inside the raw tile
    const int colo = base + ((i_first - 3) & 1);      // own column inside the old tile
    const int colm = base + ((i_first - 3) & 3);      // own column inside the tmask tile
    const SwScalars sc = prm.sc;
    CarryS ca, cb;
    carry_init(ca);
    carry_init(cb);
    double uo1 = 0.0, vo1 = 0.0, po1 = 0.0;
    bool wet0 = true, wet1 = true, wet2 = true, wet3 = true;   // warp-wide "row all wet", newest first
    long long qo = (static_cast<long long>(j0) - 5) * prm.pitch + (i - 1);
    const long long pitch = prm.pitch;
    int t_lo = active ? 4 : (1 << 30), t_hi = nsteps;
    // push boxes this lane's column and this CTA's rows can meet (0 for almost every thread)
    unsigned pmask = 0;
    if (active)
        for (int q = 0; q < prm.push.n; ++q) {
            const PushSeg2 &g = prm.push.seg[q];
            if (i >= g.i0 && i <= g.i1 && g.j1 >= j0 && g.j0 <= j1) pmask |= 1u << q;
        }
    const int pfirst = pmask ? __ffs(pmask) - 1 : 0;
    const unsigned prest = pmask & (pmask - 1);
    int jrow = j0 - 4;
    for (int c = 0; c < nchunks; ++c) {
        const int s = c % STAGES;
        mbar_wait(&full[s], (c / STAGES) & 1);
        const unsigned char *st = fused_smem + s * Cfg::STAGE_BYTES;
        const double *sp = reinterpret_cast<const double *>(st + 0 * Cfg::A) + col;
        const double *su = reinterpret_cast<const double *>(st + 1 * Cfg::A) + col;
        const double *sv = reinterpret_cast<const double *>(st + 2 * Cfg::A) + col;
        const double *suo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 0 * Cfg::AO) + colo;
        const double *svo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 1 * Cfg::AO) + colo;
        const double *spo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 2 * Cfg::AO) + colo;
        double *xb = reinterpret_cast<double *>(fused_smem + Cfg::OFF_X + (c & 1) * Cfg::XBUF_BYTES) + xo;
        const int *sm = reinterpret_cast<const int *>(st + Cfg::OFF_MASK) + colm;
        double uo_r[RY], vo_r[RY], po_r[RY];
        bool tbC[RY], tbW[RY];   // stage B's row is stage A's previous row: its masks are A's carried ones
        bool fastB[RY];
#pragma unroll
        for (int r = 0; r < RY; ++r) {
            RawS N;
            N.pW = sp[r * BW]; N.pC = sp[r * BW + 1];
            N.uC = su[r * BW + 1]; N.uE = su[r * BW + 2];
            N.vW = sv[r * BW]; N.vC = sv[r * BW + 1];
            bool tC = true, tW = true;
            Outs o;
            if (M) {
                tC = sm[r * BWM] == 1;
                tW = __shfl_up_sync(0xffffffffu, static_cast<int>(tC), 1) != 0;
                tbC[r] = ca.tC_P; tbW[r] = ca.tW_P;
                // rows N, P, PP all wet across the warp (warp-uniform): the unmasked march gives the
                // same bits without the selects -- almost every warp-row away from the islands
                wet3 = wet2; wet2 = wet1; wet1 = wet0;
                wet0 = __all_sync(0xffffffffu, tC);
                fastB[r] = wet1 && wet2 && wet3;
                if (wet0 && wet1 && wet2)
                    ingest_sh<false>(ca, N, true, true, sc, suo[r * BWO], svo[r * BWO], spo[r * BWO], o);
                else
                    ingest_sh<true>(ca, N, tC, tW, sc, suo[r * BWO], svo[r * BWO], spo[r * BWO], o);
                rotate_masks(ca, tC, tW);
            } else {
                ingest_sh<false>(ca, N, true, true, sc, suo[r * BWO], svo[r * BWO], spo[r * BWO], o);
            }
            if (a_writes) {
                xb[(0 * RY + r) * XW] = o.unew;
                xb[(1 * RY + r) * XW] = o.vnew;
                xb[(2 * RY + r) * XW] = o.pnew;
            }
            uo_r[r] = o.uold; vo_r[r] = o.vold; po_r[r] = o.pold;
        }
        __syncthreads();
        if (tid == 0 && c + STAGES < nchunks) issue(c + STAGES);
#pragma unroll
        for (int r = 0; r < RY; ++r) {
            RawS N;
            N.uC = xb[(0 * RY + r) * XW]; N.uE = xb[(0 * RY + r) * XW + 1];
            N.vW = xb[(1 * RY + r) * XW - 1]; N.vC = xb[(1 * RY + r) * XW];
            N.pW = xb[(2 * RY + r) * XW - 1]; N.pC = xb[(2 * RY + r) * XW];
            Outs o;
            if (M) {
                if (fastB[r]) ingest_sh<false>(cb, N, true, true, sc, uo1, vo1, po1, o);
                else ingest_sh<true>(cb, N, tbC[r], tbW[r], sc, uo1, vo1, po1, o);
                rotate_masks(cb, tbC[r], tbW[r]);
            } else {
                ingest_sh<false>(cb, N, true, true, sc, uo1, vo1, po1, o);
            }
            uo1 = uo_r[r]; vo1 = vo_r[r]; po1 = po_r[r];
            if (r >= t_lo && r < t_hi) {
                __stcs(prm.u2 + qo, o.unew);
                __stcs(prm.v2 + qo, o.vnew);
                __stcs(prm.p2 + qo, o.pnew);
                __stcs(prm.uold2 + qo, o.uold);
                __stcs(prm.vold2 + qo, o.vold);
                __stcs(prm.pold2 + qo, o.pold);
                if (pmask) {
                    // first box inline (an x-edge column meets it on every row; a call would save and
                    // restore ~150 live registers each time), the rest -- corners -- out of line
                    const PushSeg2 &g = prm.push.seg[pfirst];
                    if (jrow >= g.j0 && jrow <= g.j1) {
                        const size_t d = idx(g.dpitch, g.ides + (i - g.i0), g.jdes + (jrow - g.j0));
                        g.dst[0][d] = o.unew; g.dst[1][d] = o.vnew; g.dst[2][d] = o.pnew;
                        g.dst[3][d] = o.uold; g.dst[4][d] = o.vold; g.dst[5][d] = o.pold;
                    }
                    if (prest) push_row2(prm.push, prest, i, jrow, o.unew, o.vnew, o.pnew, o.uold, o.vold, o.pold);
                }
            }
            qo += pitch;
            ++jrow;
        }
        t_lo -= RY;
        t_hi -= RY;
    }
}

// ---------------------------------------------------------------------------------------
// Two steps per sweep, third form: warp-private.  ncu on the shuffle form showed 23 % of all
// warp samples waiting at the per-chunk CTA barrier of the shared-memory exchange.  Here the
// intermediate state crosses lanes by shuffle as well, so a warp depends on no other warp:
// stage A row -> shuffle -> stage B row, interleaved row by row, no barrier, no exchange
// buffer.  Lanes 2..29 store (28 columns per warp); a producer warp feeds the TMA ring as in the
// one-step sweep.
// ---------------------------------------------------------------------------------------
template <int TX, int RY, int STAGES>
struct Tma2wCfg {
    static constexpr int NW = TX / 32;
    static constexpr int OUT = 28 * NW;      // output columns per CTA
    static constexpr int BW = 28 * NW + 8;   // raw tiles, even start column
    static constexpr int BWO = 28 * NW + 6;  // old tiles, even start column
    static constexpr int A = align128(BW * RY * 8);
    static constexpr int AO = align128(BWO * RY * 8);
    static constexpr int OFF_OLD = 3 * A;
    static constexpr int STAGE_BYTES = 3 * A + 3 * AO;
    static constexpr int TX_BYTES = 3 * BW * RY * 8 + 3 * BWO * RY * 8;
    static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
    static constexpr int SMEM_BYTES = BAR_OFFSET + 2 * STAGES * 8;
    static constexpr int THREADS = TX + 32;
};

template <int TX, int RY, int STAGES, int MINB>
__global__ void __launch_bounds__(TX + 32, MINB)
fused2w_tma_kernel(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_v,
                   const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_uo,
                   const __grid_constant__ CUtensorMap tm_vo, const __grid_constant__ CUtensorMap tm_po,
                   const __grid_constant__ Tma2Params prm)
{
    using Cfg = Tma2wCfg<TX, RY, STAGES>;
    constexpr int BW = Cfg::BW, BWO = Cfg::BWO;
    extern __shared__ __align__(1024) unsigned char fused_smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(fused_smem + Cfg::BAR_OFFSET);
    uint64_t *empty = full + STAGES;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int j0 = prm.b.ys + blockIdx.y * prm.seg_rows;
    const int j1 = min(j0 + prm.seg_rows - 1, prm.b.ye);
    const int nsteps = (j1 - j0 + 1) + 4;            // raw rows j0-2 .. j1+2
    const int nchunks = (nsteps + RY - 1) / RY;
    const int i_first = prm.b.xs + blockIdx.x * Cfg::OUT;   // first output column of this CTA

    if (tid == 0) {
        if (smem_u32(fused_smem) & 127u) __trap();
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], TX / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (warp == TX / 32) {
        if (lane == 0) {
            const int x_box = (i_first - 4) & ~1;   // even 0-based column at/below lane 0's west neighbour
            const int x_old = (i_first - 3) & ~1;   // even 0-based column at/below lane 0's own column
            for (int c = 0; c < nchunks; ++c) {
                const int s = c % STAGES;
                if (c >= STAGES) mbar_wait(&empty[s], ((c / STAGES) - 1) & 1);
                unsigned char *st = fused_smem + s * Cfg::STAGE_BYTES;
                const int y_raw = (j0 - 2) + c * RY - 1;
                const int y_old = y_raw - 1;
                mbar_expect_tx(&full[s], Cfg::TX_BYTES);
                tma_load_2d(st + 0 * Cfg::A, &tm_p, x_box, y_raw, &full[s]);
                tma_load_2d(st + 1 * Cfg::A, &tm_u, x_box, y_raw, &full[s]);
                tma_load_2d(st + 2 * Cfg::A, &tm_v, x_box, y_raw, &full[s]);
                tma_load_2d(st + Cfg::OFF_OLD + 0 * Cfg::AO, &tm_uo, x_old, y_old, &full[s]);
                tma_load_2d(st + Cfg::OFF_OLD + 1 * Cfg::AO, &tm_vo, x_old, y_old, &full[s]);
                tma_load_2d(st + Cfg::OFF_OLD + 2 * Cfg::AO, &tm_po, x_old, y_old, &full[s]);
            }
        }
        return;
    }

    // ---- consumer warps: lane column = i_first + 28*warp + lane - 2 in both stages ----
    const int base = 28 * warp + lane;
    const int i = i_first + base - 2;
    const bool active = lane >= 2 && lane <= 29 && i <= prm.b.xe;
    const int col = base + ((i_first - 4) & 1);       // west neighbour inside the raw tile
    const int colo = base + ((i_first - 3) & 1);      // own column inside the old tile
    const SwScalars sc = prm.sc;
    CarryS ca, cb;
    carry_init(ca);
    carry_init(cb);
    double uo1 = 0.0, vo1 = 0.0, po1 = 0.0;   // once-filtered old fields of the row stage B emits next
    // raw step t ingests raw row j0-2+t; stage A emits row j0-3+t; stage B emits row j0-4+t
    long long qo = (static_cast<long long>(j0) - 5) * prm.pitch + (i - 1);
    const long long pitch = prm.pitch;
    int t_lo = active ? 4 : (1 << 30), t_hi = nsteps;
    for (int c = 0; c < nchunks; ++c) {
        const int s = c % STAGES;
        mbar_wait(&full[s], (c / STAGES) & 1);
        const unsigned char *st = fused_smem + s * Cfg::STAGE_BYTES;
        const double *sp = reinterpret_cast<const double *>(st + 0 * Cfg::A) + col;
        const double *su = reinterpret_cast<const double *>(st + 1 * Cfg::A) + col;
        const double *sv = reinterpret_cast<const double *>(st + 2 * Cfg::A) + col;
        const double *suo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 0 * Cfg::AO) + colo;
        const double *svo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 1 * Cfg::AO) + colo;
        const double *spo = reinterpret_cast<const double *>(st + Cfg::OFF_OLD + 2 * Cfg::AO) + colo;
#pragma unroll
        for (int r = 0; r < RY; ++r) {
            RawS N;
            N.pW = sp[r * BW]; N.pC = sp[r * BW + 1];
            N.uC = su[r * BW + 1]; N.uE = su[r * BW + 2];
            N.vW = sv[r * BW]; N.vC = sv[r * BW + 1];
            Outs oa;
            ingest_sh<false>(ca, N, true, true, sc, suo[r * BWO], svo[r * BWO], spo[r * BWO], oa);
            RawS M;   // the intermediate state's row, neighbours by shuffle (valid on lanes 1..30)
            M.uC = oa.unew; M.vC = oa.vnew; M.pC = oa.pnew;
            M.uE = __shfl_down_sync(0xffffffffu, oa.unew, 1);
            M.vW = __shfl_up_sync(0xffffffffu, oa.vnew, 1);
            M.pW = __shfl_up_sync(0xffffffffu, oa.pnew, 1);
            Outs ob;
            ingest_sh<false>(cb, M, true, true, sc, uo1, vo1, po1, ob);
            uo1 = oa.uold; vo1 = oa.vold; po1 = oa.pold;
            if (r >= t_lo && r < t_hi) {
                __stcs(prm.u2 + qo, ob.unew);
                __stcs(prm.v2 + qo, ob.vnew);
                __stcs(prm.p2 + qo, ob.pnew);
                __stcs(prm.uold2 + qo, ob.uold);
                __stcs(prm.vold2 + qo, ob.vold);
                __stcs(prm.pold2 + qo, ob.pold);
            }
            qo += pitch;
        }
        t_lo -= RY;
        t_hi -= RY;
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
}

template <class Cfg, int RY, int MINB, class Kern>
cudaError_t launch_tma2_generic(const Fused2Args &a, cudaStream_t s, Kern kern)
{
    CUtensorMap mu, mv, mp, muo, mvo, mpo;
    if (!get_map(a.u, 8, a.pitch, a.ny, Cfg::BW, RY, mu) || !get_map(a.v, 8, a.pitch, a.ny, Cfg::BW, RY, mv) ||
        !get_map(a.p, 8, a.pitch, a.ny, Cfg::BW, RY, mp) || !get_map(a.uold, 8, a.pitch, a.ny, Cfg::BWO, RY, muo) ||
        !get_map(a.vold, 8, a.pitch, a.ny, Cfg::BWO, RY, mvo) || !get_map(a.pold, 8, a.pitch, a.ny, Cfg::BWO, RY, mpo))
        return cudaErrorNotSupported;
    Tma2Params prm;
    prm.u2 = a.u2; prm.v2 = a.v2; prm.p2 = a.p2;
    prm.uold2 = a.uold2; prm.vold2 = a.vold2; prm.pold2 = a.pold2;
    const int w = a.b.xe - a.b.xs + 1, h = a.b.ye - a.b.ys + 1;
    static const int target = [] {
        // every segment recomputes two rows of the intermediate state and re-reads four raw rows,
        // so segments are longer than the one-step sweep's; DLESM_SEG_ROWS2 overrides
        const char *e = std::getenv("DLESM_SEG_ROWS2");
        const int v = e ? std::atoi(e) : 124;
        return v < 8 ? 124 : v;
    }();
    const int seg_rows = pick_seg_rows(w, h, Cfg::OUT, 148 * MINB, RY, target, 4);
    prm.pitch = a.pitch; prm.b = a.b; prm.seg_rows = seg_rows; prm.sc = a.sc;
    prm.push = a.push;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    dim3 grid((w + Cfg::OUT - 1) / Cfg::OUT, (h + seg_rows - 1) / seg_rows);
    kern<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, s>>>(mu, mv, mp, muo, mvo, mpo, prm);
    return cudaGetLastError();
}

template <int TX, int RY, int STAGES, int MINB>
cudaError_t launch_tma2(const Fused2Args &a, cudaStream_t s)
{
    return launch_tma2_generic<Tma2Cfg<TX, RY, STAGES>, RY, MINB>(a, s, fused2_tma_kernel<TX, RY, STAGES, MINB>);
}

template <int TX, int RY, int STAGES, int MINB, bool M>
cudaError_t launch_tma2s_m(const Fused2Args &a, cudaStream_t s)
{
    using Cfg = Tma2sCfg<TX, RY, STAGES, M>;
    CUtensorMap mu, mv, mp, muo, mvo, mpo, mm;
    if (!get_map(a.u, 8, a.pitch, a.ny, Cfg::BW, RY, mu) || !get_map(a.v, 8, a.pitch, a.ny, Cfg::BW, RY, mv) ||
        !get_map(a.p, 8, a.pitch, a.ny, Cfg::BW, RY, mp) || !get_map(a.uold, 8, a.pitch, a.ny, Cfg::BWO, RY, muo) ||
        !get_map(a.vold, 8, a.pitch, a.ny, Cfg::BWO, RY, mvo) || !get_map(a.pold, 8, a.pitch, a.ny, Cfg::BWO, RY, mpo))
        return cudaErrorNotSupported;
    if (M) {
        if (!get_map(a.tmask, 4, a.pitch, a.ny, Cfg::BWM, RY, mm)) return cudaErrorNotSupported;
    } else {
        mm = mp;  // unused by the unmasked kernel
    }
    Tma2Params prm;
    prm.u2 = a.u2; prm.v2 = a.v2; prm.p2 = a.p2;
    prm.uold2 = a.uold2; prm.vold2 = a.vold2; prm.pold2 = a.pold2;
    const int w = a.b.xe - a.b.xs + 1, h = a.b.ye - a.b.ys + 1;
    static const int target = [] {
        const char *e = std::getenv("DLESM_SEG_ROWS2");
        const int v = e ? std::atoi(e) : 124;
        return v < 8 ? 124 : v;
    }();
    const int seg_rows = pick_seg_rows(w, h, Cfg::OUT, 148 * MINB, RY, target, 4);
    prm.pitch = a.pitch; prm.b = a.b; prm.seg_rows = seg_rows; prm.sc = a.sc;
    prm.push = a.push;
    auto kern = fused2s_tma_kernel<TX, RY, STAGES, MINB, M>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    dim3 grid((w + Cfg::OUT - 1) / Cfg::OUT, (h + seg_rows - 1) / seg_rows);
    kern<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, s>>>(mu, mv, mp, muo, mvo, mpo, mm, prm);
    return cudaGetLastError();
}
template <int TX, int RY, int STAGES, int MINB>
cudaError_t launch_tma2s(const Fused2Args &a, cudaStream_t s)
{
    return a.tmask ? launch_tma2s_m<TX, RY, STAGES, MINB, true>(a, s)
                   : launch_tma2s_m<TX, RY, STAGES, MINB, false>(a, s);
}

template <int TX, int RY, int STAGES, int MINB>
cudaError_t launch_tma2w(const Fused2Args &a, cudaStream_t s)
{
    return launch_tma2_generic<Tma2wCfg<TX, RY, STAGES>, RY, MINB>(a, s, fused2w_tma_kernel<TX, RY, STAGES, MINB>);
}

struct Tma2Variant {
    int id;
    const char *name;
    cudaError_t (*launch)(const Fused2Args &, cudaStream_t);
};
// {threads = columns of stage A (TMA boxes are at most 256 wide: 128 only), rows per chunk,
// stages, resident CTAs per SM}
const Tma2Variant kTma2Variants[] = {
    {1, "tma2-128x4x3-occ2", launch_tma2<128, 4, 3, 2>},
    {2, "tma2-128x4x2-occ2", launch_tma2<128, 4, 2, 2>},
    {3, "tma2-128x2x3-occ3", launch_tma2<128, 2, 3, 3>},
    {4, "tma2-128x2x4-occ2", launch_tma2<128, 2, 4, 2>},
    {5, "tma2-128x2x2-occ3", launch_tma2<128, 2, 2, 3>},
    {6, "tma2-128x8x2-occ1", launch_tma2<128, 8, 2, 1>},
    {7, "tma2-128x1x4-occ3", launch_tma2<128, 1, 4, 3>},
    // ids >= 10: east / west intermediates through warp shuffles (118 output columns per CTA)
    {10, "tma2s-128x4x2-occ2", launch_tma2s<128, 4, 2, 2>},
    {11, "tma2s-128x4x3-occ2", launch_tma2s<128, 4, 3, 2>},
    {12, "tma2s-128x2x3-occ3", launch_tma2s<128, 2, 3, 3>},
    {13, "tma2s-128x2x4-occ2", launch_tma2s<128, 2, 4, 2>},
    {14, "tma2s-128x4x2-occ3", launch_tma2s<128, 4, 2, 3>},
    {15, "tma2s-256x4x2-occ1", launch_tma2s<256, 4, 2, 1>},
    {16, "tma2s-256x2x3-occ1", launch_tma2s<256, 2, 3, 1>},
    // ids >= 20: warp-private (all neighbour data by shuffle, no CTA barrier; 28 columns per warp)
    {20, "tma2w-128x4x3-occ2", launch_tma2w<128, 4, 3, 2>},
    {21, "tma2w-128x4x2-occ2", launch_tma2w<128, 4, 2, 2>},
    {22, "tma2w-128x2x4-occ2", launch_tma2w<128, 2, 4, 2>},
    {23, "tma2w-256x4x3-occ1", launch_tma2w<256, 4, 3, 1>},
    {24, "tma2w-128x8x2-occ2", launch_tma2w<128, 8, 2, 2>},
    {25, "tma2w-64x4x3-occ4", launch_tma2w<64, 4, 3, 4>},
    {26, "tma2w-96x4x3-occ3", launch_tma2w<96, 4, 3, 3>},
};
constexpr int kDefaultTma2Variant = 10;

}  // namespace

bool fused2_supported(const Fused2Args &a)
{
    if (a.tmask && a.pitch % 4 != 0) return false;   // the int32 tmask tile needs 16-byte row starts
    return a.pitch % 2 == 0 && encode_fn() != nullptr && a.b.xs >= 3 && a.b.ys >= 3 && a.b.xe + 2 <= a.pitch &&
           a.b.ye + 2 <= a.ny && !a.sc.first;
}

bool fused2_push_supported(int variant) { return variant == 0 || (variant >= 10 && variant < 20); }

cudaError_t sw_step2_fused(const Fused2Args &a, int variant, cudaStream_t s)
{
    if (a.b.xe < a.b.xs || a.b.ye < a.b.ys) return cudaSuccess;
    if (!fused2_supported(a)) return cudaErrorNotSupported;
    if ((a.push.n > 0 || a.tmask) && !fused2_push_supported(variant)) return cudaErrorNotSupported;
    if (variant == 0) variant = kDefaultTma2Variant;
    for (const auto &v : kTma2Variants)
        if (v.id == variant) {
            cudaError_t e = v.launch(a, s);
            note_launch();
            return e;
        }
    return cudaErrorInvalidValue;
}

namespace {
struct TmaVariant {
    int id;
    const char *name;
    cudaError_t (*launch)(const FusedArgs &, cudaStream_t);
};
// {columns, rows per chunk, stages, resident CTAs per SM}.  ids >= 20 start every strip on a
// 128-byte line of the row (aligned origin); ids < 20 start strips at xstart (kept for the
// comparison in DESIGN.md).
const TmaVariant kTmaVariants[] = {
    {2, "tma-128x4x2-occ3-unaligned", launch_tma<128, 4, 2, 3, false>},
    {5, "tma-128x4x3-occ2-unaligned", launch_tma<128, 4, 3, 2, false>},
    {22, "tma-128x4x2-occ3", launch_tma<128, 4, 2, 3, true>},
    {23, "tma-128x8x2-occ2", launch_tma<128, 8, 2, 2, true>},
    {24, "tma-128x4x3-occ3", launch_tma<128, 4, 3, 3, true>},
    {25, "tma-128x4x3-occ2", launch_tma<128, 4, 3, 2, true>},
    {26, "tma-128x4x4-occ2", launch_tma<128, 4, 4, 2, true>},
    {27, "tma-128x2x6-occ2", launch_tma<128, 2, 6, 2, true>},
    {28, "tma-128x8x3-occ1", launch_tma<128, 8, 3, 1, true>},
    {29, "tma-128x2x8-occ2", launch_tma<128, 2, 8, 2, true>},
    {30, "tma-128x2x4-occ3", launch_tma<128, 2, 4, 3, true>},
    {31, "tma-128x16x2-occ1", launch_tma<128, 16, 2, 1, true>},
    {32, "tma-128x8x4-occ1", launch_tma<128, 8, 4, 1, true>},
    {33, "tma-128x4x8-occ1", launch_tma<128, 4, 8, 1, true>},
    {34, "tma-128x12x3-occ1", launch_tma<128, 12, 3, 1, true>},
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
    const TmaVariant *v = find_tma(variant);
    return v ? v->name : "unknown";
}

bool fused_push_supported(const FusedArgs &a, int variant)
{
    const bool tma_ok = (a.pitch % 2 == 0) && (a.tmask == nullptr || a.pitch % 4 == 0) && encode_fn() != nullptr;
    return tma_ok && variant != 1;
}

cudaError_t sw_step_fused(const FusedArgs &a, int variant, cudaStream_t s)
{
    if (a.b.xe < a.b.xs || a.b.ye < a.b.ys) return cudaSuccess;
    if (a.push.n > 0 && !fused_push_supported(a, variant)) return cudaErrorNotSupported;
    // FP64 rows must start 16-byte aligned (even pitch); int32 tmask rows need pitch % 4 == 0
    const bool tma_ok = (a.pitch % 2 == 0) && (a.tmask == nullptr || a.pitch % 4 == 0) && encode_fn() != nullptr;
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
        const TmaVariant *v = find_tma(variant);
        if (!v) return cudaErrorInvalidValue;
        e = v->launch(a, s);
    }
    note_launch();
    return e;
}

}  // namespace k
}  // namespace dlesm
