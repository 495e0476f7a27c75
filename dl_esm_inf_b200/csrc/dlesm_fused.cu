// dlesm_fused.cu -- the whole shallow-water step as ONE sweep (96 B / point), and two steps as one.
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
// This file is compiled TWICE (dl_esm_inf_b200/build.py): with -fmad=false into namespace
// k::fma_off (one rounding per operation, bit-identical to the oracle) and with -fmad=true
// -DDLESM_CONTRACTED into k::fma_on (the compiler contracts a*b+c into DFMA: fewer FP64
// instructions, results within 1e-12 relative of the oracle).  dlesm_set_arith selects.
//
// Variants (dlesm_sw_step_fused's `variant`):
//   1     direct : raw rows come straight from global memory through L1 (any pitch, tmask ok)
//   5, 23, 25 tma : a producer warp streams row chunks of all six inputs (and the byte image of
//               the tmask) into a shared-memory ring with cp.async.bulk.tensor (TMA) + mbarrier
//               full/empty pairs; consumer warps march out of shared memory.  Needs a
//               16-byte-multiple pitch (DL_ESM_ALIGNMENT even).
//   0     auto   : tma when eligible, else direct.
#include <cuda.h>

#include <cstdint>
#include <cstdlib>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>
#include <type_traits>

#include "dlesm_kernels.cuh"

#ifdef DLESM_CONTRACTED
#define DLESM_FNS fma_on
#else
#define DLESM_FNS fma_off
#endif

namespace dlesm {
namespace k {
namespace DLESM_FNS {

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
    // cu, cv are carried DOUBLED (the 0.5 of cu = 0.5*(p+p)*u is folded into tdts8h, tdtsdxh,
    // tdtsdyh: an exact power-of-two rescaling, same bits; see SwScalars)
    double cuC_N = (N.pC + N.pW) * N.uC;
    double cuE_N = (N.pE + N.pC) * N.uE;
    double cvC_N = (N.pC + P.pC) * N.vC;
    double cvW_N = (N.pW + P.pW) * N.vW;
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
    o.unew = uo + sc.tdts8h * (zC_N + c.zC_P) * (cvC_N + cvW_N + c.cvW_P + c.cvC_P) -
             sc.tdtsdx * (hC_P - hW_P);
    o.vnew = vo - sc.tdts8h * (c.zE_P + c.zC_P) * (c.cuE_P + c.cuC_P + c.cuC_PP + c.cuE_PP) -
             sc.tdtsdy * (hC_P - c.hC_PP);
    o.pnew = po - sc.tdtsdxh * (c.cuE_P - c.cuC_P) - sc.tdtsdyh * (cvC_N - c.cvC_P);
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

// How blockIdx maps to columns and rows: the CUDA tile grid (SweepTiles; the analogue of the
// reference's OpenMP tiles, src/field_mod.f90:300-325).  x-tile `a` starts at x0 + a*tw + min(a, xextra)
// and is tw (+1 for the first xextra tiles: go_decompose's remainder rule, src/parallel_mod.f90:211-223)
// columns wide; it is cut into `strips` strips of the kernel's output width.  y-tile (segment) `by`
// starts at ys + by*th + min(by, yextra) and is th (+1) rows long; one CTA marches it.
struct TileMap {
    int strips;
    int x0, tw, xextra;
    int th, yextra;
};
__device__ __forceinline__ void tile_of_block(const TileMap &t, const Box &b, int bx, int by, int out_cols,
                                              int &i_first, int &i_last, int &j0, int &j1)
{
    const int a = bx / t.strips, sidx = bx - a * t.strips;
    const int tx0 = t.x0 + a * t.tw + min(a, t.xextra);
    i_first = tx0 + sidx * out_cols;
    i_last = min(tx0 + t.tw + (a < t.xextra ? 1 : 0) - 1, b.xe);
    j0 = b.ys + by * t.th + min(by, t.yextra);
    j1 = min(j0 + t.th + (by < t.yextra ? 1 : 0) - 1, b.ye);
}

// tile_list != null: a 1-D launch over a subset of the tile grid -- blockIdx.x indexes the list, an entry
// is by * list_gx + bx of the natural 2-D grid (masked runs: all-wet tiles go through the unmasked
// kernel, the tiles that see land through the masked one)
struct TmaStepParams {
    double *unew, *vnew, *pnew, *uold, *vold, *pold;
    int pitch;
    Box b;
    TileMap tm;
    SwScalars sc;
    PushPlan push;
    PushSync sync;
    const int *tile_list;
    int list_gx;
};
// bound of the neighbour wait in SM clock cycles (0 = unbounded); set_wait_limit, read on the slow path only
__device__ long long g_wait_limit = 0;
__device__ __forceinline__ bool wait_expired(long long t0)
{
    const long long lim = *reinterpret_cast<volatile long long *>(&g_wait_limit);
    return lim > 0 && clock64() - t0 > lim;
}
// start of an edge tile of a push sweep: wait for the neighbours' previous sweep (see PushSync); call
// before a CTA-wide barrier that precedes the first load
__device__ __forceinline__ void sync_wait_neighbours(const PushSync &sy, bool edge)
{
    if (edge && static_cast<int>(threadIdx.x) < sy.n_wait) {
        const long long t0 = clock64();
        for (;;) {
            unsigned long long v;
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(sy.local[threadIdx.x]) : "memory");
            if (v >= sy.wait_epoch) break;
            if (wait_expired(t0)) {   // DLESM_WAIT_TIMEOUT_S: the ranks are out of step -- fatal
                atomicExch(sy.err, 1);
                __threadfence_system();
                __trap();
            }
            __nanosleep(100);
        }
    }
}
// end of a push sweep, called by ONE thread per CTA after a barrier that follows every thread's
// __threadfence_system(): the CTA that draws the last ticket publishes this rank's epoch
__device__ __forceinline__ void sync_publish_if_last(const PushSync &sy)
{
    if (atomicAdd(sy.ticket, 1u) == gridDim.x * gridDim.y - 1) {
        __threadfence_system();
        for (int q = 0; q < sy.n; ++q)
            asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(sy.remote[q]), "l"(sy.epoch) : "memory");
        *sy.ticket = 0u;
    }
}
// diagnostics (dlesm_debug_cta_trace_begin/end): when set, thread 0 of every CTA of the TMA sweeps records
// its timeline -- slot 0 start, 1 end, 2 SM id, 3 after the neighbour wait (globaltimer, ns)
__device__ long long *g_cta_trace = nullptr;
__device__ __forceinline__ long long global_ns()
{
    long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// `tile`: by * (x-tiles of the natural grid) + bx -- the same entry whether the tile ran in the natural 2-D
// launch or in one of the two list launches of a split masked sweep
__device__ __forceinline__ void cta_trace(int slot, int tile)
{
    long long *tr = g_cta_trace;
    if (tr && threadIdx.x == 0) {
        long long *e = tr + 4 * static_cast<long long>(tile);
        e[slot] = global_ns();
        if (slot == 0) {
            unsigned sm;
            asm volatile("mov.u32 %0, %smid;" : "=r"(sm));
            e[2] = sm;
        }
    }
}
__device__ __forceinline__ void block_tile(const int *tile_list, int list_gx, int &bx, int &by)
{
    bx = blockIdx.x;
    by = blockIdx.y;
    if (tile_list) {
        const int t = tile_list[blockIdx.x];
        by = t / list_gx;
        bx = t - by * list_gx;
    }
}

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
constexpr int kUnalignedTmaVariant = 5;   // strips start at the x-tile's first column (tiles.nx > 1)
constexpr int align128(int v) { return (v + 127) & ~127; }
constexpr int align16(int v) { return (v + 15) & ~15; }

template <int TX, int RY, int STAGES, bool M, bool ALN>
struct TmaCfg {
    // TMA needs the first element of a box 16-byte aligned (even FP64 column; measured with
    // tools/tma_probe.cu: an odd start column raises "illegal instruction").  Every tile
    // therefore starts at the even column at or below the strip's west halo column and is
    // TX + 4 wide, which covers west halo + TX + east halo for either parity.
    // uold/vold/pold need no halo columns: with the 128-byte aligned strip origin (ALN) their
    // tiles are exactly the TX output columns (whole 32-byte sectors, nothing shared with the
    // neighbouring strip); otherwise they start at the even column at or below the first one.
    // The tmask rides the ring as its BYTE image (1 = wet; mask_to_bytes): the tile starts at the
    // multiple of 16 columns at or below the west halo column and covers west halo + TX + east halo
    // from any such start: 160 bytes per row for 128 columns (the int32 mask tile was 544).
    static constexpr int BW = TX + 4;
    static constexpr int BWO = ALN ? TX : TX + 2;
    static constexpr int BWM = align16(TX + 2 + 15);
    static constexpr int A = align128(BW * RY * 8);
    static constexpr int AO = align128(BWO * RY * 8);
    static constexpr int AM = M ? align128(BWM * RY) : 0;
    static constexpr int OFF_OLD = 3 * A;
    static constexpr int OFF_MASK = 3 * A + 3 * AO;
    static constexpr int STAGE_BYTES = 3 * A + 3 * AO + AM;
    static constexpr int TX_BYTES = 3 * BW * RY * 8 + 3 * BWO * RY * 8 + (M ? BWM * RY : 0);
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
    int i_first, i_last, j0, j1;   // first output column of this strip, last column of its x-tile, rows
    int bx_, by_;
    block_tile(prm.tile_list, prm.list_gx, bx_, by_);
    tile_of_block(prm.tm, prm.b, bx_, by_, TX, i_first, i_last, j0, j1);
    const int nsteps = (j1 - j0 + 1) + 2;            // raw rows j0-1 .. j1+1
    const int nchunks = (nsteps + RY - 1) / RY;

    if (tid == 0) {
        if (smem_u32(fused_smem) & 127u) __trap();
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], TX / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    cta_trace(0, by_ * prm.list_gx + bx_);
    // an edge tile reads halo cells the neighbours push and pushes into theirs: their previous sweep first
    sync_wait_neighbours(prm.sync, i_first - 2 < prm.b.xs || min(i_first + TX - 1, i_last) + 2 > prm.b.xe ||
                                       j0 - 2 < prm.b.ys || j1 + 2 > prm.b.ye);
    cta_trace(3, by_ * prm.list_gx + bx_);
    __syncthreads();

    if (warp == TX / 32) {
        // ---- producer warp: one lane feeds the ring ----
        if (lane == 0) {
            const int x_box = (i_first - 2) & ~1;   // even 0-based column at/below the west halo column
            const int x_box_m = (i_first - 2) & ~15;
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
    const bool active = i <= i_last && i >= prm.b.xs;
    const int col = tid + ((i_first - 2) & 1);        // west neighbour's column inside the tile
    const int colm = tid + ((i_first - 2) & 15);
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
    bool wet_prev = false;   // the previous chunk was all wet across this warp's columns and their neighbours
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
        const unsigned char *sm = st + Cfg::OFF_MASK + colm;
        // one pass of the chunk's rows; APPLY = false skips the mask selects (same bits when every
        // mask value the rows look at is wet)
        auto rows = [&](auto apply_tag) {
            constexpr bool APPLY = decltype(apply_tag)::value;
#pragma unroll
            for (int r = 0; r < RY; ++r) {
                Raw<M> N;
                N.pW = sp[r * BW]; N.pC = sp[r * BW + 1]; N.pE = sp[r * BW + 2];
                N.uW = su[r * BW]; N.uC = su[r * BW + 1]; N.uE = su[r * BW + 2];
                N.vW = sv[r * BW]; N.vC = sv[r * BW + 1]; N.vE = sv[r * BW + 2];
                if (M) {
                    if (APPLY) {
                        N.tW = sm[r * BWM] == 1; N.tC = sm[r * BWM + 1] == 1; N.tE = sm[r * BWM + 2] == 1;
                    } else {
                        N.tW = N.tC = N.tE = true;
                    }
                }
                const double uo = suo[r * BWO], vo = svo[r * BWO], po = spo[r * BWO];
                Outs out;
                ingest<M, APPLY>(cr, N, sc, uo, vo, po, out);
                if (r >= t_lo && r < t_hi) {
                    __stcs(prm.unew + qo, out.unew);
                    __stcs(prm.vnew + qo, out.vnew);
                    __stcs(prm.pnew + qo, out.pnew);
                    if (!FIRST) {
                        if (!APPLY || out.wu) __stcs(prm.uold + qo, out.uold);
                        if (!APPLY || out.wv) __stcs(prm.vold + qo, out.vold);
                        if (!APPLY || out.wt) __stcs(prm.pold + qo, out.pold);
                    }
                    if (pmask) push_row(prm.push, pmask, i, jrow, out.unew, out.vnew, out.pnew);
                }
                qo += pitch;
                ++jrow;
            }
        };
        if constexpr (M) {
            // all-wet test hoisted to the chunk (warp-uniform): a row looks back two rows, so the
            // unmasked march is taken when this chunk and the previous one are wet throughout
            bool w = true;
#pragma unroll
            for (int r = 0; r < RY; ++r) w = w && sm[r * BWM] == 1 && sm[r * BWM + 1] == 1 && sm[r * BWM + 2] == 1;
            const bool wet_now = __all_sync(0xffffffffu, w);
            if (wet_now && wet_prev) rows(std::false_type{});
            else rows(std::true_type{});
            wet_prev = wet_now;
        } else {
            rows(std::false_type{});
        }
        t_lo -= RY;
        t_hi -= RY;
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
    if (prm.sync.n > 0) {
        // the peer stores of the threads that pushed are ordered before the ticket (the other threads stored
        // into this rank's own memory only: a system fence in every thread of every CTA drains 19 k CTAs'
        // streaming stores one by one -- measured +0.7 ms per 16384^2 launch); the last CTA of a publishing
        // launch tells the neighbours that this rank's sweep is complete
        if (pmask) __threadfence_system();
        if (prm.sync.publish) {
            asm volatile("bar.sync 1, %0;" ::"r"(TX) : "memory");   // the consumer threads (the producer warp has left)
            if (tid == 0) sync_publish_if_last(prm.sync);
        }
    }
    cta_trace(1, by_ * prm.list_gx + bx_);
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

// 2-D tiled tensor map over a column-major (nx, ny) array of FP64 (elem_bytes 8) or bytes (1)
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
    CUtensorMap m;
    CUresult r = fn(&m, elem_bytes == 8 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 2,
                    const_cast<void *>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    if (g_maps.size() > 256) g_maps.clear();
    g_maps[key] = m;
    out = m;
    return true;
}

// opt a kernel into more than 48 KB of dynamic shared memory, once per kernel
cudaError_t allow_smem(const void *kern, int bytes)
{
    static std::mutex mu;
    static std::map<const void *, int> done;
    std::lock_guard<std::mutex> lock(mu);
    auto it = done.find(kern);
    if (it != done.end() && it->second >= bytes) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    // every sweep kernel asks for the SAME shared-memory carve-out (all of it): left to the driver, the masked
    // kernel (two CTAs per SM) and the unmasked one (three) got different carve-outs, an SM cannot hold both at
    // once, and the two launches of a split masked sweep ran one after the other -- the second started when the
    // first one's CTAs had all finished (CTA timelines, profiles/r03e_cta_trace_masked_4096_steady.jsonl)
    static const bool same_carveout = [] {
        const char *v = std::getenv("DLESM_SAME_CARVEOUT");   // 0: leave the carve-out to the driver (A/B runs)
        return !(v && std::atoi(v) == 0);
    }();
    if (e == cudaSuccess && same_carveout)
        e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e == cudaSuccess) done[kern] = bytes;
    return e;
}

// Pick the segment height.  Every segment costs `extra` more raw rows (its prologue) and its
// row count is rounded up to whole chunks of `ry` rows, so: segments of about `target` rows,
// the CTA count rounded up to whole waves of `slots` resident CTAs, and seg_rows + extra a
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

int env_rows(const char *name, int dflt)
{
    const char *e = std::getenv(name);
    const int v = e ? std::atoi(e) : dflt;
    return v < 8 ? dflt : v;
}

// The launch geometry: SweepTiles (the reference's tile-grid knob) -> TileMap + grid.
//   tiles.ny > 0 : that many segments along j, go_decompose's remainder rule;
//   tiles.ny = 0 : the library's choice (pick_seg_rows around `target` rows);
//   tiles.nx > 1 : that many x-tiles starting at b.xs, each cut into strips of `out_cols`;
//   else         : one x-tile starting at `x0` (the aligned origin).
void make_tile_map(const Box &b, SweepTiles tiles, int x0, int out_cols, int slots, int ry, int target, int extra,
                   TileMap &tm, dim3 &grid)
{
    const int h = b.ye - b.ys + 1;
    int ntx = tiles.nx > 1 ? tiles.nx : 1;
    if (ntx > 1) x0 = b.xs;
    const int w = b.xe - x0 + 1;
    if (ntx > w) ntx = w;
    tm.x0 = x0;
    tm.tw = w / ntx;
    tm.xextra = w % ntx;
    tm.strips = (tm.tw + (tm.xextra ? 1 : 0) + out_cols - 1) / out_cols;
    int nty;
    if (tiles.ny > 0) {
        nty = tiles.ny > h ? h : tiles.ny;
        tm.th = h / nty;
        tm.yextra = h % nty;
    } else {
        tm.th = pick_seg_rows(w, h, out_cols, slots, ry, target, extra);
        tm.yextra = 0;
        nty = (h + tm.th - 1) / tm.th;
    }
    grid = dim3(static_cast<unsigned>(tm.strips * ntx), static_cast<unsigned>(nty));
}

// ---------------------------------------------------------------------------------------
// Masked runs: which tiles of the launch grid see land at all?  A tile whose mask cells are all wet
// within `margin` cells of its outputs computes exactly what the unmasked march computes (every
// select of the masked rule passes), so it goes through the UNMASKED kernel -- full speed, small
// code -- and only the tiles near land run the masked kernel.  The classification is done once per
// (mask image, geometry) on the device and cached; the runtime calls forget_mask when it frees an image.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) classify_tiles_kernel(const unsigned char *__restrict__ m8, int pitch8, int nx,
                                                             int ny, TileMap tm, Box b, int out_cols, int margin,
                                                             unsigned char *__restrict__ all_wet)
{
    int i_first, i_last, j0, j1;
    tile_of_block(tm, b, blockIdx.x, blockIdx.y, out_cols, i_first, i_last, j0, j1);
    const int i_hi = min(i_first + out_cols - 1, i_last);
    const int ilo = max(1, i_first - margin), ihi = min(nx, i_hi + margin);
    const int jlo = max(1, j0 - margin), jhi = min(ny, j1 + margin);
    const int w = ihi - ilo + 1, h = jhi - jlo + 1;
    int wet = 1;
    for (int t = threadIdx.x; t < w * h && wet; t += blockDim.x) {
        const int jj = t / w, ii = t - jj * w;
        wet = m8[static_cast<size_t>(jlo - 1 + jj) * pitch8 + (ilo - 1 + ii)] == 1;
    }
    wet = __syncthreads_and(wet);
    if (threadIdx.x == 0) all_wet[blockIdx.y * gridDim.x + blockIdx.x] = static_cast<unsigned char>(wet);
}

struct TileLists {
    int *wet = nullptr, *land = nullptr;   // device arrays of tile ids in natural (row-major) order
    int nwet = 0, nland = 0;
    const void *mask = nullptr;
};
std::mutex &tile_mu()
{
    static std::mutex m;
    return m;
}
std::map<std::string, TileLists> &tile_cache()
{
    static std::map<std::string, TileLists> c;
    return c;
}
cudaError_t tile_lists(const unsigned char *m8, int pitch8, int nx, int ny, const TileMap &tm, const Box &b, int out_cols,
                       int margin, dim3 grid, cudaStream_t s, TileLists &out)
{
    struct Key {
        const void *m;
        int v[16];
    } key{m8, {pitch8, nx, ny, tm.strips, tm.x0, tm.tw, tm.xextra, tm.th, tm.yextra, b.xs, b.xe, b.ys, b.ye, out_cols,
               margin, static_cast<int>(grid.x)}};
    const std::string k(reinterpret_cast<const char *>(&key), sizeof(key));
    std::lock_guard<std::mutex> lock(tile_mu());
    auto it = tile_cache().find(k);
    if (it != tile_cache().end()) {
        out = it->second;
        return cudaSuccess;
    }
    const int n = static_cast<int>(grid.x * grid.y);
    unsigned char *flags = nullptr;
    cudaError_t e = cudaMalloc(&flags, n);
    if (e != cudaSuccess) return e;
    classify_tiles_kernel<<<grid, 256, 0, s>>>(m8, pitch8, nx, ny, tm, b, out_cols, margin, flags);
    std::vector<unsigned char> h(n);
    e = cudaMemcpyAsync(h.data(), flags, n, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(flags);
    if (e != cudaSuccess) return e;
    std::vector<int> wet, land;
    for (int t = 0; t < n; ++t) (h[t] ? wet : land).push_back(t);
    TileLists tl;
    tl.mask = m8;
    tl.nwet = static_cast<int>(wet.size());
    tl.nland = static_cast<int>(land.size());
    if (tl.nwet) {
        e = cudaMalloc(&tl.wet, wet.size() * sizeof(int));
        if (e == cudaSuccess) e = cudaMemcpyAsync(tl.wet, wet.data(), wet.size() * sizeof(int), cudaMemcpyHostToDevice, s);
    }
    if (e == cudaSuccess && tl.nland) {
        e = cudaMalloc(&tl.land, land.size() * sizeof(int));
        if (e == cudaSuccess) e = cudaMemcpyAsync(tl.land, land.data(), land.size() * sizeof(int), cudaMemcpyHostToDevice, s);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return e;
    tile_cache()[k] = tl;
    out = tl;
    return cudaSuccess;
}
// The two launches of a split masked sweep touch disjoint tiles: the land-tile launch runs on a side
// stream between two events so that it fills the wet-tile launch's last partial wave instead of queueing
// behind it (4096^2: 2.6 + 0.4 waves).
struct SideStream {
    cudaStream_t s = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
};
cudaError_t side_stream(SideStream &out)
{
    static SideStream ss;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    if (!ss.s) {
        cudaError_t e = cudaStreamCreateWithFlags(&ss.s, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming);
        if (e != cudaSuccess) return e;
    }
    out = ss;
    return cudaSuccess;
}

bool split_masked_launches()
{
    static const bool on = [] {
        const char *e = std::getenv("DLESM_MASK_SPLIT");
        return !(e && std::atoi(e) == 0);
    }();
    return on;
}

template <int TX, int RY, int STAGES, int MINB, bool ALN>
void tma_geometry(const FusedArgs &a, TileMap &tm, dim3 &grid)
{
    // aligned origin: strips start on 128-byte lines of the row (1-based column 1 + 16k), the
    // columns below xs are masked; otherwise strips start at xs.
    // Short segments keep the strips of a band in lockstep, so that together they sweep whole
    // contiguous rows of DRAM and re-read each other's halo rows out of L2 (measured at
    // 16384^2: 256 rows 0.93, 128 rows 0.95, 48 rows 0.995, 24-32 rows 1.01-1.02 of the
    // measured copy bandwidth).  DLESM_SEG_ROWS overrides for experiments.
    static const int target = env_rows("DLESM_SEG_ROWS", 32);
    const int x0 = ALN ? ((a.b.xs - 1) / 16) * 16 + 1 : a.b.xs;
    make_tile_map(a.b, a.tiles, x0, TX, 148 * MINB, RY, target, 2, tm, grid);
}

// one instantiation of the one-step kernel over `grid` (2-D natural grid, or 1-D over prm.tile_list)
template <int TX, int RY, int STAGES, int MINB, bool M, bool FIRST, bool ALN>
cudaError_t launch_tma_kernel(const FusedArgs &a, const TmaStepParams &prm, dim3 grid, cudaStream_t s)
{
    using Cfg = TmaCfg<TX, RY, STAGES, M, ALN>;
    CUtensorMap mu, mv, mp, muo, mvo, mpo, mm;
    if (!get_map(a.u, 8, a.pitch, a.ny, Cfg::BW, RY, mu) || !get_map(a.v, 8, a.pitch, a.ny, Cfg::BW, RY, mv) ||
        !get_map(a.p, 8, a.pitch, a.ny, Cfg::BW, RY, mp) || !get_map(a.uold, 8, a.pitch, a.ny, Cfg::BWO, RY, muo) ||
        !get_map(a.vold, 8, a.pitch, a.ny, Cfg::BWO, RY, mvo) || !get_map(a.pold, 8, a.pitch, a.ny, Cfg::BWO, RY, mpo))
        return cudaErrorNotSupported;
    if (M) {
        if (!a.mask8 || a.mask8_pitch % 16 != 0 ||
            !get_map(a.mask8, 1, a.mask8_pitch, a.ny, Cfg::BWM, RY, mm))
            return cudaErrorNotSupported;
    } else {
        mm = mp;  // unused by the unmasked kernel
    }
    auto kern = fused_tma_kernel<TX, RY, STAGES, MINB, M, FIRST, ALN>;
    cudaError_t e = allow_smem(reinterpret_cast<const void *>(kern), Cfg::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    kern<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, s>>>(mu, mv, mp, muo, mvo, mpo, mm, prm);
    return cudaGetLastError();
}

template <int TX, int RY, int STAGES, int MINB, bool M, bool FIRST, bool ALN>
cudaError_t launch_tma_mf(const FusedArgs &a, cudaStream_t s)
{
    TmaStepParams prm;
    prm.unew = a.unew; prm.vnew = a.vnew; prm.pnew = a.pnew;
    prm.uold = a.uold; prm.vold = a.vold; prm.pold = a.pold;
    dim3 grid;
    tma_geometry<TX, RY, STAGES, MINB, ALN>(a, prm.tm, grid);
    prm.pitch = a.pitch; prm.b = a.b; prm.sc = a.sc;
    prm.push = a.push;
    prm.sync = a.sync;
    prm.tile_list = nullptr;
    prm.list_gx = static_cast<int>(grid.x);
    // (with remote neighbours the two launches of a split run would both have to take tickets: the masked
    // kernel alone runs there)
    if (M && split_masked_launches() && a.mask8 && a.mask8_pitch % 16 == 0 && a.sync.n == 0 && a.sync.n_wait == 0) {
        // all-wet tiles through the unmasked kernel, the rest through the masked one (same bits)
        TileLists tl;
        cudaError_t e = tile_lists(a.mask8, a.mask8_pitch, a.pitch, a.ny, prm.tm, a.b, TX, 3, grid, s, tl);
        if (e != cudaSuccess) return e;
        SideStream ss;
        const bool both = tl.nwet && tl.nland;
        if (both) {
            e = side_stream(ss);
            if (e == cudaSuccess) e = cudaEventRecord(ss.fork, s);
            if (e == cudaSuccess) e = cudaStreamWaitEvent(ss.s, ss.fork, 0);
            if (e != cudaSuccess) return e;
        }
        // the land tiles are the slow ones (masked march, fewer CTAs per SM): they go FIRST, on the launching
        // stream, so that they start at once; the wet-tile launch follows on the side stream and fills the slots
        // around them.  (The other way round the wet launch took every slot and the land tiles ran as a tail of
        // their own: 4096^2 config 3, 380 us per two-step launch against 305 unmasked.)
        if (tl.nland) {
            prm.tile_list = tl.land;
            e = launch_tma_kernel<TX, RY, STAGES, MINB, true, FIRST, ALN>(a, prm, dim3(tl.nland), s);
            if (e != cudaSuccess) return e;
        }
        if (tl.nwet) {
            prm.tile_list = tl.wet;
            e = launch_tma_kernel<TX, RY, STAGES, MINB, false, FIRST, ALN>(a, prm, dim3(tl.nwet), both ? ss.s : s);
            if (e != cudaSuccess) return e;
        }
        if (both) {
            e = cudaEventRecord(ss.join, ss.s);
            if (e == cudaSuccess) e = cudaStreamWaitEvent(s, ss.join, 0);
        }
        return e;
    }
    return launch_tma_kernel<TX, RY, STAGES, MINB, M, FIRST, ALN>(a, prm, grid, s);
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
template <int TX, int RY, int STAGES, int MINB, bool ALN>
void grid_tma(const FusedArgs &a, int grid[2])
{
    TileMap tm;
    dim3 g;
    tma_geometry<TX, RY, STAGES, MINB, ALN>(a, tm, g);
    grid[0] = static_cast<int>(g.x);
    grid[1] = static_cast<int>(g.y);
}

// ---------------------------------------------------------------------------------------
// Two model steps per sweep (temporal blocking).  Stage A is the march above, run on a strip
// one column wider on each side and one row longer at each end, producing the intermediate
// state (u1,v1,p1 and the once-filtered old fields) of step n+1; stage B -- the same march
// again, fed by u1,v1,p1 of the west / own / east column -- produces step n+2 two rows behind.
// The intermediate state never reaches memory: 6 reads + 6 writes per point for TWO steps.
//
// Every lane computes the four intermediates (cu, cv, z, h) of its own column only and takes the
// neighbours' through warp shuffles (the first form of this sweep, where every thread recomputed
// its east / west copies, spent a third of its FP64 work on them: ncu 75 % FP64 pipe).  Warps
// overlap by two columns (lane 0 and lane 31 only feed their neighbours), so a warp yields 30
// columns: 4 warps = 120 columns of the intermediate state, 118 output columns per CTA.  The
// expression trees are unchanged -- lane l+1's cu(own) IS lane l's cu(east), operand for
// operand -- so the bits are the same.
// ---------------------------------------------------------------------------------------
struct Tma2Params {
    double *u2, *v2, *p2, *uold2, *vold2, *pold2;
    int pitch;
    Box b;
    TileMap tm;
    SwScalars sc;
    PushPlan2 push;
    PushSync sync;
    const int *tile_list;   // see TmaStepParams
    int list_gx;
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
    double cuC_N = (N.pC + N.pW) * N.uC;   // doubled, like ingest()
    double cvC_N = (N.pC + P.pC) * N.vC;
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
    o.unew = uo + sc.tdts8h * (zC_N + c.zC_P) * (cvC_N + cvW_N + c.cvW_P + c.cvC_P) -
             sc.tdtsdx * (hC_P - hW_P);
    o.vnew = vo - sc.tdts8h * (c.zE_P + c.zC_P) * (c.cuE_P + c.cuC_P + c.cuC_PP + c.cuE_PP) -
             sc.tdtsdy * (hC_P - c.hC_PP);
    o.pnew = po - sc.tdtsdxh * (c.cuE_P - c.cuC_P) - sc.tdtsdyh * (cvC_N - c.cvC_P);
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
    static constexpr int BWM = align16(VA + 2 + 15);   // byte tmask tile: lanes' columns from a multiple of 16
    static constexpr int A = align128(BW * RY * 8);
    static constexpr int AO = align128(BWO * RY * 8);
    static constexpr int AM = M ? align128(BWM * RY) : 0;
    static constexpr int OFF_OLD = 3 * A;
    static constexpr int OFF_MASK = 3 * A + 3 * AO;
    static constexpr int STAGE_BYTES = 3 * A + 3 * AO + AM;
    static constexpr int TX_BYTES = 3 * BW * RY * 8 + 3 * BWO * RY * 8 + (M ? BWM * RY : 0);
    static constexpr int XW = VA + 4;    // exchange row: index of column x = x - i_first + 3
    static constexpr int XBUF_BYTES = 3 * RY * XW * 8;
    static constexpr int OFF_X = STAGES * STAGE_BYTES;
    static constexpr int BAR_OFFSET = OFF_X + 2 * XBUF_BYTES;
    static constexpr int SMEM_BYTES = BAR_OFFSET + STAGES * 8;
    static constexpr int THREADS = TX;
};

// PIPE: software-pipelined form.  Between two CTA barriers a warp runs stage B of the PREVIOUS
// chunk and stage A of the current one -- two independent instruction streams (more ILP for the
// fixed-latency FP64 chains, and the wait for the TMA data of chunk c hides behind B(c-1)) --
// instead of A(c) | barrier | B(c).  Same exchange buffers (A(c+1) overwrites what B(c-1) read one
// barrier earlier), same arithmetic, same bits.
template <int TX, int RY, int STAGES, int MINB, bool M, bool PIPE>
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
    int i_first, i_last, j0, j1;   // first output column of this CTA, last column of its x-tile, rows
    int bx_, by_;
    block_tile(prm.tile_list, prm.list_gx, bx_, by_);
    tile_of_block(prm.tm, prm.b, bx_, by_, Cfg::OUT, i_first, i_last, j0, j1);
    const int nsteps = (j1 - j0 + 1) + 4;            // raw rows j0-2 .. j1+2
    const int nchunks = (nsteps + RY - 1) / RY;
    // lane column (both stages): i_first + 30*warp + lane - 2
    const int x_box = (i_first - 4) & ~1;   // even 0-based column at/below stage A lane 0's west neighbour
    const int x_old = (i_first - 3) & ~1;   // even 0-based column at/below stage A lane 0's own column
    const int x_msk = (i_first - 3) & ~15;  // the byte tmask tile starts on a multiple of 16 columns

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
    // an edge tile reads halo cells the neighbours push and pushes into theirs: their previous sweep first
    cta_trace(0, by_ * prm.list_gx + bx_);
    if (prm.sync.n_wait > 0) {
        sync_wait_neighbours(prm.sync, i_first - 3 < prm.b.xs || min(i_first + Cfg::OUT - 1, i_last) + 3 > prm.b.xe ||
                                           j0 - 3 < prm.b.ys || j1 + 3 > prm.b.ye);
        __syncthreads();
    }
    cta_trace(3, by_ * prm.list_gx + bx_);
    // No producer warp here: the per-chunk barrier between the two stages already tells thread 0
    // that every warp is done with a ring stage, so it refills the stage right after it.  (A
    // fifth warp would also cap the kernel at 168 registers: 10 warps on 4 schedulers.)
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
    const bool active = a_writes && i >= i_first && i < i_first + Cfg::OUT && i <= i_last;
    const int xo = base + 1;                          // exchange index of this lane's column
    const int col = base + ((i_first - 4) & 1);       // west neighbour inside the raw tile
    const int colo = base + ((i_first - 3) & 1);      // own column inside the old tile
    const int colm = base + ((i_first - 3) & 15);     // own column inside the tmask tile
    const SwScalars sc = prm.sc;
    CarryS ca, cb;
    carry_init(ca);
    carry_init(cb);
    double uo1 = 0.0, vo1 = 0.0, po1 = 0.0;
    bool wet_prev = false;    // the previous chunk was all wet across this warp's columns
    long long qo = (static_cast<long long>(j0) - 5) * prm.pitch + (i - 1);
    const long long pitch = prm.pitch;
    int t_lo = active ? 4 : (1 << 30), t_hi = nsteps;
    // push boxes this lane's column and this CTA's rows can meet (0 for almost every thread).  With a staging
    // strip the narrow boxes (x edges, corners) are deferred to the end of the CTA: `pmask` keeps the wide
    // ones (north / south rows), `dmask` (CTA-uniform) the deferred ones this CTA's columns and rows meet.
    unsigned pmask = 0;
    // deferred boxes this CTA's columns and rows meet (CTA-uniform; recomputed in the epilogue rather than
    // carried through the march: the masked form has no register to spare)
    auto deferred_boxes = [&]() {
        unsigned d = 0;
        if (prm.push.strip) {
            const int i_end = min(i_first + Cfg::OUT - 1, i_last);   // last output column of this CTA
            for (int q = 0; q < prm.push.n; ++q) {
                const PushSeg2 &g = prm.push.seg[q];
                if (g.i1 - g.i0 < 2 && g.j1 >= j0 && g.j0 <= j1 && g.i1 >= i_first && g.i0 <= i_end) d |= 1u << q;
            }
        }
        return d;
    };
    if (active)
        for (int q = 0; q < prm.push.n; ++q) {
            const PushSeg2 &g = prm.push.seg[q];
            if (!(prm.push.strip && g.i1 - g.i0 < 2) && i >= g.i0 && i <= g.i1 && g.j1 >= j0 && g.j0 <= j1)
                pmask |= 1u << q;
        }
    const int pfirst = pmask ? __ffs(pmask) - 1 : 0;
    const unsigned prest = pmask & (pmask - 1);
    // this lane's slot in the staging strip (doubles): the two columns at either x edge of the tile
    int qs = -1;
    if (active && deferred_boxes()) {
        const int h = prm.b.ye - prm.b.ys + 1;
        if (i - prm.b.xs < 2) qs = (j0 - prm.b.ys) * 12 + (i - prm.b.xs);
        else if (prm.b.xe - i < 2) qs = (h + (j0 - prm.b.ys)) * 12 + (i - (prm.b.xe - 1));
    }
    int jrow = j0 - 4;

    // state handed from stage A of a chunk to stage B of the same chunk
    double uo_r[RY], vo_r[RY], po_r[RY];
    bool tbC[RY], tbW[RY];   // stage B's row is stage A's previous row: its masks are A's carried ones
    bool fastB = false;

    // ---- stage A of chunk c: RY raw rows -> RY rows of the intermediate state ----
    auto stage_a = [&](int c) {
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
        const unsigned char *sm = st + Cfg::OFF_MASK + colm;
        auto rows = [&](auto apply_tag) {
            constexpr bool APPLY = decltype(apply_tag)::value;
#pragma unroll
            for (int r = 0; r < RY; ++r) {
                RawS N;
                N.pW = sp[r * BW]; N.pC = sp[r * BW + 1];
                N.uC = su[r * BW + 1]; N.uE = su[r * BW + 2];
                N.vW = sv[r * BW]; N.vC = sv[r * BW + 1];
                Outs o;
                if (APPLY) {
                    const bool tC = sm[r * BWM] == 1;
                    const bool tW = __shfl_up_sync(0xffffffffu, static_cast<int>(tC), 1) != 0;
                    tbC[r] = ca.tC_P; tbW[r] = ca.tW_P;
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
        };
        if constexpr (M) {
            // all-wet test hoisted to the chunk (warp-uniform).  Stage A's rows look back two raw
            // rows, stage B's three: both take the unmasked march -- same bits, no selects --
            // when this chunk and the previous one are wet throughout.
            bool w = true;
#pragma unroll
            for (int r = 0; r < RY; ++r) w = w && sm[r * BWM] == 1;
            const bool wet_now = __all_sync(0xffffffffu, w);
            fastB = wet_now && wet_prev;
            wet_prev = wet_now;
            if (fastB) {
                rows(std::false_type{});
                ca.tC_P = ca.tW_P = ca.tC_PP = true;
            } else {
                rows(std::true_type{});
            }
        } else {
            rows(std::false_type{});
        }
    };

    // ---- stage B of chunk c: those RY rows (west / own / east column) -> RY rows of step n+2 ----
    auto stage_b = [&](int c) {
        const double *xb = reinterpret_cast<const double *>(fused_smem + Cfg::OFF_X + (c & 1) * Cfg::XBUF_BYTES) + xo;
        auto rows = [&](auto apply_tag) {
            constexpr bool APPLY = decltype(apply_tag)::value;
#pragma unroll
            for (int r = 0; r < RY; ++r) {
                RawS N;
                N.uC = xb[(0 * RY + r) * XW]; N.uE = xb[(0 * RY + r) * XW + 1];
                N.vW = xb[(1 * RY + r) * XW - 1]; N.vC = xb[(1 * RY + r) * XW];
                N.pW = xb[(2 * RY + r) * XW - 1]; N.pC = xb[(2 * RY + r) * XW];
                Outs o;
                if (APPLY) {
                    ingest_sh<true>(cb, N, tbC[r], tbW[r], sc, uo1, vo1, po1, o);
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
                    if (qs >= 0) {
                        // x-edge lane: outputs of this row into the local staging strip (this lane stores
                        // rows j0.. in order: r - t_lo counts them within the chunk)
                        double *e = prm.push.strip + qs;
                        e[0] = o.unew; e[2] = o.vnew; e[4] = o.pnew;
                        e[6] = o.uold; e[8] = o.vold; e[10] = o.pold;
                        qs += 12;
                    }
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
        };
        if constexpr (M) {
            if (fastB) {
                rows(std::false_type{});
                cb.tC_P = cb.tW_P = cb.tC_PP = true;
            } else {
                rows(std::true_type{});
            }
        } else {
            rows(std::false_type{});
        }
        t_lo -= RY;
        t_hi -= RY;
    };

    if (!PIPE) {
        for (int c = 0; c < nchunks; ++c) {
            stage_a(c);
            __syncthreads();
            if (tid == 0 && c + STAGES < nchunks) issue(c + STAGES);
            stage_b(c);
        }
    } else {
        stage_a(0);
        __syncthreads();
        if (tid == 0 && STAGES < nchunks) issue(STAGES);
        for (int c = 1; c < nchunks; ++c) {
            stage_b(c - 1);   // reads the exchange rows A(c-1) wrote before the last barrier
            stage_a(c);       // overwrites the rows B(c-2) read before the last barrier
            __syncthreads();
            if (tid == 0 && c + STAGES < nchunks) issue(c + STAGES);
        }
        stage_b(nchunks - 1);
    }
    const unsigned dmask = deferred_boxes();
    if (dmask) {
        // deferred boxes: every thread of the CTA copies its share of the staged rows to the neighbours
        __syncthreads();
        const double *strip = prm.push.strip;
        const long long h = prm.b.ye - prm.b.ys + 1;
        const int i_end = min(i_first + Cfg::OUT - 1, i_last);
        for (unsigned m = dmask; m; m &= m - 1) {
            const PushSeg2 &g = prm.push.seg[__ffs(m) - 1];
            const int c0 = max(g.i0, i_first), c1 = min(g.i1, i_end);
            const int r0 = max(g.j0, j0), r1 = min(g.j1, j1);
            const int nc = c1 - c0 + 1, total = (r1 - r0 + 1) * 6 * nc;
            for (int e = tid; e < total; e += TX) {
                const int c = e % nc, f = (e / nc) % 6, r = e / (6 * nc);
                const int col = c0 + c, row = r0 + r;
                const long long srow = (col - prm.b.xs < 2) ? (row - prm.b.ys) : (h + (row - prm.b.ys));
                const int cell = (col - prm.b.xs < 2) ? (col - prm.b.xs) : (col - (prm.b.xe - 1));
                const double v = __ldcg(strip + srow * 12 + f * 2 + cell);
                if (g.dpitch == 0)   // packed: the neighbour's slot for this direction, [row][field][column]
                    g.dst[0][(static_cast<long long>(row - g.j0) * 6 + f) * (g.i1 - g.i0 + 1) + (col - g.i0)] = v;
                else
                    g.dst[f][idx(g.dpitch, g.ides + (col - g.i0), g.jdes + (row - g.j0))] = v;
            }
        }
    }
    if (prm.sync.n > 0) {
        if (pmask || dmask) __threadfence_system();   // only the threads that pushed (see the one-step sweep)
        if (prm.sync.publish) {
            __syncthreads();
            if (tid == 0) sync_publish_if_last(prm.sync);
        }
    }
    cta_trace(1, by_ * prm.list_gx + bx_);
}

// ---------------------------------------------------------------------------------------
// Two steps per sweep, warp-private form.  ncu on the shuffle form showed 23 % of all
// warp samples waiting at the per-chunk CTA barrier of the shared-memory exchange.  Here the
// intermediate state crosses lanes by shuffle as well, so a warp depends on no other warp:
// stage A row -> shuffle -> stage B row, interleaved row by row, no barrier, no exchange
// buffer.  Lanes 2..29 store (28 columns per warp); a producer warp feeds the TMA ring as in the
// one-step sweep.  No tmask, no halo push.
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
                   const __grid_constant__ CUtensorMap, const __grid_constant__ Tma2Params prm)
{
    using Cfg = Tma2wCfg<TX, RY, STAGES>;
    constexpr int BW = Cfg::BW, BWO = Cfg::BWO;
    extern __shared__ __align__(1024) unsigned char fused_smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(fused_smem + Cfg::BAR_OFFSET);
    uint64_t *empty = full + STAGES;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    int i_first, i_last, j0, j1;
    int bx_, by_;
    block_tile(prm.tile_list, prm.list_gx, bx_, by_);
    tile_of_block(prm.tm, prm.b, bx_, by_, Cfg::OUT, i_first, i_last, j0, j1);
    const int nsteps = (j1 - j0 + 1) + 4;            // raw rows j0-2 .. j1+2
    const int nchunks = (nsteps + RY - 1) / RY;

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
    const bool active = lane >= 2 && lane <= 29 && i <= i_last;
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

// every segment recomputes two rows of the intermediate state and re-reads four raw rows, so
// segments are longer than the one-step sweep's; DLESM_SEG_ROWS2 overrides
// CTAs of `kern` one SM holds (registers and shared memory decide; __launch_bounds__' MINB is only a
// floor): the wave size the segment height is rounded to.  A 4096^2 tile is 2.7 waves of 444 CTAs:
// sized for 296 slots it ran 4 waves of 98-row segments instead of 3 of 108 (+20 % time).
int resident_ctas(const void *kern, int threads, int smem, int floor_per_sm)
{
    static std::mutex mu;
    static std::map<const void *, int> cache;
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find(kern);
    if (it != cache.end()) return it->second;
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem) != cudaSuccess || per_sm < 1) {
        cudaGetLastError();
        per_sm = floor_per_sm;
    }
    int sms = 148;
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cache[kern] = per_sm * sms;
    return cache[kern];
}

template <class Cfg, int RY, int MINB>
void tma2_geometry(const Fused2Args &a, TileMap &tm, dim3 &grid, int slots = 148 * MINB)
{
    static const int target = env_rows("DLESM_SEG_ROWS2", 124);
    make_tile_map(a.b, a.tiles, a.b.xs, Cfg::OUT, slots, RY, target, 4, tm, grid);
}

// one kernel of the two-step family over `grid` (2-D natural grid, or 1-D over prm.tile_list)
template <class Cfg, int RY, bool M, class Kern>
cudaError_t launch_tma2_kernel(const Fused2Args &a, const Tma2Params &prm, dim3 grid, cudaStream_t s, Kern kern)
{
    CUtensorMap mu, mv, mp, muo, mvo, mpo, mm;
    if (!get_map(a.u, 8, a.pitch, a.ny, Cfg::BW, RY, mu) || !get_map(a.v, 8, a.pitch, a.ny, Cfg::BW, RY, mv) ||
        !get_map(a.p, 8, a.pitch, a.ny, Cfg::BW, RY, mp) || !get_map(a.uold, 8, a.pitch, a.ny, Cfg::BWO, RY, muo) ||
        !get_map(a.vold, 8, a.pitch, a.ny, Cfg::BWO, RY, mvo) || !get_map(a.pold, 8, a.pitch, a.ny, Cfg::BWO, RY, mpo))
        return cudaErrorNotSupported;
    if constexpr (M) {
        if (!a.mask8 || a.mask8_pitch % 16 != 0 || !get_map(a.mask8, 1, a.mask8_pitch, a.ny, Cfg::BWM, RY, mm))
            return cudaErrorNotSupported;
    } else {
        mm = mp;  // unused by the unmasked kernels
    }
    kern<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, s>>>(mu, mv, mp, muo, mvo, mpo, mm, prm);
    return cudaGetLastError();
}

template <class Cfg, int RY, int MINB, bool M, class Kern>
cudaError_t launch_tma2_generic(const Fused2Args &a, cudaStream_t s, Kern kern)
{
    Tma2Params prm;
    prm.u2 = a.u2; prm.v2 = a.v2; prm.p2 = a.p2;
    prm.uold2 = a.uold2; prm.vold2 = a.vold2; prm.pold2 = a.pold2;
    // per KERNEL, not per instantiation of this launcher: kernels that differ only in a template
    // argument of their own (PIPE) share this function
    cudaError_t e = allow_smem(reinterpret_cast<const void *>(kern), Cfg::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    dim3 grid;
    tma2_geometry<Cfg, RY, MINB>(a, prm.tm, grid,
                                 resident_ctas(reinterpret_cast<const void *>(kern), Cfg::THREADS, Cfg::SMEM_BYTES, MINB));
    prm.pitch = a.pitch; prm.b = a.b; prm.sc = a.sc;
    prm.push = a.push;
    prm.sync = a.sync;
    prm.tile_list = nullptr;
    prm.list_gx = static_cast<int>(grid.x);
    return launch_tma2_kernel<Cfg, RY, M>(a, prm, grid, s, kern);
}

// masked two-step sweep: all-wet tiles through the unmasked kernel `kern_wet`, tiles that see land through
// the masked kernel `kern_land` (same geometry, same bits)
template <class CfgWet, class CfgLand, int RY, int MINB, class KernWet, class KernLand>
cudaError_t launch_tma2_split(const Fused2Args &a, cudaStream_t s, KernWet kern_wet, KernLand kern_land)
{
    Tma2Params prm;
    prm.u2 = a.u2; prm.v2 = a.v2; prm.p2 = a.p2;
    prm.uold2 = a.uold2; prm.vold2 = a.vold2; prm.pold2 = a.pold2;
    cudaError_t e = allow_smem(reinterpret_cast<const void *>(kern_wet), CfgWet::SMEM_BYTES);
    if (e == cudaSuccess) e = allow_smem(reinterpret_cast<const void *>(kern_land), CfgLand::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    dim3 grid;
    const int slots = resident_ctas(reinterpret_cast<const void *>(kern_wet), CfgWet::THREADS, CfgWet::SMEM_BYTES, MINB);
    tma2_geometry<CfgWet, RY, MINB>(a, prm.tm, grid, slots);
    prm.pitch = a.pitch; prm.b = a.b; prm.sc = a.sc;
    prm.push = a.push;
    prm.sync = a.sync;       // n == 0 here: see the caller
    prm.list_gx = static_cast<int>(grid.x);
    TileLists tl;
    e = tile_lists(a.mask8, a.mask8_pitch, a.pitch, a.ny, prm.tm, a.b, CfgWet::OUT, 4, grid, s, tl);
    if (e != cudaSuccess) return e;
    SideStream ss;
    const bool both = tl.nwet && tl.nland;
    if (both) {
        e = side_stream(ss);
        if (e == cudaSuccess) e = cudaEventRecord(ss.fork, s);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(ss.s, ss.fork, 0);
        if (e != cudaSuccess) return e;
    }
    // land tiles first, on the launching stream; the wet tiles on the side stream (see launch_tma_mf)
    if (tl.nland) {
        prm.tile_list = tl.land;
        e = launch_tma2_kernel<CfgLand, RY, true>(a, prm, dim3(tl.nland), s, kern_land);
        if (e != cudaSuccess) return e;
    }
    if (tl.nwet) {
        prm.tile_list = tl.wet;
        e = launch_tma2_kernel<CfgWet, RY, false>(a, prm, dim3(tl.nwet), both ? ss.s : s, kern_wet);
        if (e != cudaSuccess) return e;
    }
    if (both) {
        e = cudaEventRecord(ss.join, ss.s);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(s, ss.join, 0);
    }
    return e;
}

template <int TX, int RY, int STAGES, int MINB, bool PIPE>
cudaError_t launch_tma2s(const Fused2Args &a, cudaStream_t s)
{
    // (with remote neighbours the two launches of a split run would both have to take tickets: the masked
    // kernel alone runs there)
    if (a.mask8 && a.mask8_pitch % 16 == 0 && split_masked_launches() && a.sync.n == 0 && a.sync.n_wait == 0)
        return launch_tma2_split<Tma2sCfg<TX, RY, STAGES, false>, Tma2sCfg<TX, RY, STAGES, true>, RY, MINB>(
            a, s, fused2s_tma_kernel<TX, RY, STAGES, MINB, false, PIPE>,
            fused2s_tma_kernel<TX, RY, STAGES, MINB, true, PIPE>);
    if (a.mask8)
        return launch_tma2_generic<Tma2sCfg<TX, RY, STAGES, true>, RY, MINB, true>(
            a, s, fused2s_tma_kernel<TX, RY, STAGES, MINB, true, PIPE>);
    return launch_tma2_generic<Tma2sCfg<TX, RY, STAGES, false>, RY, MINB, false>(
        a, s, fused2s_tma_kernel<TX, RY, STAGES, MINB, false, PIPE>);
}
template <int TX, int RY, int STAGES, int MINB, bool PIPE>
void grid_tma2s(const Fused2Args &a, int grid[2])
{
    TileMap tm;
    dim3 g;
    using Cfg = Tma2sCfg<TX, RY, STAGES, false>;
    const void *kern = reinterpret_cast<const void *>(fused2s_tma_kernel<TX, RY, STAGES, MINB, false, PIPE>);
    allow_smem(kern, Cfg::SMEM_BYTES);
    tma2_geometry<Cfg, RY, MINB>(a, tm, g, resident_ctas(kern, Cfg::THREADS, Cfg::SMEM_BYTES, MINB));
    grid[0] = static_cast<int>(g.x);
    grid[1] = static_cast<int>(g.y);
}

template <int TX, int RY, int STAGES, int MINB>
cudaError_t launch_tma2w(const Fused2Args &a, cudaStream_t s)
{
    return launch_tma2_generic<Tma2wCfg<TX, RY, STAGES>, RY, MINB, false>(a, s, fused2w_tma_kernel<TX, RY, STAGES, MINB>);
}
template <int TX, int RY, int STAGES, int MINB>
void grid_tma2w(const Fused2Args &a, int grid[2])
{
    TileMap tm;
    dim3 g;
    tma2_geometry<Tma2wCfg<TX, RY, STAGES>, RY, MINB>(a, tm, g);
    grid[0] = static_cast<int>(g.x);
    grid[1] = static_cast<int>(g.y);
}

struct Tma2Variant {
    int id;
    const char *name;
    cudaError_t (*launch)(const Fused2Args &, cudaStream_t);
    void (*grid)(const Fused2Args &, int[2]);
};
// {threads = columns of stage A (TMA boxes are at most 256 wide: 128 only), rows per chunk, stages,
// resident CTAs per SM}.  Shipped + one alternate per idea (the other shapes measured in round 1 --
// RY 2 / 8, 3-4 stages, 256 threads -- were within 1 % or slower: DESIGN.md section 4b):
//   10       shuffle form, A | barrier | B per chunk (round 1's shipped kernel)
//   17       shuffle form, software-pipelined: B(c-1) and A(c) between two barriers
//   20       warp-private (all neighbour data by shuffle, no CTA barrier; 28 columns per warp)
const Tma2Variant kTma2Variants[] = {
    {10, "tma2s-128x4x2-occ2", launch_tma2s<128, 4, 2, 2, false>, grid_tma2s<128, 4, 2, 2, false>},
    {17, "tma2s-128x4x2-occ2-pipelined", launch_tma2s<128, 4, 2, 2, true>, grid_tma2s<128, 4, 2, 2, true>},
    {20, "tma2w-128x4x3-occ2", launch_tma2w<128, 4, 3, 2>, grid_tma2w<128, 4, 3, 2>},
};
int default_tma2_variant()
{
    static const int v = [] {
        const char *e = std::getenv("DLESM_VARIANT2");
        const int x = e ? std::atoi(e) : 0;
        for (const auto &t : kTma2Variants)
            if (t.id == x) return x;
        return 10;
    }();
    return v;
}

}  // namespace

bool fused2_supported(const Fused2Args &a)
{
    return a.pitch % 2 == 0 && encode_fn() != nullptr && a.b.xs >= 3 && a.b.ys >= 3 && a.b.xe + 2 <= a.pitch &&
           a.b.ye + 2 <= a.ny && !a.sc.first && (a.mask8 == nullptr || a.mask8_pitch % 16 == 0);
}

bool fused2_push_supported(int variant) { return variant == 0 || (variant >= 10 && variant < 20); }

cudaError_t set_cta_trace(long long *buf) { return cudaMemcpyToSymbol(g_cta_trace, &buf, sizeof(buf)); }
cudaError_t set_wait_limit(long long cycles) { return cudaMemcpyToSymbol(g_wait_limit, &cycles, sizeof(cycles)); }

cudaError_t sw_step2_fused(const Fused2Args &a, int variant, cudaStream_t s)
{
    if (a.b.xe < a.b.xs || a.b.ye < a.b.ys) return cudaSuccess;
    if (!fused2_supported(a)) return cudaErrorNotSupported;
    if (variant == 0) variant = default_tma2_variant();
    if ((a.push.n > 0 || a.mask8) && !fused2_push_supported(variant)) return cudaErrorNotSupported;
    for (const auto &v : kTma2Variants)
        if (v.id == variant) {
            cudaError_t e = v.launch(a, s);
            note_launch();
            return e;
        }
    return cudaErrorInvalidValue;
}

// the runtime frees (or rebuilds) a byte mask image: drop the tile classifications made for it
void forget_mask(const unsigned char *m8)
{
    std::lock_guard<std::mutex> lock(tile_mu());
    auto &c = tile_cache();
    for (auto it = c.begin(); it != c.end();) {
        if (it->second.mask == m8) {
            cudaFree(it->second.wet);
            cudaFree(it->second.land);
            it = c.erase(it);
        } else {
            ++it;
        }
    }
}

void fused2_grid(const Fused2Args &a, int variant, int grid[2])
{
    grid[0] = grid[1] = 0;
    if (variant == 0) variant = default_tma2_variant();
    for (const auto &v : kTma2Variants)
        if (v.id == variant) v.grid(a, grid);
}

namespace {
struct TmaVariant {
    int id;
    const char *name;
    cudaError_t (*launch)(const FusedArgs &, cudaStream_t);
    void (*grid)(const FusedArgs &, int[2]);
};
// {columns, rows per chunk, stages, resident CTAs per SM}.  25 is the shipped kernel (strips on
// 128-byte lines of the row), 23 the alternate chunk shape, 5 the same kernel as 25 with strips
// starting at the first column of their x-tile (used when the tile-grid knob asks for x-tiles).
const TmaVariant kTmaVariants[] = {
    {5, "tma-128x4x3-occ2-unaligned", launch_tma<128, 4, 3, 2, false>, grid_tma<128, 4, 3, 2, false>},
    {23, "tma-128x8x2-occ2", launch_tma<128, 8, 2, 2, true>, grid_tma<128, 8, 2, 2, true>},
    {25, "tma-128x4x3-occ2", launch_tma<128, 4, 3, 2, true>, grid_tma<128, 4, 3, 2, true>},
};
const TmaVariant *find_tma(int id)
{
    for (const auto &v : kTmaVariants)
        if (v.id == id) return &v;
    return nullptr;
}
bool tma_eligible(const FusedArgs &a)
{
    // FP64 rows must start 16-byte aligned (even pitch); a masked run needs the byte image
    return (a.pitch % 2 == 0) && (a.tmask == nullptr || (a.mask8 != nullptr && a.mask8_pitch % 16 == 0)) &&
           encode_fn() != nullptr;
}
int resolve_variant(const FusedArgs &a, int variant)
{
    if (variant == 0) variant = tma_eligible(a) ? kDefaultTmaVariant : 1;
    // strips of an aligned-origin kernel must start on 128-byte lines; x-tiles start anywhere
    if (variant >= 20 && a.tiles.nx > 1) variant = kUnalignedTmaVariant;
    return variant;
}
}  // namespace

const char *fused_variant_name(int variant)
{
    if (variant == 0) return "auto";
    if (variant == 1) return "direct-128x64";
    const TmaVariant *v = find_tma(variant);
    return v ? v->name : "unknown";
}

bool fused_push_supported(const FusedArgs &a, int variant) { return tma_eligible(a) && variant != 1; }

cudaError_t sw_step_fused(const FusedArgs &a, int variant, cudaStream_t s)
{
    if (a.b.xe < a.b.xs || a.b.ye < a.b.ys) return cudaSuccess;
    if (a.push.n > 0 && !fused_push_supported(a, variant)) return cudaErrorNotSupported;
    const bool tma_ok = tma_eligible(a);
    variant = resolve_variant(a, variant);
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

void fused_grid(const FusedArgs &a, int variant, int grid[2])
{
    variant = resolve_variant(a, variant);
    if (variant == 1) {
        grid[0] = (a.b.xe - a.b.xs + 1 + 127) / 128;
        grid[1] = (a.b.ye - a.b.ys + 1 + 63) / 64;
        return;
    }
    grid[0] = grid[1] = 0;
    const TmaVariant *v = find_tma(variant);
    if (v) v->grid(a, grid);
}

}  // namespace DLESM_FNS
}  // namespace k
}  // namespace dlesm
