// dlesm_kernels.cuh -- launchers of the hand-written sm_100a kernels (FP64, HBM-bound).
// All fields are column-major (i fastest) with pitch = grid%nx; boxes are 1-based inclusive.
#pragma once

#include <cuda_runtime.h>

#include <cstddef>

namespace dlesm {
namespace k {

struct Box {
    int xs, xe, ys, ye;
};

constexpr int kMaxExchangeFields = 8;
constexpr int kMaxSegments = 16;

// one rectangular patch of every field <-> one contiguous slice of a message buffer
struct Segment {
    int i0, j0;          // 1-based start in the field
    int nx, ny;          // extent
    long long offset;    // element offset of this segment's first field in the buffer
};
struct PackPlan {
    int nseg;
    int nfields;
    int pitch;
    Segment seg[kMaxSegments];
    double *fields[kMaxExchangeFields];
};

struct SwScalars {
    double fsdx, fsdy;       // 4/dx, 4/dy
    double tdts8, tdtsdx, tdtsdy;
    double alpha;
    // the same constants halved (exact): the fused sweeps carry 2*cu, 2*cv -- they drop the 0.5 of
    // cu = 0.5*(p+p)*u and fold it into the constants cu, cv are multiplied by.  Scaling by a power
    // of two commutes with rounding, so the results keep their bits.
    double tdts8h, tdtsdxh, tdtsdyh;
    int first;               // 1 => no time filter
};
inline SwScalars make_scalars(double dx, double dy, double tdt, double alpha, int first)
{
    SwScalars sc;
    sc.fsdx = 4.0 / dx; sc.fsdy = 4.0 / dy;
    sc.tdts8 = tdt / 8.0; sc.tdtsdx = tdt / dx; sc.tdtsdy = tdt / dy;
    sc.tdts8h = 0.5 * sc.tdts8; sc.tdtsdxh = 0.5 * sc.tdtsdx; sc.tdtsdyh = 0.5 * sc.tdtsdy;
    sc.alpha = alpha; sc.first = first;
    return sc;
}

// ---- field_mod operations ----
cudaError_t copy_field(const double *in, double *out, size_t n, cudaStream_t s);
cudaError_t set_field(double *f, size_t n, double val, cudaStream_t s);
cudaError_t copy_patch(double *f, int pitch, Box src, Box dst, double *scratch, cudaStream_t s);
// partials: >= kChecksumBlocks doubles of scratch; ticket: one zeroed unsigned (left zero again);
// result: 1 double (device); host_result: null, or two mapped pinned words -- the sum, then `seq` (written
// after it, system fence between: a host that sees seq sees the sum).  One launch.
constexpr int kChecksumBlocks = 1184;  // 148 SMs x 8
// `gs` non-null (several ranks): the global sum rides in the same launch -- the folding block stores this
// rank's sum and the call's sequence number into every rank's cell for this rank (peer memory, two buffers
// alternating with the sequence), waits for every rank's cell here, adds them in rank order (the same bits
// on every rank) and publishes THAT to the host.  No collective library call, no second launch.
constexpr int kGsumMaxRanks = 64;
struct GlobalSumPlan {
    int nranks, me;
    unsigned long long seq;        // this call's number, equal on every rank
    int *err;
    // cell of rank `me` in rank r's table, and this rank's own table: cells of {double value; u64 seq}
    double *peer_cell[kGsumMaxRanks];
    const double *own_cells;
};
cudaError_t checksum(const double *f, int pitch, Box b, double *partials, unsigned *ticket, double *result,
                     double *host_result, unsigned long long seq, cudaStream_t s, const GlobalSumPlan *gs = nullptr);
cudaError_t periodic_halos(double *const *fields, int nfields, int pitch, Box internal,
                           bool periodic_x, bool periodic_y, cudaStream_t s);
cudaError_t pack(const PackPlan &plan, double *buffer, cudaStream_t s);
cudaError_t unpack(const PackPlan &plan, const double *buffer, cudaStream_t s);
cudaError_t broadcast_x(double *out, int nx, int ny, const double *xvec, cudaStream_t s);
cudaError_t broadcast_y(double *out, int nx, int ny, const double *yvec, cudaStream_t s);
cudaError_t scatter_global(double *f, int pitch, Box box, const double *global, int gnx, int dx,
                           int dy, cudaStream_t s);
cudaError_t gather_box(const double *f, int pitch, Box box, double *packed, cudaStream_t s);

// Direct halo exchange (exchange_generic over NVLink peer memory instead of pack -> NCCL -> unpack).
// Every rank owns two receive buffers (used alternately per neighbour pair) divided into one slot
// per direction tag, mapped by all its peers.  Launch 1 packs every send patch of every field
// STRAIGHT INTO THE NEIGHBOUR'S slot (contiguous remote stores), and the block that finishes last
// publishes the pair's exchange count into each neighbour's flag array (release, system scope).
// Launch 2 waits (acquire) for the neighbours' counts and unpacks this rank's own slots.
struct XSegment {
    int i0, j0, nx, ny;      // patch in the local fields
    double *buf;             // send: the neighbour's slot (peer memory); receive: this rank's slot
};
struct XPlan {
    int nseg, nfields, pitch;
    XSegment seg[kMaxSegments];
    double *fields[kMaxExchangeFields];
    int npeers;
    unsigned long long *flag[kMaxSegments];     // send: neighbour's flag cell for this rank; receive: own cell for the neighbour
    unsigned long long count[kMaxSegments];     // the pair's exchange count this launch publishes / waits for
};
// send + publish + wait + unpack in one co-resident launch (the default); the two-launch form follows
cudaError_t xdirect_exchange(const XPlan &send, const XPlan &recv, unsigned *ticket, int *err, cudaStream_t s);
cudaError_t xdirect_send(const XPlan &plan, unsigned *ticket, cudaStream_t s);
cudaError_t xdirect_recv(const XPlan &plan, int *err, cudaStream_t s);

// Tile-set exchange (several sub-domains per rank): a batch of rectangular copies, each between two
// arrays of its own pitch -- tile to tile, tile to send buffer, receive buffer to tile.  Pointers
// address the first element of the patch.  One launch per batch, blockIdx.y = copy.
struct TileOp {
    const double *src;
    double *dst;
    int spitch, dpitch, nx, ny;
};
constexpr int kTileOpsPerLaunch = 96;
struct TileOpBatch {
    TileOp op[kTileOpsPerLaunch];
    int n;
};
cudaError_t tile_ops(const TileOpBatch &b, cudaStream_t s);

// ---- shallow-water sweeps, un-fused (one launch per reference-style kernel) ----
cudaError_t compute_cu(double *cu, const double *p, const double *u, int pitch, Box b,
                       const int *tmask, cudaStream_t s);
cudaError_t compute_cv(double *cv, const double *p, const double *v, int pitch, Box b,
                       const int *tmask, cudaStream_t s);
cudaError_t compute_z(double *z, const double *p, const double *u, const double *v, int pitch,
                      Box b, double fsdx, double fsdy, const int *tmask, cudaStream_t s);
cudaError_t compute_h(double *h, const double *p, const double *u, const double *v, int pitch,
                      Box b, const int *tmask, cudaStream_t s);
cudaError_t compute_unew(double *unew, const double *uold, const double *z, const double *cv,
                         const double *h, int pitch, Box b, double tdts8, double tdtsdx,
                         const int *tmask, const double *ucur, cudaStream_t s);
cudaError_t compute_vnew(double *vnew, const double *vold, const double *z, const double *cu,
                         const double *h, int pitch, Box b, double tdts8, double tdtsdy,
                         const int *tmask, const double *vcur, cudaStream_t s);
cudaError_t compute_pnew(double *pnew, const double *pold, const double *cu, const double *cv,
                         int pitch, Box b, double tdtsdx, double tdtsdy, const int *tmask,
                         const double *pcur, cudaStream_t s);
cudaError_t time_smooth(const double *f, const double *fnew, double *fold, int pitch, Box b,
                        double alpha, const int *tmask, int points, cudaStream_t s);

// ---- halo push: the fused sweep stores its edge outputs straight into the neighbours' halo
// cells (peer-mapped device memory over NVLink, or this rank's own array for a wrap-around
// neighbour that is itself).  One entry per send-list entry of map_comms
// (src/parallel_comms_mod.f90:70-83): source box in this rank's array -> (ides, jdes) in the
// destination rank's array.
constexpr int kMaxPush = 16;
struct PushSeg {
    int i0, i1, j0, j1;   // source box, 1-based inclusive
    int ides, jdes;       // where (i0, j0) lands in the destination array
    int dpitch;           // destination pitch
    int pad_;
    double *dst[3];       // destination unew, vnew, pnew
};
struct PushPlan {
    int n;
    int pad_;
    PushSeg seg[kMaxPush];
};
// signal every remote neighbour that this rank's pushes of `epoch` are complete, then wait for
// theirs.  remote[q]: this rank's slot in neighbour q's flag array; local[q]: neighbour q's
// slot in this rank's flag array.  err (device int) is set if the wait times out.
struct PeerFlags {
    int n;
    unsigned long long *remote[kMaxPush];
    const unsigned long long *local[kMaxPush];
};
cudaError_t peer_signal_wait(const PeerFlags &pf, unsigned long long epoch, int *err, cudaStream_t s,
                             bool publish = true);
// The receiving end of the two-step sweep's packed x-edge / corner strips (PushPlan2): after the
// hand-shake (every CTA waits for the neighbours' epoch itself) the rows a neighbour left in this rank's
// slot -- [row][field 0..5][column] doubles -- go into the halo columns of the six output arrays.
struct StripScatter {
    int n;
    int pitch;
    struct Item {
        int ides, jdes, nx, ny;
        const double *slot;
    } item[8];
    double *dst[6];
};
cudaError_t peer_wait_scatter(const PeerFlags &pf, unsigned long long epoch, int *err, bool publish,
                              const StripScatter &sp, cudaStream_t s);

// The same hand-shake folded into a push sweep (no launch of its own).  Wait half: CTAs whose tile
// reads halo cells or pushes (the edge tiles) wait at their start until each of the `n_wait` neighbours of
// the PREVIOUS push sweep has published `wait_epoch` -- that sweep is complete there, so its pushes into
// this rank's halos have landed (RAW) and it no longer reads the buffer this sweep pushes into (WAR);
// interior tiles start at once and absorb the skew between ranks.  Publish half: when the last CTA of a
// launch with `publish` set finishes (ticket), it stores `epoch` into the flag cells of this sweep's `n`
// neighbours.  n_wait = n = 0: no remote neighbour, nothing happens.
struct PushSync {
    int n_wait;
    int n;
    int publish;
    const unsigned long long *local[kMaxPush];   // n_wait cells of this rank's flag array
    unsigned long long *remote[kMaxPush];        // n cells, one in each neighbour's flag array
    unsigned long long wait_epoch, epoch;
    unsigned *ticket;
    int *err;
};

// ---- the fused step ----
// Tile grid of a sweep: the CUDA analogue of the reference's OpenMP tile grid (GOCEAN_OMP_GRID="AxB",
// r2d_field(..., do_tile), src/field_mod.f90:300-325,1473-1503).  nx tiles across the swept box, each
// cut into strips of the kernel's width; ny tiles along j, one CTA marching each.  0 = the library's
// own choice (one x-tile on 128-byte lines; segment height picked for whole waves of CTAs).
struct SweepTiles {
    int nx, ny;
};
struct FusedArgs {
    const double *u, *v, *p;
    double *uold, *vold, *pold;
    double *unew, *vnew, *pnew;
    const int *tmask;            // int32 (nx, ny) as grid_init builds it; may be null (all wet)
    const unsigned char *mask8;  // byte image of tmask (1 = wet) with row pitch `mask8_pitch`, for the TMA ring
    int mask8_pitch;
    int pitch, ny;
    Box b;              // internal region (identical for U,V,T,F under SW periodic)
    SwScalars sc;
    SweepTiles tiles;
    PushPlan push;      // push.n == 0: halos of the new fields are left to the caller
    PushSync sync;      // sync.n == 0: no in-kernel hand-shake
};

// ---- two model steps in one sweep (temporal blocking): 96 B / point per TWO steps ----
// Needs a tile with halo width >= 2: u,v,p valid to depth 2 around `b`, uold,vold,pold valid to
// depth 1.  The intermediate state (step n+1) is recomputed on the depth-1 ring and never
// stored.  Outputs (step n+2) go to separate arrays: u2,v2,p2 and the twice-filtered
// uold2,vold2,pold2 (not in place: neighbouring CTAs still read the old fields' ring).
// Not the first (unfiltered) step.
// halo push of the two-step sweep: like PushSeg, six destinations (u2,v2,p2,uold2,vold2,pold2)
struct PushSeg2 {
    int i0, i1, j0, j1;
    int ides, jdes;
    int dpitch;
    int pad_;
    double *dst[6];
};
// `strip` non-null: boxes at most two columns wide (the east / west edges and the corners) are DEFERRED --
// the lanes on the tile's two x-edge column pairs store their outputs into this LOCAL staging strip
// ([side 0 = west, 1 = east][row - b.ys][field 0..5][cell 0..1] doubles) on every row, and when a CTA
// has finished its segment all its threads copy the rows of those boxes to the neighbours in one burst.
// (CTA timelines with remote x neighbours: the per-row peer stores made an edge CTA of the two-step sweep
// 3.5x longer and slowed the CTAs sharing its SM; profiles/r02q_cta_trace_n4.jsonl.)
struct PushPlan2 {
    int n;
    int pad_;
    double *strip;
    PushSeg2 seg[kMaxPush];
};
struct Fused2Args {
    const double *u, *v, *p, *uold, *vold, *pold;
    double *u2, *v2, *p2, *uold2, *vold2, *pold2;
    const unsigned char *mask8;  // may be null; byte image of the PERIODIC mask two cells deep around b
    int mask8_pitch;
    int pitch, ny;
    Box b;
    SwScalars sc;
    SweepTiles tiles;
    PushPlan2 push;     // push.n == 0: halos of the outputs are left to the caller
    PushSync sync;      // sync.n == 0: no in-kernel hand-shake
};

// The fused sweeps exist twice: namespace fma_off is compiled with -fmad=false (one rounding per
// operation: bit-identical to the oracle and to the reference's FMA-free Fortran), fma_on is the
// same source compiled with -fmad=true (contracted arithmetic: fewer FP64 instructions, results
// within 1e-12 relative of the oracle instead of bit-equal).  dlesm_set_arith selects.
#define DLESM_FUSED_API(NS)                                                                          \
    namespace NS {                                                                                   \
    /* whether sw_step_fused can push halos for these arguments (TMA variants only) */               \
    bool fused_push_supported(const FusedArgs &a, int variant);                                      \
    cudaError_t sw_step_fused(const FusedArgs &a, int variant, cudaStream_t s);                      \
    const char *fused_variant_name(int variant);                                                     \
    bool fused2_supported(const Fused2Args &a);                                                      \
    /* whether `variant` of the two-step sweep can push halos / take a mask (the shuffle form) */    \
    bool fused2_push_supported(int variant);                                                         \
    cudaError_t sw_step2_fused(const Fused2Args &a, int variant, cudaStream_t s);                    \
    /* launch geometry the sweep would use (grid.x, grid.y) -- for the tiling knob's tests */        \
    void fused_grid(const FusedArgs &a, int variant, int grid[2]);                                   \
    void fused2_grid(const Fused2Args &a, int variant, int grid[2]);                                 \
    /* a byte mask image is about to be freed: drop what was cached for it */                        \
    void forget_mask(const unsigned char *m8);                                                       \
    /* diagnostics: CTA (by * gridDim.x + bx) of the TMA sweeps records globaltimer at its start,    \
       after the neighbour wait and at its end, and its SM id: 4 long long each; null = off */       \
    cudaError_t set_cta_trace(long long *buf);                                                       \
    /* bound of the in-kernel neighbour wait, SM clock cycles; 0 = unbounded */                      \
    cudaError_t set_wait_limit(long long cycles);                                                    \
    }
DLESM_FUSED_API(fma_off)
DLESM_FUSED_API(fma_on)
#undef DLESM_FUSED_API

// periodic halo fill to `depth` cells on one rank (wrapped source index; corners included)
cudaError_t periodic_halos_deep(double *const *fields, int nfields, int pitch, Box internal, int depth,
                                cudaStream_t s);
// east / west halo columns of rows box.ys..box.ye only (box.xs..box.xe = the internal columns)
cudaError_t periodic_x_rows(double *const *fields, int nfields, int pitch, Box rows, int depth, cudaStream_t s);

// m[t] -> d[t] (to_double) or d[t] -> m[t]: the tmask rides the FP64 halo machinery once
cudaError_t mask_convert(int *m, double *d, size_t n, bool to_double, cudaStream_t s);
// byte image of an int32 tmask for the TMA ring: m8[(j-1)*pitch8 + (i-1)] = (tmask(i,j) == 1)
cudaError_t mask_to_bytes(const int *m, int pitch, int ny, unsigned char *m8, int pitch8, cudaStream_t s);

// classic initial conditions, synthetic (device sin/cos)
cudaError_t sw_init(double *u, double *v, double *p, int pitch, Box b, int gx0, int gy0, int M,
                    int N, double dx, double dy, cudaStream_t s);

// bound of every in-kernel wait for another rank (exchange, global sum, push hand-shake), SM clock cycles;
// 0 = unbounded.  Once per device.
cudaError_t set_wait_limit(long long cycles);

long long launch_count();
void note_launch(int n = 1);

}  // namespace k
}  // namespace dlesm
