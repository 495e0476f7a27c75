// dlesm_runtime.cu -- runtime context (device, stream, scratch, NCCL communicator, field
// registry) and the C ABI of include/dlesm_b200.h.  There is no CPU fallback anywhere in
// this file: a device entry point either runs its CUDA kernel or fails with DLESM_ERR_CUDA.
#include <cuda.h>
#include <cuda_runtime.h>
#include <nccl.h>

#include <sys/mman.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <atomic>
#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/dlesm_b200.h"
#include "dlesm_host.hpp"
#include "dlesm_internal.hpp"
#include "dlesm_kernels.cuh"

using namespace dlesm;

// =========================================================================================
// context
// =========================================================================================
namespace {

struct FieldRec {
    int nx, ny;
    // dirty-flag halos (halo_type%needs_update, src/halo_mod.f90:10-25): the depth to which this
    // array's halos are known to be up to date; 0 = out of date.  Every entry point that writes the
    // array resets it, halo_exchange / the periodic fill / a pushing sweep set it.
    int clean_depth = 0;
};

struct PeerGroup {
    int peer0;                 // 0-based rank; == my rank for self messages
    std::vector<int> sends;    // indices into comms lists, ordered by direction tag
    std::vector<int> recvs;
    long long send_off = 0, recv_off = 0;  // element offsets for ONE field
    long long send_n = 0, recv_n = 0;
};

// one set of map_comms tables with the per-peer message plan derived from it
struct CommTables {
    dlesm_comms t{};
    std::vector<PeerGroup> groups;
    long long send_total = 0, recv_total = 0;  // per field
    int depth = 1;                             // halo depth the tables were built for
    bool valid = false;
};

struct Context {
    std::mutex mu;
    bool device_ready = false;
    int device = -1;
    cudaStream_t stream = nullptr;
    cudaStream_t s_h2d = nullptr, s_d2h = nullptr;  // copy streams of the banded host step
    std::vector<cudaEvent_t> band_events;           // reused pool (2 per band + 1)
    double *ring_host = nullptr;                    // pinned staging of the halo ring
    size_t ring_host_n = 0;
    double *partials = nullptr;   // checksum partials + result
    double *result = nullptr;
    double *host_result = nullptr;  // pinned (and, under UVA, writable by kernels: the checksum lands here)
    unsigned *ticket = nullptr;     // last-block ticket of the one-launch checksum
    std::atomic<unsigned long long> checksum_seq{0}, gsum_seq{0};   // sequence numbers the host polls for
    double *scratch = nullptr;      // patch scratch
    size_t scratch_n = 0;
    double *sendbuf = nullptr, *recvbuf = nullptr;
    size_t xbuf_n = 0;
    std::unordered_map<const void *, FieldRec> registry;
    // communication
    bool comm_ready = false;
    ncclComm_t comm = nullptr;
    CommTables installed;   // dlesm_dev_set_comms: the module-level tables of parallel_comms_mod
    // halo push: peer-mapped neighbours (dlesm_dev_peer_register)
    struct PeerView {
        std::vector<double *> by_rank;  // this field's counterpart on rank r (0-based), or null
        std::vector<int> pitch_by_rank; // its pitch there (tile widths may differ by one)
        std::vector<std::string> handles;   // IPC handles this view holds a reference on
    };
    std::unordered_map<const void *, PeerView> peers;
    unsigned long long *flags = nullptr;              // [kMaxRanks] epochs published by neighbours
    std::vector<unsigned long long *> peer_flags;     // neighbours' flag arrays, by rank
    std::vector<void *> flag_mappings;                // flag-array mappings, closed at finalise
    unsigned long long push_epoch = 0;
    int *push_err = nullptr;                          // device flag: a neighbour wait timed out
    unsigned *push_ticket = nullptr;                  // CTA ticket of a push sweep that publishes its own epoch
    double *push_strip = nullptr;                     // local staging strip of the two-step sweep's deferred x-edge push
    size_t push_strip_cap = 0;
    // a push sweep whose hand-shake is deferred published `push_epoch` itself; nobody has waited for the
    // neighbours of THAT sweep yet.  The next push sweep's edge tiles do (PushSync), or push_settle().
    bool push_pending = false;
    k::PeerFlags pending_pf{};
    // direct halo exchange over peer memory (DLESM_EXCHANGE_DIRECT): one allocation per rank --
    // [flag cells | ticket | err | receive buffer 0 | receive buffer 1] -- mapped by every rank
    struct XDirect {
        bool ready = false;
        char *base = nullptr;
        size_t slot_cap = 0;                          // elements per direction slot
        // second region of the same allocation: 2 x 8 direction slots for the packed x-edge / corner
        // strips of the two-step halo push (k::PushPlan2), buffers alternating with the push epoch
        size_t pslot_cap = 0;
        std::vector<char *> peer_base;                // the same allocation on rank r (own for r = me)
        std::vector<unsigned long long> count;        // exchanges done with rank r (equal on both ends)
        std::vector<void *> mappings;
    } xd;
    int exchange_mode = -1;                           // -1: from DLESM_EXCHANGE at first use
    // arithmetic of the fused sweeps (dlesm_set_arith) and the tile-grid knob (GOCEAN_OMP_GRID)
    int arith = DLESM_ARITH_EXACT;
    k::SweepTiles sweep_tiles{0, 0};
    bool tiling_initialised = false;   // `tiling_initialised` of src/field_mod.f90:236
    // byte images of int32 tmasks for the TMA ring, keyed by the mask's device pointer
    struct Mask8 {
        unsigned char *p = nullptr;
        int pitch = 0;
    };
    std::unordered_map<const void *, Mask8> mask8;
    // profiling of the dominant kernel
    bool profile = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
    size_t prof_used = 0;
};

Context &ctx()
{
    static Context c;
    return c;
}

int fail_cuda(cudaError_t e, const char *what)
{
    set_error(DLESM_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
    return DLESM_ERR_CUDA;
}
int fail_nccl(ncclResult_t r, const char *what)
{
    set_error(DLESM_ERR_NCCL, std::string(what) + ": " + ncclGetErrorString(r));
    return DLESM_ERR_NCCL;
}
#define CU(call)                                               \
    do {                                                       \
        cudaError_t e_ = (call);                               \
        if (e_ != cudaSuccess) return fail_cuda(e_, #call);    \
    } while (0)
#define NC(call)                                               \
    do {                                                       \
        ncclResult_t r_ = (call);                              \
        if (r_ != ncclSuccess) return fail_nccl(r_, #call);    \
    } while (0)

int ensure_device()
{
    Context &c = ctx();
    if (c.device_ready) return DLESM_OK;
    return dlesm_device_init(-1);
}

int lookup(const void *p, FieldRec &out)
{
    Context &c = ctx();
    std::lock_guard<std::mutex> lock(c.mu);
    auto it = c.registry.find(p);
    if (it == c.registry.end()) {
        set_error(DLESM_ERR_UNKNOWN_PTR, "device pointer was not allocated by dlesm_dev_field_alloc");
        return DLESM_ERR_UNKNOWN_PTR;
    }
    out = it->second;
    return DLESM_OK;
}

// dirty-flag halos: see FieldRec::clean_depth
void mark_written(const void *p)
{
    Context &c = ctx();
    std::lock_guard<std::mutex> lock(c.mu);
    auto it = c.registry.find(p);
    if (it != c.registry.end()) it->second.clean_depth = 0;
}
void mark_clean(const void *p, int depth)
{
    Context &c = ctx();
    std::lock_guard<std::mutex> lock(c.mu);
    auto it = c.registry.find(p);
    if (it != c.registry.end()) it->second.clean_depth = depth;
}

int ensure_scratch(size_t n)
{
    Context &c = ctx();
    if (c.scratch_n >= n) return DLESM_OK;
    if (c.scratch) cudaFree(c.scratch);
    c.scratch = nullptr;
    c.scratch_n = 0;
    CU(cudaMalloc(&c.scratch, n * sizeof(double)));
    c.scratch_n = n;
    return DLESM_OK;
}

int ensure_xbuf(size_t n)
{
    Context &c = ctx();
    if (c.xbuf_n >= n) return DLESM_OK;
    if (c.sendbuf) cudaFree(c.sendbuf);
    if (c.recvbuf) cudaFree(c.recvbuf);
    c.sendbuf = c.recvbuf = nullptr;
    c.xbuf_n = 0;
    CU(cudaMalloc(&c.sendbuf, n * sizeof(double)));
    CU(cudaMalloc(&c.recvbuf, n * sizeof(double)));
    c.xbuf_n = n;
    return DLESM_OK;
}

k::Box box_of(const dlesm_region &r) { return k::Box{r.xstart, r.xstop, r.ystart, r.ystop}; }

}  // namespace

namespace dlesm {
int internal_ensure_device() { return ensure_device(); }
cudaStream_t internal_stream() { return ctx().stream; }
void internal_mark_written(const void *device_ptr) { mark_written(device_ptr); }
int internal_pitch_of(const void *device_ptr, int &pitch, int &ny)
{
    FieldRec r;
    int rc = lookup(device_ptr, r);
    if (rc != DLESM_OK) return rc;
    pitch = r.nx;
    ny = r.ny;
    return DLESM_OK;
}
}  // namespace dlesm

// =========================================================================================
// errors, lifecycle
// =========================================================================================
extern "C" const char *dlesm_last_error(void) { return last_error_cstr(); }
extern "C" int dlesm_last_status(void) { return last_status(); }

extern "C" int dlesm_parallel_init(int rank1, int nranks)
{
    if (nranks < 1 || rank1 < 1 || rank1 > nranks) {
        set_error(DLESM_ERR_NRANKS, "dlesm_parallel_init: need 1 <= rank <= nranks");
        return DLESM_ERR_NRANKS;
    }
    pstate().rank = rank1;
    pstate().nranks = nranks;
    return DLESM_OK;
}

extern "C" int dlesm_device_init(int device)
{
    Context &c = ctx();
    if (c.device_ready && (device < 0 || device == c.device)) return DLESM_OK;
    if (c.device_ready) {
        set_error(DLESM_ERR_STATE, "dlesm_device_init: already bound to another device");
        return DLESM_ERR_STATE;
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        set_error(DLESM_ERR_CUDA, std::string("no usable CUDA device (") +
                                      (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                                      "); this library has no CPU fallback");
        return DLESM_ERR_CUDA;
    }
    if (device < 0) {
        const char *lr = std::getenv("LOCAL_RANK");
        device = lr ? std::atoi(lr) % count : 0;
    }
    CU(cudaSetDevice(device));
    CU(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&c.s_h2d, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&c.s_d2h, cudaStreamNonBlocking));
    CU(cudaMalloc(&c.partials, (k::kChecksumBlocks + 16) * sizeof(double)));
    c.result = c.partials + k::kChecksumBlocks;
    c.ticket = reinterpret_cast<unsigned *>(c.partials + k::kChecksumBlocks + 8);
    CU(cudaMemset(c.ticket, 0, sizeof(unsigned)));
    CU(cudaMallocHost(&c.host_result, 8 * sizeof(double)));
    std::memset(c.host_result, 0, 8 * sizeof(double));
    {
        // How long a kernel waits for another rank before it gives up (exchange, global sum, halo push
        // hand-shake): DLESM_WAIT_TIMEOUT_S seconds, default 300, 0 = for ever like an MPI wait.  A rank that
        // is merely late (output on rank 0, an NVRTC compile, a debugger) must not end the job; a lost rank
        // must not hang it for good.  Expiry is fatal by design: carrying on would compute on stale halos.
        double secs = 300.0;
        if (const char *e = std::getenv("DLESM_WAIT_TIMEOUT_S")) {
            char *end = nullptr;
            const double v = std::strtod(e, &end);
            if (end == e || v < 0.0) {
                set_error(DLESM_ERR_ARGS, "DLESM_WAIT_TIMEOUT_S: expected a number of seconds >= 0 (0 = wait for ever)");
                return DLESM_ERR_ARGS;
            }
            secs = v;
        }
        int khz = 0;
        CU(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, device));
        const double cyc = secs * 1000.0 * (khz > 0 ? khz : 2000000);
        const long long cycles = secs == 0.0 ? 0 : (cyc < 1.0 ? 1 : (cyc > 9.0e18 ? 9000000000000000000LL : static_cast<long long>(cyc)));
        CU(k::set_wait_limit(cycles));
        CU(k::fma_off::set_wait_limit(cycles));
        CU(k::fma_on::set_wait_limit(cycles));
    }
    c.device = device;
    c.device_ready = true;
    if (const char *ar = std::getenv("DLESM_ARITH"))
        if (std::strcmp(ar, "fma") == 0 || std::strcmp(ar, "contracted") == 0) c.arith = DLESM_ARITH_CONTRACTED;
    return DLESM_OK;
}

extern "C" int dlesm_comm_unique_id(void *id128)
{
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is expected to be 128 bytes");
    ncclUniqueId id;
    NC(ncclGetUniqueId(&id));
    std::memcpy(id128, &id, sizeof(id));
    return DLESM_OK;
}

extern "C" int dlesm_comm_init(const void *id128)
{
    Context &c = ctx();
    if (c.comm_ready) return DLESM_OK;
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    NC(ncclCommInitRank(&c.comm, pstate().nranks, id, pstate().rank - 1));
    c.comm_ready = true;
    return DLESM_OK;
}

extern "C" int dlesm_parallel_finalise(void)
{
    Context &c = ctx();
    if (c.device_ready) cudaStreamSynchronize(c.stream);
    if (c.xd.ready) {
        // importers close before the exporter frees: unmap, meet, then free
        for (void *m : c.xd.mappings) cudaIpcCloseMemHandle(m);
        c.xd.mappings.clear();
        if (c.comm_ready && pstate().nranks > 1) {
            double one = 1.0;
            dlesm_global_sum(&one);
        }
        cudaFree(c.xd.base);
        c.xd = Context::XDirect{};
    }
    if (c.push_strip) {
        cudaFree(c.push_strip);
        c.push_strip = nullptr;
        c.push_strip_cap = 0;
    }
    if (c.comm_ready) {
        ncclCommDestroy(c.comm);
        c.comm = nullptr;
        c.comm_ready = false;
    }
    return DLESM_OK;
}

extern "C" void dlesm_parallel_abort(const char *msg)
{
    std::fprintf(stderr, " %s\n", msg ? msg : "");
    Context &c = ctx();
    if (c.comm_ready) ncclCommAbort(c.comm);
    std::exit(1);
}

extern "C" int dlesm_get_rank(void) { return pstate().rank; }
extern "C" int dlesm_get_num_ranks(void) { return pstate().nranks; }
extern "C" int dlesm_on_master(void) { return pstate().rank == 1; }
extern "C" void dlesm_reset_saved_state(void)
{
    reset_saved_state();
    // the SAVEd tiling of field_mod (`tiling_initialised`, `decomp`: src/field_mod.f90:236-240)
    ctx().tiling_initialised = false;
    ctx().sweep_tiles = k::SweepTiles{0, 0};
}
extern "C" int dlesm_set_extensions(int flags)
{
    pstate().extensions = flags;
    return DLESM_OK;
}

extern "C" int dlesm_global_sum(double *var)
{
    if (pstate().nranks == 1) return DLESM_OK;  // stub behaviour
    Context &c = ctx();
    if (!c.comm_ready) {
        set_error(DLESM_ERR_STATE, "global_sum: communicator not initialised (dlesm_comm_init)");
        return DLESM_ERR_STATE;
    }
    c.host_result[1] = *var;
    CU(cudaMemcpyAsync(c.result + 1, c.host_result + 1, sizeof(double), cudaMemcpyHostToDevice, c.stream));
    NC(ncclAllReduce(c.result + 1, c.result + 2, 1, ncclDouble, ncclSum, c.comm, c.stream));
    CU(cudaMemcpyAsync(c.host_result + 2, c.result + 2, sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    *var = c.host_result[2];
    return DLESM_OK;
}

// gather (src/parallel/parallel_utils_mod.f90:242-255; stub :154-161): every rank contributes n
// doubles from host memory, rank 1 receives them back to back.  Staged through the device so the
// transfer rides NCCL/NVLink like every other message of this library.
extern "C" int dlesm_gather_host(const double *send_buffer, long long n, double *recv_buffer)
{
    clear_error();
    const int nranks = pstate().nranks, me = pstate().rank;
    if (n < 0 || !send_buffer || (me == 1 && !recv_buffer)) {
        set_error(DLESM_ERR_ARGS, "gather: invalid arguments");
        return DLESM_ERR_ARGS;
    }
    if (nranks == 1) {
        std::memcpy(recv_buffer, send_buffer, static_cast<size_t>(n) * sizeof(double));
        return DLESM_OK;
    }
    Context &c = ctx();
    if (!c.comm_ready) {
        set_error(DLESM_ERR_STATE, "gather: communicator not initialised (dlesm_comm_init)");
        return DLESM_ERR_STATE;
    }
    const size_t un = static_cast<size_t>(n);
    int rc = ensure_scratch(un * (me == 1 ? nranks : 1));
    if (rc != DLESM_OK) return rc;
    CU(cudaMemcpyAsync(c.scratch, send_buffer, un * sizeof(double), cudaMemcpyHostToDevice, c.stream));
    NC(ncclGroupStart());
    if (me == 1) {
        for (int r = 1; r < nranks; ++r)
            NC(ncclRecv(c.scratch + static_cast<size_t>(r) * un, un, ncclDouble, r, c.comm, c.stream));
    } else {
        NC(ncclSend(c.scratch, un, ncclDouble, 0, c.comm, c.stream));
    }
    NC(ncclGroupEnd());
    if (me == 1)
        CU(cudaMemcpyAsync(recv_buffer, c.scratch, un * nranks * sizeof(double), cudaMemcpyDeviceToHost,
                           c.stream));
    CU(cudaStreamSynchronize(c.stream));
    return DLESM_OK;
}

// =========================================================================================
// decomposition / comm tables (host only)
// =========================================================================================
struct dlesm_decomp {
    Decomposition d;
};

extern "C" dlesm_decomp_t *dlesm_go_decompose(int domainx, int domainy, int ndomains, int ndomainx,
                                               int ndomainy, int halo_width)
{
    clear_error();
    auto *h = new dlesm_decomp;
    if (go_decompose(domainx, domainy, ndomains, ndomainx, ndomainy, halo_width, pstate().nranks, h->d) !=
        DLESM_OK) {
        delete h;
        return nullptr;
    }
    return h;
}
extern "C" int dlesm_decomp_info(const dlesm_decomp_t *h, int info[8])
{
    const Decomposition &d = h->d;
    const int v[8] = {d.global_nx, d.global_ny, d.nx, d.ny, d.ndomains, d.max_width, d.max_height, d.domperrank};
    std::memcpy(info, v, sizeof(v));
    return DLESM_OK;
}
extern "C" int dlesm_decomp_subdomain(const dlesm_decomp_t *h, int id1, dlesm_subdomain *out)
{
    if (id1 < 1 || id1 > h->d.ndomains) {
        set_error(DLESM_ERR_ARGS, "subdomain index out of range");
        return DLESM_ERR_ARGS;
    }
    *out = h->d.subdomains[id1 - 1];
    return DLESM_OK;
}
extern "C" int dlesm_decomp_proc_subdomain(const dlesm_decomp_t *h, int islot1, int irank1)
{
    const Decomposition &d = h->d;
    if (islot1 < 1 || islot1 > d.domperrank || irank1 < 1 || irank1 > d.nranks) return -1;
    return d.proc_subdomains[static_cast<size_t>(irank1 - 1) * d.domperrank + (islot1 - 1)];
}
extern "C" void dlesm_decomp_free(dlesm_decomp_t *h) { delete h; }

extern "C" int dlesm_map_comms(const dlesm_decomp_t *h, int irank1, int nranks, int pbc_x, int pbc_y,
                               int hdx, int hdy, dlesm_comms *out)
{
    clear_error();
    return build_comms(h->d, irank1, nranks, pbc_x != 0, pbc_y != 0, hdx, hdy, *out);
}
extern "C" int dlesm_iprocmap(const dlesm_decomp_t *h, int nranks, int ia, int ja)
{
    return owner_of(h->d, nranks, ia, ja);
}

// =========================================================================================
// device layer: storage + hooks
// =========================================================================================
extern "C" int dlesm_dev_field_alloc(int nx, int ny, void **device_ptr)
{
    clear_error();
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    if (nx < 1 || ny < 1) {
        set_error(DLESM_ERR_ARGS, "dlesm_dev_field_alloc: extents must be positive");
        return DLESM_ERR_ARGS;
    }
    Context &c = ctx();
    void *p = nullptr;
    const size_t bytes = static_cast<size_t>(nx) * ny * sizeof(double);
    CU(cudaMalloc(&p, bytes));
    cudaError_t e = cudaMemsetAsync(p, 0, bytes, c.stream);  // self%data(:,:) = 0 (field_mod.f90:375)
    if (e != cudaSuccess) {
        cudaFree(p);
        return fail_cuda(e, "cudaMemsetAsync");
    }
    {
        std::lock_guard<std::mutex> lock(c.mu);
        c.registry[p] = FieldRec{nx, ny};
    }
    *device_ptr = p;
    return DLESM_OK;
}

extern "C" int dlesm_dev_field_free(void *p)
{
    if (!p) return DLESM_OK;
    Context &c = ctx();
    {
        std::lock_guard<std::mutex> lock(c.mu);
        if (c.registry.erase(p) == 0) {
            set_error(DLESM_ERR_UNKNOWN_PTR, "dlesm_dev_field_free: unknown pointer");
            return DLESM_ERR_UNKNOWN_PTR;
        }
    }
    CU(cudaStreamSynchronize(c.stream));
    CU(cudaFree(p));
    return DLESM_OK;
}

namespace {
// shared by both hooks: 1-based sub-rectangle of the padded array, offset (j-1)*pitch+(i-1)
int hook_copy(void *device, void *host, bool to_device, int startx, int starty, int nx, int ny, bool blocking)
{
    FieldRec r;
    int rc = lookup(device, r);
    if (rc != DLESM_OK) return rc;
    if (startx < 1 || starty < 1 || nx < 0 || ny < 0 || startx + nx - 1 > r.nx || starty + ny - 1 > r.ny) {
        set_error(DLESM_ERR_ARGS, "device hook: sub-rectangle outside the field allocation");
        return DLESM_ERR_ARGS;
    }
    Context &c = ctx();
    const size_t off = (static_cast<size_t>(starty) - 1) * r.nx + (static_cast<size_t>(startx) - 1);
    double *d = static_cast<double *>(device) + off;
    double *h = static_cast<double *>(host) + off;
    const size_t pitch_b = static_cast<size_t>(r.nx) * sizeof(double);
    if (nx > 0 && ny > 0) {
        const cudaMemcpyKind kind = to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
        if (nx == r.nx)  // full rows: one contiguous transfer
            CU(cudaMemcpyAsync(to_device ? d : h, to_device ? h : d, pitch_b * ny, kind, c.stream));
        else
            CU(cudaMemcpy2DAsync(to_device ? d : h, pitch_b, to_device ? h : d, pitch_b,
                                 static_cast<size_t>(nx) * sizeof(double), ny, kind, c.stream));
    }
    if (blocking) CU(cudaStreamSynchronize(c.stream));
    if (to_device) mark_written(device);
    return DLESM_OK;
}
}  // namespace

extern "C" void dlesm_read_from_device_c(void *from, void *to, int startx, int starty, int nx, int ny, bool blocking)
{
    if (hook_copy(from, to, false, startx, starty, nx, ny, blocking) != DLESM_OK)
        dlesm_parallel_abort(dlesm_last_error());  // the hook signature has no status: fail loudly
}
extern "C" void dlesm_write_to_device_c(void *from, void *to, int startx, int starty, int nx, int ny, bool blocking)
{
    if (hook_copy(to, from, true, startx, starty, nx, ny, blocking) != DLESM_OK)
        dlesm_parallel_abort(dlesm_last_error());
}
extern "C" int dlesm_dev_sync(void)
{
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    Context &c = ctx();
    CU(cudaStreamSynchronize(c.stream));
    if (c.push_err) {
        int flag = 0;
        CU(cudaMemcpy(&flag, c.push_err, sizeof(int), cudaMemcpyDeviceToHost));
        if (flag) {
            set_error(DLESM_ERR_STATE, "halo push: a neighbour did not publish its epoch in time (ranks out of step?)");
            return DLESM_ERR_STATE;
        }
    }
    return DLESM_OK;
}

namespace {
// NUMA node of the GPU this process drives (sysfs), -1 if unknown
int gpu_numa_node()
{
    char bus[32] = {0};
    if (cudaDeviceGetPCIBusId(bus, sizeof(bus), ctx().device) != cudaSuccess) return -1;
    for (char *p = bus; *p; ++p) *p = static_cast<char>(std::tolower(*p));
    const std::string path = std::string("/sys/bus/pci/devices/") + bus + "/numa_node";
    FILE *f = std::fopen(path.c_str(), "r");
    if (!f) return -1;
    int node = -1;
    if (std::fscanf(f, "%d", &node) != 1) node = -1;
    std::fclose(f);
    return node;
}
std::map<void *, size_t> &numa_allocs()
{
    static std::map<void *, size_t> m;
    return m;
}
}  // namespace

// Pinned host memory for the host-state step and the hooks.  With several GPUs per node every rank
// would otherwise pin and stream from whichever node its launcher thread ran on (the round-1 scaling
// run: all eight ranks on NUMA node 0); here the pages are placed on the GPU's own node first
// (mmap + mbind(MPOL_PREFERRED) + first touch) and then registered.  Falls back to cudaMallocHost when
// the node is unknown or DLESM_HOST_NUMA=0.
extern "C" int dlesm_host_alloc(size_t bytes, void **host_ptr)
{
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    const char *e = std::getenv("DLESM_HOST_NUMA");
    const int node = (e && std::atoi(e) == 0) ? -1 : gpu_numa_node();
    if (node >= 0 && node < 1024 && bytes > 0) {
        void *p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (p != MAP_FAILED) {
            unsigned long mask[16] = {0};
            mask[node / (8 * sizeof(unsigned long))] |= 1UL << (node % (8 * sizeof(unsigned long)));
            // MPOL_PREFERRED = 1: place on `node` when the cpuset allows it, anywhere otherwise
            syscall(SYS_mbind, p, bytes, 1, mask, sizeof(mask) * 8, 0);
            madvise(p, bytes, MADV_HUGEPAGE);
            if (cudaHostRegister(p, bytes, cudaHostRegisterPortable) == cudaSuccess) {
                numa_allocs()[p] = bytes;
                *host_ptr = p;
                return DLESM_OK;
            }
            cudaGetLastError();
            munmap(p, bytes);
        }
    }
    CU(cudaMallocHost(host_ptr, bytes));
    return DLESM_OK;
}
extern "C" int dlesm_host_free(void *host_ptr)
{
    auto it = numa_allocs().find(host_ptr);
    if (it != numa_allocs().end()) {
        CU(cudaHostUnregister(host_ptr));
        munmap(host_ptr, it->second);
        numa_allocs().erase(it);
        return DLESM_OK;
    }
    CU(cudaFreeHost(host_ptr));
    return DLESM_OK;
}
extern "C" int dlesm_host_register(void *host_ptr, size_t bytes)
{
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    CU(cudaHostRegister(host_ptr, bytes, cudaHostRegisterDefault));
    return DLESM_OK;
}
extern "C" int dlesm_host_unregister(void *host_ptr)
{
    CU(cudaHostUnregister(host_ptr));
    return DLESM_OK;
}

namespace {
cudaEvent_t g_events[16] = {nullptr};
}
extern "C" int dlesm_event_record(int slot)
{
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    if (slot < 0 || slot >= 16) {
        set_error(DLESM_ERR_ARGS, "event slot out of range");
        return DLESM_ERR_ARGS;
    }
    if (!g_events[slot]) CU(cudaEventCreate(&g_events[slot]));
    CU(cudaEventRecord(g_events[slot], ctx().stream));
    return DLESM_OK;
}
extern "C" int dlesm_event_elapsed_ms(int a, int b, double *ms)
{
    if (a < 0 || a >= 16 || b < 0 || b >= 16 || !g_events[a] || !g_events[b]) {
        set_error(DLESM_ERR_ARGS, "event slot not recorded");
        return DLESM_ERR_ARGS;
    }
    CU(cudaEventSynchronize(g_events[b]));
    float f = 0.f;
    CU(cudaEventElapsedTime(&f, g_events[a], g_events[b]));
    *ms = f;
    return DLESM_OK;
}

// =========================================================================================
// device layer: field operations
// =========================================================================================
extern "C" int dlesm_dev_copy_field(const void *in, void *out)
{
    FieldRec a, b;
    int rc = lookup(in, a);
    if (rc == DLESM_OK) rc = lookup(out, b);
    if (rc != DLESM_OK) return rc;
    if (a.nx != b.nx || a.ny != b.ny) {
        set_error(DLESM_ERR_ARGS, "copy_field: fields have different shapes");
        return DLESM_ERR_ARGS;
    }
    CU(k::copy_field(static_cast<const double *>(in), static_cast<double *>(out),
                     static_cast<size_t>(a.nx) * a.ny, ctx().stream));
    mark_clean(out, a.clean_depth);   // the whole array incl. its halos is copied
    return DLESM_OK;
}

extern "C" int dlesm_dev_copy_patch(void *f, int sxs, int sxe, int sys, int sye, int dxs, int dxe, int dys, int dye)
{
    FieldRec r;
    int rc = lookup(f, r);
    if (rc != DLESM_OK) return rc;
    const int w = sxe - sxs + 1, h = sye - sys + 1;
    if (w != dxe - dxs + 1 || h != dye - dys + 1) {
        set_error(DLESM_ERR_ARGS, "copy_field: source and destination patches differ in shape");
        return DLESM_ERR_ARGS;
    }
    if (w <= 0 || h <= 0) return DLESM_OK;
    if (sxs < 1 || sys < 1 || dxs < 1 || dys < 1 || sxe > r.nx || dxe > r.nx || sye > r.ny || dye > r.ny) {
        set_error(DLESM_ERR_ARGS, "copy_field: patch outside the field allocation");
        return DLESM_ERR_ARGS;
    }
    rc = ensure_scratch(static_cast<size_t>(w) * h);
    if (rc != DLESM_OK) return rc;
    CU(k::copy_patch(static_cast<double *>(f), r.nx, k::Box{sxs, sxe, sys, sye}, k::Box{dxs, dxe, dys, dye},
                     ctx().scratch, ctx().stream));
    mark_written(f);
    return DLESM_OK;
}

extern "C" int dlesm_dev_set_field(void *f, double val)
{
    FieldRec r;
    int rc = lookup(f, r);
    if (rc != DLESM_OK) return rc;
    CU(k::set_field(static_cast<double *>(f), static_cast<size_t>(r.nx) * r.ny, val, ctx().stream));
    mark_written(f);
    return DLESM_OK;
}

extern "C" int dlesm_dev_checksum(const void *f, int xs, int xe, int ys, int ye, double *val)
{
    FieldRec r;
    int rc = lookup(f, r);
    if (rc != DLESM_OK) return rc;
    if (xs < 1 || ys < 1 || xe > r.nx || ye > r.ny) {
        set_error(DLESM_ERR_ARGS, "field_checksum: box outside the field allocation");
        return DLESM_ERR_ARGS;
    }
    Context &c = ctx();
    // one launch; the folding block stores the sum straight into pinned host memory
    const unsigned long long seq = ++c.checksum_seq;
    CU(k::checksum(static_cast<const double *>(f), r.nx, k::Box{xs, xe, ys, ye}, c.partials, c.ticket, c.result,
                   c.host_result + 4, seq, c.stream));
    // poll the sequence word instead of cudaStreamSynchronize (a driver round trip of several us); the
    // stream is queried now and then so that a failed launch ends the wait with its error
    // acquire: the value next to the sequence word is read only after the word (weakly ordered hosts too)
    const unsigned long long *flag = reinterpret_cast<const unsigned long long *>(c.host_result + 5);
    auto published = [flag]() { return __atomic_load_n(flag, __ATOMIC_ACQUIRE); };
    for (unsigned spin = 1; published() != seq; ++spin) {
        if ((spin & 0xfffff) == 0) {
            cudaError_t q = cudaStreamQuery(c.stream);
            if (q != cudaSuccess && q != cudaErrorNotReady) return fail_cuda(q, "field_checksum");
            if (q == cudaSuccess && published() != seq) {
                set_error(DLESM_ERR_CUDA, "field_checksum: the kernel finished without publishing its result");
                return DLESM_ERR_CUDA;
            }
        }
    }
    *val = c.host_result[4];
    return DLESM_OK;
}

extern "C" int dlesm_dev_periodic_halos(void *f, int xs, int xe, int ys, int ye, int px, int py)
{
    FieldRec r;
    int rc = lookup(f, r);
    if (rc != DLESM_OK) return rc;
    if (xs < 2 || ys < 2 || xe + 1 > r.nx || ye + 1 > r.ny) {
        set_error(DLESM_ERR_ARGS, "periodic halos: internal region leaves no halo inside the allocation");
        return DLESM_ERR_ARGS;
    }
    double *fl[1] = {static_cast<double *>(f)};
    CU(k::periodic_halos(fl, 1, r.nx, k::Box{xs, xe, ys, ye}, px != 0, py != 0, ctx().stream));
    if (px && py) mark_clean(f, 1);
    return DLESM_OK;
}

extern "C" int dlesm_dev_periodic_halos_deep(void *f, int xs, int xe, int ys, int ye, int depth)
{
    FieldRec r;
    int rc = lookup(f, r);
    if (rc != DLESM_OK) return rc;
    if (depth < 1 || xs <= depth || ys <= depth || xe + depth > r.nx || ye + depth > r.ny || depth > xe - xs + 1 ||
        depth > ye - ys + 1) {
        set_error(DLESM_ERR_ARGS, "periodic halos: the allocation has no room for a halo that deep");
        return DLESM_ERR_ARGS;
    }
    double *fl[1] = {static_cast<double *>(f)};
    CU(k::periodic_halos_deep(fl, 1, r.nx, k::Box{xs, xe, ys, ye}, depth, ctx().stream));
    mark_clean(f, depth);
    return DLESM_OK;
}

// =========================================================================================
// halo exchange
// =========================================================================================
extern "C" int dlesm_comm_plan_build(const dlesm_comms *t, int rank1, dlesm_comm_plan *out)
{
    std::memset(out, 0, sizeof(*out));
    if (t->nsend > DLESM_MAXCOMM || t->nrecv > DLESM_MAXCOMM || t->nsend < 0 || t->nrecv < 0) {
        set_error(DLESM_ERR_MAXCOMM, "comm plan: list longer than MaxComm");
        return DLESM_ERR_MAXCOMM;
    }
    const int me0 = rank1 - 1;
    // peers in ascending rank order
    std::vector<int> peers;
    auto note = [&](int p) {
        if (std::find(peers.begin(), peers.end(), p) == peers.end()) peers.push_back(p);
    };
    for (int s = 0; s < t->nsend; ++s)
        if (t->destination[s] >= 0 && t->nxsend[s] > 0) note(t->destination[s]);
    for (int r = 0; r < t->nrecv; ++r)
        if (t->source[r] >= 0 && t->nxrecv[r] > 0) note(t->source[r]);
    std::sort(peers.begin(), peers.end());
    if (peers.size() > DLESM_MAXCOMM) {
        set_error(DLESM_ERR_MAXCOMM, "comm plan: more peers than MaxComm");
        return DLESM_ERR_MAXCOMM;
    }
    out->npeers = static_cast<int>(peers.size());
    long long so = 0, ro = 0;
    for (int g = 0; g < out->npeers; ++g) {
        const int p0 = peers[g];
        out->peer[g] = p0;
        std::vector<int> sends, recvs;
        for (int s = 0; s < t->nsend; ++s)
            if (t->destination[s] == p0 && t->nxsend[s] > 0) sends.push_back(s);
        for (int r = 0; r < t->nrecv; ++r)
            if (t->source[r] == p0 && t->nxrecv[r] > 0) recvs.push_back(r);
        // both sides walk their messages for a peer in direction order (the tag of the
        // reference); a message sent in direction d is received as direction d
        std::stable_sort(sends.begin(), sends.end(), [&](int a, int b) { return t->dirsend[a] < t->dirsend[b]; });
        std::stable_sort(recvs.begin(), recvs.end(), [&](int a, int b) { return t->dirrecv[a] < t->dirrecv[b]; });
        out->nsend[g] = static_cast<int>(sends.size());
        out->nrecv[g] = static_cast<int>(recvs.size());
        out->send_off[g] = so;
        out->recv_off[g] = ro;
        for (size_t q = 0; q < sends.size(); ++q) {
            out->send_slot[g][q] = sends[q];
            out->send_n[g] += static_cast<long long>(t->nxsend[sends[q]]) * t->nysend[sends[q]];
        }
        for (size_t q = 0; q < recvs.size(); ++q) {
            out->recv_slot[g][q] = recvs[q];
            out->recv_n[g] += static_cast<long long>(t->nxrecv[recvs[q]]) * t->nyrecv[recvs[q]];
        }
        if (p0 == me0 && out->send_n[g] != out->recv_n[g]) {
            set_error(DLESM_ERR_ARGS, "comm plan: self-messages do not pair up");
            return DLESM_ERR_ARGS;
        }
        so += out->send_n[g];
        ro += out->recv_n[g];
    }
    out->send_total = so;
    out->recv_total = ro;
    return DLESM_OK;
}

namespace {
int build_tables(const dlesm_comms *t, int depth, CommTables &out)
{
    dlesm_comm_plan plan;
    int rc = dlesm_comm_plan_build(t, pstate().rank, &plan);
    if (rc != DLESM_OK) return rc;
    out.t = *t;
    out.groups.clear();
    for (int g = 0; g < plan.npeers; ++g) {
        PeerGroup pg;
        pg.peer0 = plan.peer[g];
        pg.sends.assign(plan.send_slot[g], plan.send_slot[g] + plan.nsend[g]);
        pg.recvs.assign(plan.recv_slot[g], plan.recv_slot[g] + plan.nrecv[g]);
        pg.send_off = plan.send_off[g];
        pg.recv_off = plan.recv_off[g];
        pg.send_n = plan.send_n[g];
        pg.recv_n = plan.recv_n[g];
        out.groups.push_back(pg);
    }
    out.send_total = plan.send_total;
    out.recv_total = plan.recv_total;
    out.depth = depth;
    out.valid = true;
    return DLESM_OK;
}

// exchange_generic on device-resident fields with the given tables.  With DLESM_EXT_DIRTY_HALOS
// fields whose halos are already up to date to the tables' depth are left out (all clean: no launch).
int xdirect_exchange(void *const *fields, int nfields, int pitch, const CommTables &ct, bool &done);

int exchange_with(void *const *fields_in, int nfields_in, const CommTables &ct)
{
    Context &c = ctx();
    if (!ct.valid || (ct.t.nsend == 0 && ct.t.nrecv == 0)) return DLESM_OK;  // nothing to do
    if (nfields_in < 1 || nfields_in > k::kMaxExchangeFields) {
        set_error(DLESM_ERR_ARGS, "halo_exchange: between 1 and 8 fields per call");
        return DLESM_ERR_ARGS;
    }
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    void *fields[k::kMaxExchangeFields];
    int nfields = 0;
    FieldRec r0{};
    // The skip is a per-rank decision, but an exchange with another rank is a contract between two: both sides
    // must agree on the field list (packed layout, slot parity) and on launching at all.  A one-sided write
    // (a hook or set_field on one rank) would break that, so fields are only skipped when every neighbour of the
    // installed tables is this rank itself; with remote neighbours every listed field is exchanged.
    bool remote_peer = false;
    for (const auto &g : ct.groups) remote_peer = remote_peer || g.peer0 != pstate().rank - 1;
    const bool skip_clean = (pstate().extensions & DLESM_EXT_DIRTY_HALOS) != 0 && !remote_peer;
    for (int q = 0; q < nfields_in; ++q) {
        FieldRec rq;
        rc = lookup(fields_in[q], rq);
        if (rc != DLESM_OK) return rc;
        if (q == 0) r0 = rq;
        if (rq.nx != r0.nx || rq.ny != r0.ny) {
            set_error(DLESM_ERR_ARGS, "halo_exchange: fields have different shapes");
            return DLESM_ERR_ARGS;
        }
        if (skip_clean && rq.clean_depth >= ct.depth) continue;
        fields[nfields++] = fields_in[q];
    }
    if (nfields == 0) return DLESM_OK;   // every halo is up to date: nothing is launched
    const dlesm_comms &t = ct.t;
    // the tables may come from another grid (module-level state): every patch must lie inside
    // the fields, or the pack / unpack kernels would run off the allocation
    for (int q = 0; q < t.nsend; ++q)
        if (t.nxsend[q] > 0 && (t.isrcsend[q] < 1 || t.jsrcsend[q] < 1 || t.isrcsend[q] + t.nxsend[q] - 1 > r0.nx ||
                                t.jsrcsend[q] + t.nysend[q] - 1 > r0.ny)) {
            set_error(DLESM_ERR_ARGS, "halo_exchange: a send patch of the installed tables lies outside the field (tables of another grid?)");
            return DLESM_ERR_ARGS;
        }
    for (int q = 0; q < t.nrecv; ++q)
        if (t.nxrecv[q] > 0 && (t.idesrecv[q] < 1 || t.jdesrecv[q] < 1 || t.idesrecv[q] + t.nxrecv[q] - 1 > r0.nx ||
                                t.jdesrecv[q] + t.nyrecv[q] - 1 > r0.ny)) {
            set_error(DLESM_ERR_ARGS, "halo_exchange: a receive patch of the installed tables lies outside the field (tables of another grid?)");
            return DLESM_ERR_ARGS;
        }
    {
        // default path: the send patches go straight into the neighbours' receive slots over NVLink
        bool done = false;
        rc = xdirect_exchange(fields, nfields, r0.nx, ct, done);
        if (rc != DLESM_OK) return rc;
        if (done) {
            for (int q = 0; q < nfields; ++q) mark_clean(fields[q], ct.depth);
            return DLESM_OK;
        }
    }
    const long long tot = std::max(ct.send_total, ct.recv_total) * nfields;
    rc = ensure_xbuf(static_cast<size_t>(tot) + 1);
    if (rc != DLESM_OK) return rc;
    const int me0 = pstate().rank - 1;

    k::PackPlan sp{}, rp{};
    sp.nfields = rp.nfields = nfields;
    sp.pitch = rp.pitch = r0.nx;
    for (int q = 0; q < nfields; ++q) sp.fields[q] = rp.fields[q] = static_cast<double *>(fields[q]);
    bool remote = false;
    for (const auto &g : ct.groups) {
        long long off = g.send_off * nfields;
        for (int s : g.sends) {
            k::Segment &sg = sp.seg[sp.nseg++];
            sg.i0 = t.isrcsend[s]; sg.j0 = t.jsrcsend[s]; sg.nx = t.nxsend[s]; sg.ny = t.nysend[s];
            sg.offset = off;
            off += static_cast<long long>(sg.nx) * sg.ny * nfields;
        }
        off = g.recv_off * nfields;
        for (int r : g.recvs) {
            k::Segment &sg = rp.seg[rp.nseg++];
            sg.i0 = t.idesrecv[r]; sg.j0 = t.jdesrecv[r]; sg.nx = t.nxrecv[r]; sg.ny = t.nyrecv[r];
            sg.offset = off;
            off += static_cast<long long>(sg.nx) * sg.ny * nfields;
        }
        if (g.peer0 != me0) remote = true;
    }
    if (remote && !c.comm_ready) {
        set_error(DLESM_ERR_STATE, "halo_exchange: communicator not initialised (dlesm_comm_init)");
        return DLESM_ERR_STATE;
    }
    CU(k::pack(sp, c.sendbuf, c.stream));
    if (remote) NC(ncclGroupStart());
    for (const auto &g : ct.groups) {
        if (g.peer0 == me0) {
            // wrap-around neighbour is this rank itself: the "message" never leaves the device
            if (g.send_n > 0)
                CU(cudaMemcpyAsync(c.recvbuf + g.recv_off * nfields, c.sendbuf + g.send_off * nfields,
                                   static_cast<size_t>(g.send_n) * nfields * sizeof(double),
                                   cudaMemcpyDeviceToDevice, c.stream));
            continue;
        }
        if (g.send_n > 0)
            NC(ncclSend(c.sendbuf + g.send_off * nfields, static_cast<size_t>(g.send_n) * nfields, ncclDouble,
                        g.peer0, c.comm, c.stream));
        if (g.recv_n > 0)
            NC(ncclRecv(c.recvbuf + g.recv_off * nfields, static_cast<size_t>(g.recv_n) * nfields, ncclDouble,
                        g.peer0, c.comm, c.stream));
    }
    if (remote) NC(ncclGroupEnd());
    CU(k::unpack(rp, c.recvbuf, c.stream));
    for (int q = 0; q < nfields; ++q) mark_clean(fields[q], ct.depth);
    return DLESM_OK;
}
}  // namespace

extern "C" int dlesm_dev_set_comms(const dlesm_comms *t)
{
    clear_error();
    Context &c = ctx();
    if (!t) {   // back to the state before any grid_init: no tables
        c.installed = CommTables{};
        return DLESM_OK;
    }
    // depth of the tables = the widest strip they receive
    int depth = 1;
    for (int q = 0; q < t->nrecv && q < DLESM_MAXCOMM; ++q)
        depth = std::max(depth, std::min(t->nxrecv[q], t->nyrecv[q]));
    CommTables ct;
    int rc = build_tables(t, depth, ct);
    if (rc != DLESM_OK) return rc;
    c.installed = ct;
    return DLESM_OK;
}

extern "C" int dlesm_dev_halo_exchange(void *const *fields, int nfields)
{
    clear_error();
    return exchange_with(fields, nfields, ctx().installed);
}

// fld%halo_exchange(depth) of a field whose values live in HOST memory (data_on_device false) in a
// build whose parallel_utils_mod has no MPI (src/field_mod.f90:1231-1256 would end in post_receive):
// the send patches of the installed tables go up to a staging array of the same shape, the device
// exchange runs, the receive patches come back.  Only the patches cross the bus.
extern "C" int dlesm_host_halo_exchange(double *host, int nx, int ny)
{
    clear_error();
    Context &c = ctx();
    const CommTables &ct = c.installed;
    if (!ct.valid || (ct.t.nsend == 0 && ct.t.nrecv == 0)) return DLESM_OK;   // one rank: a no-op, as in the reference
    if (!host || nx < 1 || ny < 1) {
        set_error(DLESM_ERR_ARGS, "host_halo_exchange: null array or empty shape");
        return DLESM_ERR_ARGS;
    }
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    static void *stage = nullptr;          // one staging array, re-used while the shape stays
    static int stage_nx = 0, stage_ny = 0;
    if (stage && (stage_nx != nx || stage_ny != ny)) {
        rc = dlesm_dev_field_free(stage);
        stage = nullptr;
        if (rc != DLESM_OK) return rc;
    }
    if (!stage) {
        rc = dlesm_dev_field_alloc(nx, ny, &stage);
        if (rc != DLESM_OK) return rc;
        stage_nx = nx; stage_ny = ny;
    }
    const dlesm_comms &t = ct.t;
    for (int q = 0; q < t.nsend; ++q)
        if (t.nxsend[q] > 0 && t.nysend[q] > 0) {
            rc = hook_copy(stage, host, true, t.isrcsend[q], t.jsrcsend[q], t.nxsend[q], t.nysend[q], false);
            if (rc != DLESM_OK) return rc;
        }
    void *ptrs[1] = {stage};
    rc = exchange_with(ptrs, 1, ct);
    if (rc != DLESM_OK) return rc;
    for (int q = 0; q < t.nrecv; ++q)
        if (t.nxrecv[q] > 0 && t.nyrecv[q] > 0) {
            rc = hook_copy(stage, host, false, t.idesrecv[q], t.jdesrecv[q], t.nxrecv[q], t.nyrecv[q], false);
            if (rc != DLESM_OK) return rc;
        }
    CU(cudaStreamSynchronize(c.stream));
    return DLESM_OK;
}

// ---------------------------------------------------------------------------------------
// Halo push set-up.  Collective over all ranks; every rank passes its fields in the same order.
// Each field's allocation is exported with cudaIpcGetMemHandle, the handles travel through one
// ncclAllGather, and the handles of this rank's neighbours (the destinations of its map_comms
// send list) are opened, which maps the neighbours' arrays into this process over NVLink.
// A flag array per rank is shared the same way for the per-step hand-shake.
// Mappings are reference-counted per IPC handle (cudaMalloc may serve several fields from one
// allocation): a field's registration can be dropped (dlesm_dev_peer_unregister) without
// unmapping what another registered field -- of this or of another model object -- still uses.
// ---------------------------------------------------------------------------------------
namespace {
constexpr int kMaxRanks = 64;
struct IpcEntry {
    cudaIpcMemHandle_t handle;
    long long offset;   // bytes from the allocation base (cudaMalloc may sub-allocate)
    int pitch, ny;
};
static_assert(sizeof(IpcEntry) == 80, "IpcEntry layout");

typedef CUresult (*GetAddressRangeFn)(CUdeviceptr *, size_t *, CUdeviceptr);
GetAddressRangeFn address_range_fn()
{
    static GetAddressRangeFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<GetAddressRangeFn>(p);
    }();
    return fn;
}

int make_entry(const void *ptr, int pitch, int ny, IpcEntry &e)
{
    std::memset(&e, 0, sizeof(e));
    GetAddressRangeFn fn = address_range_fn();
    CUdeviceptr base = reinterpret_cast<CUdeviceptr>(ptr);
    size_t size = 0;
    if (!fn || fn(&base, &size, reinterpret_cast<CUdeviceptr>(ptr)) != CUDA_SUCCESS) {
        set_error(DLESM_ERR_CUDA, "peer_register: cuMemGetAddressRange failed");
        return DLESM_ERR_CUDA;
    }
    CU(cudaIpcGetMemHandle(&e.handle, reinterpret_cast<void *>(base)));
    e.offset = static_cast<long long>(reinterpret_cast<CUdeviceptr>(ptr) - base);
    e.pitch = pitch;
    e.ny = ny;
    return DLESM_OK;
}

struct OpenedHandle {
    void *base = nullptr;
    int refs = 0;
};
std::map<std::string, OpenedHandle> &opened_handles()
{
    static std::map<std::string, OpenedHandle> m;
    return m;
}

// maps the allocation behind `e` (or reuses the mapping) and takes one reference on it; `key_out`
// identifies the mapping for release_handle
int open_entry(const IpcEntry &e, void **out, std::string &key_out)
{
    key_out.assign(reinterpret_cast<const char *>(&e.handle), sizeof(e.handle));
    auto &m = opened_handles();
    auto it = m.find(key_out);
    if (it == m.end()) {
        void *base = nullptr;
        CU(cudaIpcOpenMemHandle(&base, e.handle, cudaIpcMemLazyEnablePeerAccess));
        it = m.emplace(key_out, OpenedHandle{base, 0}).first;
    }
    ++it->second.refs;
    *out = static_cast<char *>(it->second.base) + e.offset;
    return DLESM_OK;
}
void release_handle(const std::string &key)
{
    auto &m = opened_handles();
    auto it = m.find(key);
    if (it == m.end()) return;
    if (--it->second.refs <= 0) {
        cudaIpcCloseMemHandle(it->second.base);
        m.erase(it);
    }
}

int barrier_all()
{
    if (pstate().nranks > 1 && ctx().comm_ready) {
        double one = 1.0;
        return dlesm_global_sum(&one);   // a one-double allreduce is the barrier
    }
    return DLESM_OK;
}

// neighbours (0-based ranks) named by a set of tables
void neighbours_of(const CommTables &ct, std::vector<char> &is_nb)
{
    for (const auto &g : ct.groups)
        if (g.peer0 >= 0 && g.peer0 < static_cast<int>(is_nb.size())) is_nb[g.peer0] = 1;
}

int xdirect_setup();
// slot of the packed push strip for direction tag `dir` (1..8) in the allocation at `base`
double *push_slot(char *base, unsigned long long epoch, int dir)
{
    const Context::XDirect &xd = ctx().xd;
    return reinterpret_cast<double *>(base + 4096) + 2 * 8 * xd.slot_cap + ((epoch & 1) * 8 + (dir - 1)) * xd.pslot_cap;
}

// global-sum cells of the allocation at `base`: bytes 2048..4095 of its flag page, two tables (alternating
// with the call's sequence number) of 64 cells {double value; u64 seq}
double *gsum_cells(char *base, unsigned long long seq) { return reinterpret_cast<double *>(base + 2048 + (seq & 1) * 1024); }

int peer_register_with(void *const *fields, int nfields, const CommTables &ct)
{
    Context &c = ctx();
    if (nfields < 1 || nfields > 12) {
        set_error(DLESM_ERR_ARGS, "peer_register: between 1 and 12 fields per call");
        return DLESM_ERR_ARGS;
    }
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    const int nranks = pstate().nranks, me0 = pstate().rank - 1;
    if (nranks > kMaxRanks) {
        set_error(DLESM_ERR_ARGS, "peer_register: too many ranks");
        return DLESM_ERR_ARGS;
    }
    std::vector<FieldRec> recs(nfields);
    for (int q = 0; q < nfields; ++q) {
        rc = lookup(fields[q], recs[q]);
        if (rc != DLESM_OK) return rc;
        if (c.peers.count(fields[q])) {
            set_error(DLESM_ERR_STATE, "peer_register: field is already registered (unregister it first)");
            return DLESM_ERR_STATE;
        }
    }
    if (!c.flags) {
        CU(cudaMalloc(&c.flags, kMaxRanks * sizeof(unsigned long long)));
        CU(cudaMemset(c.flags, 0, kMaxRanks * sizeof(unsigned long long)));
        CU(cudaMalloc(&c.push_err, sizeof(int)));
        CU(cudaMemset(c.push_err, 0, sizeof(int)));
        CU(cudaMalloc(&c.push_ticket, sizeof(unsigned)));
        CU(cudaMemset(c.push_ticket, 0, sizeof(unsigned)));
        c.peer_flags.assign(nranks, nullptr);
        c.peer_flags[me0] = c.flags;
    }
    // the slots of the packed push strips live in the direct exchange's allocation (collective, once)
    if (!c.xd.ready && (nranks == 1 || c.comm_ready)) {
        rc = xdirect_setup();
        if (rc != DLESM_OK) return rc;
    }
    if (nranks == 1) {
        for (int q = 0; q < nfields; ++q) {
            Context::PeerView pv;
            pv.by_rank.assign(1, static_cast<double *>(fields[q]));
            pv.pitch_by_rank.assign(1, recs[q].nx);
            c.peers[fields[q]] = pv;
        }
        return DLESM_OK;
    }
    if (!c.comm_ready || !ct.valid) {
        set_error(DLESM_ERR_STATE, "peer_register: needs dlesm_comm_init and the comm tables (dlesm_dev_set_comms) first");
        return DLESM_ERR_STATE;
    }
    // my entries: the fields, then the flag array
    const int nent = nfields + 1;
    std::vector<IpcEntry> mine(nent), all(static_cast<size_t>(nent) * nranks);
    for (int q = 0; q < nfields; ++q) {
        rc = make_entry(fields[q], recs[q].nx, recs[q].ny, mine[q]);
        if (rc != DLESM_OK) return rc;
    }
    rc = make_entry(c.flags, 0, 0, mine[nfields]);
    if (rc != DLESM_OK) return rc;
    const size_t blob = sizeof(IpcEntry) * nent;
    rc = ensure_scratch((blob * (nranks + 1) + 7) / 8);
    if (rc != DLESM_OK) return rc;
    char *dsend = reinterpret_cast<char *>(c.scratch), *drecv = dsend + blob;
    CU(cudaMemcpyAsync(dsend, mine.data(), blob, cudaMemcpyHostToDevice, c.stream));
    NC(ncclAllGather(dsend, drecv, blob, ncclChar, c.comm, c.stream));
    CU(cudaMemcpyAsync(all.data(), drecv, blob * nranks, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    std::vector<char> is_nb(nranks, 0);
    neighbours_of(ct, is_nb);
    for (int q = 0; q < nfields; ++q) {
        Context::PeerView pv;
        pv.by_rank.assign(nranks, nullptr);
        pv.pitch_by_rank.assign(nranks, 0);
        c.peers[fields[q]] = pv;
    }
    for (int r = 0; r < nranks; ++r) {
        if (!is_nb[r]) continue;
        const IpcEntry *er = &all[static_cast<size_t>(r) * nent];
        for (int q = 0; q < nfields; ++q) {
            void *p = fields[q];
            if (r != me0) {
                std::string key;
                rc = open_entry(er[q], &p, key);
                if (rc != DLESM_OK) return rc;
                c.peers[fields[q]].handles.push_back(key);
            }
            c.peers[fields[q]].by_rank[r] = static_cast<double *>(p);
            c.peers[fields[q]].pitch_by_rank[r] = er[q].pitch;
        }
        if (r != me0 && !c.peer_flags[r]) {
            // the flag arrays live as long as the library: mapped once, closed at finalise
            void *base = nullptr;
            CU(cudaIpcOpenMemHandle(&base, er[nfields].handle, cudaIpcMemLazyEnablePeerAccess));
            c.flag_mappings.push_back(base);
            c.peer_flags[r] = reinterpret_cast<unsigned long long *>(static_cast<char *>(base) + er[nfields].offset);
        }
    }
    // nobody pushes before everybody has mapped
    return barrier_all();
}

int peer_unregister_fields(void *const *fields, int nfields)
{
    Context &c = ctx();
    if (c.device_ready) CU(cudaStreamSynchronize(c.stream));
    for (int q = 0; q < nfields; ++q) {
        auto it = c.peers.find(fields[q]);
        if (it == c.peers.end()) continue;
        for (const auto &key : it->second.handles) release_handle(key);
        c.peers.erase(it);
    }
    return barrier_all();
}
}  // namespace

namespace {
constexpr size_t kXdFlagsBytes = 4096;     // flag cells [kMaxRanks] at 0, ticket at 1024, err at 1032 (push_slot assumes 4096)

// Collective, once per process (at the first direct exchange, which every rank reaches together):
// allocate this rank's flag cells + two receive buffers as ONE allocation and map every rank's.
int xdirect_setup()
{
    Context &c = ctx();
    Context::XDirect &xd = c.xd;
    const int nranks = pstate().nranks, me0 = pstate().rank - 1;
    if (nranks > kMaxRanks) {
        set_error(DLESM_ERR_ARGS, "direct exchange: too many ranks");
        return DLESM_ERR_ARGS;
    }
    size_t mb = 64;
    if (const char *e = std::getenv("DLESM_XDIRECT_MB")) mb = std::max(1L, std::atol(e));
    xd.slot_cap = mb * (1u << 20) / sizeof(double) / 8;                // 8 direction slots per buffer
    xd.pslot_cap = xd.slot_cap;
    const size_t bytes = kXdFlagsBytes + 2 * 8 * (xd.slot_cap + xd.pslot_cap) * sizeof(double);
    CU(cudaMalloc(&xd.base, bytes));
    CU(cudaMemset(xd.base, 0, kXdFlagsBytes));
    xd.peer_base.assign(nranks, nullptr);
    xd.count.assign(nranks, 0);
    xd.peer_base[me0] = xd.base;
    if (nranks > 1) {
        IpcEntry mine;
        int rc = make_entry(xd.base, 0, 0, mine);
        if (rc != DLESM_OK) return rc;
        std::vector<IpcEntry> all(nranks);
        rc = ensure_scratch((sizeof(IpcEntry) * (nranks + 1) + 7) / 8);
        if (rc != DLESM_OK) return rc;
        char *dsend = reinterpret_cast<char *>(c.scratch), *drecv = dsend + sizeof(IpcEntry);
        CU(cudaMemcpyAsync(dsend, &mine, sizeof(IpcEntry), cudaMemcpyHostToDevice, c.stream));
        NC(ncclAllGather(dsend, drecv, sizeof(IpcEntry), ncclChar, c.comm, c.stream));
        CU(cudaMemcpyAsync(all.data(), drecv, sizeof(IpcEntry) * nranks, cudaMemcpyDeviceToHost, c.stream));
        CU(cudaStreamSynchronize(c.stream));
        for (int r = 0; r < nranks; ++r) {
            if (r == me0) continue;
            void *b = nullptr;
            CU(cudaIpcOpenMemHandle(&b, all[r].handle, cudaIpcMemLazyEnablePeerAccess));
            xd.mappings.push_back(b);
            xd.peer_base[r] = static_cast<char *>(b) + all[r].offset;
        }
        int rc2 = barrier_all();     // nobody stores into a buffer before everybody has mapped
        if (rc2 != DLESM_OK) return rc2;
    }
    xd.ready = true;
    return DLESM_OK;
}

// exchange_generic (src/parallel_comms_mod.f90:1501-1855) with the neighbours' memory as the wire.
// done = false: not taken (mode NCCL) -- the caller falls through to pack -> NCCL -> unpack.
int xdirect_exchange(void *const *fields, int nfields, int pitch, const CommTables &ct, bool &done)
{
    Context &c = ctx();
    done = false;
    if (c.exchange_mode < 0) {
        const char *e = std::getenv("DLESM_EXCHANGE");
        c.exchange_mode = (e && std::strcmp(e, "nccl") == 0) ? DLESM_EXCHANGE_NCCL : DLESM_EXCHANGE_DIRECT;
    }
    if (c.exchange_mode != DLESM_EXCHANGE_DIRECT) return DLESM_OK;
    const int me0 = pstate().rank - 1;
    bool remote = false;
    for (const auto &g : ct.groups) remote = remote || g.peer0 != me0;
    if (remote && !c.comm_ready) {
        set_error(DLESM_ERR_STATE, "halo_exchange: communicator not initialised (dlesm_comm_init)");
        return DLESM_ERR_STATE;
    }
    if (!c.xd.ready) {
        int rc = xdirect_setup();
        if (rc != DLESM_OK) return rc;
    }
    Context::XDirect &xd = c.xd;
    const dlesm_comms &t = ct.t;
    k::XPlan sp{}, rp{};
    sp.nfields = rp.nfields = nfields;
    sp.pitch = rp.pitch = pitch;
    for (int q = 0; q < nfields; ++q) sp.fields[q] = rp.fields[q] = static_cast<double *>(fields[q]);
    auto slot = [&](char *base, unsigned long long cnt, int dir) {
        return reinterpret_cast<double *>(base + kXdFlagsBytes) + ((cnt & 1) * 8 + (dir - 1)) * xd.slot_cap;
    };
    // validate every group before any pair count moves: a refusal must leave this rank's counts (and with
    // them the slot parity) in step with its neighbours'
    for (const auto &g : ct.groups) {
        if (g.sends.empty() && g.recvs.empty()) continue;
        if (g.peer0 < 0 || g.peer0 >= static_cast<int>(xd.peer_base.size()) || !xd.peer_base[g.peer0]) {
            set_error(DLESM_ERR_STATE, "direct exchange: neighbour rank not mapped");
            return DLESM_ERR_STATE;
        }
        for (int s : g.sends) {
            if (t.dirsend[s] < 1 || t.dirsend[s] > 8 ||
                static_cast<size_t>(t.nxsend[s]) * t.nysend[s] * nfields > xd.slot_cap) {
                set_error(DLESM_ERR_ARGS, "direct exchange: a message exceeds its slot (raise DLESM_XDIRECT_MB or set DLESM_EXCHANGE=nccl on every rank)");
                return DLESM_ERR_ARGS;
            }
        }
    }
    for (const auto &g : ct.groups) {
        if (g.sends.empty() && g.recvs.empty()) continue;
        const unsigned long long cnt = ++xd.count[g.peer0];
        for (int s : g.sends) {
            k::XSegment &sg = sp.seg[sp.nseg++];
            sg.i0 = t.isrcsend[s]; sg.j0 = t.jsrcsend[s]; sg.nx = t.nxsend[s]; sg.ny = t.nysend[s];
            sg.buf = slot(xd.peer_base[g.peer0], cnt, t.dirsend[s]);
        }
        for (int r : g.recvs) {
            k::XSegment &sg = rp.seg[rp.nseg++];
            sg.i0 = t.idesrecv[r]; sg.j0 = t.jdesrecv[r]; sg.nx = t.nxrecv[r]; sg.ny = t.nyrecv[r];
            sg.buf = slot(xd.base, cnt, t.dirrecv[r]);
        }
        // publish into the neighbour's cell for this rank; wait on this rank's cell for the neighbour
        sp.flag[sp.npeers] = reinterpret_cast<unsigned long long *>(xd.peer_base[g.peer0]) + me0;
        sp.count[sp.npeers++] = cnt;
        rp.flag[rp.npeers] = reinterpret_cast<unsigned long long *>(xd.base) + g.peer0;
        rp.count[rp.npeers++] = cnt;
    }
    unsigned *ticket = reinterpret_cast<unsigned *>(xd.base + 1024);
    int *err = reinterpret_cast<int *>(xd.base + 1032);
    static const bool two_launches = std::getenv("DLESM_XDIRECT_TWO_LAUNCHES") != nullptr;
    if (two_launches) {
        CU(k::xdirect_send(sp, ticket, c.stream));
        CU(k::xdirect_recv(rp, err, c.stream));
    } else {
        CU(k::xdirect_exchange(sp, rp, ticket, err, c.stream));
    }
    done = true;
    return DLESM_OK;
}
}  // namespace

extern "C" int dlesm_dev_set_exchange_mode(int mode)
{
    clear_error();
    if (mode != DLESM_EXCHANGE_DIRECT && mode != DLESM_EXCHANGE_NCCL) {
        set_error(DLESM_ERR_ARGS, "set_exchange_mode: DLESM_EXCHANGE_DIRECT or DLESM_EXCHANGE_NCCL");
        return DLESM_ERR_ARGS;
    }
    ctx().exchange_mode = mode;
    return DLESM_OK;
}

// Collective.  Drops the registration of these fields: the neighbours' mappings they hold are
// released (closed when no other registered field uses the same allocation), then a barrier, after
// which each rank may free its own fields (CUDA requires importers to close before the exporter frees).
extern "C" int dlesm_dev_peer_unregister(void *const *fields, int nfields)
{
    clear_error();
    return peer_unregister_fields(fields, nfields);
}

// Collective.  Drops every registration made so far.
extern "C" int dlesm_dev_peer_unregister_all(void)
{
    clear_error();
    Context &c = ctx();
    std::vector<void *> all;
    for (const auto &kv : c.peers) all.push_back(const_cast<void *>(kv.first));
    return peer_unregister_fields(all.data(), static_cast<int>(all.size()));
}

extern "C" int dlesm_dev_peer_register(void *const *fields, int nfields)
{
    clear_error();
    return peer_register_with(fields, nfields, ctx().installed);
}

extern "C" int dlesm_dev_grid_upload(int nx, int ny, const int *tmask, const double *const metrics[16],
                                     void *device_ptrs[17])
{
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    Context &c = ctx();
    const size_t n = static_cast<size_t>(nx) * ny;
    for (int q = 0; q < 17; ++q) device_ptrs[q] = nullptr;
    if (tmask) {
        CU(cudaMalloc(&device_ptrs[0], n * sizeof(int)));
        CU(cudaMemcpyAsync(device_ptrs[0], tmask, n * sizeof(int), cudaMemcpyHostToDevice, c.stream));
    }
    for (int q = 0; q < 16; ++q) {
        if (!metrics || !metrics[q]) continue;
        rc = dlesm_dev_field_alloc(nx, ny, &device_ptrs[q + 1]);
        if (rc != DLESM_OK) return rc;
        CU(cudaMemcpyAsync(device_ptrs[q + 1], metrics[q], n * sizeof(double), cudaMemcpyHostToDevice, c.stream));
    }
    CU(cudaStreamSynchronize(c.stream));
    return DLESM_OK;
}

// Release what dlesm_dev_grid_upload handed out (grid_mod keeps the 17 values in its *_device slots): the
// tmask copy and every metric array that is not NULL; the slots are set to NULL.  Call before a second
// grid_init of the same grid and when the grid is dropped -- the reference has no grid finaliser, so
// nothing else will.
extern "C" int dlesm_dev_grid_release(void *device_ptrs[17])
{
    clear_error();
    if (!device_ptrs) return DLESM_OK;
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    CU(cudaStreamSynchronize(ctx().stream));   // no sweep may still be reading the arrays
    int first = DLESM_OK;
    if (device_ptrs[0]) {
        cudaError_t e = cudaFree(device_ptrs[0]);
        if (e != cudaSuccess) first = fail_cuda(e, "grid release: tmask");
        device_ptrs[0] = nullptr;
    }
    for (int q = 1; q < 17; ++q) {
        if (!device_ptrs[q]) continue;
        rc = dlesm_dev_field_free(device_ptrs[q]);
        if (rc != DLESM_OK && first == DLESM_OK) first = rc;
        device_ptrs[q] = nullptr;
    }
    return first;
}

// =========================================================================================
// device layer: stencils
// =========================================================================================
namespace {
int pitch_of(const void *any, int &pitch, int &ny)
{
    FieldRec r;
    int rc = lookup(any, r);
    if (rc != DLESM_OK) return rc;
    pitch = r.nx;
    ny = r.ny;
    return DLESM_OK;
}
int check_box(int pitch, int ny, int xs, int xe, int ys, int ye)
{
    // every stencil reaches one cell beyond the swept region
    if (xs < 2 || ys < 2 || xe + 1 > pitch || ye + 1 > ny) {
        set_error(DLESM_ERR_ARGS, "stencil sweep: region needs a one-cell rim inside the allocation");
        return DLESM_ERR_ARGS;
    }
    return DLESM_OK;
}
#define D(p) static_cast<double *>(p)
#define CD(p) static_cast<const double *>(p)
#define SWEEP_PROLOGUE(outptr)                                   \
    int pitch, ny_;                                              \
    int rc = pitch_of(outptr, pitch, ny_);                       \
    if (rc == DLESM_OK) rc = check_box(pitch, ny_, xs, xe, ys, ye); \
    if (rc != DLESM_OK) return rc;                               \
    mark_written(outptr);                                        \
    const k::Box b{xs, xe, ys, ye};                              \
    cudaStream_t st = ctx().stream;
}  // namespace

extern "C" int dlesm_compute_cu(void *cu, const void *p, const void *u, int xs, int xe, int ys, int ye, const void *tm)
{
    SWEEP_PROLOGUE(cu)
    CU(k::compute_cu(D(cu), CD(p), CD(u), pitch, b, static_cast<const int *>(tm), st));
    return DLESM_OK;
}
extern "C" int dlesm_compute_cv(void *cv, const void *p, const void *v, int xs, int xe, int ys, int ye, const void *tm)
{
    SWEEP_PROLOGUE(cv)
    CU(k::compute_cv(D(cv), CD(p), CD(v), pitch, b, static_cast<const int *>(tm), st));
    return DLESM_OK;
}
extern "C" int dlesm_compute_z(void *z, const void *p, const void *u, const void *v, double dx, double dy, int xs,
                               int xe, int ys, int ye, const void *tm)
{
    SWEEP_PROLOGUE(z)
    CU(k::compute_z(D(z), CD(p), CD(u), CD(v), pitch, b, 4.0 / dx, 4.0 / dy, static_cast<const int *>(tm), st));
    return DLESM_OK;
}
extern "C" int dlesm_compute_h(void *h, const void *p, const void *u, const void *v, int xs, int xe, int ys, int ye,
                               const void *tm)
{
    SWEEP_PROLOGUE(h)
    CU(k::compute_h(D(h), CD(p), CD(u), CD(v), pitch, b, static_cast<const int *>(tm), st));
    return DLESM_OK;
}
extern "C" int dlesm_compute_unew(void *unew, const void *uold, const void *z, const void *cv, const void *h,
                                  double tdt, double dx, int xs, int xe, int ys, int ye, const void *tm,
                                  const void *ucur)
{
    SWEEP_PROLOGUE(unew)
    if (tm && !ucur) {
        set_error(DLESM_ERR_ARGS, "compute_unew: masked sweep needs the current field");
        return DLESM_ERR_ARGS;
    }
    CU(k::compute_unew(D(unew), CD(uold), CD(z), CD(cv), CD(h), pitch, b, tdt / 8.0, tdt / dx,
                       static_cast<const int *>(tm), CD(ucur), st));
    return DLESM_OK;
}
extern "C" int dlesm_compute_vnew(void *vnew, const void *vold, const void *z, const void *cu, const void *h,
                                  double tdt, double dy, int xs, int xe, int ys, int ye, const void *tm,
                                  const void *vcur)
{
    SWEEP_PROLOGUE(vnew)
    if (tm && !vcur) {
        set_error(DLESM_ERR_ARGS, "compute_vnew: masked sweep needs the current field");
        return DLESM_ERR_ARGS;
    }
    CU(k::compute_vnew(D(vnew), CD(vold), CD(z), CD(cu), CD(h), pitch, b, tdt / 8.0, tdt / dy,
                       static_cast<const int *>(tm), CD(vcur), st));
    return DLESM_OK;
}
extern "C" int dlesm_compute_pnew(void *pnew, const void *pold, const void *cu, const void *cv, double tdt,
                                  double dx, double dy, int xs, int xe, int ys, int ye, const void *tm,
                                  const void *pcur)
{
    SWEEP_PROLOGUE(pnew)
    if (tm && !pcur) {
        set_error(DLESM_ERR_ARGS, "compute_pnew: masked sweep needs the current field");
        return DLESM_ERR_ARGS;
    }
    CU(k::compute_pnew(D(pnew), CD(pold), CD(cu), CD(cv), pitch, b, tdt / dx, tdt / dy,
                       static_cast<const int *>(tm), CD(pcur), st));
    return DLESM_OK;
}
extern "C" int dlesm_time_smooth(const void *f, const void *fnew, void *fold, double alpha, int xs, int xe, int ys,
                                 int ye, const void *tm, int points)
{
    SWEEP_PROLOGUE(fold)
    if (points != DLESM_GO_U_POINTS && points != DLESM_GO_V_POINTS && points != DLESM_GO_T_POINTS) {
        set_error(DLESM_ERR_POINTS, "time_smooth: points must be U, V or T");
        return DLESM_ERR_POINTS;
    }
    CU(k::time_smooth(CD(f), CD(fnew), D(fold), pitch, b, alpha, static_cast<const int *>(tm), points, st));
    return DLESM_OK;
}

namespace {
// byte image of an int32 tmask, built on first use and cached by the mask's device pointer
// (grid%tmask is immutable after grid_init); freed by release_mask8 / at finalise
int mask8_of(const void *tmask, int pitch, int ny, const unsigned char *&m8, int &pitch8)
{
    Context &c = ctx();
    auto it = c.mask8.find(tmask);
    if (it == c.mask8.end()) {
        Context::Mask8 m;
        m.pitch = (pitch + 15) & ~15;
        CU(cudaMalloc(&m.p, static_cast<size_t>(m.pitch) * ny));
        CU(k::mask_to_bytes(static_cast<const int *>(tmask), pitch, ny, m.p, m.pitch, c.stream));
        it = c.mask8.emplace(tmask, m).first;
    }
    m8 = it->second.p;
    pitch8 = it->second.pitch;
    return DLESM_OK;
}
void release_mask8(const void *tmask)
{
    Context &c = ctx();
    auto it = c.mask8.find(tmask);
    if (it == c.mask8.end()) return;
    if (c.device_ready) cudaStreamSynchronize(c.stream);
    k::fma_off::forget_mask(it->second.p);
    k::fma_on::forget_mask(it->second.p);
    cudaFree(it->second.p);
    c.mask8.erase(it);
}

struct ProfScope {
    cudaEvent_t e0 = nullptr, e1 = nullptr;
};
int prof_begin(ProfScope &ps, cudaStream_t st)
{
    Context &c = ctx();
    if (!c.profile) return DLESM_OK;
    if (c.prof_used == c.prof_events.size()) {
        cudaEvent_t x, y;
        CU(cudaEventCreate(&x));
        CU(cudaEventCreate(&y));
        c.prof_events.emplace_back(x, y);
    }
    ps.e0 = c.prof_events[c.prof_used].first;
    ps.e1 = c.prof_events[c.prof_used].second;
    ++c.prof_used;
    CU(cudaEventRecord(ps.e0, st));
    return DLESM_OK;
}
int prof_end(const ProfScope &ps, cudaStream_t st)
{
    if (ps.e1) CU(cudaEventRecord(ps.e1, st));
    return DLESM_OK;
}

// push boxes: one per send-list entry of `ct` (src/parallel_comms_mod.f90:70-83), aimed at the
// destination rank's counterparts of the `nout` output arrays
// `strips` (two-step sweep only): a box at most two columns wide is aimed at the packed slot of its
// direction tag in the neighbour's exchange allocation (dpitch = 0 marks it; dst[0] = the slot for the
// epoch the sweep will publish), which the neighbour scatters into its halo columns after the hand-shake
template <class Seg, int NOUT>
int make_push(const CommTables &ct, const void *const outs[NOUT], Seg *segs, int &nseg, k::PeerFlags &pf, const char *who,
              bool strips = false)
{
    Context &c = ctx();
    if (!ct.valid) {
        set_error(DLESM_ERR_STATE, std::string(who) + ": halo push needs the comm tables (dlesm_dev_set_comms)");
        return DLESM_ERR_STATE;
    }
    const Context::PeerView *pv[NOUT];
    for (int q = 0; q < NOUT; ++q) {
        auto it = c.peers.find(outs[q]);
        if (it == c.peers.end()) {
            set_error(DLESM_ERR_STATE, std::string(who) + ": halo push needs the output fields registered (dlesm_dev_peer_register)");
            return DLESM_ERR_STATE;
        }
        pv[q] = &it->second;
    }
    const dlesm_comms &t = ct.t;
    const int me0 = pstate().rank - 1;
    const bool single = pstate().nranks == 1;
    bool seen[64] = {false};
    nseg = 0;
    for (int sidx = 0; sidx < t.nsend; ++sidx) {
        const int dest = single ? 0 : t.destination[sidx];
        Seg &g = segs[nseg++];
        g.i0 = t.isrcsend[sidx]; g.i1 = g.i0 + t.nxsend[sidx] - 1;
        g.j0 = t.jsrcsend[sidx]; g.j1 = g.j0 + t.nysend[sidx] - 1;
        g.ides = t.idessend[sidx]; g.jdes = t.jdessend[sidx];
        if (dest < 0 || dest >= static_cast<int>(pv[0]->by_rank.size())) {
            set_error(DLESM_ERR_STATE, std::string(who) + ": send list names a rank outside the communicator");
            return DLESM_ERR_STATE;
        }
        g.dpitch = pv[0]->pitch_by_rank[dest];
        const bool packed = strips && g.i1 - g.i0 < 2 && c.xd.ready && t.dirsend[sidx] >= 1 && t.dirsend[sidx] <= 8 &&
                            dest < static_cast<int>(c.xd.peer_base.size()) && c.xd.peer_base[dest] &&
                            static_cast<size_t>(t.nxsend[sidx]) * t.nysend[sidx] * NOUT <= c.xd.pslot_cap;
        for (int q = 0; q < NOUT; ++q) {
            g.dst[q] = pv[q]->by_rank[dest];
            if (!g.dst[q]) {
                set_error(DLESM_ERR_STATE, std::string(who) + ": a neighbour's field is not mapped");
                return DLESM_ERR_STATE;
            }
        }
        if (packed) {
            g.dpitch = 0;
            g.dst[0] = push_slot(c.xd.peer_base[dest], c.push_epoch + 1, t.dirsend[sidx]);
        }
        if (dest != me0 && !seen[dest]) {
            seen[dest] = true;
            pf.remote[pf.n] = c.peer_flags[dest] + me0;
            pf.local[pf.n] = c.flags + dest;
            ++pf.n;
        }
    }
    return DLESM_OK;
}

// The fused step.  `push_tables` non-null: halo push along that send list.
// The hand-shake of a push sweep rides in the sweep itself (k::PushSync): its edge tiles wait for the
// neighbours of the previous, still unsettled push sweep, its last CTA publishes the new epoch.
// DLESM_PUSH_SYNC=kernel keeps round 1's form for A/B runs: a one-warp kernel after every push sweep
// publishes the epoch and waits for the neighbours'.  Must agree across ranks.
bool push_sync_in_sweep()
{
    static const bool in_sweep = [] {
        const char *e = std::getenv("DLESM_PUSH_SYNC");
        return !(e && std::strcmp(e, "kernel") == 0);
    }();
    return in_sweep;
}
void push_sync_fill(Context &c, const k::PeerFlags &pf, k::PushSync &sy)
{
    sy = k::PushSync{};
    if (pf.n <= 0 || !push_sync_in_sweep()) return;
    if (c.push_pending) {
        sy.n_wait = c.pending_pf.n;
        for (int q = 0; q < sy.n_wait; ++q) sy.local[q] = c.pending_pf.local[q];
        sy.wait_epoch = c.push_epoch;
    }
    sy.n = pf.n;
    for (int q = 0; q < pf.n; ++q) sy.remote[q] = pf.remote[q];
    sy.publish = 1;
    sy.epoch = c.push_epoch + 1;
    sy.ticket = c.push_ticket;
    sy.err = c.push_err;
}
// after the launch: the sweep publishes push_epoch + 1.  `deferred`: the caller sweeps again (or settles)
// before anything else reads the pushed halos; otherwise wait for the neighbours right here.
int push_sync_done(Context &c, const k::PeerFlags &pf, bool deferred, cudaStream_t st)
{
    if (pf.n <= 0) return DLESM_OK;
    ++c.push_epoch;
    if (!push_sync_in_sweep()) {
        CU(k::peer_signal_wait(pf, c.push_epoch, c.push_err, st, /*publish=*/true));
        return DLESM_OK;
    }
    c.pending_pf = pf;
    c.push_pending = true;
    if (deferred) return DLESM_OK;
    c.push_pending = false;
    CU(k::peer_signal_wait(pf, c.push_epoch, c.push_err, st, /*publish=*/false));
    return DLESM_OK;
}
int push_settle()
{
    Context &c = ctx();
    if (!c.push_pending) return DLESM_OK;
    c.push_pending = false;
    CU(k::peer_signal_wait(c.pending_pf, c.push_epoch, c.push_err, c.stream, /*publish=*/false));
    return DLESM_OK;
}

int sw_step_fused_impl(const void *u, const void *v, const void *p, void *uold, void *vold, void *pold, void *unew,
                       void *vnew, void *pnew, double dx, double dy, double tdt, double alpha, int first, int xs,
                       int xe, int ys, int ye, const void *tm, int variant, const CommTables *push_tables,
                       bool deferred = false)
{
    SWEEP_PROLOGUE(unew)
    (void)b;
    const void *all[] = {u, v, p, uold, vold, pold, vnew, pnew};
    for (const void *q : all) {
        FieldRec r;
        rc = lookup(q, r);
        if (rc != DLESM_OK) return rc;
        if (r.nx != pitch || r.ny != ny_) {
            set_error(DLESM_ERR_ARGS, "sw_step_fused: fields have different shapes");
            return DLESM_ERR_ARGS;
        }
    }
    Context &c = ctx();
    k::FusedArgs a;
    a.u = CD(u); a.v = CD(v); a.p = CD(p);
    a.uold = D(uold); a.vold = D(vold); a.pold = D(pold);
    a.unew = D(unew); a.vnew = D(vnew); a.pnew = D(pnew);
    a.tmask = static_cast<const int *>(tm);
    a.mask8 = nullptr; a.mask8_pitch = 0;
    a.pitch = pitch; a.ny = ny_;
    a.b = k::Box{xs, xe, ys, ye};
    a.sc = k::make_scalars(dx, dy, tdt, alpha, first);
    a.tiles = c.sweep_tiles;
    a.push.n = 0;
    if (tm && variant != 1 && pitch % 2 == 0) {
        rc = mask8_of(tm, pitch, ny_, a.mask8, a.mask8_pitch);
        if (rc != DLESM_OK) return rc;
    }
    const bool fma = c.arith == DLESM_ARITH_CONTRACTED;
    k::PeerFlags pf{};
    if (push_tables) {
        const bool ok = fma ? k::fma_on::fused_push_supported(a, variant) : k::fma_off::fused_push_supported(a, variant);
        if (!ok) {
            set_error(DLESM_ERR_ARGS, "sw_step_fused: halo push needs a TMA variant (even pitch)");
            return DLESM_ERR_ARGS;
        }
        const void *outs[3] = {unew, vnew, pnew};
        rc = make_push<k::PushSeg, 3>(*push_tables, outs, a.push.seg, a.push.n, pf, "sw_step_fused");
        if (rc != DLESM_OK) return rc;
    }
    push_sync_fill(c, pf, a.sync);
    ProfScope ps;
    rc = prof_begin(ps, st);
    if (rc != DLESM_OK) return rc;
    cudaError_t e = fma ? k::fma_on::sw_step_fused(a, variant, st) : k::fma_off::sw_step_fused(a, variant, st);
    if (e == cudaErrorNotSupported || e == cudaErrorInvalidValue) {
        set_error(DLESM_ERR_ARGS, "sw_step_fused: variant not available for this pitch / mask (see DESIGN.md)");
        return DLESM_ERR_ARGS;
    }
    if (e != cudaSuccess) return fail_cuda(e, "sw_step_fused");
    rc = prof_end(ps, st);
    if (rc != DLESM_OK) return rc;
    // the six arrays were written; with a push the halos of the new fields are filled as well
    const int depth_new = push_tables ? 1 : 0;
    mark_clean(unew, depth_new); mark_clean(vnew, depth_new); mark_clean(pnew, depth_new);
    if (!first) { mark_written(uold); mark_written(vold); mark_written(pold); }
    // neighbours must not start their next sweep before these pushes landed, nor push into a buffer this
    // rank is still reading: the sweep published its epoch, someone waits for theirs
    return push_sync_done(c, pf, deferred, st);
}

int sw_step2_fused_impl(const void *u, const void *v, const void *p, const void *uold, const void *vold,
                        const void *pold, void *u2, void *v2, void *p2, void *uold2, void *vold2, void *pold2,
                        double dx, double dy, double tdt, double alpha, int xs, int xe, int ys, int ye,
                        const unsigned char *mask8, int mask8_pitch, int variant, const CommTables *push_tables,
                        bool deferred = false)
{
    SWEEP_PROLOGUE(u2)
    (void)b;
    const void *all[] = {u, v, p, uold, vold, pold, v2, p2, uold2, vold2, pold2};
    for (const void *q : all) {
        FieldRec r;
        rc = lookup(q, r);
        if (rc != DLESM_OK) return rc;
        if (r.nx != pitch || r.ny != ny_) {
            set_error(DLESM_ERR_ARGS, "sw_step2_fused: fields have different shapes");
            return DLESM_ERR_ARGS;
        }
    }
    const void *ins[] = {u, v, p, uold, vold, pold};
    const void *outs[6] = {u2, v2, p2, uold2, vold2, pold2};
    for (const void *a : ins)
        for (const void *o : outs)
            if (a == o) {
                set_error(DLESM_ERR_ARGS, "sw_step2_fused: outputs must not alias inputs (neighbouring blocks re-read the ring)");
                return DLESM_ERR_ARGS;
            }
    Context &c = ctx();
    k::Fused2Args a;
    a.u = CD(u); a.v = CD(v); a.p = CD(p);
    a.uold = CD(uold); a.vold = CD(vold); a.pold = CD(pold);
    a.u2 = D(u2); a.v2 = D(v2); a.p2 = D(p2);
    a.mask8 = mask8; a.mask8_pitch = mask8_pitch;
    a.uold2 = D(uold2); a.vold2 = D(vold2); a.pold2 = D(pold2);
    a.pitch = pitch; a.ny = ny_;
    a.b = k::Box{xs, xe, ys, ye};
    a.sc = k::make_scalars(dx, dy, tdt, alpha, 0);
    a.tiles = c.sweep_tiles;
    const bool fma = c.arith == DLESM_ARITH_CONTRACTED;
    if (!(fma ? k::fma_on::fused2_supported(a) : k::fma_off::fused2_supported(a))) {
        set_error(DLESM_ERR_ARGS, "sw_step2_fused: needs an even pitch and a box with two halo cells on every side (halo_width >= 2)");
        return DLESM_ERR_ARGS;
    }
    a.push.n = 0;
    a.push.strip = nullptr;
    k::PeerFlags pf{};
    k::StripScatter sp{};
    bool strips = false;
    if (push_tables) {
        // the tables must be the depth-2 ones
        if (!k::fma_off::fused2_push_supported(variant)) {
            set_error(DLESM_ERR_ARGS, "sw_step2_fused: halo push needs the shuffle form of the sweep (variant 0 or 10..19)");
            return DLESM_ERR_ARGS;
        }
        // x-edge and corner boxes (at most two columns wide) leave through a local staging strip in ONE
        // burst per CTA into the packed slot of their direction in the neighbour's exchange allocation, and
        // the neighbour scatters them into its halo columns after the hand-shake (k::PushPlan2,
        // k::StripScatter): a peer store per row and field touched a new remote page every time and made the
        // edge CTAs 3.5x longer.  DLESM_PUSH_DEFER=0 (on every rank) keeps the per-row peer stores for A/B runs.
        static const bool defer = [] {
            const char *e = std::getenv("DLESM_PUSH_DEFER");
            return !(e && std::strcmp(e, "0") == 0);
        }();
        strips = defer && c.xd.ready;
        rc = make_push<k::PushSeg2, 6>(*push_tables, outs, a.push.seg, a.push.n, pf, "sw_step2_fused", strips);
        if (rc != DLESM_OK) return rc;
        if (strips) {
            for (int q = 0; q < a.push.n; ++q) {
                const k::PushSeg2 &g = a.push.seg[q];
                if (g.i1 - g.i0 < 2 && !((g.i0 >= xs && g.i1 <= xs + 1) || (g.i0 >= xe - 1 && g.i1 <= xe))) {
                    set_error(DLESM_ERR_STATE, "sw_step2_fused: a narrow push box away from the tile's x edges");
                    return DLESM_ERR_STATE;
                }
            }
            const size_t need = static_cast<size_t>(2) * (ye - ys + 1) * 12;
            if (c.push_strip_cap < need) {
                if (c.push_strip) {
                    CU(cudaStreamSynchronize(st));
                    CU(cudaFree(c.push_strip));
                    c.push_strip = nullptr;
                    c.push_strip_cap = 0;
                }
                CU(cudaMalloc(&c.push_strip, need * sizeof(double)));
                c.push_strip_cap = need;
            }
            a.push.strip = c.push_strip;
            // the receiving end: this rank's narrow receive boxes, from the slots of their direction tags
            const dlesm_comms &t = push_tables->t;
            sp.pitch = a.pitch;
            for (int q = 0; q < 6; ++q) sp.dst[q] = const_cast<double *>(static_cast<const double *>(outs[q]));
            for (int r = 0; r < t.nrecv && sp.n < 8; ++r) {
                if (t.nxrecv[r] > 2 || t.dirrecv[r] < 1 || t.dirrecv[r] > 8 ||
                    static_cast<size_t>(t.nxrecv[r]) * t.nyrecv[r] * 6 > c.xd.pslot_cap)
                    continue;
                k::StripScatter::Item &it = sp.item[sp.n++];
                it.ides = t.idesrecv[r]; it.jdes = t.jdesrecv[r]; it.nx = t.nxrecv[r]; it.ny = t.nyrecv[r];
                it.slot = push_slot(c.xd.base, c.push_epoch + 1, t.dirrecv[r]);
            }
        }
    }
    push_sync_fill(c, pf, a.sync);
    ProfScope ps;
    rc = prof_begin(ps, st);
    if (rc != DLESM_OK) return rc;
    cudaError_t e = fma ? k::fma_on::sw_step2_fused(a, variant, st) : k::fma_off::sw_step2_fused(a, variant, st);
    if (e == cudaErrorNotSupported || e == cudaErrorInvalidValue) {
        set_error(DLESM_ERR_ARGS, "sw_step2_fused: variant not available (see DESIGN.md)");
        return DLESM_ERR_ARGS;
    }
    if (e != cudaSuccess) return fail_cuda(e, "sw_step2_fused");
    rc = prof_end(ps, st);
    if (rc != DLESM_OK) return rc;
    for (const void *o : outs) mark_clean(o, push_tables ? 2 : 0);
    if (sp.n > 0) {
        // hand-shake and scatter in one small launch: wait for the neighbours' epoch (the sweep's last CTA
        // published this rank's; with DLESM_PUSH_SYNC=kernel this launch does), then move the strips they
        // left here into the halo columns
        ++c.push_epoch;
        c.push_pending = false;
        CU(k::peer_wait_scatter(pf, c.push_epoch, c.push_err, !push_sync_in_sweep(), sp, st));
        return DLESM_OK;
    }
    return push_sync_done(c, pf, deferred, st);
}
}  // namespace

extern "C" int dlesm_sw_step_fused(const void *u, const void *v, const void *p, void *uold, void *vold, void *pold,
                                   void *unew, void *vnew, void *pnew, double dx, double dy, double tdt,
                                   double alpha, int first, int xs, int xe, int ys, int ye, const void *tm,
                                   int variant)
{
    clear_error();
    const bool push = (variant & DLESM_FUSED_PUSH) != 0;
    return sw_step_fused_impl(u, v, p, uold, vold, pold, unew, vnew, pnew, dx, dy, tdt, alpha, first, xs, xe, ys, ye,
                              tm, variant & ~DLESM_FUSED_PUSH, push ? &ctx().installed : nullptr);
}

extern "C" int dlesm_sw_step2_fused(const void *u, const void *v, const void *p, const void *uold, const void *vold,
                                    const void *pold, void *u2, void *v2, void *p2, void *uold2, void *vold2,
                                    void *pold2, double dx, double dy, double tdt, double alpha, int xs, int xe,
                                    int ys, int ye, const void *tm, int variant)
{
    clear_error();
    const bool push = (variant & DLESM_FUSED_PUSH) != 0;
    const unsigned char *m8 = nullptr;
    int p8 = 0;
    if (tm) {
        int pitch, ny_;
        int rc = pitch_of(u2, pitch, ny_);
        if (rc == DLESM_OK) rc = mask8_of(tm, pitch, ny_, m8, p8);
        if (rc != DLESM_OK) return rc;
    }
    return sw_step2_fused_impl(u, v, p, uold, vold, pold, u2, v2, p2, uold2, vold2, pold2, dx, dy, tdt, alpha, xs, xe,
                               ys, ye, m8, p8, variant & ~DLESM_FUSED_PUSH, push ? &ctx().installed : nullptr);
}

extern "C" int dlesm_set_arith(int mode)
{
    clear_error();
    if (mode != DLESM_ARITH_EXACT && mode != DLESM_ARITH_CONTRACTED) {
        set_error(DLESM_ERR_ARGS, "dlesm_set_arith: DLESM_ARITH_EXACT or DLESM_ARITH_CONTRACTED");
        return DLESM_ERR_ARGS;
    }
    ctx().arith = mode;
    return DLESM_OK;
}
extern "C" int dlesm_get_arith(void) { return ctx().arith; }

extern "C" int dlesm_dev_set_sweep_tiles(int ntilex, int ntiley)
{
    clear_error();
    if (ntilex < 0 || ntiley < 0) {
        set_error(DLESM_ERR_ARGS, "dlesm_dev_set_sweep_tiles: tile counts must be >= 0 (0 = library default)");
        return DLESM_ERR_ARGS;
    }
    ctx().sweep_tiles = k::SweepTiles{ntilex, ntiley};
    return DLESM_OK;
}

extern "C" long long dlesm_kernel_launch_count(void) { return k::launch_count(); }
extern "C" int dlesm_profile_enable(int on)
{
    ctx().profile = on != 0;
    return DLESM_OK;
}
extern "C" int dlesm_profile_read(double *ms_total, long long *nlaunches, int reset)
{
    Context &c = ctx();
    double tot = 0.0;
    if (c.device_ready) CU(cudaStreamSynchronize(c.stream));
    for (size_t q = 0; q < c.prof_used; ++q) {
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, c.prof_events[q].first, c.prof_events[q].second));
        tot += ms;
    }
    if (ms_total) *ms_total = tot;
    if (nlaunches) *nlaunches = static_cast<long long>(c.prof_used);
    if (reset) c.prof_used = 0;
    return DLESM_OK;
}

// Diagnostics: per-CTA timeline of the TMA sweeps launched between begin and end (each launch overwrites
// the entries of its own CTAs: trace ONE launch shape at a time).  4 long long per CTA
// (by * gridDim.x + bx): start ns, end ns, SM id, ns after the neighbour wait.
namespace {
long long *g_trace_buf = nullptr;
int g_trace_cap = 0;
}  // namespace
extern "C" int dlesm_debug_cta_trace_begin(int max_ctas)
{
    clear_error();
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    if (max_ctas <= 0) {
        set_error(DLESM_ERR_ARGS, "dlesm_debug_cta_trace_begin: max_ctas must be positive");
        return DLESM_ERR_ARGS;
    }
    if (g_trace_buf) cudaFree(g_trace_buf);
    g_trace_buf = nullptr;
    CU(cudaMalloc(&g_trace_buf, static_cast<size_t>(max_ctas) * 4 * sizeof(long long)));
    CU(cudaMemset(g_trace_buf, 0, static_cast<size_t>(max_ctas) * 4 * sizeof(long long)));
    g_trace_cap = max_ctas;
    CU(k::fma_off::set_cta_trace(g_trace_buf));
    CU(k::fma_on::set_cta_trace(g_trace_buf));
    return DLESM_OK;
}
extern "C" int dlesm_debug_cta_trace_end(long long *out, int max_ctas)
{
    clear_error();
    if (!g_trace_buf) {
        set_error(DLESM_ERR_STATE, "dlesm_debug_cta_trace_end: no trace in progress");
        return DLESM_ERR_STATE;
    }
    Context &c = ctx();
    CU(cudaStreamSynchronize(c.stream));
    CU(k::fma_off::set_cta_trace(nullptr));
    CU(k::fma_on::set_cta_trace(nullptr));
    if (out)
        CU(cudaMemcpy(out, g_trace_buf, static_cast<size_t>(std::min(max_ctas, g_trace_cap)) * 4 * sizeof(long long),
                      cudaMemcpyDeviceToHost));
    cudaFree(g_trace_buf);
    g_trace_buf = nullptr;
    g_trace_cap = 0;
    return DLESM_OK;
}

// =========================================================================================
// HOST MIRROR: grid_type
// =========================================================================================
struct dlesm_grid {
    dlesm_grid_info info{};
    dlesm_decomp dh;                  // handed out by dlesm_grid_decomp
    Decomposition &decomp = dh.d;
    bool decomposed = false;
    bool user_tmask = false;
    std::vector<int> tmask;           // materialised lazily when no mask was supplied
    std::vector<double> xt, yt;       // 1-D coordinate vectors (xt(ji,:) and yt(:,jj))
    dlesm_comms comms{};
    bool have_comms = false;
    bool pbc_x = false, pbc_y = false;
    void *dev[17] = {nullptr};
    int tile_id = 0;                  // > 0: one of several sub-domains of this rank (dlesm_tileset), 1-based id
};

extern "C" dlesm_grid_t *dlesm_grid_create(int grid_name, const int bcs[3], int grid_offsets)
{
    clear_error();
    if (grid_offsets < 0) {
        set_error(DLESM_ERR_GRID_OFFSET, "ERROR: grid offset not specified in call to grid_constructor.");
        return nullptr;
    }
    if (grid_name != DLESM_GO_ARAKAWA_C && grid_name != DLESM_GO_ARAKAWA_B) {
        set_error(DLESM_ERR_GRID_NAME, "grid_constructor: ERROR: unsupported grid type");
        return nullptr;
    }
    if (grid_offsets != DLESM_GO_OFFSET_NE && grid_offsets != DLESM_GO_OFFSET_NW &&
        grid_offsets != DLESM_GO_OFFSET_SE && grid_offsets != DLESM_GO_OFFSET_SW) {
        set_error(DLESM_ERR_GRID_OFFSET, "grid_constructor: ERROR: unsupported relative offsets of grid types");
        return nullptr;
    }
    auto *g = new dlesm_grid;
    g->info.name = grid_name;
    g->info.offset = grid_offsets;
    for (int q = 0; q < 3; ++q) g->info.boundary_conditions[q] = bcs[q];
    return g;
}

extern "C" int dlesm_grid_decompose(dlesm_grid_t *g, int domainx, int domainy, int ndomains, int ndomainx,
                                    int ndomainy, int halo_width)
{
    clear_error();
    int rc = go_decompose(domainx, domainy, ndomains, ndomainx, ndomainy, halo_width, pstate().nranks, g->decomp);
    if (rc != DLESM_OK) return rc;
    if (pstate().rank > g->decomp.ndomains) {
        set_error(DLESM_ERR_ARGS, "grid%decompose: this rank has no sub-domain");
        return DLESM_ERR_ARGS;
    }
    g->info.subdomain = g->decomp.subdomains[pstate().rank - 1];  // src/grid_mod.f90:207
    g->info.global_nx = g->decomp.global_nx;
    g->info.global_ny = g->decomp.global_ny;
    g->decomposed = true;
    return DLESM_OK;
}

extern "C" int dlesm_grid_init(dlesm_grid_t *g, double dxarg, double dyarg, const int *tmask, int tmask_nx,
                               int tmask_ny)
{
    clear_error();
    if (!g->decomposed) {
        set_error(DLESM_ERR_STATE, "grid_init: grid%decompose must be called first");
        return DLESM_ERR_STATE;
    }
    int alignment = 1;
    int rc = parse_alignment(std::getenv("DL_ESM_ALIGNMENT"), alignment);
    if (rc != DLESM_OK) return rc;
    GridDims gd;
    rc = grid_dims(g->info.subdomain, alignment, gd);
    if (rc != DLESM_OK) return rc;
    const dlesm_subdomain &sub = g->info.subdomain;
    const bool px = g->info.boundary_conditions[0] == DLESM_GO_BC_PERIODIC;
    const bool py = g->info.boundary_conditions[1] == DLESM_GO_BC_PERIODIC;
    // a tile of a tile set gets its halos from the set's own exchange plan (periodic wrap included)
    const bool dist = pstate().nranks > 1 && g->tile_id == 0;
    const bool ext = (pstate().extensions & DLESM_EXT_PERIODIC_DISTMEM) != 0;
    if (dist && (px || py) && !ext) {
        // src/grid_mod.f90:437-442 and :559-564
        set_error(DLESM_ERR_PBC_DISTMEM, "grid_init: ERROR: Periodic boundary conditions are not yet supported.");
        return DLESM_ERR_PBC_DISTMEM;
    }
    if (tmask) {
        if (tmask_nx < sub.internal.xstop + 1 || tmask_ny < sub.internal.ystop + 1) {
            set_error(DLESM_ERR_ARGS, "grid_init: tmask smaller than the sub-domain it must cover");
            return DLESM_ERR_ARGS;
        }
        fill_tmask(sub, gd.nx, gd.ny, tmask, tmask_nx, g->tmask);
        g->user_tmask = true;
    }
    g->info.nx = gd.nx;
    g->info.ny = gd.ny;
    g->info.alignment = alignment;
    g->info.padding = gd.padding;
    g->info.dx = dxarg;
    g->info.dy = dyarg;
    coord_vectors(sub, gd.nx, gd.ny, dxarg, dyarg, g->xt, g->yt);
    g->pbc_x = px && dist;
    g->pbc_y = py && dist;
    if (dist) {
        // map_comms(decomp, tmask, .false., (/1,1/), ierr) -- src/grid_mod.f90:566
        rc = build_comms(g->decomp, pstate().rank, pstate().nranks, g->pbc_x, g->pbc_y, 1, 1, g->comms);
        if (rc != DLESM_OK) {
            set_error(rc, "Set-up of communication tables (call to map_comms()) failed.");
            return rc;
        }
        g->have_comms = true;
        rc = dlesm_dev_set_comms(&g->comms);  // module-level state, as in parallel_comms_mod
        if (rc != DLESM_OK) return rc;
    } else if (g->tile_id == 0) {
        dlesm_dev_set_comms(nullptr);         // one rank: map_comms finds no neighbour (stub build)
    }
    g->info.initialised = 1;
    return DLESM_OK;
}

extern "C" int dlesm_grid_get_info(const dlesm_grid_t *g, dlesm_grid_info *out)
{
    *out = g->info;
    return DLESM_OK;
}
extern "C" const dlesm_decomp_t *dlesm_grid_decomp(const dlesm_grid_t *g)
{
    return &g->dh;
}
extern "C" int dlesm_grid_comms(const dlesm_grid_t *g, dlesm_comms *out)
{
    if (!g->have_comms) {
        dlesm_comms empty;
        std::memset(&empty, 0, sizeof(empty));
        *out = empty;
        return DLESM_OK;
    }
    *out = g->comms;
    return DLESM_OK;
}

namespace {
double metric_constant(const dlesm_grid &g, int which, bool &is_const)
{
    is_const = true;
    switch (which) {
    case DLESM_GRID_DX_T: case DLESM_GRID_DX_U: case DLESM_GRID_DX_V: case DLESM_GRID_DX_F: return g.info.dx;
    case DLESM_GRID_DY_T: case DLESM_GRID_DY_U: case DLESM_GRID_DY_V: case DLESM_GRID_DY_F: return g.info.dy;
    case DLESM_GRID_AREA_T: case DLESM_GRID_AREA_U: case DLESM_GRID_AREA_V: return g.info.dx * g.info.dy;
    case DLESM_GRID_GPHIU: case DLESM_GRID_GPHIV: case DLESM_GRID_GPHIF: return 50.0;
    default: is_const = false; return 0.0;
    }
}
void ensure_tmask(dlesm_grid &g)
{
    if (g.tmask.empty()) fill_tmask(g.info.subdomain, g.info.nx, g.info.ny, nullptr, 0, g.tmask);
}
}  // namespace

extern "C" int dlesm_grid_host_array(dlesm_grid_t *g, int which, void *out)
{
    clear_error();
    if (!g->info.initialised) {
        set_error(DLESM_ERR_STATE, "grid arrays exist only after grid_init");
        return DLESM_ERR_STATE;
    }
    const int nx = g->info.nx, ny = g->info.ny;
    const size_t n = static_cast<size_t>(nx) * ny;
    if (which == DLESM_GRID_TMASK) {
        ensure_tmask(*g);
        std::memcpy(out, g->tmask.data(), n * sizeof(int));
        return DLESM_OK;
    }
    double *o = static_cast<double *>(out);
    bool is_const;
    const double cval = metric_constant(*g, which, is_const);
    if (is_const) {
        std::fill(o, o + n, cval);
    } else if (which == DLESM_GRID_XT) {
        for (int j = 0; j < ny; ++j) std::copy(g->xt.begin(), g->xt.end(), o + static_cast<size_t>(j) * nx);
    } else if (which == DLESM_GRID_YT) {
        for (int j = 0; j < ny; ++j) std::fill(o + static_cast<size_t>(j) * nx, o + static_cast<size_t>(j + 1) * nx, g->yt[j]);
    } else {
        set_error(DLESM_ERR_ARGS, "unknown grid array id");
        return DLESM_ERR_ARGS;
    }
    return DLESM_OK;
}

extern "C" int dlesm_grid_device_array(dlesm_grid_t *g, int which, void **device_ptr)
{
    clear_error();
    if (!g->info.initialised || which < 0 || which > 16) {
        set_error(DLESM_ERR_STATE, "grid device arrays exist only after grid_init");
        return DLESM_ERR_STATE;
    }
    if (g->dev[which]) {
        *device_ptr = g->dev[which];
        return DLESM_OK;
    }
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    Context &c = ctx();
    const int nx = g->info.nx, ny = g->info.ny;
    const size_t n = static_cast<size_t>(nx) * ny;
    if (which == DLESM_GRID_TMASK) {
        ensure_tmask(*g);
        void *p = nullptr;
        CU(cudaMalloc(&p, n * sizeof(int)));
        CU(cudaMemcpyAsync(p, g->tmask.data(), n * sizeof(int), cudaMemcpyHostToDevice, c.stream));
        CU(cudaStreamSynchronize(c.stream));
        g->dev[which] = p;
    } else {
        void *p = nullptr;
        rc = dlesm_dev_field_alloc(nx, ny, &p);
        if (rc != DLESM_OK) return rc;
        bool is_const;
        const double cval = metric_constant(*g, which, is_const);
        if (is_const) {
            CU(k::set_field(static_cast<double *>(p), n, cval, c.stream));
        } else {
            const std::vector<double> &vec = which == DLESM_GRID_XT ? g->xt : g->yt;
            double *dv = nullptr;
            CU(cudaMalloc(&dv, vec.size() * sizeof(double)));
            CU(cudaMemcpyAsync(dv, vec.data(), vec.size() * sizeof(double), cudaMemcpyHostToDevice, c.stream));
            if (which == DLESM_GRID_XT) CU(k::broadcast_x(static_cast<double *>(p), nx, ny, dv, c.stream));
            else CU(k::broadcast_y(static_cast<double *>(p), nx, ny, dv, c.stream));
            CU(cudaStreamSynchronize(c.stream));
            CU(cudaFree(dv));
        }
        g->dev[which] = p;
    }
    *device_ptr = g->dev[which];
    return DLESM_OK;
}

extern "C" void dlesm_grid_free(dlesm_grid_t *g)
{
    if (!g) return;
    if (g->dev[0]) {
        release_mask8(g->dev[0]);
        cudaFree(g->dev[0]);
    }
    for (int q = 1; q < 17; ++q)
        if (g->dev[q]) dlesm_dev_field_free(g->dev[q]);
    delete g;
}

// =========================================================================================
// HOST MIRROR: r2d_field
// =========================================================================================
struct dlesm_field {
    dlesm_grid *grid = nullptr;
    FieldBounds fb;
    int nx = 0, ny = 0;
    int ntiles = 0;
    std::vector<dlesm_region> tile_internal, tile_whole;   // tile(:)%internal / %whole (src/tile_mod.f90:36-42)
    void *dptr = nullptr;
};

namespace {
// get_grid_dims (src/field_mod.f90:1473-1503): GOCEAN_OMP_GRID = "AxB"
bool get_grid_dims(int &nx, int &ny)
{
    const char *e = std::getenv("GOCEAN_OMP_GRID");
    if (!e) return false;
    const char *x = std::strchr(e, 'x');
    if (!x) {
        std::fprintf(stderr, "\ndlesm get_grid_dims: failed to parse GOCEAN_OMP_GRID string: %s\n"
                             "   -  will use defaults for dimensions of tiling grid\n", e);
        return false;
    }
    char *end = nullptr;
    const long a = std::strtol(e, &end, 10);
    if (end == e || end != x || a < 1) return false;
    const long b = std::strtol(x + 1, &end, 10);
    if (end == x + 1 || b < 1) return false;
    nx = static_cast<int>(a);
    ny = static_cast<int>(b);
    return true;
}

// do_tile (src/field_mod.f90:300-325).  The reference cuts the field's internal region into one tile
// per OpenMP thread with go_decompose and keeps that decomposition for every later field (SAVEd
// `decomp`, `tiling_initialised`).  Here a tile is the unit of work of one CTA column-strip march:
// GOCEAN_OMP_GRID="AxB" gives the tile grid (A x-tiles, each cut into strips of the sweep's width; B
// tiles = segments along j); without it the library's own choice stands (one x-tile, the segment
// count the fused sweep picks for whole waves of CTAs).  The first tiled field installs the grid for
// every sweep (module-level state, like the reference's), dlesm_dev_set_sweep_tiles overrides.
int setup_tiles(dlesm_field *f)
{
    Context &c = ctx();
    const dlesm_region &in = f->fb.internal;
    if (!c.tiling_initialised) {
        int a = 1, b = 0;
        if (get_grid_dims(a, b)) c.sweep_tiles = k::SweepTiles{a, b};
        c.tiling_initialised = true;
    }
    int ntx = c.sweep_tiles.nx > 0 ? c.sweep_tiles.nx : 1, nty = c.sweep_tiles.ny;
    if (nty <= 0) {
        // the default segment count of the one-step fused sweep over this box
        k::FusedArgs a{};
        a.pitch = f->nx - (f->nx % 2); a.ny = f->ny;   // geometry only
        a.b = box_of(in);
        a.tiles = k::SweepTiles{0, 0};
        int grid[2] = {1, 1};
        k::fma_off::fused_grid(a, 25, grid);
        nty = grid[1] > 0 ? grid[1] : 1;
    }
    ntx = std::min(ntx, std::max(in.nx, 1));
    nty = std::min(nty, std::max(in.ny, 1));
    // tile boxes by go_decompose's remainder rule (src/parallel_mod.f90:211-223), in the field's
    // own index space: internal = the tile's points, whole = one cell more on every side
    f->ntiles = ntx * nty;
    f->tile_internal.resize(f->ntiles);
    f->tile_whole.resize(f->ntiles);
    const int tw = in.nx / ntx, xr = in.nx % ntx, th = in.ny / nty, yr = in.ny % nty;
    for (int b = 0; b < nty; ++b)
        for (int a = 0; a < ntx; ++a) {
            dlesm_region r;
            r.xstart = in.xstart + a * tw + std::min(a, xr);
            r.nx = tw + (a < xr ? 1 : 0);
            r.xstop = r.xstart + r.nx - 1;
            r.ystart = in.ystart + b * th + std::min(b, yr);
            r.ny = th + (b < yr ? 1 : 0);
            r.ystop = r.ystart + r.ny - 1;
            f->tile_internal[b * ntx + a] = r;
            dlesm_region w = r;
            w.xstart -= 1; w.xstop += 1; w.ystart -= 1; w.ystop += 1;
            w.nx += 2; w.ny += 2;
            f->tile_whole[b * ntx + a] = w;
        }
    return DLESM_OK;
}
}  // namespace

extern "C" dlesm_field_t *dlesm_r2d_field_create(dlesm_grid_t *g, int grid_points, int do_tile,
                                                 const double *init_global_data, int gnx, int gny)
{
    clear_error();
    if (!g || !g->info.initialised) {
        set_error(DLESM_ERR_STATE, "r2d_field: the grid must be initialised (grid_init) first");
        return nullptr;
    }
    auto *f = new dlesm_field;
    f->grid = g;
    f->nx = g->info.nx;
    f->ny = g->info.ny;
    if (field_bounds(g->info.subdomain, g->info.nx, g->info.ny, g->info.offset, g->info.boundary_conditions,
                     grid_points, f->fb) != DLESM_OK) {
        delete f;
        return nullptr;
    }
    f->ntiles = 0;   // "We're not tiling this field so set ntiles to zero" (src/field_mod.f90:324)
    if (do_tile && f->fb.internal.nx > 0 && f->fb.internal.ny > 0) setup_tiles(f);
    if (dlesm_dev_field_alloc(f->nx, f->ny, &f->dptr) != DLESM_OK) {
        delete f;
        return nullptr;
    }
    if (init_global_data) {
        // src/field_mod.f90:378-389: data(ji,jj) = init_global_data(ji+dx, jj+dy) over the GRID's
        // internal region
        const dlesm_subdomain &sub = g->info.subdomain;
        if (gnx < sub.global.xstop || gny < sub.global.ystop) {
            set_error(DLESM_ERR_ARGS, "r2d_field: init_global_data smaller than the global domain");
            dlesm_free_field(f);
            return nullptr;
        }
        Context &c = ctx();
        double *dg = nullptr;
        const size_t gn = static_cast<size_t>(gnx) * gny;
        cudaError_t e = cudaMalloc(&dg, gn * sizeof(double));
        if (e == cudaSuccess) e = cudaMemcpyAsync(dg, init_global_data, gn * sizeof(double), cudaMemcpyHostToDevice, c.stream);
        if (e == cudaSuccess)
            e = k::scatter_global(static_cast<double *>(f->dptr), f->nx, box_of(sub.internal), dg, gnx,
                                  sub.global.xstart - f->fb.internal.xstart,
                                  sub.global.ystart - f->fb.internal.ystart, c.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
        if (dg) cudaFree(dg);
        if (e != cudaSuccess) {
            fail_cuda(e, "r2d_field scatter");
            dlesm_free_field(f);
            return nullptr;
        }
    }
    return f;
}

extern "C" dlesm_grid_t *dlesm_field_grid(const dlesm_field_t *f) { return f ? f->grid : nullptr; }
extern "C" int dlesm_field_get_info(const dlesm_field_t *f, dlesm_field_info *out)
{
    out->defined_on = f->fb.defined_on;
    out->internal = f->fb.internal;
    out->whole = f->fb.whole;
    out->num_halos = f->fb.num_halos;
    for (int q = 0; q < 4; ++q) out->halo[q] = f->fb.halo[q];
    out->nx = f->nx;
    out->ny = f->ny;
    out->ntiles = f->ntiles;
    out->data_on_device = 1;
    out->device_ptr = f->dptr;
    // halo_type%needs_update (src/halo_mod.f90:10-12) from the array's dirty flag
    FieldRec r;
    const bool dirty = lookup(f->dptr, r) != DLESM_OK || r.clean_depth < 1;
    for (int q = 0; q < 4; ++q) out->halo[q].needs_update = dirty ? 1 : 0;
    clear_error();
    return DLESM_OK;
}

extern "C" int dlesm_field_tile(const dlesm_field_t *f, int itile1, dlesm_region *internal, dlesm_region *whole)
{
    clear_error();
    if (!f || itile1 < 1 || itile1 > f->ntiles) {
        set_error(DLESM_ERR_ARGS, "field tile index out of range (1..ntiles)");
        return DLESM_ERR_ARGS;
    }
    if (internal) *internal = f->tile_internal[itile1 - 1];
    if (whole) *whole = f->tile_whole[itile1 - 1];
    return DLESM_OK;
}

extern "C" int dlesm_field_halo_is_dirty(const dlesm_field_t *f, int depth)
{
    FieldRec r;
    if (!f || lookup(f->dptr, r) != DLESM_OK) {
        clear_error();
        return 1;
    }
    return r.clean_depth < (depth < 1 ? 1 : depth) ? 1 : 0;
}
extern "C" int dlesm_field_set_halo_dirty(dlesm_field_t *f)
{
    if (f) mark_written(f->dptr);
    return DLESM_OK;
}

extern "C" int dlesm_grid_field_bounds(const dlesm_grid_t *g, int grid_points, dlesm_field_info *out)
{
    clear_error();
    if (!g || !g->info.initialised) {
        set_error(DLESM_ERR_STATE, "r2d_field: the grid must be initialised (grid_init) first");
        return DLESM_ERR_STATE;
    }
    FieldBounds fb;
    int rc = field_bounds(g->info.subdomain, g->info.nx, g->info.ny, g->info.offset,
                          g->info.boundary_conditions, grid_points, fb);
    if (rc != DLESM_OK) return rc;
    out->defined_on = fb.defined_on;
    out->internal = fb.internal;
    out->whole = fb.whole;
    out->num_halos = fb.num_halos;
    for (int q = 0; q < 4; ++q) out->halo[q] = fb.halo[q];
    out->nx = g->info.nx;
    out->ny = g->info.ny;
    out->ntiles = 0;
    out->data_on_device = 0;
    out->device_ptr = nullptr;
    return DLESM_OK;
}

extern "C" int dlesm_field_write_to_device(dlesm_field_t *f, const double *host, int startx, int starty, int nx,
                                           int ny, bool blocking)
{
    clear_error();
    return hook_copy(f->dptr, const_cast<double *>(host), true, startx, starty, nx, ny, blocking);
}
extern "C" int dlesm_field_read_from_device(dlesm_field_t *f, double *host, int startx, int starty, int nx, int ny,
                                            bool blocking)
{
    clear_error();
    return hook_copy(f->dptr, host, false, startx, starty, nx, ny, blocking);
}
extern "C" int dlesm_field_set_data(dlesm_field_t *f, const double *host)
{
    return dlesm_field_write_to_device(f, host, 1, 1, f->nx, f->ny, true);
}
extern "C" int dlesm_field_get_data(dlesm_field_t *f, double *host)
{
    return dlesm_field_read_from_device(f, host, 1, 1, f->nx, f->ny, true);
}

extern "C" int dlesm_fields_halo_exchange(dlesm_field_t *const *f, int nfields, int depth)
{
    clear_error();
    if (nfields < 1 || nfields > k::kMaxExchangeFields) {
        set_error(DLESM_ERR_ARGS, "halo_exchange: between 1 and 8 fields per call");
        return DLESM_ERR_ARGS;
    }
    void *ptrs[k::kMaxExchangeFields];
    for (int q = 0; q < nfields; ++q) ptrs[q] = f[q]->dptr;
    // The reference ignores `depth` ("currently ignored and hardwired to 1",
    // src/field_mod.f90:1226-1229; MAX_HALO_DEPTH = 1, src/parallel_comms_mod.f90:48).  With the
    // DLESM_EXT_HALO_DEPTH extension a deeper exchange is served from tables built for that
    // depth, up to the halo width the grid was decomposed with.
    dlesm_grid *g = f[0]->grid;
    if (depth > 1 && (pstate().extensions & DLESM_EXT_HALO_DEPTH) && g->have_comms) {
        // tables of that depth, used for this call only: the module-level (depth-1) state stays put
        dlesm_comms deep;
        int rc = build_comms(g->decomp, pstate().rank, pstate().nranks, g->pbc_x, g->pbc_y, depth, depth, deep);
        if (rc != DLESM_OK) return rc;
        CommTables ct;
        rc = build_tables(&deep, depth, ct);
        if (rc != DLESM_OK) return rc;
        return exchange_with(ptrs, nfields, ct);
    }
    return dlesm_dev_halo_exchange(ptrs, nfields);
}
extern "C" int dlesm_field_halo_exchange(dlesm_field_t *f, int depth)
{
    dlesm_field_t *one[1] = {f};
    return dlesm_fields_halo_exchange(one, 1, depth);
}

extern "C" int dlesm_field_apply_periodic_halos(dlesm_field_t *f)
{
    clear_error();
    if (f->fb.num_halos <= 0) return DLESM_OK;
    const int *bcs = f->grid->info.boundary_conditions;
    double *fl[1] = {static_cast<double *>(f->dptr)};
    const bool px = bcs[0] == DLESM_GO_BC_PERIODIC, py = bcs[1] == DLESM_GO_BC_PERIODIC;
    if ((pstate().extensions & DLESM_EXT_DIRTY_HALOS) && dlesm_field_halo_is_dirty(f, 1) == 0) return DLESM_OK;
    CU(k::periodic_halos(fl, 1, f->nx, box_of(f->fb.internal), px, py, ctx().stream));
    mark_clean(f->dptr, 1);   // every halo in the field's list (the periodic directions) is filled
    return DLESM_OK;
}

extern "C" int dlesm_field_gather_inner_data(dlesm_field_t *f, double *global_data)
{
    clear_error();
    Context &c = ctx();
    const Decomposition &d = f->grid->decomp;
    const dlesm_region &in = f->fb.internal;
    const int nranks = pstate().nranks, me = pstate().rank;
    if (nranks == 1) {
        // src/field_mod.f90:1333-1343
        CU(cudaMemcpy2DAsync(global_data, static_cast<size_t>(d.global_nx) * sizeof(double),
                             static_cast<const double *>(f->dptr) + (static_cast<size_t>(in.ystart) - 1) * f->nx +
                                 (in.xstart - 1),
                             static_cast<size_t>(f->nx) * sizeof(double), static_cast<size_t>(in.nx) * sizeof(double),
                             in.ny, cudaMemcpyDeviceToHost, c.stream));
        CU(cudaStreamSynchronize(c.stream));
        return DLESM_OK;
    }
    if (!c.comm_ready) {
        set_error(DLESM_ERR_STATE, "gather_inner_data: communicator not initialised (dlesm_comm_init)");
        return DLESM_ERR_STATE;
    }
    // :1348-1351: every rank contributes a fixed-size slot
    const int halo_x = in.xstart - 1, halo_y = in.ystart - 1;
    const size_t n = static_cast<size_t>(d.max_width - 2 * halo_x) * (d.max_height - 2 * halo_y);
    int rc = ensure_scratch(n * (me == 1 ? nranks : 1));
    if (rc != DLESM_OK) return rc;
    CU(k::gather_box(static_cast<const double *>(f->dptr), f->nx, box_of(in), c.scratch, c.stream));
    NC(ncclGroupStart());
    if (me == 1) {
        for (int r = 1; r < nranks; ++r)
            NC(ncclRecv(c.scratch + static_cast<size_t>(r) * n, n, ncclDouble, r, c.comm, c.stream));
    } else {
        NC(ncclSend(c.scratch, n, ncclDouble, 0, c.comm, c.stream));
    }
    NC(ncclGroupEnd());
    if (me == 1) {
        std::vector<double> recv(n * nranks);
        CU(cudaMemcpyAsync(recv.data(), c.scratch, recv.size() * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
        CU(cudaStreamSynchronize(c.stream));
        // :1373-1388
        for (int r = 1; r <= nranks; ++r) {
            const dlesm_region &gr = d.subdomains[r - 1].global;
            size_t q = static_cast<size_t>(r - 1) * n;
            for (int jj = gr.ystart; jj <= gr.ystop; ++jj)
                for (int ji = gr.xstart; ji <= gr.xstop; ++ji)
                    global_data[(static_cast<size_t>(jj) - 1) * d.global_nx + (ji - 1)] = recv[q++];
        }
    } else {
        CU(cudaStreamSynchronize(c.stream));
    }
    return DLESM_OK;
}

extern "C" int dlesm_copy_field(const dlesm_field_t *in, dlesm_field_t *out)
{
    clear_error();
    return dlesm_dev_copy_field(in->dptr, out->dptr);
}
extern "C" int dlesm_copy_field_patch(dlesm_field_t *f, const dlesm_region *src, const dlesm_region *dest)
{
    clear_error();
    return dlesm_dev_copy_patch(f->dptr, src->xstart, src->xstop, src->ystart, src->ystop, dest->xstart,
                                dest->xstop, dest->ystart, dest->ystop);
}
extern "C" int dlesm_set_field(dlesm_field_t *f, double val)
{
    clear_error();
    return dlesm_dev_set_field(f->dptr, val);
}
extern "C" int dlesm_field_checksum(dlesm_field_t *f, double *val)
{
    clear_error();
    const dlesm_region &in = f->fb.internal;
    if (pstate().nranks == 1) return dlesm_dev_checksum(f->dptr, in.xstart, in.xstop, in.ystart, in.ystop, val);
    // several ranks: local sum -> all-reduce -> host, chained on the stream with one synchronise
    Context &c = ctx();
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    if (!c.comm_ready) {
        set_error(DLESM_ERR_STATE, "field_checksum: communicator not initialised (dlesm_comm_init)");
        return DLESM_ERR_STATE;
    }
    const int nranks = pstate().nranks, me0 = pstate().rank - 1;
    if (c.exchange_mode < 0) {
        const char *e = std::getenv("DLESM_EXCHANGE");
        c.exchange_mode = (e && std::strcmp(e, "nccl") == 0) ? DLESM_EXCHANGE_NCCL : DLESM_EXCHANGE_DIRECT;
    }
    if (nranks <= k::kGsumMaxRanks && c.exchange_mode != DLESM_EXCHANGE_NCCL) {
        // ONE launch: local sum, global sum through the ranks' peer-mapped cells (rank order: the same bits
        // everywhere), result into pinned host memory which the host polls
        if (!c.xd.ready) {
            rc = xdirect_setup();      // collective; every rank reaches its first checksum together
            if (rc != DLESM_OK) return rc;
        }
        const unsigned long long gseq = ++c.gsum_seq;
        k::GlobalSumPlan gs{};
        gs.nranks = nranks; gs.me = me0; gs.seq = gseq;
        gs.err = reinterpret_cast<int *>(c.xd.base + 1032);
        for (int r = 0; r < nranks; ++r)
            gs.peer_cell[r] = gsum_cells(c.xd.peer_base[r], gseq) + 2 * me0;
        gs.own_cells = gsum_cells(c.xd.base, gseq);
        CU(k::checksum(static_cast<const double *>(f->dptr), f->nx, box_of(in), c.partials, c.ticket, c.result,
                       c.host_result + 6, gseq, c.stream, &gs));
        const unsigned long long *flag = reinterpret_cast<const unsigned long long *>(c.host_result + 7);
        auto published = [flag]() { return __atomic_load_n(flag, __ATOMIC_ACQUIRE); };
        for (unsigned spin = 1; published() != gseq; ++spin) {
            if ((spin & 0xfffff) == 0) {
                cudaError_t q = cudaStreamQuery(c.stream);
                if (q != cudaSuccess && q != cudaErrorNotReady) return fail_cuda(q, "field_checksum");
                if (q == cudaSuccess && published() != gseq) {
                    set_error(DLESM_ERR_CUDA, "field_checksum: the kernel finished without publishing its result");
                    return DLESM_ERR_CUDA;
                }
            }
        }
        *val = c.host_result[6];
        return DLESM_OK;
    }
    CU(k::checksum(static_cast<const double *>(f->dptr), f->nx, box_of(in), c.partials, c.ticket, c.result, nullptr, 0,
                   c.stream));
    NC(ncclAllReduce(c.result, c.result + 2, 1, ncclDouble, ncclSum, c.comm, c.stream));
    CU(cudaMemcpyAsync(c.host_result + 2, c.result + 2, sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    *val = c.host_result[2];
    return DLESM_OK;
}
extern "C" int dlesm_free_field(dlesm_field_t *f)
{
    if (!f) return DLESM_OK;
    int rc = f->dptr ? dlesm_dev_field_free(f->dptr) : DLESM_OK;
    delete f;
    return rc;
}

// =========================================================================================
// HOST MIRROR: shallow-water application layer
// =========================================================================================
struct dlesm_sw {
    dlesm_grid *grid = nullptr;
    dlesm_field *f[DLESM_SW_NFIELDS] = {nullptr};
    double dt = 0, tdt = 0, alpha = 0;
    int mode = DLESM_SW_FUSED;
    int variant = 0;
    long long ncycle = 0;
    const int *tmask_dev = nullptr;
    bool periodic_local = true;  // halos by local copies (1 rank) or by exchange (several)
    bool push = false;           // halos stored by the fused sweep itself (DLESM_HALO_PUSH)
    // halo push state of THIS object: the send list the sweeps push along (depth 1, or 2 with the
    // two-step sweep) and the buffers the neighbours have mapped.  Nothing of it lives in the
    // module-level comm state, so other fields' halo_exchange calls are unaffected.
    CommTables push_tables;
    std::vector<void *> registered;
    // two steps per sweep (dlesm_sw_set_two_step): second set of old fields (the two-step sweep
    // cannot filter in place), and whether u,v,p / uold,vold,pold currently hold deep halos
    bool two_step = false;
    int variant2 = 0;
    dlesm_field *old2[3] = {nullptr, nullptr, nullptr};
    bool deep_valid = false;
    unsigned char *mask2 = nullptr;   // byte image of the tmask with a PERIODIC ring two cells deep
    int mask2_pitch = 0;
    bool tiled = false;               // model of one tile of a tile set: halos come from dlesm_tileset_*
};

extern "C" dlesm_sw_t *dlesm_sw_create(dlesm_grid_t *g, double dt, double alpha, int mode, int use_tmask)
{
    clear_error();
    if (!g || !g->info.initialised) {
        set_error(DLESM_ERR_STATE, "sw_create: the grid must be initialised first");
        return nullptr;
    }
    const int *bcs = g->info.boundary_conditions;
    if (g->info.offset != DLESM_GO_OFFSET_SW || bcs[0] != DLESM_GO_BC_PERIODIC || bcs[1] != DLESM_GO_BC_PERIODIC) {
        set_error(DLESM_ERR_ARGS, "sw_create: the shallow-water layer needs GO_OFFSET_SW and periodic BCs in x and y");
        return nullptr;
    }
    auto *s = new dlesm_sw;
    s->grid = g;
    s->dt = s->tdt = dt;
    s->alpha = alpha;
    s->mode = mode;
    s->tiled = g->tile_id > 0;
    s->periodic_local = pstate().nranks == 1 && !s->tiled;
    static const int pts[DLESM_SW_NFIELDS] = {
        DLESM_GO_U_POINTS, DLESM_GO_V_POINTS, DLESM_GO_T_POINTS, DLESM_GO_U_POINTS, DLESM_GO_V_POINTS,
        DLESM_GO_T_POINTS, DLESM_GO_U_POINTS, DLESM_GO_V_POINTS, DLESM_GO_T_POINTS, DLESM_GO_U_POINTS,
        DLESM_GO_V_POINTS, DLESM_GO_F_POINTS, DLESM_GO_T_POINTS};
    const int nf = mode == DLESM_SW_FUSED ? DLESM_SW_CU : DLESM_SW_NFIELDS;  // cu,cv,z,h only un-fused
    // an application built for tiling asks for tiled fields; here: whenever the knob is set
    const int do_tile = std::getenv("GOCEAN_OMP_GRID") != nullptr;
    for (int q = 0; q < nf; ++q) {
        s->f[q] = dlesm_r2d_field_create(g, pts[q], do_tile, nullptr, 0, 0);
        if (!s->f[q]) {
            dlesm_sw_free(s);
            return nullptr;
        }
    }
    if (use_tmask) {
        void *p = nullptr;
        if (dlesm_grid_device_array(g, DLESM_GRID_TMASK, &p) != DLESM_OK) {
            dlesm_sw_free(s);
            return nullptr;
        }
        s->tmask_dev = static_cast<const int *>(p);
    }
    return s;
}

namespace {
int sw_fill_halos(dlesm_sw *s, int first_field, int count)
{
    if (s->tiled) return DLESM_OK;   // deferred: the tile set exchanges every tile's fields at once
    if (s->periodic_local) {
        double *fl[k::kMaxExchangeFields];
        for (int q = 0; q < count; ++q) fl[q] = static_cast<double *>(s->f[first_field + q]->dptr);
        // every SW periodic field shares the T-point internal region
        CU(k::periodic_halos(fl, count, s->f[first_field]->nx, box_of(s->f[first_field]->fb.internal), true, true,
                             ctx().stream));
        for (int q = 0; q < count; ++q) mark_clean(fl[q], 1);
        return DLESM_OK;
    }
    return dlesm_fields_halo_exchange(&s->f[first_field], count, 1);
}

// Collective.  Drops this object's registrations (the neighbours unmap its buffers and it unmaps theirs).
int sw_release_peers(dlesm_sw *s)
{
    if (s->registered.empty()) return DLESM_OK;
    int rc = peer_unregister_fields(s->registered.data(), static_cast<int>(s->registered.size()));
    s->registered.clear();
    return rc;
}

// Halo push set-up of the model (collective): build the send list at the depth the sweeps push
// (1, or 2 with the two-step sweep) and map the neighbours' counterparts of every buffer a sweep
// stores into -- both rotating sets of u,v,p and, with the two-step sweep, of the old fields.
int sw_configure_push(dlesm_sw *s)
{
    const int depth = s->two_step ? 2 : 1;
    dlesm_grid &g = *s->grid;
    dlesm_comms t;
    int rc;
    if (pstate().nranks == 1)
        rc = dlesm_map_comms(&g.dh, 1, 1, 1, 1, depth, depth, &t);   // wrap-around neighbours = this rank
    else
        rc = build_comms(g.decomp, pstate().rank, pstate().nranks, g.pbc_x, g.pbc_y, depth, depth, t);
    if (rc == DLESM_OK) rc = build_tables(&t, depth, s->push_tables);
    if (rc != DLESM_OK) return rc;
    void *bufs[12];
    int n = 0;
    for (int q = 0; q < 3; ++q) bufs[n++] = s->f[DLESM_SW_U + q]->dptr;
    for (int q = 0; q < 3; ++q) bufs[n++] = s->f[DLESM_SW_UNEW + q]->dptr;
    if (s->two_step) {
        for (int q = 0; q < 3; ++q) bufs[n++] = s->f[DLESM_SW_UOLD + q]->dptr;
        for (int q = 0; q < 3; ++q) bufs[n++] = s->old2[q]->dptr;
    }
    rc = sw_release_peers(s);                                   // the registered set is replaced (collective)
    if (rc == DLESM_OK) rc = peer_register_with(bufs, n, s->push_tables);   // collective
    if (rc != DLESM_OK) return rc;
    s->registered.assign(bufs, bufs + n);
    return DLESM_OK;
}
}  // namespace

extern "C" int dlesm_sw_set_halo_mode(dlesm_sw_t *s, int mode)
{
    clear_error();
    if (mode == DLESM_HALO_EXCHANGE) {
        s->push = false;
        s->deep_valid = false;
        return sw_release_peers(s);      // collective, like the call that mapped them
    }
    if (mode != DLESM_HALO_PUSH || s->mode != DLESM_SW_FUSED) {
        set_error(DLESM_ERR_ARGS, "sw_set_halo_mode: DLESM_HALO_PUSH needs the fused mode");
        return DLESM_ERR_ARGS;
    }
    if (s->f[0]->nx % 2 != 0) {
        set_error(DLESM_ERR_ARGS, "sw_set_halo_mode: halo push needs a TMA-eligible pitch (DL_ESM_ALIGNMENT even)");
        return DLESM_ERR_ARGS;
    }
    s->push = true;
    return sw_configure_push(s);
}

namespace {
// halos of u,v,p,uold,vold,pold two cells deep: wrapped copies on one rank, the depth-2
// exchange (DLESM_EXT_HALO_DEPTH tables) on several
int sw_fill_deep_halos(dlesm_sw *s)
{
    dlesm_field *fl[6];
    for (int q = 0; q < 6; ++q) fl[q] = s->f[DLESM_SW_U + q];
    if (s->periodic_local) {
        double *dp[6];
        for (int q = 0; q < 6; ++q) dp[q] = static_cast<double *>(fl[q]->dptr);
        CU(k::periodic_halos_deep(dp, 6, fl[0]->nx, box_of(s->f[DLESM_SW_P]->fb.internal), 2, ctx().stream));
        for (int q = 0; q < 6; ++q) mark_clean(dp[q], 2);
        return DLESM_OK;
    }
    if (s->push && s->push_tables.valid && s->push_tables.depth == 2) {
        void *ptrs[6];
        for (int q = 0; q < 6; ++q) ptrs[q] = fl[q]->dptr;
        return exchange_with(ptrs, 6, s->push_tables);
    }
    return dlesm_fields_halo_exchange(fl, 6, 2);
}
}  // namespace

extern "C" int dlesm_sw_set_two_step(dlesm_sw_t *s, int on, int variant2)
{
    clear_error();
    if (!on) {
        s->two_step = false;
        if (s->push) return sw_configure_push(s);   // back to the depth-1 send list
        return DLESM_OK;
    }
    const dlesm_region &in = s->f[DLESM_SW_P]->fb.internal;
    const int nx = s->f[0]->nx, ny = s->f[0]->ny;
    if (s->mode != DLESM_SW_FUSED) {
        set_error(DLESM_ERR_ARGS, "sw_set_two_step: needs the fused mode");
        return DLESM_ERR_ARGS;
    }
    if (nx % 2 != 0 || in.xstart < 3 || in.ystart < 3 || in.xstop + 2 > nx || in.ystop + 2 > ny ||
        in.xstop - in.xstart < 1 || in.ystop - in.ystart < 1) {
        set_error(DLESM_ERR_ARGS, "sw_set_two_step: needs an even pitch and a grid decomposed with halo_width >= 2");
        return DLESM_ERR_ARGS;
    }
    if (!s->periodic_local && !(pstate().extensions & DLESM_EXT_HALO_DEPTH)) {
        set_error(DLESM_ERR_STATE, "sw_set_two_step: several ranks need DLESM_EXT_HALO_DEPTH (depth-2 exchange)");
        return DLESM_ERR_STATE;
    }
    static const int pts[3] = {DLESM_GO_U_POINTS, DLESM_GO_V_POINTS, DLESM_GO_T_POINTS};
    for (int q = 0; q < 3; ++q)
        if (!s->old2[q]) {
            s->old2[q] = dlesm_r2d_field_create(s->grid, pts[q], s->f[0]->ntiles > 0, nullptr, 0, 0);
            if (!s->old2[q]) return last_status();
        }
    if (s->tmask_dev && !s->mask2) {
        // grid_init replicates the mask outward beyond the first halo ring (src/grid_mod.f90:400-431);
        // the two-step sweep reads the ring two cells deep, where the periodic model needs the
        // wrapped mask.  Build it once: mask -> FP64 image -> depth-2 halo fill -> mask -> byte image.
        const size_t n = static_cast<size_t>(nx) * ny;
        dlesm_field *tmp = dlesm_r2d_field_create(s->grid, DLESM_GO_T_POINTS, 0, nullptr, 0, 0);
        if (!tmp) return last_status();
        int *m32 = nullptr;
        CU(cudaMalloc(&m32, n * sizeof(int)));
        CU(cudaMemcpyAsync(m32, s->tmask_dev, n * sizeof(int), cudaMemcpyDeviceToDevice, ctx().stream));
        CU(k::mask_convert(m32, static_cast<double *>(tmp->dptr), n, true, ctx().stream));
        int rc = DLESM_OK;
        if (s->periodic_local) {
            double *dp[1] = {static_cast<double *>(tmp->dptr)};
            CU(k::periodic_halos_deep(dp, 1, nx, box_of(in), 2, ctx().stream));
        } else {
            mark_written(tmp->dptr);
            rc = dlesm_fields_halo_exchange(&tmp, 1, 2);
        }
        if (rc == DLESM_OK) {
            CU(k::mask_convert(m32, static_cast<double *>(tmp->dptr), n, false, ctx().stream));
            s->mask2_pitch = (nx + 15) & ~15;
            CU(cudaMalloc(&s->mask2, static_cast<size_t>(s->mask2_pitch) * ny));
            CU(k::mask_to_bytes(m32, nx, ny, s->mask2, s->mask2_pitch, ctx().stream));
        }
        CU(cudaStreamSynchronize(ctx().stream));
        cudaFree(m32);
        dlesm_free_field(tmp);
        if (rc != DLESM_OK) return rc;
    }
    s->two_step = true;
    s->variant2 = variant2;
    s->deep_valid = false;
    if (s->push) return sw_configure_push(s);   // send list two cells deep, the old fields mapped too
    return DLESM_OK;
}

extern "C" int dlesm_sw_init_device(dlesm_sw_t *s)
{
    clear_error();
    const dlesm_subdomain &sub = s->grid->info.subdomain;
    const dlesm_region &in = s->f[DLESM_SW_P]->fb.internal;
    const int n = s->f[0]->nx;
    CU(k::sw_init(static_cast<double *>(s->f[DLESM_SW_U]->dptr), static_cast<double *>(s->f[DLESM_SW_V]->dptr),
                  static_cast<double *>(s->f[DLESM_SW_P]->dptr), n, box_of(in), sub.global.xstart,
                  sub.global.ystart, s->grid->info.global_nx, s->grid->info.global_ny, s->grid->info.dx,
                  s->grid->info.dy, ctx().stream));
    // the internal regions were just overwritten: whatever the halos held is stale (dirty-flag halos)
    for (int q = 0; q < 3; ++q) mark_written(s->f[DLESM_SW_U + q]->dptr);
    int rc = sw_fill_halos(s, DLESM_SW_U, 3);
    if (rc != DLESM_OK) return rc;
    for (int q = 0; q < 3; ++q) {
        rc = dlesm_copy_field(s->f[DLESM_SW_U + q], s->f[DLESM_SW_UOLD + q]);
        if (rc != DLESM_OK) return rc;
    }
    s->ncycle = 0;
    s->deep_valid = false;
    s->tdt = s->dt;
    return DLESM_OK;
}

extern "C" int dlesm_sw_set_state(dlesm_sw_t *s, const double *u, const double *v, const double *p)
{
    clear_error();
    const double *src[3] = {u, v, p};
    for (int q = 0; q < 3; ++q) {
        int rc = dlesm_field_set_data(s->f[DLESM_SW_U + q], src[q]);
        if (rc != DLESM_OK) return rc;
        rc = dlesm_copy_field(s->f[DLESM_SW_U + q], s->f[DLESM_SW_UOLD + q]);
        if (rc != DLESM_OK) return rc;
    }
    s->ncycle = 0;
    s->deep_valid = false;
    s->tdt = s->dt;
    return DLESM_OK;
}

extern "C" int dlesm_sw_step(dlesm_sw_t *s, int nsteps)
{
    clear_error();
    if (s->tiled) {
        set_error(DLESM_ERR_STATE, "sw_step: this model is one tile of a tile set; step the set (dlesm_tileset_sw_step)");
        return DLESM_ERR_STATE;
    }
    dlesm_grid &g = *s->grid;
    const double dx = g.info.dx, dy = g.info.dy;
    for (int it = 0; it < nsteps; ++it) {
        const int first = s->ncycle == 0;
        dlesm_field **f = s->f;
        if (s->two_step && !first && nsteps - it >= 2) {
            // two steps in one sweep: 96 B / point for both
            const dlesm_region &rt = f[DLESM_SW_P]->fb.internal;
            int rc = DLESM_OK;
            if (!s->deep_valid) {
                rc = push_settle();
                if (rc == DLESM_OK) rc = sw_fill_deep_halos(s);
            }
            if (rc != DLESM_OK) return rc;
            rc = sw_step2_fused_impl(f[DLESM_SW_U]->dptr, f[DLESM_SW_V]->dptr, f[DLESM_SW_P]->dptr,
                                     f[DLESM_SW_UOLD]->dptr, f[DLESM_SW_VOLD]->dptr, f[DLESM_SW_POLD]->dptr,
                                     f[DLESM_SW_UNEW]->dptr, f[DLESM_SW_VNEW]->dptr, f[DLESM_SW_PNEW]->dptr,
                                     s->old2[0]->dptr, s->old2[1]->dptr, s->old2[2]->dptr, dx, dy, s->tdt, s->alpha,
                                     rt.xstart, rt.xstop, rt.ystart, rt.ystop, s->mask2, s->mask2_pitch, s->variant2,
                                     s->push ? &s->push_tables : nullptr, /*deferred=*/true);
            if (rc != DLESM_OK) return rc;
            for (int q = 0; q < 3; ++q) {
                std::swap(f[DLESM_SW_U + q]->dptr, f[DLESM_SW_UNEW + q]->dptr);
                std::swap(f[DLESM_SW_UOLD + q]->dptr, s->old2[q]->dptr);
            }
            if (!s->push) rc = sw_fill_deep_halos(s);   // with push the sweep stored the halos itself
            if (rc != DLESM_OK) return rc;
            s->deep_valid = true;
            s->ncycle += 2;
            ++it;
            continue;
        }
        s->deep_valid = false;
        const dlesm_region &ru = f[DLESM_SW_U]->fb.internal, &rv = f[DLESM_SW_V]->fb.internal;
        const dlesm_region &rt = f[DLESM_SW_P]->fb.internal;
        int rc;
#define P(q) f[q]->dptr
        if (s->mode == DLESM_SW_FUSED) {
            rc = sw_step_fused_impl(P(DLESM_SW_U), P(DLESM_SW_V), P(DLESM_SW_P), P(DLESM_SW_UOLD), P(DLESM_SW_VOLD),
                                    P(DLESM_SW_POLD), P(DLESM_SW_UNEW), P(DLESM_SW_VNEW), P(DLESM_SW_PNEW), dx, dy,
                                    s->tdt, s->alpha, first, rt.xstart, rt.xstop, rt.ystart, rt.ystop,
                                    s->tmask_dev, s->variant, s->push ? &s->push_tables : nullptr,
                                    /*deferred=*/true);
            if (rc != DLESM_OK) return rc;
            if (!s->push) {
                rc = sw_fill_halos(s, DLESM_SW_UNEW, 3);
                if (rc != DLESM_OK) return rc;
            }
            // u <- unew by rotating the device buffers (copy_field semantics without the copy)
            for (int q = 0; q < 3; ++q) std::swap(f[DLESM_SW_U + q]->dptr, f[DLESM_SW_UNEW + q]->dptr);
        } else {
            const dlesm_region &rf = f[DLESM_SW_Z]->fb.internal;
            const void *tm = s->tmask_dev;
            rc = dlesm_compute_cu(P(DLESM_SW_CU), P(DLESM_SW_P), P(DLESM_SW_U), ru.xstart, ru.xstop, ru.ystart, ru.ystop, tm);
            if (rc == DLESM_OK) rc = dlesm_compute_cv(P(DLESM_SW_CV), P(DLESM_SW_P), P(DLESM_SW_V), rv.xstart, rv.xstop, rv.ystart, rv.ystop, tm);
            if (rc == DLESM_OK) rc = dlesm_compute_z(P(DLESM_SW_Z), P(DLESM_SW_P), P(DLESM_SW_U), P(DLESM_SW_V), dx, dy, rf.xstart, rf.xstop, rf.ystart, rf.ystop, tm);
            if (rc == DLESM_OK) rc = dlesm_compute_h(P(DLESM_SW_H), P(DLESM_SW_P), P(DLESM_SW_U), P(DLESM_SW_V), rt.xstart, rt.xstop, rt.ystart, rt.ystop, tm);
            if (rc == DLESM_OK) rc = sw_fill_halos(s, DLESM_SW_CU, 4);
            if (rc == DLESM_OK) rc = dlesm_compute_unew(P(DLESM_SW_UNEW), P(DLESM_SW_UOLD), P(DLESM_SW_Z), P(DLESM_SW_CV), P(DLESM_SW_H), s->tdt, dx, ru.xstart, ru.xstop, ru.ystart, ru.ystop, tm, P(DLESM_SW_U));
            if (rc == DLESM_OK) rc = dlesm_compute_vnew(P(DLESM_SW_VNEW), P(DLESM_SW_VOLD), P(DLESM_SW_Z), P(DLESM_SW_CU), P(DLESM_SW_H), s->tdt, dy, rv.xstart, rv.xstop, rv.ystart, rv.ystop, tm, P(DLESM_SW_V));
            if (rc == DLESM_OK) rc = dlesm_compute_pnew(P(DLESM_SW_PNEW), P(DLESM_SW_POLD), P(DLESM_SW_CU), P(DLESM_SW_CV), s->tdt, dx, dy, rt.xstart, rt.xstop, rt.ystart, rt.ystop, tm, P(DLESM_SW_P));
            if (rc == DLESM_OK) rc = sw_fill_halos(s, DLESM_SW_UNEW, 3);
            if (rc == DLESM_OK && !first) {
                rc = dlesm_time_smooth(P(DLESM_SW_U), P(DLESM_SW_UNEW), P(DLESM_SW_UOLD), s->alpha, ru.xstart, ru.xstop, ru.ystart, ru.ystop, tm, DLESM_GO_U_POINTS);
                if (rc == DLESM_OK) rc = dlesm_time_smooth(P(DLESM_SW_V), P(DLESM_SW_VNEW), P(DLESM_SW_VOLD), s->alpha, rv.xstart, rv.xstop, rv.ystart, rv.ystop, tm, DLESM_GO_V_POINTS);
                if (rc == DLESM_OK) rc = dlesm_time_smooth(P(DLESM_SW_P), P(DLESM_SW_PNEW), P(DLESM_SW_POLD), s->alpha, rt.xstart, rt.xstop, rt.ystart, rt.ystop, tm, DLESM_GO_T_POINTS);
            }
            for (int q = 0; q < 3 && rc == DLESM_OK; ++q) rc = dlesm_copy_field(f[DLESM_SW_UNEW + q], f[DLESM_SW_U + q]);
            if (rc != DLESM_OK) return rc;
        }
#undef P
        if (first) s->tdt = s->tdt + s->tdt;
        ++s->ncycle;
    }
    // the last push sweep published its epoch; the neighbours' must be in before the caller reads halos
    return push_settle();
}

// One model time step driven with HOST state (the `e2e` path of bench.py): instead of
// write_to_device x6 -> step -> read_from_device x6 back to back (H2D, compute and D2H
// serialised: 2 x 12.9 GB over one PCIe direction at a time at 16384^2), the tile is streamed
// through the device in row bands.  Band k+1 uploads while band k is swept while band k-1
// downloads: three streams, both PCIe directions busy, the sweep hidden behind the copies.
// Each row of each field crosses the bus exactly once per direction.
extern "C" int dlesm_sw_step_host(dlesm_sw_t *s, const double *const host_in[6], double *const host_out[6],
                                  int band_rows)
{
    clear_error();
    if (!s || s->mode != DLESM_SW_FUSED) {
        set_error(DLESM_ERR_ARGS, "sw_step_host: needs a fused-mode shallow-water object");
        return DLESM_ERR_ARGS;
    }
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    Context &c = ctx();
    dlesm_grid &g = *s->grid;
    dlesm_field **f = s->f;
    const dlesm_region &rt = f[DLESM_SW_P]->fb.internal;
    const int nx = f[0]->nx, ny = f[0]->ny;
    const size_t row_bytes = static_cast<size_t>(nx) * sizeof(double);
    if (band_rows <= 0) band_rows = std::max(64, ((rt.ny / 128) + 31) / 32 * 32);
    const int nbands = (rt.ny + band_rows - 1) / band_rows;
    while (c.band_events.size() < static_cast<size_t>(2 * nbands + 2)) {
        cudaEvent_t e;
        CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        c.band_events.push_back(e);
    }
    const int first = s->ncycle == 0;
    // inputs: device slots U,V,P,UOLD,VOLD,POLD; outputs: UNEW,VNEW,PNEW + the filtered *OLD
    static const int in_slot[6] = {DLESM_SW_U, DLESM_SW_V, DLESM_SW_P, DLESM_SW_UOLD, DLESM_SW_VOLD, DLESM_SW_POLD};
    static const int out_slot[6] = {DLESM_SW_UNEW, DLESM_SW_VNEW, DLESM_SW_PNEW,
                                    DLESM_SW_UOLD, DLESM_SW_VOLD, DLESM_SW_POLD};
    // whatever is queued on the library stream (a previous step) must finish before we overwrite
    cudaEvent_t e_start = c.band_events[2 * nbands];
    CU(cudaEventRecord(e_start, c.stream));
    CU(cudaStreamWaitEvent(c.s_h2d, e_start, 0));
    CU(cudaStreamWaitEvent(c.s_d2h, e_start, 0));
    int up_next = 1;  // next row (1-based) not yet uploaded
    for (int b = 0; b < nbands; ++b) {
        const int j0 = rt.ystart + b * band_rows;
        const int j1 = std::min(rt.ystop, j0 + band_rows - 1);
        // rows the sweep of [j0, j1] reads: j0-1 .. j1+1 of u,v,p; j0..j1 of the old fields
        const int up_to = (b == nbands - 1) ? ny : j1 + 1;
        if (up_to >= up_next) {
            const size_t off = static_cast<size_t>(up_next - 1) * nx;
            const size_t bytes = static_cast<size_t>(up_to - up_next + 1) * row_bytes;
            for (int q = 0; q < 6; ++q)
                CU(cudaMemcpyAsync(static_cast<double *>(f[in_slot[q]]->dptr) + off, host_in[q] + off, bytes,
                                   cudaMemcpyHostToDevice, c.s_h2d));
            up_next = up_to + 1;
        }
        cudaEvent_t e_up = c.band_events[2 * b], e_done = c.band_events[2 * b + 1];
        CU(cudaEventRecord(e_up, c.s_h2d));
        CU(cudaStreamWaitEvent(c.stream, e_up, 0));
        rc = dlesm_sw_step_fused(f[DLESM_SW_U]->dptr, f[DLESM_SW_V]->dptr, f[DLESM_SW_P]->dptr,
                                 f[DLESM_SW_UOLD]->dptr, f[DLESM_SW_VOLD]->dptr, f[DLESM_SW_POLD]->dptr,
                                 f[DLESM_SW_UNEW]->dptr, f[DLESM_SW_VNEW]->dptr, f[DLESM_SW_PNEW]->dptr,
                                 g.info.dx, g.info.dy, s->tdt, s->alpha, first, rt.xstart, rt.xstop, j0, j1,
                                 s->tmask_dev, s->variant);
        if (rc != DLESM_OK) return rc;
        CU(cudaEventRecord(e_done, c.stream));
        CU(cudaStreamWaitEvent(c.s_d2h, e_done, 0));
        const size_t off = static_cast<size_t>(j0 - 1) * nx;
        const size_t bytes = static_cast<size_t>(j1 - j0 + 1) * row_bytes;
        for (int q = 0; q < 6; ++q)
            CU(cudaMemcpyAsync(host_out[q] + off, static_cast<const double *>(f[out_slot[q]]->dptr) + off, bytes,
                               cudaMemcpyDeviceToHost, c.s_d2h));
    }
    // halo ring of the three new fields: periodic copies (one rank) or the NCCL exchange
    rc = sw_fill_halos(s, DLESM_SW_UNEW, 3);
    if (rc != DLESM_OK) return rc;
    // ring -> host: the two halo rows are contiguous; the two halo columns are gathered on the
    // device, shipped as one block and scattered by the host
    const size_t col_n = static_cast<size_t>(ny);
    rc = ensure_scratch(6 * col_n);
    if (rc != DLESM_OK) return rc;
    if (c.ring_host_n < 6 * col_n) {
        if (c.ring_host) cudaFreeHost(c.ring_host);
        c.ring_host = nullptr;
        c.ring_host_n = 0;
        CU(cudaMallocHost(&c.ring_host, 6 * col_n * sizeof(double)));
        c.ring_host_n = 6 * col_n;
    }
    const int xw = rt.xstart - 1, xe = rt.xstop + 1;
    for (int q = 0; q < 3; ++q) {
        const double *d = static_cast<const double *>(f[DLESM_SW_UNEW + q]->dptr);
        CU(k::gather_box(d, nx, k::Box{xw, xw, 1, ny}, c.scratch + (2 * q) * col_n, c.stream));
        CU(k::gather_box(d, nx, k::Box{xe, xe, 1, ny}, c.scratch + (2 * q + 1) * col_n, c.stream));
    }
    CU(cudaMemcpyAsync(c.ring_host, c.scratch, 6 * col_n * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    // all band downloads must have landed before the ring rows/columns overwrite their edges
    cudaEvent_t e_dl = c.band_events[2 * nbands + 1];
    CU(cudaEventRecord(e_dl, c.s_d2h));
    CU(cudaStreamWaitEvent(c.stream, e_dl, 0));
    for (int q = 0; q < 3; ++q) {
        const double *d = static_cast<const double *>(f[DLESM_SW_UNEW + q]->dptr);
        const int rows[2] = {rt.ystart - 1, rt.ystop + 1};
        for (int r : rows) {
            const size_t off = static_cast<size_t>(r - 1) * nx;
            CU(cudaMemcpyAsync(host_out[q] + off, d + off, row_bytes, cudaMemcpyDeviceToHost, c.stream));
        }
    }
    CU(cudaStreamSynchronize(c.stream));
    for (int q = 0; q < 3; ++q) {
        double *h = host_out[q];
        const double *w = c.ring_host + (2 * q) * col_n, *e = c.ring_host + (2 * q + 1) * col_n;
        for (int j = rt.ystart - 1; j <= rt.ystop + 1; ++j) {
            h[static_cast<size_t>(j - 1) * nx + (xw - 1)] = w[j - 1];
            h[static_cast<size_t>(j - 1) * nx + (xe - 1)] = e[j - 1];
        }
    }
    for (int q = 0; q < 3; ++q) std::swap(f[DLESM_SW_U + q]->dptr, f[DLESM_SW_UNEW + q]->dptr);
    if (first) s->tdt = s->tdt + s->tdt;
    ++s->ncycle;
    s->deep_valid = false;
    return DLESM_OK;
}

namespace {
// 2 * npass model steps per PCIe round trip on one rank (periodic wrap is local).  The bands of pass 1
// follow the upload as in the two-step form; pass p+1 of band b needs pass p of bands b-1, b, b+1, so it
// runs two bands behind -- skewed, on the same stream, while later bands are still uploading -- and the
// last pass's bands download as they finish.  Only the first and last band of a pass wait for the whole
// previous pass (they read the wrapped halo rows).  Both PCIe directions stay busy; each row crosses the
// bus once per direction for 2 * npass steps.  Outputs ping-pong between the two buffer sets; a band's
// east / west halo columns are filled right after its sweep (periodic_x_rows), the north / south rows
// when a pass is complete.
int sw_step_host_passes(dlesm_sw *s, const double *const host_in[6], double *const host_out[6], int band_rows,
                        int npass)
{
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    Context &c = ctx();
    dlesm_grid &g = *s->grid;
    dlesm_field **f = s->f;
    const dlesm_region &rt = f[DLESM_SW_P]->fb.internal;
    const int nx = f[0]->nx, ny = f[0]->ny;
    const size_t row_bytes = static_cast<size_t>(nx) * sizeof(double);
    if (band_rows <= 0) band_rows = std::max(64, ((rt.ny / 128) + 31) / 32 * 32);
    band_rows = std::max(band_rows, 4);
    const int nbands = (rt.ny + band_rows - 1) / band_rows;
    if (rt.ny - (nbands - 1) * band_rows < 2 && nbands > 1) {
        set_error(DLESM_ERR_ARGS, "sw_step_host_n: the last row band must hold at least two rows (change band_rows)");
        return DLESM_ERR_ARGS;
    }
    while (c.band_events.size() < static_cast<size_t>(2 * nbands + 2)) {
        cudaEvent_t e;
        CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        c.band_events.push_back(e);
    }
    cudaEvent_t e_start = c.band_events[2 * nbands];
    CU(cudaEventRecord(e_start, c.stream));
    CU(cudaStreamWaitEvent(c.s_h2d, e_start, 0));
    CU(cudaStreamWaitEvent(c.s_d2h, e_start, 0));
    // set 0 = the model's current fields (uploaded into), set 1 = the spare ones; pass p reads set p % 2
    double *set[2][6];
    for (int q = 0; q < 6; ++q) set[0][q] = static_cast<double *>(f[DLESM_SW_U + q]->dptr);
    for (int q = 0; q < 3; ++q) {
        set[1][q] = static_cast<double *>(f[DLESM_SW_UNEW + q]->dptr);
        set[1][3 + q] = static_cast<double *>(s->old2[q]->dptr);
    }
    auto band_j0 = [&](int b) { return rt.ystart + b * band_rows; };
    auto band_j1 = [&](int b) { return std::min(rt.ystop, band_j0(b) + band_rows - 1); };
    std::vector<std::vector<char>> done(npass, std::vector<char>(nbands, 0));
    std::vector<int> ndone(npass, 0);
    std::vector<char> filled(npass, 0);
    auto issue = [&](int p, int b) -> int {
        double **in = set[p % 2], **out = set[(p + 1) % 2];
        const int j0 = band_j0(b), j1 = band_j1(b);
        int r2 = sw_step2_fused_impl(in[0], in[1], in[2], in[3], in[4], in[5], out[0], out[1], out[2], out[3], out[4],
                                     out[5], g.info.dx, g.info.dy, s->tdt, s->alpha, rt.xstart, rt.xstop, j0, j1,
                                     s->mask2, s->mask2_pitch, s->variant2, nullptr);
        if (r2 != DLESM_OK) return r2;
        CU(k::periodic_x_rows(out, 6, nx, k::Box{rt.xstart, rt.xstop, j0, j1}, 2, c.stream));
        if (p == npass - 1) {
            cudaEvent_t e_done = c.band_events[2 * b + 1];
            CU(cudaEventRecord(e_done, c.stream));
            CU(cudaStreamWaitEvent(c.s_d2h, e_done, 0));
            const size_t off = static_cast<size_t>(j0 - 1) * nx;
            const size_t bytes = static_cast<size_t>(j1 - j0 + 1) * row_bytes;
            for (int q = 0; q < 6; ++q)
                CU(cudaMemcpyAsync(host_out[q] + off, out[q] + off, bytes, cudaMemcpyDeviceToHost, c.s_d2h));
        }
        done[p][b] = 1;
        ++ndone[p];
        return DLESM_OK;
    };
    auto ready = [&](int p, int b) {
        if (done[p][b] || !done[p - 1][b]) return false;
        const bool south = b > 0 ? done[p - 1][b - 1] != 0 : filled[p - 1] != 0;
        const bool north = b < nbands - 1 ? done[p - 1][b + 1] != 0 : filled[p - 1] != 0;
        return south && north;
    };
    auto advance = [&]() -> int {     // everything the finished work makes ready, in pass order
        for (bool progress = true; progress;) {
            progress = false;
            for (int p = 0; p < npass; ++p) {
                if (ndone[p] == nbands && !filled[p]) {
                    // north / south halo rows (and the corners) of this pass's outputs
                    CU(k::periodic_halos_deep(set[(p + 1) % 2], 6, nx, box_of(rt), 2, c.stream));
                    filled[p] = 1;
                    progress = true;
                }
                if (p == 0) continue;
                for (int b = 0; b < nbands; ++b)
                    if (ready(p, b)) {
                        int r2 = issue(p, b);
                        if (r2 != DLESM_OK) return r2;
                        progress = true;
                    }
            }
        }
        return DLESM_OK;
    };
    int up_next = 1;
    for (int b = 0; b < nbands; ++b) {
        const int j1 = band_j1(b);
        const int up_to = (b == nbands - 1) ? ny : std::min(ny, j1 + 2);
        if (up_to >= up_next) {
            const size_t off = static_cast<size_t>(up_next - 1) * nx;
            const size_t bytes = static_cast<size_t>(up_to - up_next + 1) * row_bytes;
            for (int q = 0; q < 6; ++q)
                CU(cudaMemcpyAsync(set[0][q] + off, host_in[q] + off, bytes, cudaMemcpyHostToDevice, c.s_h2d));
            up_next = up_to + 1;
        }
        cudaEvent_t e_up = c.band_events[2 * b];
        CU(cudaEventRecord(e_up, c.s_h2d));
        CU(cudaStreamWaitEvent(c.stream, e_up, 0));
        rc = issue(0, b);
        if (rc == DLESM_OK) rc = advance();
        if (rc != DLESM_OK) return rc;
    }
    rc = advance();
    if (rc != DLESM_OK) return rc;
    if (ndone[npass - 1] != nbands || !filled[npass - 1]) {
        set_error(DLESM_ERR_STATE, "sw_step_host_n: the band schedule did not complete");
        return DLESM_ERR_STATE;
    }
    double **fin = set[npass % 2];
    // the two + two halo rows, full width, once every band's download has landed
    cudaEvent_t e_dl = c.band_events[2 * nbands + 1];
    CU(cudaEventRecord(e_dl, c.s_d2h));
    CU(cudaStreamWaitEvent(c.stream, e_dl, 0));
    for (int q = 0; q < 6; ++q) {
        const int rows[2] = {rt.ystart - 2, rt.ystop + 1};
        for (int r : rows) {
            const size_t off = static_cast<size_t>(r - 1) * nx;
            CU(cudaMemcpyAsync(host_out[q] + off, fin[q] + off, 2 * row_bytes, cudaMemcpyDeviceToHost, c.stream));
        }
    }
    CU(cudaStreamSynchronize(c.stream));
    if (npass % 2 == 1)
        for (int q = 0; q < 3; ++q) {
            std::swap(f[DLESM_SW_U + q]->dptr, f[DLESM_SW_UNEW + q]->dptr);
            std::swap(f[DLESM_SW_UOLD + q]->dptr, s->old2[q]->dptr);
        }
    s->deep_valid = true;
    s->ncycle += 2 * npass;
    return DLESM_OK;
}
}  // namespace

// Two model steps per PCIe round trip: the bands go through the two-step sweep (temporal blocking,
// tile with two halo cells), so each row of each field crosses the bus once per direction for
// TWO steps -- half the bytes per model step of dlesm_sw_step_host.  nsteps = 1 is that call.
// Needs dlesm_sw_set_two_step(on), the exchange halo mode and a model past its first step (the
// forward start-up step has no two-step form); host arrays carry the two-cell halo ring in and out.
extern "C" int dlesm_sw_step_host_n(dlesm_sw_t *s, const double *const host_in[6], double *const host_out[6],
                                    int band_rows, int nsteps)
{
    if (nsteps == 1) return dlesm_sw_step_host(s, host_in, host_out, band_rows);
    clear_error();
    if (s && nsteps >= 4 && nsteps % 2 == 0 && s->two_step && !s->push && s->ncycle > 0 && s->periodic_local)
        return sw_step_host_passes(s, host_in, host_out, band_rows, nsteps / 2);
    if (!s || nsteps != 2 || !s->two_step || s->push || s->ncycle == 0) {
        set_error(DLESM_ERR_ARGS, "sw_step_host_n: nsteps is 1, 2 or (one rank) an even number; more than 1 needs "
                                  "set_two_step(on), the exchange halo mode and a model past its first step");
        return DLESM_ERR_ARGS;
    }
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    Context &c = ctx();
    dlesm_grid &g = *s->grid;
    dlesm_field **f = s->f;
    const dlesm_region &rt = f[DLESM_SW_P]->fb.internal;
    const int nx = f[0]->nx, ny = f[0]->ny;
    const size_t row_bytes = static_cast<size_t>(nx) * sizeof(double);
    if (band_rows <= 0) band_rows = std::max(64, ((rt.ny / 128) + 31) / 32 * 32);
    const int nbands = (rt.ny + band_rows - 1) / band_rows;
    while (c.band_events.size() < static_cast<size_t>(2 * nbands + 2)) {
        cudaEvent_t e;
        CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        c.band_events.push_back(e);
    }
    cudaEvent_t e_start = c.band_events[2 * nbands];
    CU(cudaEventRecord(e_start, c.stream));
    CU(cudaStreamWaitEvent(c.s_h2d, e_start, 0));
    CU(cudaStreamWaitEvent(c.s_d2h, e_start, 0));
    dlesm_field *in_f[6], *out_f[6];
    for (int q = 0; q < 6; ++q) in_f[q] = f[DLESM_SW_U + q];                    // u,v,p,uold,vold,pold
    for (int q = 0; q < 3; ++q) { out_f[q] = f[DLESM_SW_UNEW + q]; out_f[3 + q] = s->old2[q]; }
    int up_next = 1;
    for (int b = 0; b < nbands; ++b) {
        const int j0 = rt.ystart + b * band_rows;
        const int j1 = std::min(rt.ystop, j0 + band_rows - 1);
        // rows the two-step sweep of [j0, j1] reads: j0-2 .. j1+2 of u,v,p; j0-1 .. j1+1 of the old fields
        const int up_to = (b == nbands - 1) ? ny : std::min(ny, j1 + 2);
        if (up_to >= up_next) {
            const size_t off = static_cast<size_t>(up_next - 1) * nx;
            const size_t bytes = static_cast<size_t>(up_to - up_next + 1) * row_bytes;
            for (int q = 0; q < 6; ++q)
                CU(cudaMemcpyAsync(static_cast<double *>(in_f[q]->dptr) + off, host_in[q] + off, bytes,
                                   cudaMemcpyHostToDevice, c.s_h2d));
            up_next = up_to + 1;
        }
        cudaEvent_t e_up = c.band_events[2 * b], e_done = c.band_events[2 * b + 1];
        CU(cudaEventRecord(e_up, c.s_h2d));
        CU(cudaStreamWaitEvent(c.stream, e_up, 0));
        rc = sw_step2_fused_impl(in_f[0]->dptr, in_f[1]->dptr, in_f[2]->dptr, in_f[3]->dptr, in_f[4]->dptr,
                                 in_f[5]->dptr, out_f[0]->dptr, out_f[1]->dptr, out_f[2]->dptr, out_f[3]->dptr,
                                 out_f[4]->dptr, out_f[5]->dptr, g.info.dx, g.info.dy, s->tdt, s->alpha, rt.xstart,
                                 rt.xstop, j0, j1, s->mask2, s->mask2_pitch, s->variant2, nullptr);
        if (rc != DLESM_OK) return rc;
        CU(cudaEventRecord(e_done, c.stream));
        CU(cudaStreamWaitEvent(c.s_d2h, e_done, 0));
        const size_t off = static_cast<size_t>(j0 - 1) * nx;
        const size_t bytes = static_cast<size_t>(j1 - j0 + 1) * row_bytes;
        for (int q = 0; q < 6; ++q)
            CU(cudaMemcpyAsync(host_out[q] + off, static_cast<const double *>(out_f[q]->dptr) + off, bytes,
                               cudaMemcpyDeviceToHost, c.s_d2h));
    }
    for (int q = 0; q < 3; ++q) {
        std::swap(f[DLESM_SW_U + q]->dptr, f[DLESM_SW_UNEW + q]->dptr);
        std::swap(f[DLESM_SW_UOLD + q]->dptr, s->old2[q]->dptr);
    }
    rc = sw_fill_deep_halos(s);            // two-cell ring of all six fields (wrapped copies / depth-2 exchange)
    if (rc != DLESM_OK) return rc;
    s->deep_valid = true;
    s->ncycle += 2;
    // ring -> host: 2 + 2 halo rows are contiguous; the 2 + 2 halo columns are gathered on the
    // device, shipped as one block and scattered by the host
    const size_t col_n = 2 * static_cast<size_t>(ny);
    rc = ensure_scratch(12 * col_n);
    if (rc != DLESM_OK) return rc;
    if (c.ring_host_n < 12 * col_n) {
        if (c.ring_host) cudaFreeHost(c.ring_host);
        c.ring_host = nullptr;
        c.ring_host_n = 0;
        CU(cudaMallocHost(&c.ring_host, 12 * col_n * sizeof(double)));
        c.ring_host_n = 12 * col_n;
    }
    const int xw = rt.xstart - 2, xe = rt.xstop + 1;      // first column of the west / east halo pair
    for (int q = 0; q < 6; ++q) {
        const double *d = static_cast<const double *>(f[DLESM_SW_U + q]->dptr);
        CU(k::gather_box(d, nx, k::Box{xw, xw + 1, 1, ny}, c.scratch + (2 * q) * col_n, c.stream));
        CU(k::gather_box(d, nx, k::Box{xe, xe + 1, 1, ny}, c.scratch + (2 * q + 1) * col_n, c.stream));
    }
    CU(cudaMemcpyAsync(c.ring_host, c.scratch, 12 * col_n * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    cudaEvent_t e_dl = c.band_events[2 * nbands + 1];
    CU(cudaEventRecord(e_dl, c.s_d2h));
    CU(cudaStreamWaitEvent(c.stream, e_dl, 0));
    for (int q = 0; q < 6; ++q) {
        const double *d = static_cast<const double *>(f[DLESM_SW_U + q]->dptr);
        const int rows[2] = {rt.ystart - 2, rt.ystop + 1};
        for (int r : rows) {
            const size_t off = static_cast<size_t>(r - 1) * nx;
            CU(cudaMemcpyAsync(host_out[q] + off, d + off, 2 * row_bytes, cudaMemcpyDeviceToHost, c.stream));
        }
    }
    CU(cudaStreamSynchronize(c.stream));
    for (int q = 0; q < 6; ++q) {
        double *h = host_out[q];
        const double *w = c.ring_host + (2 * q) * col_n, *e = c.ring_host + (2 * q + 1) * col_n;
        for (int j = rt.ystart - 2; j <= rt.ystop + 2; ++j) {
            double *row = h + static_cast<size_t>(j - 1) * nx;
            row[xw - 1] = w[2 * (j - 1)];
            row[xw] = w[2 * (j - 1) + 1];
            row[xe - 1] = e[2 * (j - 1)];
            row[xe] = e[2 * (j - 1) + 1];
        }
    }
    return DLESM_OK;
}

extern "C" dlesm_field_t *dlesm_sw_field(dlesm_sw_t *s, int which)
{
    if (which < 0 || which >= DLESM_SW_NFIELDS) return nullptr;
    return s->f[which];
}
extern "C" int dlesm_sw_set_variant(dlesm_sw_t *s, int variant)
{
    s->variant = variant;
    return DLESM_OK;
}
// Collective when the object pushes halos on several ranks (its buffers are mapped by the
// neighbours: they unmap, then a barrier, then the memory is freed) -- call it on every rank.
extern "C" void dlesm_sw_free(dlesm_sw_t *s)
{
    if (!s) return;
    sw_release_peers(s);
    for (auto *fl : s->f)
        if (fl) dlesm_free_field(fl);
    for (auto *fl : s->old2)
        if (fl) dlesm_free_field(fl);
    if (s->mask2) {
        k::fma_off::forget_mask(s->mask2);
        k::fma_on::forget_mask(s->mask2);
        cudaFree(s->mask2);
    }
    delete s;
}

// Non-collective release for finalisers (a garbage collector must never enter a collective: ranks
// would have to agree on its order).  Buffers the neighbours still have mapped are NOT freed --
// they stay allocated until dlesm_dev_peer_unregister_all / process exit -- everything else is.
extern "C" void dlesm_sw_free_local(dlesm_sw_t *s)
{
    if (!s) return;
    const bool mapped = !s->registered.empty() && pstate().nranks > 1;
    auto is_mapped = [&](void *p) {
        return mapped && std::find(s->registered.begin(), s->registered.end(), p) != s->registered.end();
    };
    auto drop = [&](dlesm_field *fl) {
        if (!fl) return;
        if (is_mapped(fl->dptr)) {
            std::fprintf(stderr, "dlesm: shallow-water object finalised while its buffers are mapped by the neighbours; "
                                 "call free() on every rank instead -- leaking %p\n", fl->dptr);
            delete fl;
        } else {
            dlesm_free_field(fl);
        }
    };
    if (!mapped) sw_release_peers(s);   // one rank: local bookkeeping only
    for (auto *fl : s->f) drop(fl);
    for (auto *fl : s->old2) drop(fl);
    if (s->mask2) {
        k::fma_off::forget_mask(s->mask2);
        k::fma_on::forget_mask(s->mask2);
        cudaFree(s->mask2);
    }
    delete s;
}

// the launch grid (grid.x, grid.y) the one-step (two_step = 0) or two-step sweep of this model uses:
// what the tile-grid knob (GOCEAN_OMP_GRID / dlesm_dev_set_sweep_tiles) changes
extern "C" int dlesm_sw_launch_grid(dlesm_sw_t *s, int two_step, int grid[2])
{
    clear_error();
    const dlesm_region &rt = s->f[DLESM_SW_P]->fb.internal;
    const bool fma = ctx().arith == DLESM_ARITH_CONTRACTED;
    if (two_step) {
        k::Fused2Args a{};
        a.pitch = s->f[0]->nx; a.ny = s->f[0]->ny;
        a.b = box_of(rt);
        a.tiles = ctx().sweep_tiles;
        if (fma) k::fma_on::fused2_grid(a, s->variant2, grid);
        else k::fma_off::fused2_grid(a, s->variant2, grid);
    } else {
        k::FusedArgs a{};
        a.pitch = s->f[0]->nx; a.ny = s->f[0]->ny;
        a.b = box_of(rt);
        a.tiles = ctx().sweep_tiles;
        a.tmask = s->tmask_dev;
        static const unsigned char dummy = 0;
        if (a.tmask) { a.mask8 = &dummy; a.mask8_pitch = 16; }   // geometry only
        if (fma) k::fma_on::fused_grid(a, s->variant, grid);
        else k::fma_off::fused_grid(a, s->variant, grid);
    }
    return DLESM_OK;
}

// =========================================================================================
// Several sub-domains per rank (SURVEY F3 extension).  The reference carries the bookkeeping
// (decomposition_type%proc_subdomains, src/decomposition_mod.f90:64-67; go_decompose(ndomains > nranks),
// src/parallel_mod.f90:141-152) and stops at map_comms ("\TODO support multiple subdomains per rank",
// src/parallel_comms_mod.f90:295).  Here: a tile set owns one ordinary grid per sub-domain of this rank;
// fields and models are created on those grids as usual; ONE exchange fills the halos of every tile --
// same-rank neighbours by device copies, the others packed per peer rank (build_tileset_plan).
// =========================================================================================
struct dlesm_tileset {
    dlesm_grid *parent = nullptr;          // name, offset, BCs and the decomposition
    std::vector<dlesm_grid *> tiles;       // one grid per owned sub-domain, in proc_subdomains order
    std::vector<TilesetPlan> plans;        // index = depth - 1, built on first use
    bool pbc_x = false, pbc_y = false;
};

extern "C" dlesm_tileset_t *dlesm_tileset_create(int grid_name, const int bcs[3], int grid_offsets, int domainx,
                                                 int domainy, int ndomains, int ndomainx, int ndomainy,
                                                 int halo_width)
{
    clear_error();
    dlesm_grid *parent = dlesm_grid_create(grid_name, bcs, grid_offsets);
    if (!parent) return nullptr;
    int rc = go_decompose(domainx, domainy, ndomains, ndomainx, ndomainy, halo_width, pstate().nranks, parent->decomp);
    if (rc != DLESM_OK) {
        delete parent;
        return nullptr;
    }
    const Decomposition &d = parent->decomp;
    if (d.ndomains % pstate().nranks != 0) {
        set_error(DLESM_ERR_ARGS, "tile set: the number of sub-domains must be a multiple of the number of ranks");
        delete parent;
        return nullptr;
    }
    auto *ts = new dlesm_tileset;
    ts->parent = parent;
    ts->pbc_x = bcs[0] == DLESM_GO_BC_PERIODIC;
    ts->pbc_y = bcs[1] == DLESM_GO_BC_PERIODIC;
    parent->info.global_nx = d.global_nx;
    parent->info.global_ny = d.global_ny;
    const int me = pstate().rank;
    for (int s = 0; s < d.domperrank; ++s) {
        const int id = d.proc_subdomains[static_cast<size_t>(me - 1) * d.domperrank + s];
        if (id < 1) continue;
        auto *g = new dlesm_grid;
        g->info = parent->info;
        g->decomp = d;
        g->info.subdomain = d.subdomains[id - 1];
        g->decomposed = true;
        g->tile_id = id;
        ts->tiles.push_back(g);
    }
    parent->info.subdomain = ts->tiles.empty() ? dlesm_subdomain{} : ts->tiles[0]->info.subdomain;
    parent->decomposed = true;
    return ts;
}

extern "C" int dlesm_tileset_ntiles(const dlesm_tileset_t *ts) { return static_cast<int>(ts->tiles.size()); }
extern "C" dlesm_grid_t *dlesm_tileset_grid(dlesm_tileset_t *ts, int k)
{
    if (k < 0 || k >= static_cast<int>(ts->tiles.size())) return nullptr;
    return ts->tiles[k];
}
extern "C" int dlesm_tileset_tile_id(const dlesm_tileset_t *ts, int k)
{
    if (k < 0 || k >= static_cast<int>(ts->tiles.size())) return -1;
    return ts->tiles[k]->tile_id;
}
extern "C" const dlesm_decomp_t *dlesm_tileset_decomp(const dlesm_tileset_t *ts) { return &ts->parent->dh; }

extern "C" void dlesm_tileset_free(dlesm_tileset_t *ts)
{
    if (!ts) return;
    for (auto *g : ts->tiles) dlesm_grid_free(g);
    dlesm_grid_free(ts->parent);
    delete ts;
}

// the plan as rows of 14 integers (CPU-testable: no device involved):
// kind, peer0, src_tile, dst_tile, si0, sj0, di0, dj0, nx, ny, offset, dst_id, src_id, dir
extern "C" int dlesm_tileset_plan(const dlesm_decomp_t *d, int rank1, int nranks, int pbc_x, int pbc_y, int depth,
                                  long long *rows, int max_rows, int *nrows, long long totals[2])
{
    clear_error();
    TilesetPlan plan;
    int rc = build_tileset_plan(d->d, rank1, nranks, pbc_x != 0, pbc_y != 0, depth, plan);
    if (rc != DLESM_OK) return rc;
    *nrows = static_cast<int>(plan.segs.size());
    totals[0] = plan.send_total;
    totals[1] = plan.recv_total;
    if (*nrows > max_rows) {
        set_error(DLESM_ERR_ARGS, "tileset_plan: more rows than the caller's array holds");
        return DLESM_ERR_ARGS;
    }
    for (int q = 0; q < *nrows; ++q) {
        const TileSeg &g = plan.segs[q];
        const long long v[14] = {g.kind, g.peer0, g.src_tile, g.dst_tile, g.si0, g.sj0, g.di0, g.dj0,
                                 g.nx, g.ny, g.offset, g.dst_id, g.src_id, g.dir};
        std::copy(v, v + 14, rows + static_cast<size_t>(q) * 14);
    }
    return DLESM_OK;
}

namespace {
int tileset_plan_for(dlesm_tileset *ts, int depth, const TilesetPlan *&out)
{
    if (depth < 1 || depth > ts->parent->decomp.hwidth) {
        set_error(DLESM_ERR_HALO_DEPTH, "tile set: halo depth exceeds the halo width of the decomposition");
        return DLESM_ERR_HALO_DEPTH;
    }
    if (static_cast<int>(ts->plans.size()) < depth) ts->plans.resize(depth);
    TilesetPlan &p = ts->plans[depth - 1];
    if (p.owned.empty()) {
        int rc = build_tileset_plan(ts->parent->decomp, pstate().rank, pstate().nranks, ts->pbc_x, ts->pbc_y, depth, p);
        if (rc != DLESM_OK) return rc;
    }
    out = &p;
    return DLESM_OK;
}

// fields[q * ntiles + k] = device array of field q on tile k
int tileset_exchange(dlesm_tileset *ts, void *const *fields, int nfields, int depth)
{
    const TilesetPlan *pp;
    int rc = tileset_plan_for(ts, depth, pp);
    if (rc != DLESM_OK) return rc;
    const TilesetPlan &plan = *pp;
    const int nt = static_cast<int>(ts->tiles.size());
    if (nfields < 1 || nfields > k::kMaxExchangeFields) {
        set_error(DLESM_ERR_ARGS, "tile set exchange: between 1 and 8 fields per call");
        return DLESM_ERR_ARGS;
    }
    rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    Context &c = ctx();
    std::vector<int> pitch(nt);
    for (int k = 0; k < nt; ++k)
        for (int q = 0; q < nfields; ++q) {
            FieldRec r;
            rc = lookup(fields[q * nt + k], r);
            if (rc != DLESM_OK) return rc;
            if (r.nx != ts->tiles[k]->info.nx || r.ny != ts->tiles[k]->info.ny) {
                set_error(DLESM_ERR_ARGS, "tile set exchange: a field does not have its tile's shape");
                return DLESM_ERR_ARGS;
            }
            pitch[k] = r.nx;
        }
    rc = ensure_xbuf(static_cast<size_t>(std::max(plan.send_total, plan.recv_total)) * nfields + 1);
    if (rc != DLESM_OK) return rc;
    bool remote = false;
    for (const TileSeg &g : plan.segs) remote = remote || g.kind != 0;
    if (remote && !c.comm_ready) {
        set_error(DLESM_ERR_STATE, "tile set exchange: communicator not initialised (dlesm_comm_init)");
        return DLESM_ERR_STATE;
    }
    auto at = [&](int q, int k, int i0, int j0) {
        return static_cast<double *>(fields[q * nt + k]) + (static_cast<size_t>(j0) - 1) * pitch[k] + (i0 - 1);
    };
    k::TileOpBatch batch;
    batch.n = 0;
    auto flush = [&]() -> int {
        if (batch.n > 0) CU(k::tile_ops(batch, c.stream));
        batch.n = 0;
        return DLESM_OK;
    };
    auto add = [&](const double *src, int sp, double *dst, int dp, int nx, int ny) -> int {
        if (batch.n == k::kTileOpsPerLaunch) {
            int r2 = flush();
            if (r2 != DLESM_OK) return r2;
        }
        batch.op[batch.n++] = k::TileOp{src, dst, sp, dp, nx, ny};
        return DLESM_OK;
    };
    // phase 1: same-rank halos directly, everything that leaves the rank into the send buffer
    // (segment-major, field-minor: both ends agree on it)
    for (const TileSeg &g : plan.segs) {
        for (int q = 0; q < nfields && rc == DLESM_OK; ++q) {
            if (g.kind == 0)
                rc = add(at(q, g.src_tile, g.si0, g.sj0), pitch[g.src_tile], at(q, g.dst_tile, g.di0, g.dj0),
                         pitch[g.dst_tile], g.nx, g.ny);
            else if (g.kind == 1)
                rc = add(at(q, g.src_tile, g.si0, g.sj0), pitch[g.src_tile],
                         c.sendbuf + g.offset * nfields + static_cast<long long>(q) * g.nx * g.ny, g.nx, g.nx, g.ny);
        }
        if (rc != DLESM_OK) return rc;
    }
    rc = flush();
    if (rc != DLESM_OK) return rc;
    if (remote) {
        // one message per peer rank and direction of travel, whatever the number of tile pairs behind it
        std::map<int, std::pair<long long, long long>> so, ro;   // peer -> (first offset, elements), one field
        for (const TileSeg &g : plan.segs) {
            if (g.kind == 0) continue;
            auto &m = g.kind == 1 ? so : ro;
            auto it = m.find(g.peer0);
            const long long n = static_cast<long long>(g.nx) * g.ny;
            if (it == m.end()) m[g.peer0] = {g.offset, n};
            else it->second.second += n;
        }
        NC(ncclGroupStart());
        for (const auto &kv : so)
            NC(ncclSend(c.sendbuf + kv.second.first * nfields, static_cast<size_t>(kv.second.second) * nfields,
                        ncclDouble, kv.first, c.comm, c.stream));
        for (const auto &kv : ro)
            NC(ncclRecv(c.recvbuf + kv.second.first * nfields, static_cast<size_t>(kv.second.second) * nfields,
                        ncclDouble, kv.first, c.comm, c.stream));
        NC(ncclGroupEnd());
        for (const TileSeg &g : plan.segs) {
            if (g.kind != 2) continue;
            for (int q = 0; q < nfields && rc == DLESM_OK; ++q)
                rc = add(c.recvbuf + g.offset * nfields + static_cast<long long>(q) * g.nx * g.ny, g.nx,
                         at(q, g.dst_tile, g.di0, g.dj0), pitch[g.dst_tile], g.nx, g.ny);
            if (rc != DLESM_OK) return rc;
        }
        rc = flush();
        if (rc != DLESM_OK) return rc;
    }
    for (int q = 0; q < nfields * nt; ++q) mark_clean(fields[q], depth);
    return DLESM_OK;
}
}  // namespace

// halo_exchange(depth) of `nfields` fields on every tile of the set; fields[q * ntiles + k] = field q of tile k
extern "C" int dlesm_tileset_halo_exchange(dlesm_tileset_t *ts, dlesm_field_t *const *fields, int nfields, int depth)
{
    clear_error();
    const int nt = static_cast<int>(ts->tiles.size());
    std::vector<void *> ptrs(static_cast<size_t>(nfields) * nt);
    for (size_t q = 0; q < ptrs.size(); ++q) {
        if (fields[q]->grid != ts->tiles[q % nt]) {
            set_error(DLESM_ERR_ARGS, "tile set exchange: fields must be listed field-major, one per tile, on the set's grids");
            return DLESM_ERR_ARGS;
        }
        ptrs[q] = fields[q]->dptr;
    }
    return tileset_exchange(ts, ptrs.data(), nfields, depth);
}

// gather_inner_data (src/field_mod.f90:1311-1391) of one field per tile into the global array on rank 1
extern "C" int dlesm_tileset_gather_inner_data(dlesm_tileset_t *ts, dlesm_field_t *const *fields, double *global_data)
{
    clear_error();
    Context &c = ctx();
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    const Decomposition &d = ts->parent->decomp;
    const int nt = static_cast<int>(ts->tiles.size());
    const int nranks = pstate().nranks, me = pstate().rank;
    const int hw = d.hwidth;
    const size_t slot = static_cast<size_t>(d.max_width - 2 * hw) * (d.max_height - 2 * hw);   // one tile
    const size_t per_rank = slot * d.domperrank;
    rc = ensure_scratch(per_rank * (me == 1 ? nranks : 1));
    if (rc != DLESM_OK) return rc;
    for (int k = 0; k < nt; ++k)
        CU(k::gather_box(static_cast<const double *>(fields[k]->dptr), fields[k]->nx,
                         box_of(ts->tiles[k]->info.subdomain.internal), c.scratch + slot * k, c.stream));
    if (nranks > 1) {
        if (!c.comm_ready) {
            set_error(DLESM_ERR_STATE, "tile set gather: communicator not initialised (dlesm_comm_init)");
            return DLESM_ERR_STATE;
        }
        NC(ncclGroupStart());
        if (me == 1) {
            for (int r = 1; r < nranks; ++r)
                NC(ncclRecv(c.scratch + per_rank * r, per_rank, ncclDouble, r, c.comm, c.stream));
        } else {
            NC(ncclSend(c.scratch, per_rank, ncclDouble, 0, c.comm, c.stream));
        }
        NC(ncclGroupEnd());
    }
    if (me != 1) {
        CU(cudaStreamSynchronize(c.stream));
        return DLESM_OK;
    }
    std::vector<double> recv(per_rank * nranks);
    CU(cudaMemcpyAsync(recv.data(), c.scratch, recv.size() * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    for (int r = 0; r < nranks; ++r)
        for (int s = 0; s < d.domperrank; ++s) {
            const int id = d.proc_subdomains[static_cast<size_t>(r) * d.domperrank + s];
            if (id < 1) continue;
            // T-point internal region of the tile == its piece of the global domain; the packed box of
            // any staggering has the field's own internal extents, which for interior tiles equal it
            const dlesm_region &gr = d.subdomains[id - 1].global;
            const dlesm_region &in = d.subdomains[id - 1].internal;
            size_t q = per_rank * r + slot * s;
            for (int jj = 0; jj < in.ny; ++jj)
                for (int ji = 0; ji < in.nx; ++ji)
                    global_data[(static_cast<size_t>(gr.ystart) - 1 + jj) * d.global_nx + (gr.xstart - 1 + ji)] = recv[q++];
        }
    return DLESM_OK;
}

// nsteps steps of the fused shallow-water model on every tile: per step one sweep per tile, then
// ONE exchange of unew, vnew, pnew over all tiles, then the buffer rotation.
extern "C" int dlesm_tileset_sw_step(dlesm_tileset_t *ts, dlesm_sw_t *const *sws, int nsteps)
{
    clear_error();
    const int nt = static_cast<int>(ts->tiles.size());
    for (int k = 0; k < nt; ++k)
        if (!sws[k] || sws[k]->grid != ts->tiles[k] || sws[k]->mode != DLESM_SW_FUSED ||
            sws[k]->ncycle != sws[0]->ncycle) {
            set_error(DLESM_ERR_ARGS, "tileset_sw_step: one fused-mode model per tile, in tile order, all at the same step");
            return DLESM_ERR_ARGS;
        }
    int rc = ensure_device();
    if (rc != DLESM_OK) return rc;
    std::vector<void *> ptrs(static_cast<size_t>(3) * nt);
    for (int it = 0; it < nsteps; ++it) {
        for (int k = 0; k < nt; ++k) {
            dlesm_sw *s = sws[k];
            dlesm_field **f = s->f;
            const dlesm_region &rt = f[DLESM_SW_P]->fb.internal;
            const int first = s->ncycle == 0;
            rc = sw_step_fused_impl(f[DLESM_SW_U]->dptr, f[DLESM_SW_V]->dptr, f[DLESM_SW_P]->dptr,
                                    f[DLESM_SW_UOLD]->dptr, f[DLESM_SW_VOLD]->dptr, f[DLESM_SW_POLD]->dptr,
                                    f[DLESM_SW_UNEW]->dptr, f[DLESM_SW_VNEW]->dptr, f[DLESM_SW_PNEW]->dptr,
                                    s->grid->info.dx, s->grid->info.dy, s->tdt, s->alpha, first, rt.xstart, rt.xstop,
                                    rt.ystart, rt.ystop, s->tmask_dev, s->variant, nullptr);
            if (rc != DLESM_OK) return rc;
            for (int q = 0; q < 3; ++q) ptrs[static_cast<size_t>(q) * nt + k] = f[DLESM_SW_UNEW + q]->dptr;
        }
        rc = tileset_exchange(ts, ptrs.data(), 3, 1);
        if (rc != DLESM_OK) return rc;
        for (int k = 0; k < nt; ++k) {
            dlesm_sw *s = sws[k];
            for (int q = 0; q < 3; ++q) std::swap(s->f[DLESM_SW_U + q]->dptr, s->f[DLESM_SW_UNEW + q]->dptr);
            if (s->ncycle == 0) s->tdt = s->tdt + s->tdt;
            ++s->ncycle;
            s->deep_valid = false;
        }
    }
    return DLESM_OK;
}

// halos of u, v, p after dlesm_sw_init_device / dlesm_sw_set_state on every tile, and the old fields from them
extern "C" int dlesm_tileset_sw_sync_state(dlesm_tileset_t *ts, dlesm_sw_t *const *sws)
{
    clear_error();
    const int nt = static_cast<int>(ts->tiles.size());
    std::vector<void *> ptrs(static_cast<size_t>(3) * nt);
    for (int k = 0; k < nt; ++k)
        for (int q = 0; q < 3; ++q) ptrs[static_cast<size_t>(q) * nt + k] = sws[k]->f[DLESM_SW_U + q]->dptr;
    int rc = tileset_exchange(ts, ptrs.data(), 3, 1);
    for (int k = 0; k < nt && rc == DLESM_OK; ++k)
        for (int q = 0; q < 3 && rc == DLESM_OK; ++q)
            rc = dlesm_copy_field(sws[k]->f[DLESM_SW_U + q], sws[k]->f[DLESM_SW_UOLD + q]);
    return rc;
}
