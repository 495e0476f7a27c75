// dlesm_host.cpp -- host-side bookkeeping (no CUDA).  See dlesm_host.hpp.
//
// Contract: bit-identical integer results to the reference's Fortran for every
// configuration the reference supports; extensions (periodic wrap across ranks, halo
// depth > 1) are flagged where they start.
#include "dlesm_host.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace dlesm {

// ---------------------------------------------------------------------------
// error state
// ---------------------------------------------------------------------------
namespace {
thread_local int t_status = DLESM_OK;
thread_local std::string t_message;
// go_decompose's `hwidth` is a SAVEd local in the reference (src/parallel_mod.f90:103-105):
// a halo_width passed once sticks for later calls that omit it.  Mirrored for parity.
int g_saved_hwidth = 1;
}  // namespace

ParallelState &pstate()
{
    static ParallelState s;
    return s;
}

void set_error(int code, const std::string &msg)
{
    t_status = code;
    t_message = msg;
}
int last_status() { return t_status; }
const char *last_error_cstr() { return t_message.c_str(); }
void clear_error()
{
    t_status = DLESM_OK;
    t_message.clear();
}
void reset_saved_state() { g_saved_hwidth = 1; }

// ---------------------------------------------------------------------------
// go_decompose -- src/parallel_mod.f90:70-332
// ---------------------------------------------------------------------------
namespace {
// Split `len` points over `n` tiles: len/n each, the first (len - n*(len/n)) tiles get one
// more (src/parallel_mod.f90:204-223, 251-270).  Returns the 1-based global start of every
// tile plus one-past-the-end.
std::vector<int> split_axis(int len, int n)
{
    const int base = len / n;
    int extra = (n * base < len) ? len - n * base : 0;
    std::vector<int> cuts(n + 1);
    int pos = 1;
    for (int t = 0; t < n; ++t) {
        cuts[t] = pos;
        pos += base + (extra > 0 ? 1 : 0);
        if (extra > 0) --extra;
    }
    cuts[n] = pos;
    return cuts;
}
}  // namespace

int go_decompose(int domainx, int domainy, int ndomains, int ndomainx, int ndomainy,
                 int halo_width, int nranks, Decomposition &d)
{
    int ndom;
    bool auto_tile;
    if (ndomains < 0) {
        if (ndomainx < 0 && ndomainy < 0) {
            ndom = nranks;  // one subdomain per rank (:112-116)
            auto_tile = true;
        } else if (ndomainx >= 0 && ndomainy >= 0) {
            ndom = ndomainx * ndomainy;
            auto_tile = false;
        } else {
            set_error(DLESM_ERR_ARGS, "go_decompose: invalid arguments supplied");
            return DLESM_ERR_ARGS;
        }
    } else {
        ndom = ndomains;
        auto_tile = true;
    }
    if (nranks < 1) {
        set_error(DLESM_ERR_NRANKS, "go_decompose: nranks must be >= 1 ");
        return DLESM_ERR_NRANKS;
    }
    if (halo_width >= 0) {
        if (halo_width < 1 && nranks > 1) {
            set_error(DLESM_ERR_HALO_WIDTH,
                      "go_decompose: halo width must be > 0 if running on more than one process");
            return DLESM_ERR_HALO_WIDTH;
        }
        g_saved_hwidth = halo_width;
    }
    if (ndom < 1) {
        set_error(DLESM_ERR_ARGS, "go_decompose: number of sub-domains must be >= 1");
        return DLESM_ERR_ARGS;
    }
    const int hw = g_saved_hwidth;

    // tile grid (:167-194): largest divisor of ndom not above sqrt(ndom), the longer domain
    // side gets the larger factor.  REAL() is single precision in the reference.
    int ntilex, ntiley;
    if (auto_tile) {
        ntilex = static_cast<int>(std::sqrt(static_cast<float>(ndom)));
        while (ndom % ntilex != 0) --ntilex;
        ntiley = ndom / ntilex;
        const bool want_x_larger = domainx > domainy;
        if ((want_x_larger && ntilex < ntiley) || (!want_x_larger && ntiley < ntilex))
            std::swap(ntilex, ntiley);
    } else {
        ntilex = ndomainx;
        ntiley = ndomainy;
    }
    if (ntilex < 1 || ntiley < 1) {
        set_error(DLESM_ERR_ARGS, "go_decompose: tiling grid dimensions must be >= 1");
        return DLESM_ERR_ARGS;
    }

    d = Decomposition();
    d.global_nx = domainx;
    d.global_ny = domainy;
    d.nx = ntilex;
    d.ny = ntiley;
    d.ndomains = ndom;
    d.nranks = nranks;
    d.hwidth = hw;
    // ceiling(real(ndom)/real(nranks)) (:142), single precision like the reference
    d.domperrank = static_cast<int>(std::ceil(static_cast<float>(ndom) / static_cast<float>(nranks)));
    d.proc_subdomains.assign(static_cast<size_t>(d.domperrank) * nranks, -1);
    {
        int next = 0;
        for (int r = 0; r < nranks; ++r)
            for (int s = 0; s < d.domperrank; ++s) {
                ++next;
                if (next > ndom) break;
                d.proc_subdomains[static_cast<size_t>(r) * d.domperrank + s] = next;
            }
    }

    d.xcuts = split_axis(domainx, ntilex);
    d.ycuts = split_axis(domainy, ntiley);
    d.subdomains.resize(static_cast<size_t>(ntilex) * ntiley);
    for (int ty = 0; ty < ntiley; ++ty) {
        const int height = d.ycuts[ty + 1] - d.ycuts[ty];
        for (int tx = 0; tx < ntilex; ++tx) {
            const int width = d.xcuts[tx + 1] - d.xcuts[tx];
            dlesm_subdomain &s = d.subdomains[static_cast<size_t>(ty) * ntilex + tx];
            s.internal.xstart = hw + 1;
            s.internal.xstop = hw + width;
            s.internal.nx = width;
            s.internal.ystart = hw + 1;
            s.internal.ystop = hw + height;
            s.internal.ny = height;
            s.global.xstart = d.xcuts[tx];
            s.global.xstop = d.xcuts[tx] + width - 1;
            s.global.nx = 2 * hw + width;  // full width incl. halos (:281)
            s.global.ystart = d.ycuts[ty];
            s.global.ystop = d.ycuts[ty] + height - 1;
            s.global.ny = 2 * hw + height;
            d.max_width = std::max(d.max_width, s.global.nx);
            d.max_height = std::max(d.max_height, s.global.ny);
        }
    }
    // ndom may differ from ntilex*ntiley only through explicit ndomainx/y, where they agree
    d.subdomains.resize(static_cast<size_t>(ndom));
    return DLESM_OK;
}

// ---------------------------------------------------------------------------
// iprocmap -- src/parallel_comms_mod.f90:1365-1398, by bisection on the tile cuts
// ---------------------------------------------------------------------------
int owner_of(const Decomposition &d, int nranks, int ia, int ja)
{
    if (ia < 1 || ia > d.global_nx || ja < 1 || ja > d.global_ny) return 0;
    const int tx = static_cast<int>(std::upper_bound(d.xcuts.begin(), d.xcuts.end(), ia) -
                                    d.xcuts.begin()) - 1;
    const int ty = static_cast<int>(std::upper_bound(d.ycuts.begin(), d.ycuts.end(), ja) -
                                    d.ycuts.begin()) - 1;
    if (tx < 0 || tx >= d.nx || ty < 0 || ty >= d.ny) return 0;
    const int id = ty * d.nx + tx + 1;
    // the reference only searches the first nprocp sub-domains
    return (id <= nranks && id <= d.ndomains) ? id : 0;
}

// ---------------------------------------------------------------------------
// map_comms -- src/parallel_comms_mod.f90:178-1172
//
// go_decompose only produces regular tensor-product tilings, so across every edge there
// is at most one neighbour whose extent along the edge equals ours and the diagonal
// owner never coincides with an edge owner: one message per edge, one depth x depth
// message per existing diagonal neighbour, in the list order the reference produces
// (W, E, S, N edges, then SW, NE, NW, SE corners).
// ---------------------------------------------------------------------------
namespace {
struct Msg {
    int dir, peer0, isrc, jsrc, ides, jdes, nx, ny;
};
int push_send(dlesm_comms &c, const Msg &m)
{
    if (c.nsend >= DLESM_MAXCOMM) return DLESM_ERR_MAXCOMM;
    const int k = c.nsend++;
    c.dirsend[k] = m.dir; c.destination[k] = m.peer0;
    c.isrcsend[k] = m.isrc; c.jsrcsend[k] = m.jsrc;
    c.idessend[k] = m.ides; c.jdessend[k] = m.jdes;
    c.nxsend[k] = m.nx; c.nysend[k] = m.ny; c.nsendp[k] = m.nx * m.ny;
    return DLESM_OK;
}
int push_recv(dlesm_comms &c, const Msg &m)
{
    if (c.nrecv >= DLESM_MAXCOMM) return DLESM_ERR_MAXCOMM;
    const int k = c.nrecv++;
    c.dirrecv[k] = m.dir; c.source[k] = m.peer0;
    c.isrcrecv[k] = m.isrc; c.jsrcrecv[k] = m.jsrc;
    c.idesrecv[k] = m.ides; c.jdesrecv[k] = m.jdes;
    c.nxrecv[k] = m.nx; c.nyrecv[k] = m.ny; c.nrecvp[k] = m.nx * m.ny;
    return DLESM_OK;
}
int wrap(int v, int n, bool periodic)
{
    if (!periodic) return v;
    if (v < 1) return v + n;
    if (v > n) return v - n;
    return v;
}
}  // namespace

int build_comms(const Decomposition &d, int irank1, int nranks, bool pbc_x, bool pbc_y,
                int ddx, int ddy, dlesm_comms &c)
{
    int *fields[] = {c.dirsend, c.destination, c.isrcsend, c.jsrcsend, c.idessend, c.jdessend,
                     c.nxsend, c.nysend, c.nsendp, c.dirrecv, c.source, c.isrcrecv, c.jsrcrecv,
                     c.idesrecv, c.jdesrecv, c.nxrecv, c.nyrecv, c.nrecvp};
    for (int *f : fields) std::fill(f, f + DLESM_MAXCOMM, -999);
    c.nsend = c.nrecv = 0;

    if (irank1 < 1 || irank1 > d.ndomains) {
        set_error(DLESM_ERR_ARGS, "map_comms: rank has no sub-domain");
        return DLESM_ERR_ARGS;
    }
    if (ddx < 1 || ddy < 1 || ddx > d.hwidth || ddy > d.hwidth) {
        set_error(DLESM_ERR_HALO_DEPTH, "map_comms: specified halo depth exceeds the halo width "
                                        "the decomposition was built with");
        return DLESM_ERR_HALO_DEPTH;
    }
    const dlesm_subdomain &me = d.subdomains[irank1 - 1];
    const dlesm_region &mi = me.internal;
    const int ielb = me.global.xstart, ieub = me.global.xstop;
    const int jelb = me.global.ystart, jeub = me.global.ystop;
    const int w = mi.nx, h = mi.ny;

    auto owner = [&](int ia, int ja) {
        return owner_of(d, nranks, wrap(ia, d.global_nx, pbc_x), wrap(ja, d.global_ny, pbc_y));
    };
    auto sub = [&](int id) -> const dlesm_region & { return d.subdomains[id - 1].internal; };
    int rc = DLESM_OK;
    auto add = [&](bool is_send, const Msg &m) {
        if (rc == DLESM_OK) rc = is_send ? push_send(c, m) : push_recv(c, m);
    };

    // West border: send towards -x (Iplus), receive what the neighbour sent towards +x (:285-471)
    if (int nb = owner(ielb - 1, jelb)) {
        const dlesm_region &o = sub(nb);
        add(true, {DLESM_IPLUS, nb - 1, mi.xstart, mi.ystart, o.xstop + 1, o.ystart, ddx, h});
        add(false, {DLESM_IMINUS, nb - 1, o.xstop - ddx + 1, o.ystart, mi.xstart - ddx, mi.ystart, ddx, h});
    }
    // East border (:473-636).  isrcrecv follows the reference's "internal%xstart + ihalo"
    // (:521); it is informational only -- the exchange never reads it.
    if (int nb = owner(ieub + 1, jelb)) {
        const dlesm_region &o = sub(nb);
        add(true, {DLESM_IMINUS, nb - 1, mi.xstop - ddx + 1, mi.ystart, o.xstart - ddx, o.ystart, ddx, h});
        add(false, {DLESM_IPLUS, nb - 1, o.xstart + ddx, o.ystart, mi.xstop + 1, mi.ystart, ddx, h});
    }
    // South border (:638-818)
    if (int nb = owner(ielb, jelb - 1)) {
        const dlesm_region &o = sub(nb);
        add(true, {DLESM_JPLUS, nb - 1, mi.xstart, mi.ystart, o.xstart, o.ystop + 1, w, ddy});
        add(false, {DLESM_JMINUS, nb - 1, o.xstart, o.ystop - ddy + 1, mi.xstart, mi.ystart - ddy, w, ddy});
    }
    // North border (:820-986)
    if (int nb = owner(ielb, jeub + 1)) {
        const dlesm_region &o = sub(nb);
        add(true, {DLESM_JMINUS, nb - 1, mi.xstart, mi.ystop - ddy + 1, o.xstart, o.ystart - ddy, w, ddy});
        add(false, {DLESM_JPLUS, nb - 1, o.xstart, o.ystart, mi.xstart, mi.ystop + 1, w, ddy});
    }
    // Diagonals (:988-1170): idirn 5..8 = SW, NE, NW, SE; the receive carries the opposite tag
    static const int west[9] = {0, 1, 0, 0, 0, 1, 0, 1, 0}, east[9] = {0, 0, 1, 0, 0, 0, 1, 0, 1};
    static const int south[9] = {0, 0, 0, 1, 0, 1, 0, 0, 1}, north[9] = {0, 0, 0, 0, 1, 0, 1, 1, 0};
    static const int opposite[9] = {0, 2, 1, 4, 3, 6, 5, 8, 7};
    const bool any_pbc = pbc_x || pbc_y;
    for (int dir = 5; dir <= 8; ++dir) {
        const int wx = west[dir], ex = east[dir], sy = south[dir], ny_ = north[dir];
        const int iout = wx * (ielb - 1) + ex * (ieub + 1);
        const int iin = wx * ielb + ex * ieub;
        const int jout = sy * (jelb - 1) + ny_ * (jeub + 1);
        const int jin = sy * jelb + ny_ * jeub;
        const int px = owner(iout, jin), py = owner(iin, jout), pd = owner(iout, jout);
        bool need = pd > 0 && px > 0 && py > 0;
        // the reference sends a diagonal message only when its owner takes part in neither
        // edge exchange (:1039-1040); with wrap-around neighbours (extension) the diagonal
        // owner can coincide with an edge owner and still gets its own message
        if (!any_pbc) need = need && pd != px && pd != py;
        if (!need) continue;
        const dlesm_region &o = sub(pd);
        add(true, {dir, pd - 1, ex * (mi.xstop - ddx + 1) + wx * mi.xstart,
                   ny_ * (mi.ystop - ddy + 1) + sy * mi.ystart,
                   wx * (o.xstop + 1) + ex * (o.xstart - ddx),
                   sy * (o.ystop + 1) + ny_ * (o.ystart - ddy), ddx, ddy});
        add(false, {opposite[dir], pd - 1, wx * (o.xstop - ddx + 1) + ex * o.xstart,
                    sy * (o.ystop - ddy + 1) + ny_ * o.ystart,
                    ex * (mi.xstop + 1) + wx * (mi.xstart - ddx),
                    ny_ * (mi.ystop + 1) + sy * (mi.ystart - ddy), ddx, ddy});
    }
    if (rc != DLESM_OK)
        set_error(rc, "map_comms: number of separate communications exceeds maximum of 16");
    return rc;
}

// ---------------------------------------------------------------------------
// several sub-domains per rank: one exchange plan over the tiles a rank owns
// ---------------------------------------------------------------------------
int build_tileset_plan(const Decomposition &d, int rank1, int nranks, bool pbc_x, bool pbc_y, int depth,
                       TilesetPlan &out)
{
    out = TilesetPlan{};
    if (d.nranks != nranks || d.domperrank < 1 || rank1 < 1 || rank1 > nranks) {
        set_error(DLESM_ERR_ARGS, "tile set: the decomposition was built for another number of ranks");
        return DLESM_ERR_ARGS;
    }
    // tile id -> (owner rank, index in the owner's list)
    std::vector<int> owner0(d.ndomains + 1, -1), local(d.ndomains + 1, -1);
    for (int r = 0; r < nranks; ++r) {
        int k = 0;
        for (int s = 0; s < d.domperrank; ++s) {
            const int id = d.proc_subdomains[static_cast<size_t>(r) * d.domperrank + s];
            if (id < 1) continue;
            owner0[id] = r;
            local[id] = k++;
            if (r == rank1 - 1) out.owned.push_back(id);
        }
    }
    if (out.owned.empty()) {
        set_error(DLESM_ERR_ARGS, "tile set: this rank owns no sub-domain");
        return DLESM_ERR_ARGS;
    }
    std::vector<TileSeg> copies, sends, recvs;
    for (size_t k = 0; k < out.owned.size(); ++k) {
        const int id = out.owned[k];
        dlesm_comms t;
        int rc = build_comms(d, id, d.ndomains, pbc_x, pbc_y, depth, depth, t);
        if (rc != DLESM_OK) return rc;
        const dlesm_region &in = d.subdomains[id - 1].internal;
        for (int s = 0; s < t.nsend; ++s) {
            if (t.nxsend[s] <= 0 || t.nysend[s] <= 0) continue;
            // a same-rank copy reads while other copies write: sends must read internal points only
            if (t.isrcsend[s] < in.xstart || t.isrcsend[s] + t.nxsend[s] - 1 > in.xstop ||
                t.jsrcsend[s] < in.ystart || t.jsrcsend[s] + t.nysend[s] - 1 > in.ystop) {
                set_error(DLESM_ERR_ARGS, "tile set: a send patch reaches outside the internal region");
                return DLESM_ERR_ARGS;
            }
            TileSeg g;
            const int dst = t.destination[s] + 1;
            g.peer0 = owner0[dst];
            g.src_tile = static_cast<int>(k);
            g.si0 = t.isrcsend[s]; g.sj0 = t.jsrcsend[s];
            g.di0 = t.idessend[s]; g.dj0 = t.jdessend[s];
            g.nx = t.nxsend[s]; g.ny = t.nysend[s];
            g.dst_id = dst; g.src_id = id; g.dir = t.dirsend[s];
            if (g.peer0 == rank1 - 1) {
                g.kind = 0;
                g.dst_tile = local[dst];
                copies.push_back(g);
            } else {
                g.kind = 1;
                sends.push_back(g);
            }
        }
        for (int r = 0; r < t.nrecv; ++r) {
            if (t.nxrecv[r] <= 0 || t.nyrecv[r] <= 0) continue;
            const int src = t.source[r] + 1;
            if (owner0[src] == rank1 - 1) continue;   // arrives as the sender's same-rank copy
            TileSeg g;
            g.kind = 2;
            g.peer0 = owner0[src];
            g.dst_tile = static_cast<int>(k);
            g.di0 = t.idesrecv[r]; g.dj0 = t.jdesrecv[r];
            g.nx = t.nxrecv[r]; g.ny = t.nyrecv[r];
            g.dst_id = id; g.src_id = src; g.dir = t.dirrecv[r];   // received with the sender's tag
            recvs.push_back(g);
        }
    }
    auto by_key = [](const TileSeg &a, const TileSeg &b) {
        if (a.peer0 != b.peer0) return a.peer0 < b.peer0;
        if (a.dst_id != b.dst_id) return a.dst_id < b.dst_id;
        if (a.src_id != b.src_id) return a.src_id < b.src_id;
        return a.dir < b.dir;
    };
    std::sort(sends.begin(), sends.end(), by_key);
    std::sort(recvs.begin(), recvs.end(), by_key);
    for (TileSeg &g : sends) { g.offset = out.send_total; out.send_total += static_cast<long long>(g.nx) * g.ny; }
    for (TileSeg &g : recvs) { g.offset = out.recv_total; out.recv_total += static_cast<long long>(g.nx) * g.ny; }
    out.segs = copies;
    out.segs.insert(out.segs.end(), sends.begin(), sends.end());
    out.segs.insert(out.segs.end(), recvs.begin(), recvs.end());
    return DLESM_OK;
}

// ---------------------------------------------------------------------------
// grid sizing -- src/grid_mod.f90:344-385
// ---------------------------------------------------------------------------
int parse_alignment(const char *env, int &alignment)
{
    alignment = 1;
    if (!env) return DLESM_OK;  // not present: no padding beyond the mandatory column
    const size_t len = std::strlen(env);
    if (len > 3) {
        set_error(DLESM_ERR_ALIGN_DIGITS, "Error: Only numbers of up to 3 digits are supported in "
                                          "the DL_ESM_ALIGNMENT environment variable.");
        return DLESM_ERR_ALIGN_DIGITS;
    }
    // Fortran list of an I3 field: optional sign, digits, blanks ignored
    long v = 0;
    int sign = 1;
    bool digits = false, bad = false;
    for (size_t k = 0; k < len && !bad; ++k) {
        const char ch = env[k];
        if (ch == ' ') continue;
        if ((ch == '+' || ch == '-') && !digits && v == 0) { if (ch == '-') sign = -1; continue; }
        if (ch < '0' || ch > '9') { bad = true; break; }
        v = v * 10 + (ch - '0');
        digits = true;
    }
    v *= sign;
    if (bad || v < 1) {
        set_error(DLESM_ERR_ALIGN_VALUE, std::string("Error: Cannot convert DL_ESM_ALIGNMENT value (") +
                                             env + ") into a positive integer.");
        return DLESM_ERR_ALIGN_VALUE;
    }
    alignment = static_cast<int>(v);
    return DLESM_OK;
}

int grid_dims(const dlesm_subdomain &sub, int alignment, GridDims &out)
{
    // always at least one extra column, up to the next multiple of ALIGNMENT (:364-369)
    out.padding = alignment - (sub.global.nx % alignment);
    out.nx = sub.global.nx + out.padding;
    out.ny = sub.global.ny + 1;  // exactly one extra row (:385)
    if (out.nx % alignment != 0) {
        set_error(DLESM_ERR_ALIGN_VALUE, "Error: Could not satisfy alignment requierements.");
        return DLESM_ERR_ALIGN_VALUE;
    }
    return DLESM_OK;
}

void fill_tmask(const dlesm_subdomain &sub, int nx, int ny, const int *tin, int tin_nx,
                std::vector<int> &tmask)
{
    // The box [xstart-1:xstop+1] x [ystart-1:ystop+1] is copied from the caller's mask (or
    // set wet); the N, S, low-x, high-x fills of :416-431 then leave every other cell equal
    // to the nearest box cell.  Without a caller mask the reference leaves those cells
    // uninitialised (:444-453); the same nearest-cell rule is used here (all wet).
    const int x0 = sub.internal.xstart - 1, x1 = sub.internal.xstop + 1;
    const int y0 = sub.internal.ystart - 1, y1 = sub.internal.ystop + 1;
    tmask.assign(static_cast<size_t>(nx) * ny, 1);
    if (!tin) return;
    for (int j = 1; j <= ny; ++j) {
        const int js = std::min(std::max(j, y0), y1);
        for (int i = 1; i <= nx; ++i) {
            const int is = std::min(std::max(i, x0), x1);
            tmask[static_cast<size_t>(j - 1) * nx + (i - 1)] =
                tin[static_cast<size_t>(js - 1) * tin_nx + (is - 1)];
        }
    }
}

void coord_vectors(const dlesm_subdomain &sub, int nx, int ny, double dx, double dy,
                   std::vector<double> &xt, std::vector<double> &yt)
{
    // running sums, not multiplications, so that every bit matches :536-556
    xt.assign(nx, 0.0);
    yt.assign(ny, 0.0);
    const int xs = sub.internal.xstart, ys = sub.internal.ystart;
    xt[xs - 1] = sub.global.xstart * dx;
    for (int i = xs + 1; i <= nx; ++i) xt[i - 1] = xt[i - 2] + dx;
    for (int i = xs - 1; i >= 1; --i) xt[i - 1] = xt[i] - dx;
    yt[ys - 1] = sub.global.ystart * dy;
    for (int j = ys + 1; j <= ny; ++j) yt[j - 1] = yt[j - 2] + dy;
    for (int j = ys - 1; j >= 1; --j) yt[j - 1] = yt[j] - dy;
}

// ---------------------------------------------------------------------------
// field bounds -- src/field_mod.f90:563-1122, 1394-1464
// ---------------------------------------------------------------------------
namespace {
dlesm_region make_region(int xs, int xe, int ys, int ye)
{
    dlesm_region r;
    r.xstart = xs; r.xstop = xe; r.ystart = ys; r.ystop = ye;
    r.nx = xe - xs + 1; r.ny = ye - ys + 1;
    return r;
}
// the halo list of init_periodic_bc_halos (:1394-1464): E, W columns over the internal rows,
// then N, S rows over xstart-1:xstop+1.  The reference never sets the regions' nx/ny.
void periodic_halos(const int bcs[3], FieldBounds &f)
{
    const dlesm_region &in = f.internal;
    int n = 0;
    auto put = [&](dlesm_region dst, dlesm_region src) {
        dst.nx = dst.ny = src.nx = src.ny = 0;
        f.halo[n].needs_update = 0;
        f.halo[n].dest = dst;
        f.halo[n].source = src;
        ++n;
    };
    if (bcs[0] == DLESM_GO_BC_PERIODIC) {
        put(make_region(in.xstop + 1, in.xstop + 1, in.ystart, in.ystop),
            make_region(in.xstart, in.xstart, in.ystart, in.ystop));
        put(make_region(in.xstart - 1, in.xstart - 1, in.ystart, in.ystop),
            make_region(in.xstop, in.xstop, in.ystart, in.ystop));
    }
    if (bcs[1] == DLESM_GO_BC_PERIODIC) {
        put(make_region(in.xstart - 1, in.xstop + 1, in.ystop + 1, in.ystop + 1),
            make_region(in.xstart - 1, in.xstop + 1, in.ystart, in.ystart));
        put(make_region(in.xstart - 1, in.xstop + 1, in.ystart - 1, in.ystart - 1),
            make_region(in.xstart - 1, in.xstop + 1, in.ystop, in.ystop));
    }
    f.num_halos = n;
}
int field_stop(const char *msg)
{
    set_error(DLESM_ERR_FIELD_STOP, msg);
    return DLESM_ERR_FIELD_STOP;
}
}  // namespace

int field_bounds(const dlesm_subdomain &sub, int grid_nx, int grid_ny, int offset,
                 const int bcs[3], int points, FieldBounds &f)
{
    f = FieldBounds();
    const dlesm_region &si = sub.internal;
    const bool px = bcs[0] == DLESM_GO_BC_PERIODIC, py = bcs[1] == DLESM_GO_BC_PERIODIC;
    int xs = si.xstart, xe = si.xstop, ys = si.ystart, ye = si.ystop;

    if (points == DLESM_GO_ALL_POINTS) {  // field_init :628-648
        f.defined_on = points;
        xs = 1; xe = grid_nx; ys = 1; ye = grid_ny;
        f.num_halos = 0;
    } else if (points >= DLESM_GO_U_POINTS && points <= DLESM_GO_F_POINTS) {
        f.defined_on = points;
        if (offset == DLESM_GO_OFFSET_NE) {
            // c{u,v,t,f}_ne_init (:755-786, 872-895, 965-1000, 1088-1122): every stagger
            // shares the T-point region; periodic BCs are not implemented
            if (px || py) return field_stop("ERROR: c?_ne_init: implement periodic boundary conditions!");
            if (points == DLESM_GO_U_POINTS) f.num_halos = 0;  // only cu_ne_init sets it (:784)
        } else if (offset == DLESM_GO_OFFSET_SW) {
            switch (points) {
            case DLESM_GO_U_POINTS:  // cu_sw_init :675-751
                if (!px) xs = si.xstart + 1;
                break;
            case DLESM_GO_V_POINTS:  // cv_sw_init :813-868
                if (!py) return field_stop("cv_sw_init: IMPLEMENT non-periodic BCs!");
                break;
            case DLESM_GO_T_POINTS:  // ct_sw_init :922-961
                break;
            case DLESM_GO_F_POINTS:  // cf_sw_init :1027-1084
                if (!px || !py) return field_stop("cf_sw_init: CHECK non-periodic BCs!");
                break;
            }
            f.internal = make_region(xs, xe, ys, ye);
            periodic_halos(bcs, f);
        } else {
            return field_stop("c?_field_init: ERROR - unsupported grid offset!");
        }
    } else {
        set_error(DLESM_ERR_POINTS, "r2d_field_constructor: ERROR: invalid specifier for type of mesh points");
        return DLESM_ERR_POINTS;
    }
    f.internal = make_region(xs, xe, ys, ye);
    // whole = internal +- NBOUNDARY (=1) in both dimensions whatever the BCs (:606-622)
    f.whole = make_region(xs - 1, xe + 1, ys - 1, ye + 1);
    return DLESM_OK;
}

}  // namespace dlesm
