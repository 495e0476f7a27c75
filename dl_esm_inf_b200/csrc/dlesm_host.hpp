// dlesm_host.hpp -- host-side bookkeeping of the drop-in: regions, decomposition, grid
// sizing, field bounds and communication tables.  Pure C++17, no CUDA: usable on a
// CPU-only box.  The arithmetic contract is the reference's Fortran (cited per function);
// the implementation is this project's own.
#pragma once

#include <string>
#include <vector>

#include "../../include/dlesm_b200.h"

namespace dlesm {

// ---- process-wide parallel state (parallel_utils_mod's module variables) ----
struct ParallelState {
    int rank = 1;    // 1-based
    int nranks = 1;
    int extensions = 0;
};
ParallelState &pstate();

void set_error(int code, const std::string &msg);
int last_status();
const char *last_error_cstr();
void clear_error();

// ---- decomposition (src/parallel_mod.f90:70-332) ----
struct Decomposition {
    int global_nx = 0, global_ny = 0;
    int nx = 0, ny = 0;          // tile grid
    int ndomains = 0;
    int max_width = 0, max_height = 0;
    int domperrank = 0, nranks = 0;
    int hwidth = 1;              // halo width the subdomains were built with
    std::vector<dlesm_subdomain> subdomains;   // 0-based storage of 1-based ids
    std::vector<int> proc_subdomains;          // (domperrank, nranks) column-major
    // x / y cut positions of the regular tile grid (global 1-based start of each tile
    // column / row, plus one-past-the-end), for O(log n) owner lookup
    std::vector<int> xcuts, ycuts;
};

int go_decompose(int domainx, int domainy, int ndomains, int ndomainx, int ndomainy,
                 int halo_width, int nranks, Decomposition &out);
void reset_saved_state();

// owner (1-based) of global point (ia, ja), 0 if none among the first `nranks` domains
int owner_of(const Decomposition &d, int nranks, int ia, int ja);

// ---- map_comms (src/parallel_comms_mod.f90:178-1172) ----
int build_comms(const Decomposition &d, int irank1, int nranks, bool pbc_x, bool pbc_y,
                int depth_x, int depth_y, dlesm_comms &out);

// ---- several sub-domains per rank (extension F3; the reference carries proc_subdomains,
// src/decomposition_mod.f90:64-67, and marks map_comms "\TODO support multiple subdomains per
// rank", src/parallel_comms_mod.f90:295) ----
// One halo exchange over every tile this rank owns.  Tables are map_comms' own, built per tile with
// the tile id in the place of the rank; a message whose two ends live on this rank becomes a device
// copy, the others are packed per peer RANK in the order (destination tile, source tile, direction)
// that both ends can compute.
struct TileSeg {
    int kind = 0;                   // 0 same-rank copy, 1 send (pack), 2 receive (unpack)
    int peer0 = 0;                  // 0-based rank at the other end (this rank for kind 0)
    int src_tile = -1, dst_tile = -1;   // 0-based indices into this rank's tile list (-1: the buffer)
    int si0 = 0, sj0 = 0, di0 = 0, dj0 = 0, nx = 0, ny = 0;   // 1-based corners in the tiles' local arrays
    long long offset = 0;           // element offset of the patch in the send / receive buffer, one field
    int dst_id = 0, src_id = 0, dir = 0;   // the ordering key (1-based tile ids, direction tag)
};
struct TilesetPlan {
    std::vector<int> owned;         // 1-based ids of this rank's tiles, proc_subdomains(:, rank)
    std::vector<TileSeg> segs;      // copies, then sends grouped by peer, then receives grouped by peer
    long long send_total = 0, recv_total = 0;
};
int build_tileset_plan(const Decomposition &d, int rank1, int nranks, bool pbc_x, bool pbc_y, int depth,
                       TilesetPlan &out);

// ---- grid sizing (src/grid_mod.f90:344-385) ----
int parse_alignment(const char *envval, int &alignment);
struct GridDims {
    int nx = 0, ny = 0, padding = 0;
};
int grid_dims(const dlesm_subdomain &sub, int alignment, GridDims &out);
// tmask fill rules of src/grid_mod.f90:400-453 into a (nx,ny) column-major int array
void fill_tmask(const dlesm_subdomain &sub, int nx, int ny, const int *tmask_in, int tmask_in_nx,
                std::vector<int> &tmask);
// 1-D coordinate recurrences of src/grid_mod.f90:536-556
void coord_vectors(const dlesm_subdomain &sub, int nx, int ny, double dx, double dy,
                   std::vector<double> &xt, std::vector<double> &yt);

// ---- field bounds (src/field_mod.f90:563-1122, 1394-1464) ----
struct FieldBounds {
    int defined_on = 0;
    dlesm_region internal{};
    dlesm_region whole{};
    int num_halos = -1;
    dlesm_halo halo[4]{};
};
int field_bounds(const dlesm_subdomain &sub, int grid_nx, int grid_ny, int grid_offset,
                 const int bcs[3], int grid_points, FieldBounds &out);

}  // namespace dlesm
