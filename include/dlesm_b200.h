/*
 * dlesm_b200.h -- C ABI of libdlesm_b200.so, the B200-native drop-in for the hot path of
 * stfc/dl_esm_inf's fd_lib / dm_fd_lib: finite-difference stencil sweeps over r2d_field
 * data on an Arakawa C-grid grid_type, and the halo exchange between them.
 *
 * Two layers, both plain C (pointers, ints, doubles; no C++/torch types):
 *
 *   1. DEVICE LAYER (dlesm_dev_*, dlesm_read_from_device_c, ...): what a Fortran build of
 *      the reference binds through ISO_C_BINDING.  The Fortran side keeps its own
 *      grid_type / r2d_field bookkeeping; these entry points own device storage, copies,
 *      reductions, the exchange and the stencils.  INTEGRATION.md shows the bind(C) stubs.
 *
 *   2. HOST MIRROR (dlesm_grid_*, dlesm_field_*, dlesm_sw_*): the same public API as the
 *      reference's grid_mod / field_mod / parallel_mod, implemented in C++ above layer 1,
 *      for hosts without a Fortran toolchain (this image).  Same names, argument meaning,
 *      1-based inclusive regions and error behaviour as the Fortran.
 *
 * Every int-returning function returns DLESM_OK (0) or a negative DLESM_ERR_* code and
 * records a message for dlesm_last_error(); a Fortran shim turns non-zero into
 * gocean_stop(msg) (reference: src/gocean_mod.F90:50-57 has no status returns).
 * There is NO CPU fallback: device entry points fail with DLESM_ERR_CUDA when no GPU
 * is usable.
 *
 * Citations "src/X:a-b" are to /root/reference/finite_difference/src/X.
 */
#ifndef DLESM_B200_H
#define DLESM_B200_H

#include <stdbool.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------ */
/* constants (values identical to the reference's Fortran parameters)                   */
/* ------------------------------------------------------------------------------------ */
/* src/grid_mod.f90:45-69 */
#define DLESM_GO_ARAKAWA_C 0
#define DLESM_GO_ARAKAWA_B 1
#define DLESM_GO_OFFSET_SW 0
#define DLESM_GO_OFFSET_SE 1
#define DLESM_GO_OFFSET_NW 2
#define DLESM_GO_OFFSET_NE 3
#define DLESM_GO_OFFSET_ANY 4
#define DLESM_GO_BC_PERIODIC 0
#define DLESM_GO_BC_EXTERNAL 1
#define DLESM_GO_BC_NONE 2
/* src/field_mod.f90:47-52 */
#define DLESM_GO_U_POINTS 0
#define DLESM_GO_V_POINTS 1
#define DLESM_GO_T_POINTS 2
#define DLESM_GO_F_POINTS 3
#define DLESM_GO_ALL_POINTS 4
/* src/parallel_comms_mod.f90:70,101-110 */
#define DLESM_MAXCOMM 16
#define DLESM_IPLUS 1
#define DLESM_IMINUS 2
#define DLESM_JPLUS 3
#define DLESM_JMINUS 4
#define DLESM_IPLUSJPLUS 5
#define DLESM_IMINUSJMINUS 6
#define DLESM_IPLUSJMINUS 7
#define DLESM_IMINUSJPLUS 8

#define DLESM_OK 0
#define DLESM_ERR_ARGS (-1)         /* go_decompose: invalid arguments supplied        */
#define DLESM_ERR_NRANKS (-2)
#define DLESM_ERR_HALO_WIDTH (-3)
#define DLESM_ERR_ALIGN_DIGITS (-4) /* DL_ESM_ALIGNMENT longer than 3 characters       */
#define DLESM_ERR_ALIGN_VALUE (-5)  /* DL_ESM_ALIGNMENT not a positive integer         */
#define DLESM_ERR_PBC_DISTMEM (-6)  /* periodic BCs with >1 rank and extension not on  */
#define DLESM_ERR_GRID_NAME (-7)
#define DLESM_ERR_GRID_OFFSET (-8)
#define DLESM_ERR_FIELD_STOP (-9)   /* a c?_{sw,ne}_init gocean_stop branch            */
#define DLESM_ERR_POINTS (-10)
#define DLESM_ERR_MAXCOMM (-12)     /* map_comms list overflow (reference ierr = -12)  */
#define DLESM_ERR_HALO_DEPTH (-13)
#define DLESM_ERR_CUDA (-20)        /* no device / CUDA failure: fail loudly           */
#define DLESM_ERR_NCCL (-21)
#define DLESM_ERR_STATE (-22)       /* call order (e.g. field before grid_init)        */
#define DLESM_ERR_UNKNOWN_PTR (-23) /* device pointer not allocated by this library    */

/* last error message of the calling thread ("" if none) */
const char *dlesm_last_error(void);
/* library version string, e.g. "dlesm_b200 0.2 (sm_100a, ...) src <hash>" */
const char *dlesm_version(void);
/* hash of every source, header and compiler flag this library was built from (32 hex digits);
 * dl_esm_inf_b200/build.py compares it with the sources on disk and rebuilds on mismatch */
const char *dlesm_source_hash(void);

/* ------------------------------------------------------------------------------------ */
/* index boxes -- src/region_mod.f90:7-12, src/decomposition_mod.f90:44-50,              */
/*                src/halo_mod.f90:9-25.  1-based, inclusive, exactly as in Fortran.     */
/* ------------------------------------------------------------------------------------ */
typedef struct {
    int nx, ny;
    int xstart, xstop;
    int ystart, ystop;
} dlesm_region;

typedef struct {
    dlesm_region global;   /* where the internal part sits in the global domain; nx,ny =
                              full extent incl. halos (src/parallel_mod.f90:272-290) */
    dlesm_region internal; /* internal region in local indices */
} dlesm_subdomain;

typedef struct {
    int needs_update;
    dlesm_region source;
    dlesm_region dest;
} dlesm_halo;

/* ------------------------------------------------------------------------------------ */
/* lifecycle -- replaces parallel_utils_mod / parallel_utils_stub_mod                    */
/* (src/parallel/parallel_utils_mod.f90:77-125, gocean_initialise src/gocean_mod.F90:22) */
/* ------------------------------------------------------------------------------------ */
/* Set this process's 1-based rank and the number of ranks.  Host only: touches neither
 * CUDA nor NCCL, so decomposition / bookkeeping entry points work on a CPU-only box.
 * A process that never calls it is rank 1 of 1 (the serial stub's behaviour). */
int dlesm_parallel_init(int rank1, int nranks);
/* Bind the process to a GPU (cudaSetDevice, streams, scratch).  device < 0 => LOCAL_RANK
 * env or 0.  Called implicitly (device 0 / LOCAL_RANK) by the first device operation. */
int dlesm_device_init(int device);
/* Multi-rank bring-up without MPI: rank 1 creates a 128-byte id, the launcher ships it to
 * all ranks (torch.distributed, a file, ...), every rank calls dlesm_comm_init. */
int dlesm_comm_unique_id(void *id128);
int dlesm_comm_init(const void *id128);
/* parallel_finalise (parallel_utils_mod.f90:94-98): destroys communicator, frees pools */
int dlesm_parallel_finalise(void);
/* parallel_abort (:104-111): message to stderr, abort communicator, exit(1) */
void dlesm_parallel_abort(const char *msg);
int dlesm_get_rank(void);      /* 1-based, parallel_utils_mod.f90:115-118 */
int dlesm_get_num_ranks(void); /* :122-125 */
int dlesm_on_master(void);     /* src/parallel_mod.f90:337-340 */
/* global_sum (parallel_utils_mod.f90:230-238): in-place sum over ranks of one double
 * (no-op for one rank, as the stub :148-150) */
int dlesm_global_sum(double *var);
/* gather (parallel_utils_mod.f90:242-255, stub :154-161): every rank contributes n doubles
 * (host memory); rank 1's recv_buffer (n * nranks doubles) receives them in rank order. */
int dlesm_gather_host(const double *send_buffer, long long n, double *recv_buffer);
/* go_decompose keeps halo_width between calls (SAVEd local, src/parallel_mod.f90:103-105);
 * this resets it to 1 as at program start. */
void dlesm_reset_saved_state(void);
/* Extensions the reference marks TODO.  bit 0: allow periodic BCs with
 * more than one rank (wrap-around neighbours); default 0 = refuse as the reference. */
int dlesm_set_extensions(int flags);
#define DLESM_EXT_PERIODIC_DISTMEM 1
/* bit 1: honour halo_exchange(depth) for depth > 1, up to the halo_width the grid was
 * decomposed with (the reference hard-wires depth 1: MAX_HALO_DEPTH, src/parallel_comms_mod.f90:48,
 * src/field_mod.f90:1226-1229). */
#define DLESM_EXT_HALO_DEPTH 2
/* bit 2: dirty-flag halos.  halo_type%needs_update exists in the reference but its set_dirty /
 * is_dirty methods are commented out (src/halo_mod.f90:10-40) and every halo_exchange runs.  With
 * this bit the library tracks, per device array, whether its halos are up to date (every entry point
 * that writes an array marks it dirty; halo_exchange, the periodic fill and a pushing sweep mark it
 * clean) and halo_exchange / apply_periodic_halos skip fields whose halos are clean: no launch.
 * An exchange that involves another rank is never shortened: the skip would be one rank's decision
 * about a message both sides must agree on, so it applies to self-neighbour (one-rank periodic)
 * exchanges and periodic fills only. */
#define DLESM_EXT_DIRTY_HALOS 4

/* ------------------------------------------------------------------------------------ */
/* HOST MIRROR: decomposition -- src/parallel_mod.f90:70-332, decomposition_mod.f90      */
/* ------------------------------------------------------------------------------------ */
typedef struct dlesm_decomp dlesm_decomp_t;

/* Optional Fortran arguments: pass a negative value for "not present".  nranks comes from
 * dlesm_parallel_init.  Returns NULL on error (see dlesm_last_error / dlesm_last_status). */
dlesm_decomp_t *dlesm_go_decompose(int domainx, int domainy, int ndomains, int ndomainx,
                                   int ndomainy, int halo_width);
int dlesm_last_status(void);
/* info[8] = global_nx, global_ny, nx, ny, ndomains, max_width, max_height, domperrank */
int dlesm_decomp_info(const dlesm_decomp_t *d, int info[8]);
int dlesm_decomp_subdomain(const dlesm_decomp_t *d, int idomain1, dlesm_subdomain *out);
/* proc_subdomains(islot1, irank1), -1 where empty */
int dlesm_decomp_proc_subdomain(const dlesm_decomp_t *d, int islot1, int irank1);
void dlesm_decomp_free(dlesm_decomp_t *d);

/* map_comms tables -- module variables of src/parallel_comms_mod.f90:70-83 */
typedef struct {
    int nsend, nrecv;
    int dirsend[DLESM_MAXCOMM], destination[DLESM_MAXCOMM];
    int isrcsend[DLESM_MAXCOMM], jsrcsend[DLESM_MAXCOMM];
    int idessend[DLESM_MAXCOMM], jdessend[DLESM_MAXCOMM];
    int nxsend[DLESM_MAXCOMM], nysend[DLESM_MAXCOMM], nsendp[DLESM_MAXCOMM];
    int dirrecv[DLESM_MAXCOMM], source[DLESM_MAXCOMM];
    int isrcrecv[DLESM_MAXCOMM], jsrcrecv[DLESM_MAXCOMM];
    int idesrecv[DLESM_MAXCOMM], jdesrecv[DLESM_MAXCOMM];
    int nxrecv[DLESM_MAXCOMM], nyrecv[DLESM_MAXCOMM], nrecvp[DLESM_MAXCOMM];
} dlesm_comms;

/* map_comms (src/parallel_comms_mod.f90:178-1172) for rank `irank1` of `nranks`.
 * pbc_x / pbc_y: wrap-around neighbours (extension; the reference aborts if pbc).
 * Unused slots hold -999 like the reference's initialisation (:246-261). */
int dlesm_map_comms(const dlesm_decomp_t *d, int irank1, int nranks, int pbc_x, int pbc_y,
                    int halo_depth_x, int halo_depth_y, dlesm_comms *out);
/* iprocmap (:1365-1398): owner (1-based) of global point (ia, ja) or 0 */
int dlesm_iprocmap(const dlesm_decomp_t *d, int nranks, int ia, int ja);

/* The message plan the device exchange derives from a rank's map_comms tables.  The reference
 * posts one MPI message per list entry and matches them by tag = direction
 * (src/parallel_comms_mod.f90:1606,1647); NCCL and NVLink peer copies have no tags, so all
 * patches for one peer travel as ONE buffer whose segments are ordered by direction on both
 * sides.  Host-only (no CUDA): tests/dist_host_worker.py drives a gloo exchange from it. */
typedef struct {
    int npeers;
    int peer[DLESM_MAXCOMM];                    /* 0-based rank, ascending; may be this rank */
    int nsend[DLESM_MAXCOMM], nrecv[DLESM_MAXCOMM];             /* segments per peer         */
    int send_slot[DLESM_MAXCOMM][DLESM_MAXCOMM];  /* 0-based indices into the send lists ... */
    int recv_slot[DLESM_MAXCOMM][DLESM_MAXCOMM];  /* ... and recv lists, in buffer order      */
    long long send_off[DLESM_MAXCOMM], recv_off[DLESM_MAXCOMM];  /* element offsets, 1 field */
    long long send_n[DLESM_MAXCOMM], recv_n[DLESM_MAXCOMM];      /* elements per field        */
    long long send_total, recv_total;
} dlesm_comm_plan;
int dlesm_comm_plan_build(const dlesm_comms *tables, int rank1, dlesm_comm_plan *out);

/* ------------------------------------------------------------------------------------ */
/* HOST MIRROR: grid_type -- src/grid_mod.f90:75-157                                     */
/* ------------------------------------------------------------------------------------ */
typedef struct dlesm_grid dlesm_grid_t;

typedef struct {
    int name, offset;
    int global_nx, global_ny;
    int nx, ny;                 /* padded local allocation (src/grid_mod.f90:368-385) */
    double dx, dy;
    int boundary_conditions[3];
    dlesm_subdomain subdomain;
    int alignment, padding;     /* DL_ESM_ALIGNMENT value used and columns added */
    int initialised;            /* grid_init done */
} dlesm_grid_info;

/* grid_constructor (src/grid_mod.f90:217-304).  grid_offsets < 0 means "not present",
 * which is an error exactly as in the reference (:241-246). */
dlesm_grid_t *dlesm_grid_create(int grid_name, const int boundary_conditions[3], int grid_offsets);
/* grid%decompose (src/grid_mod.f90:183-211) */
int dlesm_grid_decompose(dlesm_grid_t *g, int domainx, int domainy, int ndomains, int ndomainx,
                         int ndomainy, int halo_width);
/* grid_init (src/grid_mod.f90:330-570).  tmask may be NULL (all-wet); otherwise column-major
 * tmask(tmask_nx, tmask_ny), indexed with the LOCAL indices of the subdomain like the
 * Fortran dummy.  Reads DL_ESM_ALIGNMENT from the environment like the reference. */
int dlesm_grid_init(dlesm_grid_t *g, double dxarg, double dyarg, const int *tmask, int tmask_nx,
                    int tmask_ny);
int dlesm_grid_get_info(const dlesm_grid_t *g, dlesm_grid_info *out);
const dlesm_decomp_t *dlesm_grid_decomp(const dlesm_grid_t *g);
int dlesm_grid_comms(const dlesm_grid_t *g, dlesm_comms *out);
/* Grid arrays.  `which` is one of DLESM_GRID_*.  Host copies are written into caller
 * memory (column-major nx*ny); device copies are materialised on first request and cached
 * (the reference leaves the 17 *_device c_ptr slots NULL for the user to fill,
 * src/grid_mod.f90:286-302). */
#define DLESM_GRID_TMASK 0
#define DLESM_GRID_DX_T 1
#define DLESM_GRID_DY_T 2
#define DLESM_GRID_DX_U 3
#define DLESM_GRID_DY_U 4
#define DLESM_GRID_DX_V 5
#define DLESM_GRID_DY_V 6
#define DLESM_GRID_DX_F 7
#define DLESM_GRID_DY_F 8
#define DLESM_GRID_AREA_T 9
#define DLESM_GRID_AREA_U 10
#define DLESM_GRID_AREA_V 11
#define DLESM_GRID_GPHIU 12
#define DLESM_GRID_GPHIV 13
#define DLESM_GRID_GPHIF 14
#define DLESM_GRID_XT 15
#define DLESM_GRID_YT 16
int dlesm_grid_host_array(dlesm_grid_t *g, int which, void *out /* int32 or double nx*ny */);
int dlesm_grid_device_array(dlesm_grid_t *g, int which, void **device_ptr);
void dlesm_grid_free(dlesm_grid_t *g);

/* ------------------------------------------------------------------------------------ */
/* HOST MIRROR: r2d_field -- src/field_mod.f90:110-166                                   */
/* ------------------------------------------------------------------------------------ */
typedef struct dlesm_field dlesm_field_t;

typedef struct {
    int defined_on;
    dlesm_region internal;
    dlesm_region whole;
    int num_halos;        /* -1 where the reference leaves it unset (cv/ct/cf_ne_init) */
    dlesm_halo halo[4];
    int nx, ny;           /* allocation = grid%nx, grid%ny (src/field_mod.f90:327-333) */
    int ntiles;
    int data_on_device;   /* always 1: storage is device-resident */
    void *device_ptr;
} dlesm_field_info;

/* r2d_field(grid, grid_points [, do_tile] [, init_global_data]) -- src/field_mod.f90:242-390.
 * Device storage is allocated and zero-filled; init_global_data (column-major
 * (gnx, gny), may be NULL) is scattered as :378-389. */
dlesm_field_t *dlesm_r2d_field_create(dlesm_grid_t *g, int grid_points, int do_tile,
                                      const double *init_global_data, int gnx, int gny);
int dlesm_field_get_info(const dlesm_field_t *f, dlesm_field_info *out);
/* tile(itile)%internal / %whole (src/field_mod.f90:316-319, src/tile_mod.f90:36-42), 1 <= itile <= ntiles.
 * r2d_field(..., do_tile=.true.) cuts the field's internal region into the tile grid the sweeps use:
 * GOCEAN_OMP_GRID="AxB" (src/field_mod.f90:1473-1503) = A x-tiles (each cut into strips of the sweep's
 * width) by B tiles (segments) along j, go_decompose's remainder rule; without it the library's choice. */
int dlesm_field_tile(const dlesm_field_t *f, int itile1, dlesm_region *internal, dlesm_region *whole);
/* halo_type%needs_update as a query / set_dirty (src/halo_mod.f90:18-40): is the halo out of date to
 * `depth` cells?  (tracked always; acted upon under DLESM_EXT_DIRTY_HALOS) */
int dlesm_field_halo_is_dirty(const dlesm_field_t *f, int depth);
int dlesm_field_set_halo_dirty(dlesm_field_t *f);
/* fld%grid (src/field_mod.f90:114) */
dlesm_grid_t *dlesm_field_grid(const dlesm_field_t *f);
/* set_field_bounds only (src/field_mod.f90:563-624): the regions and halo list a field of
 * this point type would get; no storage is allocated, so it works without a GPU
 * (device_ptr = NULL, data_on_device = 0 in the result). */
int dlesm_grid_field_bounds(const dlesm_grid_t *g, int grid_points, dlesm_field_info *out);
/* set_data / get_data (:530-559): whole-array host <-> device */
int dlesm_field_set_data(dlesm_field_t *f, const double *host_array);
int dlesm_field_get_data(dlesm_field_t *f, double *host_array);
/* write_to_device / read_from_device (:407-525): 1-based sub-rectangle of the padded array;
 * `host_array` is the full (nx, ny) host array.  Pass 1,1,nx,ny for the defaults. */
int dlesm_field_write_to_device(dlesm_field_t *f, const double *host_array, int startx, int starty,
                                int nx, int ny, bool blocking);
int dlesm_field_read_from_device(dlesm_field_t *f, double *host_array, int startx, int starty,
                                 int nx, int ny, bool blocking);
/* fld%halo_exchange(depth) (:1231-1256): device-resident, all 8 directions */
int dlesm_field_halo_exchange(dlesm_field_t *f, int depth);
/* several fields in one grouped exchange (one message per neighbour) */
int dlesm_fields_halo_exchange(dlesm_field_t *const *f, int nfields, int depth);
/* serial periodic BCs: copy_field(fld, halo%source, halo%dest) for ihalo = 1..num_halos */
int dlesm_field_apply_periodic_halos(dlesm_field_t *f);
/* gather_inner_data (:1313-1390): global_data (global_nx, global_ny) filled on the master */
int dlesm_field_gather_inner_data(dlesm_field_t *f, double *global_data);
/* copy_field overloads (:1126-1187) */
int dlesm_copy_field(const dlesm_field_t *field_in, dlesm_field_t *field_out);
int dlesm_copy_field_patch(dlesm_field_t *field, const dlesm_region *src, const dlesm_region *dest);
/* set_field (:1191-1202) */
int dlesm_set_field(dlesm_field_t *fld, double val);
/* field_checksum (:1209-1219): SUM(ABS(internal region)) then global_sum */
int dlesm_field_checksum(dlesm_field_t *field, double *val);
/* free_field (:395-403); also releases the device storage */
int dlesm_free_field(dlesm_field_t *fld);

/* ------------------------------------------------------------------------------------ */
/* DEVICE LAYER: storage and the data_on_device hooks                                    */
/* ------------------------------------------------------------------------------------ */
/* Allocate a zero-filled (nx, ny) FP64 device array; registers its pitch so that the
 * hook signature below (which carries no pitch) can address sub-rectangles. */
int dlesm_dev_field_alloc(int nx, int ny, void **device_ptr);
int dlesm_dev_field_free(void *device_ptr);
/* EXACTLY the reference's C-flavoured hook signature (src/field_mod.f90:66-72, 87-93):
 * read:  from = device_ptr, to = C_LOC(data);  write: from = C_LOC(data), to = device_ptr.
 * (startx, starty, nx, ny) is a 1-based sub-rectangle of the padded array; element (i,j)
 * lives at (j-1)*size(data,1)+(i-1).  Non-blocking copies are ordered on one stream; a
 * blocking call synchronises it (the reference only blocks on the last of four). */
void dlesm_read_from_device_c(void *from, void *to, int startx, int starty, int nx, int ny,
                              bool blocking);
void dlesm_write_to_device_c(void *from, void *to, int startx, int starty, int nx, int ny,
                             bool blocking);
int dlesm_dev_sync(void);
/* Page-locked host memory for the host mirror of fld%data (fast, truly asynchronous hook
 * copies); or pin an existing allocation (a Fortran allocatable) in place. */
int dlesm_host_alloc(size_t bytes, void **host_ptr);
int dlesm_host_free(void *host_ptr);
int dlesm_host_register(void *host_ptr, size_t bytes);
int dlesm_host_unregister(void *host_ptr);
/* CUDA-event stopwatch on the library's stream (slots 0..15): record, then read the
 * elapsed milliseconds between two recorded slots (synchronises on the later one). */
int dlesm_event_record(int slot);
int dlesm_event_elapsed_ms(int slot_start, int slot_stop, double *ms);

/* ------------------------------------------------------------------------------------ */
/* DEVICE LAYER: field operations on device pointers from dlesm_dev_field_alloc          */
/* ------------------------------------------------------------------------------------ */
int dlesm_dev_copy_field(const void *field_in, void *field_out);                  /* :1152 */
int dlesm_dev_copy_patch(void *field, int sxstart, int sxstop, int systart, int systop,
                         int dxstart, int dxstop, int dystart, int dystop);       /* :1179 */
int dlesm_dev_set_field(void *field, double val);                                 /* :1191 */
/* dlesm_dev_copy_patch for callers that cannot allocate device memory.  No counterpart in the
 * reference (it has no device code); same result as dlesm_dev_copy_patch, i.e. :1179-1187 with the
 * right-hand side read first.  Meant for hosts where a cudaMalloc inside the library is not an
 * option: the application's own pool already owns all of the GPU's memory (the call would fail), the
 * time-step loop must not take the device-wide synchronisation an allocation implies, or the memory
 * budget is fixed up front.  An overlapping copy is therefore staged through an area that is part of
 * the library image -- DLESM_STATIC_SCRATCH_DOUBLES doubles (2 MiB) on every GPU that loads the
 * library, held until the process exits -- and the entry point calls no cudaMalloc / cudaFree and
 * does not touch the library's context.
 *   field        device memory of the caller, nx x ny doubles, column-major (need not come from
 *                dlesm_dev_field_alloc); boxes are 1-based inclusive, inside the allocation
 *   stream       a cudaStream_t of the CURRENT device (NULL = default stream); blocking != 0 waits
 *                for it before returning
 * Limit: overlapping patches of more than DLESM_STATIC_SCRATCH_DOUBLES cells return
 * DLESM_ERR_ARGS (split the copy); disjoint patches are copied directly and have no limit.
 * Sharing: there is ONE staging area per GPU.  It keeps the previous call's contents, which is
 * harmless (every cell read is written first by the same call; nothing needs to start at zero), but
 * calls that can run at the same time on one GPU -- other streams, other threads -- race on it.
 * Set-up: the first call on each device looks up the area's device address; make that call during
 * initialisation, before any stream capture on that device. */
#define DLESM_STATIC_SCRATCH_DOUBLES 262144
int dlesm_dev_copy_patch_static(void *field, int nx, int ny, int sxstart, int sxstop, int systart,
                                int systop, int dxstart, int dxstop, int dystart, int dystop,
                                void *stream, int blocking);
/* local SUM(ABS(field(xstart:xstop, ystart:ystop))) -- :1289-1302; deterministic two-pass
 * tree (no atomics).  Follow with dlesm_global_sum for :1305. */
int dlesm_dev_checksum(const void *field, int xstart, int xstop, int ystart, int ystop,
                       double *val);
/* periodic halo fill for a field whose internal region is the given box: E, W columns then
 * N, S rows over xstart-1:xstop+1 (src/field_mod.f90:1394-1464) */
int dlesm_dev_periodic_halos(void *field, int xstart, int xstop, int ystart, int ystop,
                             int periodic_x, int periodic_y);
/* Install the comm tables computed by the (unchanged Fortran or mirrored) map_comms for this
 * rank; later dlesm_dev_halo_exchange calls use them (module-level state, like parallel_comms_mod's).
 * NULL removes them (the state before any grid_init: halo_exchange is a no-op). */
int dlesm_dev_set_comms(const dlesm_comms *tables);
/* exchange_generic (src/parallel_comms_mod.f90:1501-1855) on device-resident fields: fused
 * pack kernel -> grouped NCCL send/recv (self-neighbours: on-device) -> fused unpack kernel */
int dlesm_dev_halo_exchange(void *const *fields, int nfields);
/* How dlesm_dev_halo_exchange moves the patches between ranks.  DIRECT (default): one kernel packs every
 * send patch straight into the neighbour's receive slot over NVLink peer memory and publishes a flag, a
 * second waits for the neighbours' flags and unpacks -- two launches, no library call.  NCCL: pack kernel,
 * grouped ncclSend/ncclRecv, unpack kernel.  Same result, bit for bit.  The mode must be the same on every
 * rank; the environment variable DLESM_EXCHANGE=nccl|direct sets the initial value. */
#define DLESM_EXCHANGE_DIRECT 0
#define DLESM_EXCHANGE_NCCL 1
int dlesm_dev_set_exchange_mode(int mode);
/* fld%halo_exchange(depth) (src/field_mod.f90:1231-1256) for a field whose values are in HOST memory
 * (data_on_device false): send patches up to a staging array, device exchange, receive patches back.
 * host = fld%data (nx*ny, column-major).  A no-op on one rank, like the reference. */
int dlesm_host_halo_exchange(double *host, int nx, int ny);
/* Upload grid arrays once (grid_mod would call this at the end of grid_init): tmask int32
 * (nx,ny) and 16 FP64 metric arrays in the DLESM_GRID_* order (NULL entries are skipped);
 * device_ptrs[17] receives the 17 *_device values of src/grid_mod.f90:104-150. */
int dlesm_dev_grid_upload(int nx, int ny, const int *tmask, const double *const metrics[16],
                          void *device_ptrs[17]);
/* Free what dlesm_dev_grid_upload returned (NULL entries are skipped) and set the 17 slots to NULL.
 * The reference's grid_mod has no finaliser and grid_init may run again on the same grid
 * (src/grid_mod.f90:104-150 only declares the slots): the patched grid_init calls this before it
 * uploads again, and an application calls it when it drops a grid. */
int dlesm_dev_grid_release(void *device_ptrs[17]);

/* ------------------------------------------------------------------------------------ */
/* DEVICE LAYER: shallow-water stencils (SURVEY.md a-S).  One launcher per kernel of the  */
/* un-fused formulation; each sweeps the OUTPUT field's internal region                  */
/* (xstart:xstop, ystart:ystop), like the PSy-layer loop nest of SURVEY CS-3.            */
/* tmask: device int32 (nx,ny) or NULL (all wet); masked rule in DESIGN.md.              */
/* ------------------------------------------------------------------------------------ */
int dlesm_compute_cu(void *cu, const void *p, const void *u, int xstart, int xstop, int ystart,
                     int ystop, const void *tmask);
int dlesm_compute_cv(void *cv, const void *p, const void *v, int xstart, int xstop, int ystart,
                     int ystop, const void *tmask);
int dlesm_compute_z(void *z, const void *p, const void *u, const void *v, double dx, double dy,
                    int xstart, int xstop, int ystart, int ystop, const void *tmask);
int dlesm_compute_h(void *h, const void *p, const void *u, const void *v, int xstart, int xstop,
                    int ystart, int ystop, const void *tmask);
int dlesm_compute_unew(void *unew, const void *uold, const void *z, const void *cv, const void *h,
                       double tdt, double dx, int xstart, int xstop, int ystart, int ystop,
                       const void *tmask, const void *ucur);
int dlesm_compute_vnew(void *vnew, const void *vold, const void *z, const void *cu, const void *h,
                       double tdt, double dy, int xstart, int xstop, int ystart, int ystop,
                       const void *tmask, const void *vcur);
int dlesm_compute_pnew(void *pnew, const void *pold, const void *cu, const void *cv, double tdt,
                       double dx, double dy, int xstart, int xstop, int ystart, int ystop,
                       const void *tmask, const void *pcur);
/* points = DLESM_GO_U/V/T_POINTS selects the cells the mask rule looks at */
int dlesm_time_smooth(const void *f, const void *fnew, void *fold, double alpha, int xstart,
                      int xstop, int ystart, int ystop, const void *tmask, int points);
/* The whole step fused into one sweep: reads u,v,p (with valid depth-1 halos) and
 * uold,vold,pold; writes unew,vnew,pnew on the internal region and, unless `first`, the
 * time-filtered uold,vold,pold in place.  96 B / point.  Halos of the new fields are NOT
 * filled (follow with dlesm_dev_periodic_halos or dlesm_dev_halo_exchange).  `variant`
 * selects the kernel (0 = default/best; others documented in DESIGN.md, for benchmarking). */
/* Halo push (multi-GPU form of the fused step).  OR DLESM_FUSED_PUSH into `variant`: the sweep
 * then also stores its edge outputs straight into the neighbours' halo cells of unew, vnew,
 * pnew -- peer-mapped device memory over NVLink (or this rank's own array where the wrap-around
 * neighbour is itself) -- following the send list installed by dlesm_dev_set_comms, and ends
 * with a flag hand-shake between neighbours.  No pack / NCCL / unpack, no separate halo kernel:
 * on return (stream order) the halos of the new fields are filled, corners included.
 * Needs unew, vnew, pnew registered on every rank with dlesm_dev_peer_register (collective,
 * same order everywhere; register both buffer sets when they rotate), a TMA-eligible pitch,
 * and all ranks stepping in lockstep. */
#define DLESM_FUSED_PUSH 0x1000
int dlesm_dev_peer_register(void *const *fields, int nfields);
/* Collective: drop the registration of these fields (the neighbours' mappings they hold are released,
 * closed when no other registered field shares the allocation), then barrier.  Call before freeing
 * registered fields (importers must close a CUDA IPC mapping before the exporter frees). */
int dlesm_dev_peer_unregister(void *const *fields, int nfields);
/* Collective: the same for every field registered so far. */
int dlesm_dev_peer_unregister_all(void);
/* Arithmetic of the fused sweeps.  DLESM_ARITH_EXACT (default): one rounding per operation
 * (-fmad=false), bit-identical to the reference's FMA-free Fortran loops.  DLESM_ARITH_CONTRACTED: the
 * same kernels compiled with FMA contraction -- fewer FP64 instructions -- results within north_star's
 * 1e-12 relative tolerance of the exact ones.  The environment variable DLESM_ARITH=fma selects the
 * contracted form at start-up. */
#define DLESM_ARITH_EXACT 0
#define DLESM_ARITH_CONTRACTED 1
int dlesm_set_arith(int mode);
int dlesm_get_arith(void);
/* The tile grid of the sweeps: ntilex x-tiles by ntiley segments (0 = the library's choice).  What
 * r2d_field(..., do_tile) + GOCEAN_OMP_GRID="AxB" set (src/field_mod.f90:300-325); a Fortran build calls
 * this from the field constructor. */
int dlesm_dev_set_sweep_tiles(int ntilex, int ntiley);
int dlesm_sw_step_fused(const void *u, const void *v, const void *p, void *uold, void *vold,
                        void *pold, void *unew, void *vnew, void *pnew, double dx, double dy,
                        double tdt, double alpha, int first, int xstart, int xstop, int ystart,
                        int ystop, const void *tmask, int variant);

/* TWO model steps in one sweep (temporal blocking; extension, no reference counterpart: the
 * reference's formulation is one loop nest per kernel).  The intermediate state
 * of step n+1 is recomputed on a one-cell ring around each CTA's block and never stored, so the
 * two steps cost the 96 B / point of one.  Results are bit-identical to two dlesm_sw_step_fused
 * calls with a halo fill in between.  Requirements: a tile decomposed with halo_width >= 2
 * (xstart, ystart >= 3); u,v,p valid to depth 2 and uold,vold,pold valid to depth 1 around the
 * box; `tdt` = 2*dt (not the first step); even pitch.  `tmask` (may be NULL; int32, pitch a
 * multiple of 4) must hold the periodic mask two cells deep around the box -- grid_init
 * replicates outward beyond the first ring, dlesm_sw_set_two_step builds the wrapped copy.  Writes u2,v2,p2 (step n+2)
 * and the twice-filtered uold2,vold2,pold2 on the box -- separate arrays, not in place.  Halos
 * of the outputs are NOT filled (dlesm_dev_periodic_halos_deep / halo_exchange(depth 2)) unless
 * DLESM_FUSED_PUSH is ORed into `variant`: then the sweep pushes them into the neighbours' arrays
 * as dlesm_sw_step_fused does, following the installed send list (install the depth-2 tables
 * with dlesm_dev_set_comms; register all six outputs with dlesm_dev_peer_register).
 * `variant` 0 = default.  Timed by dlesm_profile_enable like dlesm_sw_step_fused. */
int dlesm_sw_step2_fused(const void *u, const void *v, const void *p, const void *uold,
                         const void *vold, const void *pold, void *u2, void *v2, void *p2,
                         void *uold2, void *vold2, void *pold2, double dx, double dy, double tdt,
                         double alpha, int xstart, int xstop, int ystart, int ystop, const void *tmask,
                         int variant);
/* periodic halo fill of one field to `depth` cells on one rank (the reference's lists,
 * src/field_mod.f90:1394-1464, are one cell deep; depth 1 == dlesm_dev_periodic_halos) */
int dlesm_dev_periodic_halos_deep(void *f, int xstart, int xstop, int ystart, int ystop, int depth);

/* ------------------------------------------------------------------------------------ */
/* HOST MIRROR: the application ("PSy") layer for the shallow-water model -- not part of  */
/* the reference library; provided so that tests/bench drive the fields the way a GOcean   */
/* application would.                                                                     */
/* ------------------------------------------------------------------------------------ */
typedef struct dlesm_sw dlesm_sw_t;
#define DLESM_SW_UNFUSED 0 /* one launch per reference-style kernel, copy_field rotation */
#define DLESM_SW_FUSED 1   /* single fused sweep, pointer rotation                       */
/* field ids for dlesm_sw_field */
enum { DLESM_SW_U = 0, DLESM_SW_V, DLESM_SW_P, DLESM_SW_UOLD, DLESM_SW_VOLD, DLESM_SW_POLD,
       DLESM_SW_UNEW, DLESM_SW_VNEW, DLESM_SW_PNEW, DLESM_SW_CU, DLESM_SW_CV, DLESM_SW_Z,
       DLESM_SW_H, DLESM_SW_NFIELDS };

dlesm_sw_t *dlesm_sw_create(dlesm_grid_t *g, double dt, double alpha, int mode, int use_tmask);
/* classic "shallow" initial conditions computed on the device (synthetic data) */
int dlesm_sw_init_device(dlesm_sw_t *s);
/* initial u,v,p from host arrays (nx,ny) incl. valid halos; uold,vold,pold <- copies */
int dlesm_sw_set_state(dlesm_sw_t *s, const double *u, const double *v, const double *p);
int dlesm_sw_step(dlesm_sw_t *s, int nsteps);
/* One time step with HOST state, streamed through the device in row bands (upload of band
 * k+1, sweep of band k and download of band k-1 overlap; both PCIe directions busy).
 * host_in[6]  = u, v, p, uold, vold, pold  (nx*ny each, column-major, u/v/p with valid halos);
 * host_out[6] = the new u, v, p (halos filled) and the time-filtered uold, vold, pold.
 * host_out[q] may be host_in[q] (in place).  Pinned memory (dlesm_host_alloc/register) makes
 * the copies asynchronous.  Equivalent to write_to_device x6, dlesm_sw_step(s, 1),
 * read_from_device x6 of the reference's hooks; the device state ends up the same.
 * band_rows <= 0 selects the default.  Fused mode only. */
int dlesm_sw_step_host(dlesm_sw_t *s, const double *const host_in[6], double *const host_out[6],
                       int band_rows);
/* nsteps model steps per PCIe round trip.  nsteps = 1 is dlesm_sw_step_host.  nsteps = 2 streams the
 * bands through the two-step sweep (dlesm_sw_set_two_step(on), tile decomposed with halo_width 2,
 * exchange halo mode, model past its first step): each row crosses the bus once per direction for
 * two steps.  On one rank nsteps may be any even number: nsteps / 2 skewed passes of the two-step sweep over the
 * bands (pass p+1 runs two bands behind pass p while later bands still upload; the last pass's bands download
 * as they finish), one bus crossing per direction per CALL.  Host arrays carry a valid two-cell halo ring of all
 * six fields in, and get it back. */
int dlesm_sw_step_host_n(dlesm_sw_t *s, const double *const host_in[6], double *const host_out[6],
                         int band_rows, int nsteps);
dlesm_field_t *dlesm_sw_field(dlesm_sw_t *s, int which);
/* ------------------------------------------------------------------------------------ */
/* Several sub-domains per rank (extension).  The reference carries the bookkeeping --     */
/* decomposition_type%proc_subdomains (src/decomposition_mod.f90:64-67), go_decompose with  */
/* ndomains > nranks (src/parallel_mod.f90:141-152) -- and stops at map_comms ("\TODO      */
/* support multiple subdomains per rank", src/parallel_comms_mod.f90:295).                  */
/* A tile set owns one ordinary grid per sub-domain of this rank (grid_init each with its   */
/* local T-mask, create fields / models on them as usual); one exchange fills the halos of  */
/* every tile: same-rank neighbours by device copies, the others packed per peer rank.      */
/* Periodic wrap is part of the plan (no DLESM_EXT_PERIODIC_DISTMEM needed).                */
/* ------------------------------------------------------------------------------------ */
typedef struct dlesm_tileset dlesm_tileset_t;
dlesm_tileset_t *dlesm_tileset_create(int grid_name, const int boundary_conditions[3], int grid_offsets,
                                      int domainx, int domainy, int ndomains, int ndomainx, int ndomainy,
                                      int halo_width);
int dlesm_tileset_ntiles(const dlesm_tileset_t *ts);            /* tiles of THIS rank (domperrank) */
dlesm_grid_t *dlesm_tileset_grid(dlesm_tileset_t *ts, int k);   /* 0-based; owned by the set */
int dlesm_tileset_tile_id(const dlesm_tileset_t *ts, int k);    /* 1-based sub-domain id, proc_subdomains(k+1, rank) */
const dlesm_decomp_t *dlesm_tileset_decomp(const dlesm_tileset_t *ts);
void dlesm_tileset_free(dlesm_tileset_t *ts);
/* host only: this rank's exchange plan, rows of 14 integers
 * (kind 0 copy / 1 send / 2 receive, peer rank (0-based), src tile, dst tile, si0, sj0, di0, dj0, nx, ny,
 *  buffer offset, destination tile id, source tile id, direction tag); totals = send / receive elements */
int dlesm_tileset_plan(const dlesm_decomp_t *d, int rank1, int nranks, int pbc_x, int pbc_y, int depth,
                       long long *rows, int max_rows, int *nrows, long long totals[2]);
/* fields[q * ntiles + k] = field q on tile k; collective over the ranks */
int dlesm_tileset_halo_exchange(dlesm_tileset_t *ts, dlesm_field_t *const *fields, int nfields, int depth);
/* fields[k] = one field per tile -> global array (global_nx * global_ny) on rank 1; collective */
int dlesm_tileset_gather_inner_data(dlesm_tileset_t *ts, dlesm_field_t *const *fields, double *global_data);
/* shallow-water models created on the set's grids (fused mode): step them together */
int dlesm_tileset_sw_step(dlesm_tileset_t *ts, dlesm_sw_t *const *sws, int nsteps);
/* after dlesm_sw_init_device / dlesm_sw_set_state on every tile: halos of u,v,p, then the old fields */
int dlesm_tileset_sw_sync_state(dlesm_tileset_t *ts, dlesm_sw_t *const *sws);

int dlesm_sw_set_variant(dlesm_sw_t *s, int variant);
/* How the halos of unew, vnew, pnew get filled after the fused sweep:
 * DLESM_HALO_EXCHANGE (default): periodic copies on one rank, pack -> NCCL -> unpack on several;
 * DLESM_HALO_PUSH: the sweep stores them itself into the neighbours' arrays (DLESM_FUSED_PUSH).
 * Collective: every rank must make the same call. */
#define DLESM_HALO_EXCHANGE 0
#define DLESM_HALO_PUSH 1
int dlesm_sw_set_halo_mode(dlesm_sw_t *s, int mode);
/* on != 0: dlesm_sw_step advances two steps per sweep (dlesm_sw_step2_fused) whenever at least
 * two steps remain and the first step is behind; an odd remainder runs the one-step sweep.
 * Needs the fused mode, an even pitch (a multiple of 4 with a tmask) and a grid decomposed with
 * halo_width >= 2;
 * halos are then kept two cells deep (periodic copies on one rank, halo_exchange(depth 2) on
 * several -- DLESM_EXT_HALO_DEPTH).  Allocates three more fields.  With DLESM_HALO_PUSH (either
 * call order; both collective) the two-step sweep stores the neighbours' halo cells of its six
 * outputs itself, two cells deep, and no exchange follows it.  `variant2` selects the two-step
 * kernel (0 = default). */
int dlesm_sw_set_two_step(dlesm_sw_t *s, int on, int variant2);
/* kernels launched by this library since process start (all streams) */
long long dlesm_kernel_launch_count(void);
/* CUDA-event time (ms) spent in dominant-kernel launches since the last call with reset=1,
 * and their count; enabled by dlesm_profile_enable(1) (adds event records around each
 * fused-step launch; used by bench.py for the roofline figure) */
int dlesm_profile_enable(int on);
int dlesm_profile_read(double *ms_total, long long *nlaunches, int reset);
/* Diagnostics: per-CTA timeline of the fused TMA sweeps launched between begin and end.  `out` gets 4
 * long long per CTA (index by * grid.x + bx of dlesm_sw_launch_grid): start ns, end ns, SM id, ns after
 * the wait for the neighbours' previous sweep (halo push).  Later launches overwrite earlier ones. */
int dlesm_debug_cta_trace_begin(int max_ctas);
int dlesm_debug_cta_trace_end(long long *out, int max_ctas);
/* Collective when the object pushes halos on several ranks (call it on every rank). */
void dlesm_sw_free(dlesm_sw_t *s);
/* Non-collective release for finalisers: frees everything except buffers the neighbours still have
 * mapped (those leak until dlesm_dev_peer_unregister_all / process exit; a warning is printed). */
void dlesm_sw_free_local(dlesm_sw_t *s);
/* launch grid (grid.x, grid.y) of the model's one-step (two_step = 0) or two-step sweep */
int dlesm_sw_launch_grid(dlesm_sw_t *s, int two_step, int grid[2]);

/* ------------------------------------------------------------------------------------ */
/* PSy LAYER: metadata-driven kernel dispatch -- kernel_mod / argument_mod                */
/* (src/kernel_mod.f90:30-50, src/argument_mod.f90:39-112, the calling convention of      */
/* src/infrastructure_mod.f90:13-41: `type, extends(kernel_type)` with meta_args,          */
/* ITERATES_OVER, index_offset and a per-point `code(ji, jj, args...)`).                   */
/*                                                                                      */
/* The application supplies its per-point kernel as a CUDA C++ device function with the   */
/* reference's calling convention,                                                       */
/*     __device__ void <name>_code(int ji, int jj, <one parameter per meta_arg>)           */
/* and the same metadata it would write in Fortran.  The library generates the sweep      */
/* (the `do jj / do ji` nest of SURVEY CS-3 as a coalesced 2-D grid), compiles it for      */
/* sm_100a with NVRTC (-fmad=false, like the rest of the library) and launches it over    */
/* device-resident fields; grid-property arguments are resolved from the grid.             */
/* Parameter types, by meta_arg:                                                         */
/*   field, GO_READ                 -> dlesm_cfield  (a(ji,jj), 1-based, read-only)        */
/*   field, GO_WRITE/READWRITE/INC  -> dlesm_field   (a(ji,jj) is an lvalue)               */
/*   GO_R_SCALAR / GO_I_SCALAR      -> double / int                                       */
/*   GO_GRID_MASK_T                 -> dlesm_imask   (int tmask(ji,jj))                    */
/*   other GO_GRID_* arrays         -> dlesm_cfield                                       */
/*   GO_GRID_DX_CONST / DY_CONST    -> double;  GO_GRID_{X,Y}_{MIN,MAX}_INDEX -> int       */
/* ------------------------------------------------------------------------------------ */
/* src/argument_mod.f90:39-46 */
enum { DLESM_GO_READ = 0, DLESM_GO_WRITE, DLESM_GO_READWRITE, DLESM_GO_INC, DLESM_GO_MIN, DLESM_GO_MAX,
       DLESM_GO_SUM };
/* :68-71 */
#define DLESM_GO_R_SCALAR 0
#define DLESM_GO_I_SCALAR 1
#define DLESM_GO_EVERY 1
#define DLESM_GO_CU 1
#define DLESM_GO_CV 2
#define DLESM_GO_CT 3
#define DLESM_GO_CF 4
/* :75-112 */
enum { DLESM_GO_TIME_STEP = 1, DLESM_GO_GRID_AREA_T, DLESM_GO_GRID_AREA_U, DLESM_GO_GRID_AREA_V,
       DLESM_GO_GRID_MASK_T, DLESM_GO_GRID_DX_T, DLESM_GO_GRID_DX_U, DLESM_GO_GRID_DX_V,
       DLESM_GO_GRID_DY_T, DLESM_GO_GRID_DY_U, DLESM_GO_GRID_DY_V, DLESM_GO_GRID_LAT_U,
       DLESM_GO_GRID_LAT_V, DLESM_GO_GRID_DX_CONST, DLESM_GO_GRID_DY_CONST,
       DLESM_GO_GRID_X_MIN_INDEX, DLESM_GO_GRID_X_MAX_INDEX, DLESM_GO_GRID_Y_MIN_INDEX,
       DLESM_GO_GRID_Y_MAX_INDEX };
/* src/kernel_mod.f90:35-37 */
#define DLESM_GO_INTERNAL_PTS 0
#define DLESM_GO_EXTERNAL_PTS 1
#define DLESM_GO_ALL_PTS 2
/* go_arg (src/argument_mod.f90:57-61).  In Fortran the element names carry their meaning; the
 * integer values overlap (GO_CU == GO_TIME_STEP == GO_I_SCALAR == 1), so `kind` says which
 * family `element` belongs to. */
#define DLESM_ARG_FIELD 0  /* element = GO_CU / GO_CV / GO_CT / GO_CF / GO_EVERY */
#define DLESM_ARG_SCALAR 1 /* element = GO_R_SCALAR / GO_I_SCALAR                 */
#define DLESM_ARG_GRID 2   /* element = GO_GRID_*                                  */
typedef struct {
    int arg_intent; /* DLESM_GO_READ ...                                         */
    int kind;       /* DLESM_ARG_*                                               */
    int element;
    int stencil[3]; /* go_stencil rows, e.g. {000, 010, 000} = GO_POINTWISE (informational) */
} dlesm_go_arg;

typedef struct dlesm_kernel dlesm_kernel_t;
/* Build a kernel object.  Host only (NVRTC needs no GPU): compiles at once and reports
 * compile errors through dlesm_last_error().  iterates_over = DLESM_GO_INTERNAL_PTS or
 * DLESM_GO_ALL_PTS (GO_EXTERNAL_PTS is not supported, as in PSyclone); index_offset is checked
 * against the grid at launch (DLESM_GO_OFFSET_ANY matches every grid). */
dlesm_kernel_t *dlesm_kernel_create(const char *name, const char *cuda_source, const dlesm_go_arg *meta_args,
                                    int nargs, int iterates_over, int index_offset);
/* the generated translation unit (prelude + the application's source + the sweep), for inspection */
const char *dlesm_kernel_generated_source(const dlesm_kernel_t *k);
/* Sweep the kernel.  fields / rscalars / iscalars are consumed in metadata order (grid-property
 * arguments take nothing from them).  The sweep covers the `internal` (GO_INTERNAL_PTS) or `whole`
 * (GO_ALL_PTS) region of the first field argument that is written (GO_WRITE / READWRITE / INC). */
int dlesm_kernel_launch(dlesm_kernel_t *k, dlesm_field_t *const *fields, int nfields, const double *rscalars,
                        int nrscalars, const int *iscalars, int niscalars);
/* Device-layer form for a Fortran PSy layer: arrays = device pointers of the field AND grid-array
 * arguments in metadata order (fld%device_ptr, grid%tmask_device, ...); reals = GO_R_SCALARs and real
 * grid scalars in order; ints = GO_I_SCALARs and index properties in order; explicit sweep bounds. */
int dlesm_kernel_launch_dev(dlesm_kernel_t *k, void *const *arrays, int narrays, const double *reals, int nreals,
                            const int *ints, int nints, int xstart, int xstop, int ystart, int ystop);
void dlesm_kernel_free(dlesm_kernel_t *k);

#ifdef __cplusplus
}
#endif
#endif /* DLESM_B200_H */
