/*
 * dlesm_oracle.h -- CPU oracle for the dl_esm_inf hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is a plain-C restatement (1-based, column-major, same loop order) of the
 * reference's Fortran for the hot path named in BASELINE.json `north_star`.  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it; the product library (dl_esm_inf_b200/) never does.
 *
 * Parity status:
 *   - bookkeeping (decomposition, padding, regions, halos, comm tables, pack/unpack,
 *     scatter/gather, device-hook sub-rectangles, checksum-of-ones): PINNED against
 *     every golden vector the reference's own tests hold (see tests/test_oracle_golden.py,
 *     SURVEY.md 8c): test_gsum, test_halos, test_reduction, test_device_io, example/model.
 *   - shallow-water stencil arithmetic: "parity unpinned" -- the reference ships no
 *     stencil and no test computes one; this file is the normative
 *     definition of those loops.
 *
 * Citations: src/X = /root/reference/finite_difference/src/X.
 */
#ifndef DLESM_ORACLE_H
#define DLESM_ORACLE_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* src/region_mod.f90:7-12 -- inclusive 1-based index box */
typedef struct {
    int nx, ny;
    int xstart, xstop;
    int ystart, ystop;
} orc_region;

/* src/decomposition_mod.f90:44-50 */
typedef struct {
    orc_region global;
    orc_region internal;
} orc_subdomain;

#define ORC_MAX_DOMAINS 4096

/* src/decomposition_mod.f90:54-68 */
typedef struct {
    int global_nx, global_ny;
    int nx, ny;
    int ndomains;
    int max_width, max_height;
    int domperrank;                 /* first extent of proc_subdomains(:,:) */
    int nranks;                     /* second extent of proc_subdomains(:,:) */
    orc_subdomain subdomains[ORC_MAX_DOMAINS];
    int proc_subdomains[ORC_MAX_DOMAINS]; /* (domperrank, nranks) column-major */
} orc_decomp;

/* src/grid_mod.f90:45-69 */
enum { ORC_ARAKAWA_C = 0, ORC_ARAKAWA_B = 1 };
enum { ORC_OFFSET_SW = 0, ORC_OFFSET_SE = 1, ORC_OFFSET_NW = 2, ORC_OFFSET_NE = 3, ORC_OFFSET_ANY = 4 };
enum { ORC_BC_PERIODIC = 0, ORC_BC_EXTERNAL = 1, ORC_BC_NONE = 2 };
/* src/field_mod.f90:47-52 */
enum { ORC_U_POINTS = 0, ORC_V_POINTS = 1, ORC_T_POINTS = 2, ORC_F_POINTS = 3, ORC_ALL_POINTS = 4 };

/* error codes standing in for gocean_stop / parallel_abort (the reference has no
 * status returns on this path: src/gocean_mod.F90:50-57) */
enum {
    ORC_OK = 0,
    ORC_ERR_ARGS = -1,          /* go_decompose: invalid arguments supplied */
    ORC_ERR_NRANKS = -2,
    ORC_ERR_HALO_WIDTH = -3,
    ORC_ERR_ALIGN_DIGITS = -4,  /* DL_ESM_ALIGNMENT longer than 3 characters */
    ORC_ERR_ALIGN_VALUE = -5,   /* DL_ESM_ALIGNMENT not a positive integer */
    ORC_ERR_PBC_DISTMEM = -6,   /* periodic BCs with distributed memory */
    ORC_ERR_GRID_NAME = -7,
    ORC_ERR_GRID_OFFSET = -8,
    ORC_ERR_FIELD_STOP = -9,    /* a c?_{sw,ne}_init gocean_stop branch */
    ORC_ERR_POINTS = -10,
    ORC_ERR_MAXCOMM = -12,      /* addsend/addrecv list overflow (ierr=-12) */
    ORC_ERR_HALO_DEPTH = -13,
    ORC_ERR_TOO_MANY = -14      /* oracle capacity, not a reference error */
};

/* ---- parallel_mod ---------------------------------------------------------- */
/* src/parallel_mod.f90:70-332.  Optional arguments: pass a negative value for "absent". */
int orc_go_decompose(int domainx, int domainy, int ndomains, int ndomainx, int ndomainy,
                     int halo_width, int nranks, orc_decomp *decomp);
/* `hwidth` is an initialised (hence SAVEd) local: src/parallel_mod.f90:103-105 */
void orc_reset_saved_state(void);

/* ---- grid_mod --------------------------------------------------------------- */
/* src/grid_mod.f90:344-363: parse DL_ESM_ALIGNMENT (NULL => variable not present) */
int orc_parse_alignment(const char *envval, int *alignment);
/* src/grid_mod.f90:217-304 validation part of grid_constructor */
int orc_grid_constructor_check(int grid_name, int grid_offset);

typedef struct {
    int nx, ny;              /* padded local extents: src/grid_mod.f90:368-385 */
    int padding;
    double dx, dy;
} orc_grid_dims;

int orc_grid_dims_compute(const orc_subdomain *sub, int alignment, orc_grid_dims *out);

/* src/grid_mod.f90:330-570.  All output arrays are (nx,ny) column-major and may be NULL
 * to skip.  tmask_in may be NULL (=> all-wet box, rest left untouched = "uninitialised").
 * metrics[16] order: dx_t,dy_t,dx_u,dy_u,dx_v,dy_v,dx_f,dy_f,area_t,area_u,area_v,
 *                    gphiu,gphiv,gphif,xt,yt */
int orc_grid_init(const orc_subdomain *sub, const int bcs[3], int nranks, int dist_mem_enabled,
                  int alignment, double dxarg, double dyarg,
                  const int *tmask_in, int tmask_in_nx,
                  orc_grid_dims *dims, int *tmask_out, double *metrics[16]);

/* ---- field_mod -------------------------------------------------------------- */
typedef struct {
    orc_region source;
    orc_region dest;
} orc_halo;

typedef struct {
    int defined_on;
    orc_region internal;
    orc_region whole;
    int num_halos;          /* -1 where the reference leaves it unset (cv/ct/cf_ne_init) */
    orc_halo halo[4];
} orc_field_bounds;

/* src/field_mod.f90:563-1122, 1394-1464 */
int orc_set_field_bounds(const orc_subdomain *sub, int grid_nx, int grid_ny, int grid_offset,
                         const int bcs[3], int grid_points, orc_field_bounds *out);

/* r2d_field constructor data part: zero fill + optional scatter, src/field_mod.f90:375-389 */
void orc_field_construct(const orc_subdomain *sub, const orc_field_bounds *fb, int nx, int ny,
                         const double *init_global_data, int gnx, double *data);

/* src/field_mod.f90:1126-1133 / 1152-1174 (ntiles == 0) */
void orc_copy_field(const double *field_in, double *field_out, int nx, int ny);
/* src/field_mod.f90:1140-1148 / 1179-1187 (RHS evaluated before assignment) */
void orc_copy_patch(double *field, int nx, const orc_region *src, const orc_region *dest);
/* src/field_mod.f90:1191-1202 */
void orc_set_field(double *field, int nx, int ny, double val);
/* src/field_mod.f90:1289-1302: SUM(ABS(box)), sequential, i fastest */
double orc_array_checksum(const double *field, int nx, int xstart, int xstop, int ystart, int ystop);
/* same sum in long double: the accuracy yardstick of SURVEY.md 8d */
double orc_array_checksum_ld(const double *field, int nx, int xstart, int xstop, int ystart, int ystop);
/* serial periodic-BC fill: copy_field over halo(1..num_halos), SURVEY CS-6 */
void orc_apply_periodic_halos(double *field, int nx, const orc_field_bounds *fb);

/* device hooks, sub-rectangle semantics: tests/device_computation/test_device_io.f90:50-90 */
void orc_write_to_device(const double *host, double *device, int pitch,
                         int startx, int starty, int nx, int ny);
void orc_read_from_device(const double *device, double *host, int pitch,
                          int startx, int starty, int nx, int ny);

/* src/field_mod.f90:1313-1390.  fields[r] / bounds[r] are rank r+1's; global_data is
 * (global_nx, global_ny) on the master. */
int orc_gather_inner_data(const orc_decomp *decomp, int nranks, const double *const *fields,
                          const int *pitches, const orc_field_bounds *bounds,
                          double *global_data);

/* ---- parallel_comms_mod ------------------------------------------------------ */
#define ORC_MAXCOMM 16
enum { ORC_NONE = 0, ORC_IPLUS = 1, ORC_IMINUS = 2, ORC_JPLUS = 3, ORC_JMINUS = 4,
       ORC_IPLUSJPLUS = 5, ORC_IMINUSJMINUS = 6, ORC_IPLUSJMINUS = 7, ORC_IMINUSJPLUS = 8 };

/* module-level lists of src/parallel_comms_mod.f90:70-83 (one instance per rank here) */
typedef struct {
    int nsend, nrecv;
    int dirsend[ORC_MAXCOMM], destination[ORC_MAXCOMM];
    int isrcsend[ORC_MAXCOMM], jsrcsend[ORC_MAXCOMM];
    int idessend[ORC_MAXCOMM], jdessend[ORC_MAXCOMM];
    int nxsend[ORC_MAXCOMM], nysend[ORC_MAXCOMM], nsendp[ORC_MAXCOMM];
    int dirrecv[ORC_MAXCOMM], source[ORC_MAXCOMM];
    int isrcrecv[ORC_MAXCOMM], jsrcrecv[ORC_MAXCOMM];
    int idesrecv[ORC_MAXCOMM], jdesrecv[ORC_MAXCOMM];
    int nxrecv[ORC_MAXCOMM], nyrecv[ORC_MAXCOMM], nrecvp[ORC_MAXCOMM];
} orc_comms;

/* src/parallel_comms_mod.f90:1365-1398 */
int orc_iprocmap(const orc_decomp *decomp, int nprocp, int ia, int ja);
/* src/parallel_comms_mod.f90:178-1172; irank is 1-based */
int orc_map_comms(const orc_decomp *decomp, int irank, int nprocp, int pbc,
                  const int halo_depths[2], orc_comms *out);
/* src/parallel_comms_mod.f90:1501-1855 for all ranks at once: pack every enabled send
 * (j outer, i inner), deliver by (source, direction tag, FIFO), unpack.  comm1..4 as in
 * field_mod.f90:1247-1248. */
int orc_exchange_generic_all(int nranks, double *const *b2, const int *pitches,
                             const orc_comms *comms, int comm1, int comm2, int comm3, int comm4);

/* ---- shallow-water stencils (NOT in the reference: parity unpinned) ----------- */
/* Per-point formulas of SURVEY.md a-S (Sadourny scheme as used by the GOcean "shallow"
 * benchmark), swept as the PSy layer of SURVEY CS-3 would: jj outer, ji inner over the
 * OUTPUT field's internal region.  `tmask` may be NULL (all wet).  Masked rule (project
 * defined): a point is advanced only when every T cell it straddles has
 * tmask == 1 (U: (i-1,j),(i,j); V: (i,j-1),(i,j); F: the four around it; T: itself);
 * otherwise diagnostic fields (cu,cv,z,h) are set to 0, prognostic "new" fields copy the
 * current value (`cur`) and the time filter leaves `fold` untouched. */
typedef struct {
    int nx, ny;            /* allocation: pitch and rows */
    double dx, dy;         /* grid%dx, grid%dy */
    double dt, tdt, alpha;
} orc_sw_params;

void orc_compute_cu(double *cu, const double *p, const double *u, int nx, const orc_region *r,
                    const int *tmask);
void orc_compute_cv(double *cv, const double *p, const double *v, int nx, const orc_region *r,
                    const int *tmask);
void orc_compute_z(double *z, const double *p, const double *u, const double *v, int nx,
                   const orc_region *r, double dx, double dy, const int *tmask);
void orc_compute_h(double *h, const double *p, const double *u, const double *v, int nx,
                   const orc_region *r, const int *tmask);
void orc_compute_unew(double *unew, const double *uold, const double *z, const double *cv,
                      const double *h, int nx, const orc_region *r, double tdt, double dx,
                      const int *tmask, const double *ucur);
void orc_compute_vnew(double *vnew, const double *vold, const double *z, const double *cu,
                      const double *h, int nx, const orc_region *r, double tdt, double dy,
                      const int *tmask, const double *vcur);
void orc_compute_pnew(double *pnew, const double *pold, const double *cu, const double *cv,
                      int nx, const orc_region *r, double tdt, double dx, double dy,
                      const int *tmask, const double *pcur);
/* points: ORC_U_POINTS / V / T selects which cells the mask rule looks at */
void orc_time_smooth(const double *f, const double *fnew, double *fold, int nx,
                     const orc_region *r, double alpha, const int *tmask, int points);

/* classic "shallow" initial conditions on the T-point internal region `t`, then the
 * periodic halos of fb_u/fb_v/fb_t; uold,vold,pold are not touched here */
void orc_sw_init(double *u, double *v, double *p, int nx, int ny, const orc_region *t,
                 double dx, double dy, const orc_field_bounds *fb_u,
                 const orc_field_bounds *fb_v, const orc_field_bounds *fb_t);

/* one full time step in the reference-style un-fused formulation (SURVEY a-S):
 * fields[13] = u,v,p,uold,vold,pold,unew,vnew,pnew,cu,cv,z,h.  `first` => no smoothing
 * (the caller doubles tdt afterwards).  u,v,p must have valid periodic halos on entry
 * and have them again on exit. */
void orc_sw_step(double *fields[13], const orc_sw_params *prm, const orc_field_bounds *fb_u,
                 const orc_field_bounds *fb_v, const orc_field_bounds *fb_t,
                 const orc_field_bounds *fb_f, const int *tmask, int first);

/* Multi-threaded dm_fd_lib-equivalent CPU baseline (bench only).  nthreads workers, one
 * go_decompose subdomain each (SW offset, local arrays sized as grid_init would), the
 * un-fused sweeps above on each subdomain, and between sweep phases an exchange that
 * follows exchange_generic's pack -> copy -> unpack order with the periodic wrap added
 * (extension: the reference refuses periodic BCs under MPI).  nthreads == 1 is
 * the serial fd_lib path (copy_field periodic halos, no exchange).
 * Returns wall seconds for `nsteps` timed steps (after `warmup` untimed ones);
 * checksums[3] = field_checksum of p,u,v after the last step (summed over workers). */
void orc_sw_last_checksums_ld(double out[3]);
double orc_sw_run_threads(int global_nx, int global_ny, int nthreads, int nsteps, int warmup,
                          double *checksums);

#ifdef __cplusplus
}
#endif
#endif
