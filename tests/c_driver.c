/* Plain-C driver of libdlesm_b200.so: the call sequence a Fortran application makes through the B200
 * build of field_mod / grid_mod (fortran/patches/), with no Python and no ctypes in between.
 *
 *   c_driver model [nx ny]     example/model.f90:54-109: four fields = rank, halo_exchange(1) on each,
 *                              field_checksum of each.  Run twice: fields host-resident (the UNMODIFIED
 *                              program: fld%data written directly, data_on_device false) and
 *                              device-resident (data_on_device true; copy/set/checksum/exchange on the GPU).
 *                              One rank: 4 x 40.0; two ranks: 4 x 60.0 (tests/golden/reference_known_answers.json).
 *   c_driver halos nx ny       tests/dist_mem/test_halos.f90:60-265: hill fields, halo_exchange, every halo
 *                              cell that has a neighbour equals hill() at its own position.
 *
 * The host-side bookkeeping the Fortran library does itself (go_decompose, grid_init metrics, field
 * bounds) comes from the library's host entry points; everything after that is the DEVICE layer the
 * patched modules call, in their order:
 *   grid_init  ->  dlesm_dev_grid_upload, dlesm_dev_set_comms
 *   r2d_field  ->  dlesm_dev_field_alloc, hooks = dlesm_{read_from,write_to}_device_c
 *   halo_exchange  -> dlesm_host_halo_exchange | dlesm_dev_halo_exchange
 *   set_field / copy_field / field_checksum -> dlesm_dev_set_field / dlesm_dev_copy_field /
 *                                              dlesm_dev_checksum + dlesm_global_sum
 *   free_field ->  dlesm_dev_field_free
 *
 * Ranks: DLESM_RANK / DLESM_NRANKS (else RANK / WORLD_SIZE) and LOCAL_RANK from the environment (one process per GPU); the 128-byte NCCL id
 * travels through the file named by DLESM_ID_FILE (rank 0 writes it, the others poll).
 * Exit status 0 = every check passed. */
#include <math.h>
#include <stdbool.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "dlesm_b200.h"

static int g_rank = 1, g_nranks = 1, g_fail = 0;

#define CHECK(call)                                                                                   \
    do {                                                                                              \
        int rc_ = (call);                                                                             \
        if (rc_ != DLESM_OK) {                                                                        \
            fprintf(stderr, "[rank %d] %s -> %d: %s\n", g_rank, #call, rc_, dlesm_last_error());      \
            exit(2);                                                                                  \
        }                                                                                             \
    } while (0)

static int env_int(const char *name, int dflt)
{
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

/* gocean_initialise(): parallel_init of the NCCL variant of parallel_utils_mod */
static void gocean_initialise(void)
{
    /* the names fortran/parallel_utils_nccl_mod.f90 reads, else torchrun's */
    g_rank = env_int("DLESM_RANK", env_int("RANK", 0)) + 1;
    g_nranks = env_int("DLESM_NRANKS", env_int("WORLD_SIZE", 1));
    CHECK(dlesm_parallel_init(g_rank, g_nranks));
    CHECK(dlesm_device_init(env_int("LOCAL_RANK", 0)));
    if (g_nranks > 1) {
        unsigned char id[128];
        const char *path = getenv("DLESM_ID_FILE");
        if (!path) { fprintf(stderr, "several ranks need DLESM_ID_FILE\n"); exit(2); }
        if (g_rank == 1) {
            char tmp[4096];
            CHECK(dlesm_comm_unique_id(id));
            snprintf(tmp, sizeof tmp, "%s.tmp", path);
            FILE *f = fopen(tmp, "wb");
            if (!f || fwrite(id, 1, 128, f) != 128) { perror("id file"); exit(2); }
            fclose(f);
            rename(tmp, path);
        } else {
            FILE *f = NULL;
            for (int t = 0; t < 600 && !(f = fopen(path, "rb")); ++t) usleep(100000);
            if (!f || fread(id, 1, 128, f) != 128) { fprintf(stderr, "no NCCL id in %s\n", path); exit(2); }
            fclose(f);
        }
        CHECK(dlesm_comm_init(id));
    }
}

/* what type(r2d_field) carries (src/field_mod.f90:103-166) */
typedef struct {
    dlesm_field_info fb;  /* defined_on, internal, whole: set_field_bounds */
    int nx, ny;
    double *data;         /* fld%data(1:nx, 1:ny), column-major */
    void *device_ptr;
    bool data_on_device;
    void (*read_from_device_c)(void *, void *, int, int, int, int, bool);
    void (*write_to_device_c)(void *, void *, int, int, int, int, bool);
} r2d_field;

typedef struct {
    dlesm_grid_t *g;
    dlesm_grid_info info;
    double *xt, *yt, *dx_u, *dy_v, *dx_f, *dy_f;
} grid_t;

#define AT(f, i, j) ((f)->data[(size_t)((j) - 1) * (f)->nx + ((i) - 1)])

/* grid_type(...) + decompose + grid_init, ending with the upload the patched grid_init adds */
static void make_grid(grid_t *G, int nx, int ny)
{
    const int bcs[3] = {DLESM_GO_BC_EXTERNAL, DLESM_GO_BC_EXTERNAL, DLESM_GO_BC_NONE};
    G->g = dlesm_grid_create(DLESM_GO_ARAKAWA_C, bcs, DLESM_GO_OFFSET_NE);
    if (!G->g) { fprintf(stderr, "grid_create: %s\n", dlesm_last_error()); exit(2); }
    CHECK(dlesm_grid_decompose(G->g, nx, ny, -1, -1, -1, -1));
    CHECK(dlesm_grid_get_info(G->g, &G->info));
    const int tnx = G->info.subdomain.global.nx, tny = G->info.subdomain.global.ny;
    int *tmask = malloc(sizeof(int) * tnx * tny);
    for (int q = 0; q < tnx * tny; ++q) tmask[q] = 1;
    CHECK(dlesm_grid_init(G->g, 1.0, 1.0, tmask, tnx, tny));
    free(tmask);
    CHECK(dlesm_grid_get_info(G->g, &G->info));
    /* the arrays Fortran's grid_init computed on the host ... */
    const size_t n = (size_t)G->info.nx * G->info.ny;
    double *m[16];
    int *tm = malloc(sizeof(int) * n);
    CHECK(dlesm_grid_host_array(G->g, DLESM_GRID_TMASK, tm));
    for (int q = 0; q < 16; ++q) {
        m[q] = malloc(sizeof(double) * n);
        CHECK(dlesm_grid_host_array(G->g, DLESM_GRID_DX_T + q, m[q]));
    }
    /* ... go to the device once: the *_device slots (src/grid_mod.f90:104-150) */
    void *dev[17];
    CHECK(dlesm_dev_grid_upload(G->info.nx, G->info.ny, tm, (const double *const *)m, dev));
    for (int q = 0; q < 17; ++q)
        if (!dev[q]) { fprintf(stderr, "grid upload left slot %d null\n", q); g_fail++; }
    /* and this rank's map_comms tables go to the device exchange */
    if (g_nranks > 1) {
        dlesm_comms t;
        CHECK(dlesm_grid_comms(G->g, &t));
        CHECK(dlesm_dev_set_comms(&t));
    }
    G->dx_u = m[DLESM_GRID_DX_U - 1]; G->dy_v = m[DLESM_GRID_DY_V - 1];
    G->dx_f = m[DLESM_GRID_DX_F - 1]; G->dy_f = m[DLESM_GRID_DY_F - 1];
    G->xt = m[DLESM_GRID_XT - 1];     G->yt = m[DLESM_GRID_YT - 1];
    free(tm);
}

/* r2d_field(grid, points): host array + (B200 build) device storage and the two C hooks */
static void make_field(r2d_field *f, grid_t *G, int points)
{
    memset(f, 0, sizeof *f);
    CHECK(dlesm_grid_field_bounds(G->g, points, &f->fb));
    f->nx = G->info.nx; f->ny = G->info.ny;
    f->data = calloc((size_t)f->nx * f->ny, sizeof(double));
    CHECK(dlesm_dev_field_alloc(f->nx, f->ny, &f->device_ptr));
    f->read_from_device_c = dlesm_read_from_device_c;
    f->write_to_device_c = dlesm_write_to_device_c;
    f->data_on_device = false;
}

static void field_to_device(r2d_field *f)   /* fortran/field_device_mod.f90 */
{
    CHECK(dlesm_host_register(f->data, (size_t)f->nx * f->ny * sizeof(double)));
    f->data_on_device = true;
    f->write_to_device_c(f->data, f->device_ptr, 1, 1, f->nx, f->ny, true);
}

static void halo_exchange(r2d_field *f)     /* fld%halo_exchange(1) */
{
    if (f->data_on_device) {
        void *ptrs[1] = {f->device_ptr};
        CHECK(dlesm_dev_halo_exchange(ptrs, 1));
    } else {
        CHECK(dlesm_host_halo_exchange(f->data, f->nx, f->ny));
    }
}

static double field_checksum(r2d_field *f)  /* src/field_mod.f90:1209-1219, 1289-1307 */
{
    const dlesm_region *r = &f->fb.internal;
    double val = 0.0;
    if (f->data_on_device) {
        CHECK(dlesm_dev_checksum(f->device_ptr, r->xstart, r->xstop, r->ystart, r->ystop, &val));
    } else {
        for (int j = r->ystart; j <= r->ystop; ++j)
            for (int i = r->xstart; i <= r->xstop; ++i) val += fabs(AT(f, i, j));
    }
    CHECK(dlesm_global_sum(&val));
    return val;
}

static void free_field(r2d_field *f)
{
    if (f->data_on_device) CHECK(dlesm_host_unregister(f->data));
    CHECK(dlesm_dev_field_free(f->device_ptr));
    free(f->data);
    memset(f, 0, sizeof *f);
}

static const int kPoints[4] = {DLESM_GO_U_POINTS, DLESM_GO_V_POINTS, DLESM_GO_T_POINTS, DLESM_GO_F_POINTS};
static const char *kNames[4] = {"U", "V", "T", "F"};

static void expect(double got, double want, const char *what)
{
    if (fabs(got - want) > 1e-12 * fabs(want)) {
        fprintf(stderr, "[rank %d] %s: got %.17g, want %.17g\n", g_rank, what, got, want);
        g_fail++;
    }
}

/* example/model.f90:54-109 */
static void run_model(int nx, int ny)
{
    grid_t G;
    make_grid(&G, nx, ny);
    /* sum over ranks of rank * internal points of that rank = what the program prints */
    double want = (double)g_rank * G.info.subdomain.internal.nx * G.info.subdomain.internal.ny;
    CHECK(dlesm_global_sum(&want));
    for (int pass = 0; pass < 2; ++pass) {
        r2d_field f[4], copy;
        double sums[4];
        for (int q = 0; q < 4; ++q) {
            make_field(&f[q], &G, kPoints[q]);
            if (pass == 0) {
                /* init_field_by_rank: field%data(:,:) = real(my_rank), written on the host */
                for (size_t k = 0; k < (size_t)f[q].nx * f[q].ny; ++k) f[q].data[k] = (double)g_rank;
            } else {
                /* device-resident: set_field dispatches to the device */
                field_to_device(&f[q]);
                CHECK(dlesm_dev_set_field(f[q].device_ptr, (double)g_rank));
            }
        }
        for (int q = 0; q < 4; ++q) halo_exchange(&f[q]);
        for (int q = 0; q < 4; ++q) sums[q] = field_checksum(&f[q]);
        if (pass == 1) {
            /* copy_field(field_in, field_out) on the device, then get_data(): hook read-back */
            make_field(&copy, &G, DLESM_GO_T_POINTS);
            field_to_device(&copy);
            CHECK(dlesm_dev_copy_field(f[2].device_ptr, copy.device_ptr));
            copy.read_from_device_c(copy.device_ptr, copy.data, 1, 1, copy.nx, copy.ny, true);
            const dlesm_region *r = &copy.fb.internal;
            for (int j = r->ystart; j <= r->ystop; ++j)
                for (int i = r->xstart; i <= r->xstop; ++i)
                    if (AT(&copy, i, j) != (double)g_rank) { g_fail++; break; }
            expect(field_checksum(&copy), want, "checksum of the device copy");
            free_field(&copy);
        }
        if (dlesm_on_master()) {
            printf("\n%s fields:\n", pass == 0 ? "host-resident" : "device-resident");
            for (int q = 0; q < 4; ++q) printf("%s checksum = %15.8E\n", kNames[q], sums[q]);
        }
        /* the T field's internal region is the sub-domain: its checksum is the known answer; U, V and F
         * on the NE-offset grid have the same internal extents except at the domain's north/east edge */
        expect(sums[2], want, "T checksum");
        if (g_nranks == 1 && nx == 4 && ny == 10)
            for (int q = 0; q < 4; ++q) expect(sums[q], 40.0, kNames[q]);
        if (g_nranks == 2 && nx == 4 && ny == 10)
            for (int q = 0; q < 4; ++q) expect(sums[q], 60.0, kNames[q]);
        for (int q = 0; q < 4; ++q) free_field(&f[q]);
    }
    dlesm_grid_free(G.g);
}

/* tests/dist_mem/test_halos.f90:142-189 */
static double hill(const grid_t *G, const r2d_field *f, int i, int j)
{
    const size_t k = (size_t)(j - 1) * f->nx + (i - 1);
    double x = G->xt[k], y = G->yt[k];
    switch (f->fb.defined_on) {   /* GO_OFFSET_NE */
    case DLESM_GO_U_POINTS: x += 0.5 * G->dx_u[k]; break;
    case DLESM_GO_V_POINTS: y += 0.5 * G->dy_v[k]; break;
    case DLESM_GO_F_POINTS: x += 0.5 * G->dx_f[k]; y += 0.5 * G->dy_f[k]; break;
    default: break;
    }
    return (double)(float)(10000.0 * x + y);   /* real(...) of default kind, as the reference writes it */
}

static void init_field_hill(const grid_t *G, r2d_field *f)
{
    const dlesm_region *r = &f->fb.internal;
    memset(f->data, 0, sizeof(double) * f->nx * f->ny);
    for (int j = r->ystart; j <= r->ystop; ++j)
        for (int i = r->xstart; i <= r->xstop; ++i) AT(f, i, j) = hill(G, f, i, j);
    for (int j = 1; j <= f->ny; ++j) {
        for (int i = 1; i < r->xstart; ++i) AT(f, i, j) = AT(f, r->xstart, j);
        for (int i = r->xstop + 1; i <= f->nx; ++i) AT(f, i, j) = AT(f, r->xstop, j);
    }
    for (int i = 1; i <= f->nx; ++i) {
        for (int j = 1; j < r->ystart; ++j) AT(f, i, j) = AT(f, i, r->ystart);
        for (int j = r->ystop + 1; j <= f->ny; ++j) AT(f, i, j) = AT(f, i, r->ystop);
    }
}

static int check_hill_halos(const grid_t *G, const r2d_field *f, const char *name)
{
    const dlesm_region *r = &f->fb.internal;
    const dlesm_subdomain *sb = &G->info.subdomain;
    int bad = 0, checked = 0;
#define CELL(i, j)                                                                                     \
    do {                                                                                               \
        ++checked;                                                                                     \
        if (fabs(AT(f, (i), (j)) - hill(G, f, (i), (j))) > 1.0e-8) {                                    \
            if (bad++ < 5)                                                                             \
                fprintf(stderr, "[rank %d] ERROR in halo value for %s: %d %d %.1f %.1f\n", g_rank, name, \
                        (i), (j), AT(f, (i), (j)), hill(G, f, (i), (j)));                               \
        }                                                                                              \
    } while (0)
    if (sb->global.xstart > 1)
        for (int j = r->ystart; j <= r->ystop; ++j) CELL(r->xstart - 1, j);
    if (sb->global.xstop < G->info.global_nx)
        for (int j = r->ystart; j <= r->ystop; ++j) CELL(r->xstop + 1, j);
    if (sb->global.ystart > 1)
        for (int i = r->xstart; i <= r->xstop; ++i) CELL(i, r->ystart - 1);
    if (sb->global.ystop < G->info.global_ny)
        for (int i = r->xstart; i <= r->xstop; ++i) CELL(i, r->ystop + 1);
#undef CELL
    g_fail += bad;
    return checked;
}

static void run_halos(int nx, int ny)
{
    grid_t G;
    make_grid(&G, nx, ny);
    for (int pass = 0; pass < 2; ++pass) {
        r2d_field f[4];
        long checked = 0;
        for (int q = 0; q < 4; ++q) {
            make_field(&f[q], &G, kPoints[q]);
            init_field_hill(&G, &f[q]);
            if (pass == 1) field_to_device(&f[q]);
        }
        for (int q = 0; q < 4; ++q) halo_exchange(&f[q]);
        for (int q = 0; q < 4; ++q) {
            if (pass == 1)   /* get_data() */
                f[q].read_from_device_c(f[q].device_ptr, f[q].data, 1, 1, f[q].nx, f[q].ny, true);
            checked += check_hill_halos(&G, &f[q], kNames[q]);
        }
        printf("[rank %d] %s fields: %ld halo cells checked on a %dx%d domain\n", g_rank,
               pass == 0 ? "host-resident" : "device-resident", checked, nx, ny);
        if (g_nranks > 1 && checked == 0) g_fail++;
        for (int q = 0; q < 4; ++q) free_field(&f[q]);
    }
    dlesm_grid_free(G.g);
}

int main(int argc, char **argv)
{
    const char *what = argc > 1 ? argv[1] : "model";
    int nx = argc > 3 ? atoi(argv[2]) : 4, ny = argc > 3 ? atoi(argv[3]) : 10;
    gocean_initialise();
    if (strcmp(what, "model") == 0) run_model(nx, ny);
    else if (strcmp(what, "halos") == 0) run_halos(nx, ny);
    else { fprintf(stderr, "usage: c_driver model|halos [nx ny]\n"); return 2; }
    CHECK(dlesm_dev_sync());
    if (dlesm_on_master() && g_fail == 0) printf("\nc_driver %s: all checks passed on %d rank(s)\n", what, g_nranks);
    const int fail = g_fail;
    CHECK(dlesm_parallel_finalise());   /* gocean_finalise() */
    return fail ? 1 : 0;
}
