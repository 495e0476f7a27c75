/*
 * dlesm_oracle_sw.c -- shallow-water sweeps over fd_lib fields.  TEST INFRASTRUCTURE ONLY.
 *
 * PARITY UNPINNED: the reference ships no stencil kernel; the formulas are
 * those of SURVEY.md a-S (the public GOcean/PSyclone "shallow" benchmark, Sadourny scheme)
 * and this file is the project's normative definition of them.  What IS taken from the
 * reference is everything around the arithmetic: the per-point kernel convention
 * (src/infrastructure_mod.f90:32-41), the PSy-layer loop nest over the output field's
 * internal region with jj outer / ji inner (SURVEY CS-3), the periodic halo fill by
 * copy_field over fld%halo(1..num_halos) in list order (src/field_mod.f90:1179-1187,
 * 1394-1464) and the whole-array copy_field (src/field_mod.f90:1152-1159).
 *
 * Build with -ffp-contract=off: one rounding per operation, as gfortran -O1 on x86-64.
 */
#include "dlesm_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define IDX(nx, i, j) (((size_t)(j) - 1) * (size_t)(nx) + ((size_t)(i) - 1))

/* ---- mask predicates (project-defined rule, see header) ---- */
static int wet_t(const int *tm, int nx, int i, int j) { return !tm || tm[IDX(nx, i, j)] == 1; }
static int wet_u(const int *tm, int nx, int i, int j)
{
    return !tm || (tm[IDX(nx, i, j)] == 1 && tm[IDX(nx, i - 1, j)] == 1);
}
static int wet_v(const int *tm, int nx, int i, int j)
{
    return !tm || (tm[IDX(nx, i, j)] == 1 && tm[IDX(nx, i, j - 1)] == 1);
}
static int wet_f(const int *tm, int nx, int i, int j)
{
    return !tm || (tm[IDX(nx, i, j)] == 1 && tm[IDX(nx, i - 1, j)] == 1 &&
                   tm[IDX(nx, i, j - 1)] == 1 && tm[IDX(nx, i - 1, j - 1)] == 1);
}

/* cu(i,j) = 0.5*(p(i,j)+p(i-1,j))*u(i,j) on U points */
void orc_compute_cu(double *cu, const double *p, const double *u, int nx, const orc_region *r,
                    const int *tmask)
{
    int ji, jj;
    for (jj = r->ystart; jj <= r->ystop; ++jj)
        for (ji = r->xstart; ji <= r->xstop; ++ji) {
            if (!wet_u(tmask, nx, ji, jj)) { cu[IDX(nx, ji, jj)] = 0.0; continue; }
            cu[IDX(nx, ji, jj)] = 0.5 * (p[IDX(nx, ji, jj)] + p[IDX(nx, ji - 1, jj)]) * u[IDX(nx, ji, jj)];
        }
}

/* cv(i,j) = 0.5*(p(i,j)+p(i,j-1))*v(i,j) on V points */
void orc_compute_cv(double *cv, const double *p, const double *v, int nx, const orc_region *r,
                    const int *tmask)
{
    int ji, jj;
    for (jj = r->ystart; jj <= r->ystop; ++jj)
        for (ji = r->xstart; ji <= r->xstop; ++ji) {
            if (!wet_v(tmask, nx, ji, jj)) { cv[IDX(nx, ji, jj)] = 0.0; continue; }
            cv[IDX(nx, ji, jj)] = 0.5 * (p[IDX(nx, ji, jj)] + p[IDX(nx, ji, jj - 1)]) * v[IDX(nx, ji, jj)];
        }
}

/* z(i,j) = (fsdx*(v(i,j)-v(i-1,j)) - fsdy*(u(i,j)-u(i,j-1))) /
 *          (p(i-1,j-1)+p(i,j-1)+p(i,j)+p(i-1,j))   on F points */
void orc_compute_z(double *z, const double *p, const double *u, const double *v, int nx,
                   const orc_region *r, double dx, double dy, const int *tmask)
{
    const double fsdx = 4.0 / dx, fsdy = 4.0 / dy;
    int ji, jj;
    for (jj = r->ystart; jj <= r->ystop; ++jj)
        for (ji = r->xstart; ji <= r->xstop; ++ji) {
            if (!wet_f(tmask, nx, ji, jj)) { z[IDX(nx, ji, jj)] = 0.0; continue; }
            z[IDX(nx, ji, jj)] =
                (fsdx * (v[IDX(nx, ji, jj)] - v[IDX(nx, ji - 1, jj)]) -
                 fsdy * (u[IDX(nx, ji, jj)] - u[IDX(nx, ji, jj - 1)])) /
                (p[IDX(nx, ji - 1, jj - 1)] + p[IDX(nx, ji, jj - 1)] + p[IDX(nx, ji, jj)] +
                 p[IDX(nx, ji - 1, jj)]);
        }
}

/* h(i,j) = p(i,j)+0.25*(u(i+1,j)**2+u(i,j)**2+v(i,j+1)**2+v(i,j)**2) on T points */
void orc_compute_h(double *h, const double *p, const double *u, const double *v, int nx,
                   const orc_region *r, const int *tmask)
{
    int ji, jj;
    for (jj = r->ystart; jj <= r->ystop; ++jj)
        for (ji = r->xstart; ji <= r->xstop; ++ji) {
            if (!wet_t(tmask, nx, ji, jj)) { h[IDX(nx, ji, jj)] = 0.0; continue; }
            h[IDX(nx, ji, jj)] =
                p[IDX(nx, ji, jj)] +
                0.25 * (u[IDX(nx, ji + 1, jj)] * u[IDX(nx, ji + 1, jj)] +
                        u[IDX(nx, ji, jj)] * u[IDX(nx, ji, jj)] +
                        v[IDX(nx, ji, jj + 1)] * v[IDX(nx, ji, jj + 1)] +
                        v[IDX(nx, ji, jj)] * v[IDX(nx, ji, jj)]);
        }
}

/* unew(i,j) = uold(i,j) + tdts8*(z(i,j+1)+z(i,j))*(cv(i,j+1)+cv(i-1,j+1)+cv(i-1,j)+cv(i,j))
 *             - tdtsdx*(h(i,j)-h(i-1,j)) */
void orc_compute_unew(double *unew, const double *uold, const double *z, const double *cv,
                      const double *h, int nx, const orc_region *r, double tdt, double dx,
                      const int *tmask, const double *ucur)
{
    const double tdts8 = tdt / 8.0, tdtsdx = tdt / dx;
    int ji, jj;
    for (jj = r->ystart; jj <= r->ystop; ++jj)
        for (ji = r->xstart; ji <= r->xstop; ++ji) {
            if (!wet_u(tmask, nx, ji, jj)) { unew[IDX(nx, ji, jj)] = ucur[IDX(nx, ji, jj)]; continue; }
            unew[IDX(nx, ji, jj)] =
                uold[IDX(nx, ji, jj)] +
                tdts8 * (z[IDX(nx, ji, jj + 1)] + z[IDX(nx, ji, jj)]) *
                    (cv[IDX(nx, ji, jj + 1)] + cv[IDX(nx, ji - 1, jj + 1)] +
                     cv[IDX(nx, ji - 1, jj)] + cv[IDX(nx, ji, jj)]) -
                tdtsdx * (h[IDX(nx, ji, jj)] - h[IDX(nx, ji - 1, jj)]);
        }
}

/* vnew(i,j) = vold(i,j) - tdts8*(z(i+1,j)+z(i,j))*(cu(i+1,j)+cu(i,j)+cu(i,j-1)+cu(i+1,j-1))
 *             - tdtsdy*(h(i,j)-h(i,j-1)) */
void orc_compute_vnew(double *vnew, const double *vold, const double *z, const double *cu,
                      const double *h, int nx, const orc_region *r, double tdt, double dy,
                      const int *tmask, const double *vcur)
{
    const double tdts8 = tdt / 8.0, tdtsdy = tdt / dy;
    int ji, jj;
    for (jj = r->ystart; jj <= r->ystop; ++jj)
        for (ji = r->xstart; ji <= r->xstop; ++ji) {
            if (!wet_v(tmask, nx, ji, jj)) { vnew[IDX(nx, ji, jj)] = vcur[IDX(nx, ji, jj)]; continue; }
            vnew[IDX(nx, ji, jj)] =
                vold[IDX(nx, ji, jj)] -
                tdts8 * (z[IDX(nx, ji + 1, jj)] + z[IDX(nx, ji, jj)]) *
                    (cu[IDX(nx, ji + 1, jj)] + cu[IDX(nx, ji, jj)] + cu[IDX(nx, ji, jj - 1)] +
                     cu[IDX(nx, ji + 1, jj - 1)]) -
                tdtsdy * (h[IDX(nx, ji, jj)] - h[IDX(nx, ji, jj - 1)]);
        }
}

/* pnew(i,j) = pold(i,j) - tdtsdx*(cu(i+1,j)-cu(i,j)) - tdtsdy*(cv(i,j+1)-cv(i,j)) */
void orc_compute_pnew(double *pnew, const double *pold, const double *cu, const double *cv,
                      int nx, const orc_region *r, double tdt, double dx, double dy,
                      const int *tmask, const double *pcur)
{
    const double tdtsdx = tdt / dx, tdtsdy = tdt / dy;
    int ji, jj;
    for (jj = r->ystart; jj <= r->ystop; ++jj)
        for (ji = r->xstart; ji <= r->xstop; ++ji) {
            if (!wet_t(tmask, nx, ji, jj)) { pnew[IDX(nx, ji, jj)] = pcur[IDX(nx, ji, jj)]; continue; }
            pnew[IDX(nx, ji, jj)] = pold[IDX(nx, ji, jj)] -
                                    tdtsdx * (cu[IDX(nx, ji + 1, jj)] - cu[IDX(nx, ji, jj)]) -
                                    tdtsdy * (cv[IDX(nx, ji, jj + 1)] - cv[IDX(nx, ji, jj)]);
        }
}

/* fold(i,j) = f(i,j) + alpha*(fnew(i,j) - 2*f(i,j) + fold(i,j)) */
void orc_time_smooth(const double *f, const double *fnew, double *fold, int nx,
                     const orc_region *r, double alpha, const int *tmask, int points)
{
    int ji, jj;
    for (jj = r->ystart; jj <= r->ystop; ++jj)
        for (ji = r->xstart; ji <= r->xstop; ++ji) {
            int wet = points == ORC_U_POINTS ? wet_u(tmask, nx, ji, jj)
                    : points == ORC_V_POINTS ? wet_v(tmask, nx, ji, jj)
                                             : wet_t(tmask, nx, ji, jj);
            if (!wet) continue;
            fold[IDX(nx, ji, jj)] =
                f[IDX(nx, ji, jj)] +
                alpha * (fnew[IDX(nx, ji, jj)] - 2.0 * f[IDX(nx, ji, jj)] + fold[IDX(nx, ji, jj)]);
        }
}

/* ---- initial conditions (classic "shallow": SURVEY 8d C2) ---- */
/* gi,gj: 1-based global T-point indices; M,N: global extents */
static void sw_ic_point(int gi, int gj, int M, int N, double dx, double dy, double *u, double *v,
                        double *p)
{
    const double a = 1.0e6, pi = 3.14159265358979323846;
    const double di = 2.0 * pi / (double)M, dj = 2.0 * pi / (double)N;
    const double el = (double)N * dx;
    const double pcf = pi * pi * a * a / (el * el);
    /* psi(I,J) = a*sin((I-.5)di)*sin((J-.5)dj) */
    const double s_i = sin(((double)gi - 0.5) * di), s_i1 = sin(((double)gi + 0.5) * di);
    const double s_j = sin(((double)gj - 0.5) * dj), s_j1 = sin(((double)gj + 0.5) * dj);
    const double psi_ij = a * s_i * s_j, psi_ij1 = a * s_i * s_j1, psi_i1j = a * s_i1 * s_j;
    *p = pcf * (cos(2.0 * ((double)gi - 1.0) * di) + cos(2.0 * ((double)gj - 1.0) * dj)) + 50000.0;
    *u = -(psi_ij1 - psi_ij) / dy;
    *v = (psi_i1j - psi_ij) / dx;
}

void orc_sw_init(double *u, double *v, double *p, int nx, int ny, const orc_region *t,
                 double dx, double dy, const orc_field_bounds *fb_u,
                 const orc_field_bounds *fb_v, const orc_field_bounds *fb_t)
{
    int ji, jj;
    int M = t->xstop - t->xstart + 1, N = t->ystop - t->ystart + 1;
    (void)ny;
    for (jj = t->ystart; jj <= t->ystop; ++jj)
        for (ji = t->xstart; ji <= t->xstop; ++ji)
            sw_ic_point(ji - t->xstart + 1, jj - t->ystart + 1, M, N, dx, dy,
                        &u[IDX(nx, ji, jj)], &v[IDX(nx, ji, jj)], &p[IDX(nx, ji, jj)]);
    orc_apply_periodic_halos(u, nx, fb_u);
    orc_apply_periodic_halos(v, nx, fb_v);
    orc_apply_periodic_halos(p, nx, fb_t);
}

/* one step, un-fused, serial periodic BCs */
void orc_sw_step(double *f[13], const orc_sw_params *prm, const orc_field_bounds *fb_u,
                 const orc_field_bounds *fb_v, const orc_field_bounds *fb_t,
                 const orc_field_bounds *fb_f, const int *tmask, int first)
{
    double *u = f[0], *v = f[1], *p = f[2], *uold = f[3], *vold = f[4], *pold = f[5];
    double *unew = f[6], *vnew = f[7], *pnew = f[8], *cu = f[9], *cv = f[10], *z = f[11], *h = f[12];
    const int nx = prm->nx, ny = prm->ny;

    orc_compute_cu(cu, p, u, nx, &fb_u->internal, tmask);
    orc_compute_cv(cv, p, v, nx, &fb_v->internal, tmask);
    orc_compute_z(z, p, u, v, nx, &fb_f->internal, prm->dx, prm->dy, tmask);
    orc_compute_h(h, p, u, v, nx, &fb_t->internal, tmask);

    orc_apply_periodic_halos(cu, nx, fb_u);
    orc_apply_periodic_halos(cv, nx, fb_v);
    orc_apply_periodic_halos(z, nx, fb_f);
    orc_apply_periodic_halos(h, nx, fb_t);

    orc_compute_unew(unew, uold, z, cv, h, nx, &fb_u->internal, prm->tdt, prm->dx, tmask, u);
    orc_compute_vnew(vnew, vold, z, cu, h, nx, &fb_v->internal, prm->tdt, prm->dy, tmask, v);
    orc_compute_pnew(pnew, pold, cu, cv, nx, &fb_t->internal, prm->tdt, prm->dx, prm->dy, tmask, p);

    orc_apply_periodic_halos(unew, nx, fb_u);
    orc_apply_periodic_halos(vnew, nx, fb_v);
    orc_apply_periodic_halos(pnew, nx, fb_t);

    if (!first) {
        orc_time_smooth(u, unew, uold, nx, &fb_u->internal, prm->alpha, tmask, ORC_U_POINTS);
        orc_time_smooth(v, vnew, vold, nx, &fb_v->internal, prm->alpha, tmask, ORC_V_POINTS);
        orc_time_smooth(p, pnew, pold, nx, &fb_t->internal, prm->alpha, tmask, ORC_T_POINTS);
    }
    orc_copy_field(unew, u, nx, ny);
    orc_copy_field(vnew, v, nx, ny);
    orc_copy_field(pnew, p, nx, ny);
}

/* ======================================================================== */
/* threaded dm_fd_lib-equivalent runner (bench only)                        */
/* ======================================================================== */

typedef struct sw_shared sw_shared;

typedef struct {
    sw_shared *sh;
    int rank; /* 1-based */
    orc_subdomain sub;
    int nx, ny;
    orc_field_bounds fb_u, fb_v, fb_t, fb_f;
    double *f[13];
    double *sendW, *sendE, *sendS, *sendN; /* packed edges, read by the neighbours */
    double cks[3];
    long double cks_ld[3]; /* the same sums accumulated in long double (exact to ~1e-19 relative) */
} sw_worker;

struct sw_shared {
    int nthreads, nsteps, warmup;
    int gnx, gny;
    orc_decomp *decomp;
    sw_worker *w;
    pthread_barrier_t bar;
    double t0, t1;
};

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* neighbour tile (1-based rank) with periodic wrap on the tile grid; tiles are numbered
 * row-major with x fastest (src/parallel_mod.f90:262-315) */
static int neighbour(const orc_decomp *d, int rank, int dxt, int dyt)
{
    int ix = (rank - 1) % d->nx, iy = (rank - 1) / d->nx;
    ix = (ix + dxt + d->nx) % d->nx;
    iy = (iy + dyt + d->ny) % d->ny;
    return iy * d->nx + ix + 1;
}

/* exchange_generic order (pack j outer / i inner -> copy -> unpack) with the periodic wrap:
 * E/W columns over the internal rows first, then N/S rows over xstart-1..xstop+1 so the
 * corners travel with the rows (same result as the 8-message scheme of map_comms) */
static void exchange_fields(sw_worker *me, double *const *flds, int nf)
{
    sw_shared *sh = me->sh;
    const orc_region *in = &me->sub.internal;
    int nx = me->nx, k, i, j;
    int hgt = in->ystop - in->ystart + 1, wid = in->xstop - in->xstart + 3;
    sw_worker *wW = &sh->w[neighbour(sh->decomp, me->rank, -1, 0) - 1];
    sw_worker *wE = &sh->w[neighbour(sh->decomp, me->rank, +1, 0) - 1];
    sw_worker *wS = &sh->w[neighbour(sh->decomp, me->rank, 0, -1) - 1];
    sw_worker *wN = &sh->w[neighbour(sh->decomp, me->rank, 0, +1) - 1];

    for (k = 0; k < nf; ++k)
        for (j = 0; j < hgt; ++j) {
            me->sendW[(size_t)k * hgt + j] = flds[k][IDX(nx, in->xstart, in->ystart + j)];
            me->sendE[(size_t)k * hgt + j] = flds[k][IDX(nx, in->xstop, in->ystart + j)];
        }
    pthread_barrier_wait(&sh->bar);
    for (k = 0; k < nf; ++k)
        for (j = 0; j < hgt; ++j) {
            flds[k][IDX(nx, in->xstop + 1, in->ystart + j)] = wE->sendW[(size_t)k * hgt + j];
            flds[k][IDX(nx, in->xstart - 1, in->ystart + j)] = wW->sendE[(size_t)k * hgt + j];
        }
    for (k = 0; k < nf; ++k)
        for (i = 0; i < wid; ++i) {
            me->sendS[(size_t)k * wid + i] = flds[k][IDX(nx, in->xstart - 1 + i, in->ystart)];
            me->sendN[(size_t)k * wid + i] = flds[k][IDX(nx, in->xstart - 1 + i, in->ystop)];
        }
    pthread_barrier_wait(&sh->bar);
    for (k = 0; k < nf; ++k)
        for (i = 0; i < wid; ++i) {
            flds[k][IDX(nx, in->xstart - 1 + i, in->ystop + 1)] = wN->sendS[(size_t)k * wid + i];
            flds[k][IDX(nx, in->xstart - 1 + i, in->ystart - 1)] = wS->sendN[(size_t)k * wid + i];
        }
    /* no third barrier: the next pack of sendS/sendN happens after the next call's first
     * barrier, by which time every neighbour has finished reading them */
}

static void *sw_worker_main(void *arg)
{
    sw_worker *me = (sw_worker *)arg;
    sw_shared *sh = me->sh;
    double **f = me->f;
    const int nx = me->nx, ny = me->ny;
    const double dx = 1.0e5, dy = 1.0e5, dt = 90.0, alpha = 0.001;
    double tdt = dt;
    int step, ji, jj, k;
    const orc_region *in = &me->sub.internal;
    const int serial = (sh->nthreads == 1);

    /* first-touch allocation + initial conditions from the global analytic fields */
    for (k = 0; k < 13; ++k) f[k] = (double *)calloc((size_t)nx * ny, sizeof(double));
    for (jj = in->ystart; jj <= in->ystop; ++jj)
        for (ji = in->xstart; ji <= in->xstop; ++ji)
            sw_ic_point(me->sub.global.xstart + ji - in->xstart,
                        me->sub.global.ystart + jj - in->ystart, sh->gnx, sh->gny, dx, dy,
                        &f[0][IDX(nx, ji, jj)], &f[1][IDX(nx, ji, jj)], &f[2][IDX(nx, ji, jj)]);
    if (serial) {
        orc_apply_periodic_halos(f[0], nx, &me->fb_u);
        orc_apply_periodic_halos(f[1], nx, &me->fb_v);
        orc_apply_periodic_halos(f[2], nx, &me->fb_t);
    } else {
        pthread_barrier_wait(&sh->bar);
        exchange_fields(me, f, 3);
    }
    orc_copy_field(f[0], f[3], nx, ny);
    orc_copy_field(f[1], f[4], nx, ny);
    orc_copy_field(f[2], f[5], nx, ny);

    for (step = 0; step < sh->warmup + sh->nsteps; ++step) {
        if (step == sh->warmup) {
            pthread_barrier_wait(&sh->bar);
            if (me->rank == 1) sh->t0 = now_s();
        }
        if (serial) {
            orc_sw_params prm;
            prm.nx = nx; prm.ny = ny; prm.dx = dx; prm.dy = dy; prm.dt = dt; prm.tdt = tdt;
            prm.alpha = alpha;
            orc_sw_step(f, &prm, &me->fb_u, &me->fb_v, &me->fb_t, &me->fb_f, NULL, step == 0);
        } else {
            orc_compute_cu(f[9], f[2], f[0], nx, &me->fb_u.internal, NULL);
            orc_compute_cv(f[10], f[2], f[1], nx, &me->fb_v.internal, NULL);
            orc_compute_z(f[11], f[2], f[0], f[1], nx, &me->fb_f.internal, dx, dy, NULL);
            orc_compute_h(f[12], f[2], f[0], f[1], nx, &me->fb_t.internal, NULL);
            exchange_fields(me, &f[9], 4);
            orc_compute_unew(f[6], f[3], f[11], f[10], f[12], nx, &me->fb_u.internal, tdt, dx, NULL, f[0]);
            orc_compute_vnew(f[7], f[4], f[11], f[9], f[12], nx, &me->fb_v.internal, tdt, dy, NULL, f[1]);
            orc_compute_pnew(f[8], f[5], f[9], f[10], nx, &me->fb_t.internal, tdt, dx, dy, NULL, f[2]);
            exchange_fields(me, &f[6], 3);
            if (step != 0) {
                orc_time_smooth(f[0], f[6], f[3], nx, &me->fb_u.internal, alpha, NULL, ORC_U_POINTS);
                orc_time_smooth(f[1], f[7], f[4], nx, &me->fb_v.internal, alpha, NULL, ORC_V_POINTS);
                orc_time_smooth(f[2], f[8], f[5], nx, &me->fb_t.internal, alpha, NULL, ORC_T_POINTS);
            }
            orc_copy_field(f[6], f[0], nx, ny);
            orc_copy_field(f[7], f[1], nx, ny);
            orc_copy_field(f[8], f[2], nx, ny);
        }
        if (step == 0) tdt = tdt + tdt;
    }
    pthread_barrier_wait(&sh->bar);
    if (me->rank == 1) sh->t1 = now_s();

    me->cks[0] = orc_array_checksum(f[2], nx, in->xstart, in->xstop, in->ystart, in->ystop);
    me->cks[1] = orc_array_checksum(f[0], nx, in->xstart, in->xstop, in->ystart, in->ystop);
    me->cks[2] = orc_array_checksum(f[1], nx, in->xstart, in->xstop, in->ystart, in->ystop);
    {
        static const int which[3] = {2, 0, 1};
        int q, i, j;
        for (q = 0; q < 3; ++q) {
            long double acc = 0.0L;
            for (j = in->ystart; j <= in->ystop; ++j)
                for (i = in->xstart; i <= in->xstop; ++i)
                    acc += fabsl((long double)f[which[q]][(size_t)(j - 1) * nx + (i - 1)]);
            me->cks_ld[q] = acc;
        }
    }
    pthread_barrier_wait(&sh->bar);
    for (k = 0; k < 13; ++k) free(f[k]);
    return NULL;
}

/* checksums of the last orc_sw_run_threads call accumulated in long double: at 16384^2 the
 * reference's own double-precision sum(abs(field)) is several 1e-12 off the exact value, more than
 * the tolerance the GPU result is held to, so full-size comparisons use these */
static long double g_last_cks_ld[3];
void orc_sw_last_checksums_ld(double out[3])
{
    out[0] = (double)g_last_cks_ld[0];
    out[1] = (double)g_last_cks_ld[1];
    out[2] = (double)g_last_cks_ld[2];
}

double orc_sw_run_threads(int global_nx, int global_ny, int nthreads, int nsteps, int warmup,
                          double *checksums)
{
    static const int bcs[3] = {ORC_BC_PERIODIC, ORC_BC_PERIODIC, ORC_BC_NONE};
    sw_shared sh;
    pthread_t *tid;
    orc_decomp *decomp = (orc_decomp *)malloc(sizeof(orc_decomp));
    int r, maxw = 0, maxh = 0;
    double secs = -1.0;

    if (nthreads < 1 || !decomp) { free(decomp); return -1.0; }
    orc_reset_saved_state();
    if (orc_go_decompose(global_nx, global_ny, nthreads, -1, -1, -1, nthreads, decomp) != ORC_OK) {
        free(decomp);
        return -1.0;
    }
    memset(&sh, 0, sizeof(sh));
    sh.nthreads = nthreads; sh.nsteps = nsteps; sh.warmup = warmup;
    sh.gnx = global_nx; sh.gny = global_ny; sh.decomp = decomp;
    sh.w = (sw_worker *)calloc((size_t)nthreads, sizeof(sw_worker));
    tid = (pthread_t *)calloc((size_t)nthreads, sizeof(pthread_t));
    pthread_barrier_init(&sh.bar, NULL, (unsigned)nthreads);
    for (r = 0; r < nthreads; ++r) {
        sw_worker *w = &sh.w[r];
        orc_grid_dims gd;
        w->sh = &sh;
        w->rank = r + 1;
        w->sub = decomp->subdomains[r];
        orc_grid_dims_compute(&w->sub, 1, &gd);
        w->nx = gd.nx; w->ny = gd.ny;
        orc_set_field_bounds(&w->sub, gd.nx, gd.ny, ORC_OFFSET_SW, bcs, ORC_U_POINTS, &w->fb_u);
        orc_set_field_bounds(&w->sub, gd.nx, gd.ny, ORC_OFFSET_SW, bcs, ORC_V_POINTS, &w->fb_v);
        orc_set_field_bounds(&w->sub, gd.nx, gd.ny, ORC_OFFSET_SW, bcs, ORC_T_POINTS, &w->fb_t);
        orc_set_field_bounds(&w->sub, gd.nx, gd.ny, ORC_OFFSET_SW, bcs, ORC_F_POINTS, &w->fb_f);
        if (w->sub.global.nx > maxw) maxw = w->sub.global.nx;
        if (w->sub.global.ny > maxh) maxh = w->sub.global.ny;
    }
    for (r = 0; r < nthreads; ++r) {
        sw_worker *w = &sh.w[r];
        w->sendW = (double *)calloc((size_t)4 * maxh, sizeof(double));
        w->sendE = (double *)calloc((size_t)4 * maxh, sizeof(double));
        w->sendS = (double *)calloc((size_t)4 * maxw, sizeof(double));
        w->sendN = (double *)calloc((size_t)4 * maxw, sizeof(double));
    }
    for (r = 0; r < nthreads; ++r) pthread_create(&tid[r], NULL, sw_worker_main, &sh.w[r]);
    for (r = 0; r < nthreads; ++r) pthread_join(tid[r], NULL);
    secs = sh.t1 - sh.t0;
    g_last_cks_ld[0] = g_last_cks_ld[1] = g_last_cks_ld[2] = 0.0L;
    for (r = 0; r < nthreads; ++r) {
        g_last_cks_ld[0] += sh.w[r].cks_ld[0];
        g_last_cks_ld[1] += sh.w[r].cks_ld[1];
        g_last_cks_ld[2] += sh.w[r].cks_ld[2];
    }
    if (checksums) {
        checksums[0] = checksums[1] = checksums[2] = 0.0;
        for (r = 0; r < nthreads; ++r) {
            checksums[0] += sh.w[r].cks[0];
            checksums[1] += sh.w[r].cks[1];
            checksums[2] += sh.w[r].cks[2];
        }
    }
    for (r = 0; r < nthreads; ++r) {
        free(sh.w[r].sendW); free(sh.w[r].sendE); free(sh.w[r].sendS); free(sh.w[r].sendN);
    }
    pthread_barrier_destroy(&sh.bar);
    free(sh.w); free(tid); free(decomp);
    return secs;
}
