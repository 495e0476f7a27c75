/*
 * dlesm_oracle.c -- CPU oracle for the dl_esm_inf hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's Fortran (1-based, column-major, same loop
 * order), each function citing the reference file:line it follows.  See dlesm_oracle.h
 * for the parity status (bookkeeping pinned by the reference's own tests; shallow-water
 * stencil arithmetic "parity unpinned" -- the reference ships no stencil).
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fPIC -shared (no FMA contraction, so
 * it behaves like gfortran -O1 on baseline x86-64, compiler_setup/gnu.sh:7-8).
 *
 * src/X = /root/reference/finite_difference/src/X
 */
#include "dlesm_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* Fortran (i,j) with lower bounds 1 and leading extent nx */
#define IDX(nx, i, j) (((size_t)(j) - 1) * (size_t)(nx) + ((size_t)(i) - 1))

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* ======================================================================== */
/* parallel_mod                                                             */
/* ======================================================================== */

/* src/parallel_mod.f90:103-105: "integer :: hwidth = 1" is initialised in its
 * declaration, so it has the SAVE attribute and persists across calls. */
static int saved_hwidth = 1;

void orc_reset_saved_state(void) { saved_hwidth = 1; }

/* src/parallel_mod.f90:70-332 */
int orc_go_decompose(int domainx, int domainy, int ndomains, int ndomainx, int ndomainy,
                     int halo_width, int nranks, orc_decomp *decomp)
{
    int ndom, auto_tile;
    int ntilex, ntiley, ierr;
    int internal_width, internal_height, nwidth;
    int junder, iunder, iunder_row, height, width;
    int ith, ival, jval, ji, jj;
    int xlen, ylen, domperrank;
    orc_subdomain *subdomain = NULL;

    /* :112-126 optional arguments */
    if (ndomains < 0) {
        if (ndomainx < 0 && ndomainy < 0) {
            ndom = nranks;
            auto_tile = 1;
        } else if (ndomainx >= 0 && ndomainy >= 0) {
            ndom = ndomainx * ndomainy;
            auto_tile = 0;
        } else {
            return ORC_ERR_ARGS;
        }
    } else {
        ndom = ndomains;
        auto_tile = 1;
    }

    /* :128-131 */
    if (nranks < 1) return ORC_ERR_NRANKS;

    /* :133-139 */
    if (halo_width >= 0) {
        if (halo_width < 1 && nranks > 1) return ORC_ERR_HALO_WIDTH;
        saved_hwidth = halo_width;
    }

    if (ndom > ORC_MAX_DOMAINS || ndom < 1) return ORC_ERR_TOO_MANY;

    /* :141-152 rank -> subdomain map; ceiling(real(ndom)/real(nranks)) */
    domperrank = (int)ceilf((float)ndom / (float)nranks);
    if ((long)domperrank * nranks > ORC_MAX_DOMAINS) return ORC_ERR_TOO_MANY;
    decomp->domperrank = domperrank;
    decomp->nranks = nranks;
    for (ji = 0; ji < domperrank * nranks; ++ji) decomp->proc_subdomains[ji] = -1;
    ji = 0;
    for (ith = 1; ith <= nranks; ++ith) {
        for (jj = 1; jj <= domperrank; ++jj) {
            ji = ji + 1;
            if (ji > ndom) break;
            decomp->proc_subdomains[(size_t)(ith - 1) * domperrank + (jj - 1)] = ji;
        }
    }

    /* :154-165 */
    decomp->global_nx = domainx;
    decomp->global_ny = domainy;
    decomp->ndomains = ndom;
    xlen = domainx;
    ylen = domainy;

    /* :167-194 */
    if (auto_tile) {
        ntilex = (int)sqrtf((float)ndom);
        while (ndom % ntilex != 0) ntilex = ntilex - 1;
        ntiley = ndom / ntilex;
        if (xlen > ylen) {
            if (ntilex < ntiley) {
                ierr = ntiley;
                ntiley = ntilex;
                ntilex = ierr;
            }
        } else {
            if (ntiley < ntilex) {
                ierr = ntiley;
                ntiley = ntilex;
                ntilex = ierr;
            }
        }
    } else {
        ntilex = ndomainx;
        ntiley = ndomainy;
    }
    decomp->nx = ntilex;
    decomp->ny = ntiley;

    /* :204-223 */
    internal_width = xlen / ntilex;
    internal_height = ylen / ntiley;
    nwidth = ntiley * internal_height;
    junder = (nwidth < ylen) ? ylen - nwidth : 0;
    nwidth = ntilex * internal_width;
    iunder = (nwidth < xlen) ? xlen - nwidth : 0;

    /* :231-317 */
    ith = 1;
    jval = 1;
    decomp->max_width = 0;
    decomp->max_height = 0;
    for (jj = 1; jj <= ntiley; ++jj) {
        if (junder > 0) {
            height = internal_height + 1;
            junder = junder - 1;
        } else {
            height = internal_height;
        }
        ival = 1;
        iunder_row = iunder;
        for (ji = 1; ji <= ntilex; ++ji) {
            if (iunder_row > 0) {
                width = internal_width + 1;
                iunder_row = iunder_row - 1;
            } else {
                width = internal_width;
            }
            subdomain = &decomp->subdomains[ith - 1];
            subdomain->internal.xstart = saved_hwidth + 1;
            subdomain->internal.xstop = subdomain->internal.xstart + width - 1;
            subdomain->internal.nx = width;
            subdomain->global.xstart = ival;
            subdomain->global.xstop = subdomain->global.xstart + width - 1;
            subdomain->global.nx = 2 * saved_hwidth + width;

            subdomain->internal.ystart = saved_hwidth + 1;
            subdomain->internal.ystop = subdomain->internal.ystart + height - 1;
            subdomain->internal.ny = height;
            subdomain->global.ystart = jval;
            subdomain->global.ystop = subdomain->global.ystart + height - 1;
            subdomain->global.ny = 2 * saved_hwidth + subdomain->internal.ny;

            decomp->max_width = imax(decomp->max_width, subdomain->global.nx);
            decomp->max_height = imax(decomp->max_height, subdomain->global.ny);

            ival = subdomain->global.xstop + 1;
            ith = ith + 1;
        }
        jval = subdomain->global.ystop + 1;
    }
    return ORC_OK;
}

/* ======================================================================== */
/* grid_mod                                                                 */
/* ======================================================================== */

/* src/grid_mod.f90:344-363: get_environment_variable into character(len=3), then
 * read(strvalue,"(i3)").  Status 1 = not present, -1 = truncated. */
int orc_parse_alignment(const char *envval, int *alignment)
{
    size_t len, k;
    int value = 0, seen_digit = 0, sign = 1;
    *alignment = 1;
    if (envval == NULL) return ORC_OK;            /* not present => ALIGNMENT = 1 */
    len = strlen(envval);
    if (len > 3) return ORC_ERR_ALIGN_DIGITS;     /* did not fit in strvalue */
    /* Fortran I3 edit descriptor on a blank-padded 3-char field: blanks ignored */
    for (k = 0; k < len; ++k) {
        char c = envval[k];
        if (c == ' ') continue;
        if ((c == '+' || c == '-') && !seen_digit && value == 0) {
            if (c == '-') sign = -1;
            continue;
        }
        if (c < '0' || c > '9') return ORC_ERR_ALIGN_VALUE;
        value = value * 10 + (c - '0');
        seen_digit = 1;
    }
    /* an all-blank field reads as 0 under the default BLANK=NULL mode => < 1 => stop */
    value *= sign;
    if (value < 1) return ORC_ERR_ALIGN_VALUE;
    *alignment = value;
    return ORC_OK;
}

/* src/grid_mod.f90:250-278 */
int orc_grid_constructor_check(int grid_name, int grid_offset)
{
    if (grid_name != ORC_ARAKAWA_C && grid_name != ORC_ARAKAWA_B) return ORC_ERR_GRID_NAME;
    if (grid_offset != ORC_OFFSET_NE && grid_offset != ORC_OFFSET_NW &&
        grid_offset != ORC_OFFSET_SE && grid_offset != ORC_OFFSET_SW)
        return ORC_ERR_GRID_OFFSET;
    return ORC_OK;
}

/* src/grid_mod.f90:364-385 */
int orc_grid_dims_compute(const orc_subdomain *sub, int alignment, orc_grid_dims *out)
{
    int padding = alignment - (sub->global.nx % alignment);
    out->padding = padding;
    out->nx = sub->global.nx + padding;
    if (out->nx % alignment != 0) return ORC_ERR_ALIGN_VALUE;
    out->ny = sub->global.ny + 1;
    return ORC_OK;
}

/* src/grid_mod.f90:330-570 */
int orc_grid_init(const orc_subdomain *sub, const int bcs[3], int nranks, int dist_mem_enabled,
                  int alignment, double dxarg, double dyarg,
                  const int *tmask_in, int tmask_in_nx,
                  orc_grid_dims *dims, int *tmask, double *m[16])
{
    int ji, jj, k, rc;
    int xstart, xstop, ystart, ystop;
    int nx, ny;
    double *dx_t, *dy_t, *dx_u, *dy_u, *dx_v, *dy_v, *dx_f, *dy_f;
    double *area_t, *area_u, *area_v, *gphiu, *gphiv, *gphif, *xt, *yt;

    rc = orc_grid_dims_compute(sub, alignment, dims);
    if (rc != ORC_OK) return rc;
    nx = dims->nx;
    ny = dims->ny;

    /* :387-391 */
    xstart = sub->internal.xstart;
    xstop = sub->internal.xstop;
    ystart = sub->internal.ystart;
    ystop = sub->internal.ystop;

    if (tmask_in != NULL) {
        if (tmask) {
            /* :407-412 copy-in */
            for (jj = ystart - 1; jj <= ystop + 1; ++jj)
                for (ji = xstart - 1; ji <= xstop + 1; ++ji)
                    tmask[IDX(nx, ji, jj)] = tmask_in[IDX(tmask_in_nx, ji, jj)];
            /* :416-419 North */
            for (jj = ystop + 2; jj <= ny; ++jj)
                for (ji = 1; ji <= nx; ++ji)
                    tmask[IDX(nx, ji, jj)] = tmask[IDX(nx, ji, ystop + 1)];
            /* :420-423 South */
            for (jj = 1; jj <= ystart - 2; ++jj)
                for (ji = 1; ji <= nx; ++ji)
                    tmask[IDX(nx, ji, jj)] = tmask[IDX(nx, ji, ystart - 1)];
            /* :424-427 "East" (low x) */
            for (ji = 1; ji <= xstart - 2; ++ji)
                for (jj = 1; jj <= ny; ++jj)
                    tmask[IDX(nx, ji, jj)] = tmask[IDX(nx, xstart - 1, jj)];
            /* :428-431 "West" (high x; fills the alignment padding) */
            for (ji = xstop + 2; ji <= nx; ++ji)
                for (jj = 1; jj <= ny; ++jj)
                    tmask[IDX(nx, ji, jj)] = tmask[IDX(nx, xstop + 1, jj)];
        }
    } else {
        /* :437-442 */
        if (nranks > 1 && (bcs[0] == ORC_BC_PERIODIC || bcs[1] == ORC_BC_PERIODIC))
            return ORC_ERR_PBC_DISTMEM;
        /* :448-452 all-wet box only; everything outside stays as the caller left it */
        if (tmask) {
            for (jj = ystart - 1; jj <= ystop + 1; ++jj)
                for (ji = xstart - 1; ji <= xstop + 1; ++ji)
                    tmask[IDX(nx, ji, jj)] = 1;
        }
    }

    /* :457-459 */
    dims->dx = dxarg;
    dims->dy = dyarg;

    if (m != NULL) {
        dx_t = m[0]; dy_t = m[1]; dx_u = m[2]; dy_u = m[3]; dx_v = m[4]; dy_v = m[5];
        dx_f = m[6]; dy_f = m[7]; area_t = m[8]; area_u = m[9]; area_v = m[10];
        gphiu = m[11]; gphiv = m[12]; gphif = m[13]; xt = m[14]; yt = m[15];

        /* :481-495 */
        for (jj = 1; jj <= ny; ++jj)
            for (ji = 1; ji <= nx; ++ji) {
                size_t q = IDX(nx, ji, jj);
                if (dx_t) dx_t[q] = dxarg;
                if (dy_t) dy_t[q] = dyarg;
                if (dx_u) dx_u[q] = dxarg;
                if (dy_u) dy_u[q] = dyarg;
                if (dx_v) dx_v[q] = dxarg;
                if (dy_v) dy_v[q] = dyarg;
                if (dx_f) dx_f[q] = dxarg;
                if (dy_f) dy_f[q] = dyarg;
            }
        /* :501-509 (dx_* and dy_* are the constants just stored) */
        for (jj = 1; jj <= ny; ++jj)
            for (ji = 1; ji <= nx; ++ji) {
                size_t q = IDX(nx, ji, jj);
                if (area_t) area_t[q] = dxarg * dyarg;
                if (area_u) area_u[q] = dxarg * dyarg;
                if (area_v) area_v[q] = dxarg * dyarg;
            }
        /* :516-522 */
        for (jj = 1; jj <= ny; ++jj)
            for (ji = 1; ji <= nx; ++ji) {
                size_t q = IDX(nx, ji, jj);
                if (gphiu) gphiu[q] = 50.0;
                if (gphiv) gphiv[q] = 50.0;
                if (gphif) gphif[q] = 50.0;
            }
        /* :529-556 T-point coordinates by running sums */
        if (xt) {
            for (k = 0; k < nx * ny; ++k) xt[k] = 0.0;
            for (jj = 1; jj <= ny; ++jj)
                xt[IDX(nx, xstart, jj)] = sub->global.xstart * dxarg;
            for (ji = xstart + 1; ji <= nx; ++ji)
                for (jj = 1; jj <= ny; ++jj)
                    xt[IDX(nx, ji, jj)] = xt[IDX(nx, ji - 1, jj)] + dxarg;
            for (ji = xstart - 1; ji >= 1; --ji)
                for (jj = 1; jj <= ny; ++jj)
                    xt[IDX(nx, ji, jj)] = xt[IDX(nx, ji + 1, jj)] - dxarg;
        }
        if (yt) {
            for (k = 0; k < nx * ny; ++k) yt[k] = 0.0;
            for (ji = 1; ji <= nx; ++ji)
                yt[IDX(nx, ji, ystart)] = sub->global.ystart * dyarg;
            for (jj = ystart + 1; jj <= ny; ++jj)
                for (ji = 1; ji <= nx; ++ji)
                    yt[IDX(nx, ji, jj)] = yt[IDX(nx, ji, jj - 1)] + dyarg;
            for (jj = ystart - 1; jj >= 1; --jj)
                for (ji = 1; ji <= nx; ++ji)
                    yt[IDX(nx, ji, jj)] = yt[IDX(nx, ji, jj + 1)] - dyarg;
        }
    }

    /* :559-564 */
    if (dist_mem_enabled && (bcs[0] == ORC_BC_PERIODIC || bcs[1] == ORC_BC_PERIODIC))
        return ORC_ERR_PBC_DISTMEM;
    /* :566 map_comms is called separately through orc_map_comms */
    return ORC_OK;
}

/* ======================================================================== */
/* field_mod                                                                */
/* ======================================================================== */

#define NBOUNDARY 1 /* src/field_mod.f90:227 */

/* src/field_mod.f90:1394-1464 */
static void init_periodic_bc_halos(const int bcs[3], orc_field_bounds *fld)
{
    int ihalo = 0;
    fld->num_halos = 0;
    if (bcs[0] == ORC_BC_PERIODIC) fld->num_halos += 2;
    if (bcs[1] == ORC_BC_PERIODIC) fld->num_halos += 2;

    if (bcs[0] == ORC_BC_PERIODIC) {
        /* E-most column set to W-most internal column */
        orc_halo *h = &fld->halo[ihalo++];
        h->dest.xstart = fld->internal.xstop + 1;
        h->dest.xstop = fld->internal.xstop + 1;
        h->dest.ystart = fld->internal.ystart;
        h->dest.ystop = fld->internal.ystop;
        h->source.xstart = fld->internal.xstart;
        h->source.xstop = fld->internal.xstart;
        h->source.ystart = fld->internal.ystart;
        h->source.ystop = fld->internal.ystop;
        /* W-most column set to E-most internal column */
        h = &fld->halo[ihalo++];
        h->dest.xstart = fld->internal.xstart - 1;
        h->dest.xstop = fld->internal.xstart - 1;
        h->dest.ystart = fld->internal.ystart;
        h->dest.ystop = fld->internal.ystop;
        h->source.xstart = fld->internal.xstop;
        h->source.xstop = fld->internal.xstop;
        h->source.ystart = fld->internal.ystart;
        h->source.ystop = fld->internal.ystop;
    }
    if (bcs[1] == ORC_BC_PERIODIC) {
        /* N-most row set to S-most internal row */
        orc_halo *h = &fld->halo[ihalo++];
        h->dest.xstart = fld->internal.xstart - 1;
        h->dest.xstop = fld->internal.xstop + 1;
        h->dest.ystart = fld->internal.ystop + 1;
        h->dest.ystop = fld->internal.ystop + 1;
        h->source.xstart = fld->internal.xstart - 1;
        h->source.xstop = fld->internal.xstop + 1;
        h->source.ystart = fld->internal.ystart;
        h->source.ystop = fld->internal.ystart;
        /* S-most row set to N-most internal row */
        h = &fld->halo[ihalo++];
        h->dest.xstart = fld->internal.xstart - 1;
        h->dest.xstop = fld->internal.xstop + 1;
        h->dest.ystart = fld->internal.ystart - 1;
        h->dest.ystop = fld->internal.ystart - 1;
        h->source.xstart = fld->internal.xstart - 1;
        h->source.xstop = fld->internal.xstop + 1;
        h->source.ystart = fld->internal.ystop;
        h->source.ystop = fld->internal.ystop;
    }
    /* the halo_type region components nx, ny are never set by the reference */
}

/* src/field_mod.f90:563-1122 */
int orc_set_field_bounds(const orc_subdomain *sub, int grid_nx, int grid_ny, int grid_offset,
                         const int bcs[3], int grid_points, orc_field_bounds *fld)
{
    const orc_region *si = &sub->internal;
    memset(fld, 0, sizeof(*fld));
    fld->num_halos = -1; /* "unset" marker */

    switch (grid_points) {
    case ORC_U_POINTS: /* cu_field_init :652-671 */
        fld->defined_on = ORC_U_POINTS;
        if (grid_offset == ORC_OFFSET_SW) { /* cu_sw_init :675-751 */
            if (bcs[0] == ORC_BC_PERIODIC) {
                fld->internal.xstart = si->xstart;
                fld->internal.xstop = si->xstop;
            } else {
                fld->internal.xstart = si->xstart + 1;
                fld->internal.xstop = si->xstop;
            }
            fld->internal.ystart = si->ystart;
            fld->internal.ystop = si->ystop;
            init_periodic_bc_halos(bcs, fld);
        } else if (grid_offset == ORC_OFFSET_NE) { /* cu_ne_init :755-786 */
            if (bcs[0] != ORC_BC_PERIODIC) {
                fld->internal.xstart = si->xstart;
                fld->internal.xstop = si->xstop;
            } else {
                return ORC_ERR_FIELD_STOP;
            }
            if (bcs[1] != ORC_BC_PERIODIC) {
                fld->internal.ystart = si->ystart;
                fld->internal.ystop = si->ystop;
            } else {
                return ORC_ERR_FIELD_STOP;
            }
            fld->num_halos = 0;
        } else {
            return ORC_ERR_FIELD_STOP;
        }
        break;
    case ORC_V_POINTS: /* cv_field_init :790-809 */
        fld->defined_on = ORC_V_POINTS;
        if (grid_offset == ORC_OFFSET_SW) { /* cv_sw_init :813-868 */
            if (bcs[1] == ORC_BC_PERIODIC) {
                fld->internal.xstart = si->xstart;
                fld->internal.xstop = si->xstop;
                fld->internal.ystart = si->ystart;
                fld->internal.ystop = si->ystop;
            } else {
                return ORC_ERR_FIELD_STOP; /* :844 */
            }
            init_periodic_bc_halos(bcs, fld);
        } else if (grid_offset == ORC_OFFSET_NE) { /* cv_ne_init :872-895 */
            if (bcs[0] != ORC_BC_PERIODIC) {
                fld->internal.xstart = si->xstart;
                fld->internal.xstop = si->xstop;
            } else {
                return ORC_ERR_FIELD_STOP;
            }
            if (bcs[1] != ORC_BC_PERIODIC) {
                fld->internal.ystart = si->ystart;
                fld->internal.ystop = si->ystop;
            } else {
                return ORC_ERR_FIELD_STOP;
            }
        } else {
            return ORC_ERR_FIELD_STOP;
        }
        break;
    case ORC_T_POINTS: /* ct_field_init :899-918 */
        fld->defined_on = ORC_T_POINTS;
        if (grid_offset == ORC_OFFSET_SW) { /* ct_sw_init :922-961 */
            fld->internal.xstart = si->xstart;
            fld->internal.xstop = si->xstop;
            fld->internal.ystart = si->ystart;
            fld->internal.ystop = si->ystop;
            init_periodic_bc_halos(bcs, fld);
        } else if (grid_offset == ORC_OFFSET_NE) { /* ct_ne_init :965-1000 */
            if (bcs[0] != ORC_BC_PERIODIC) {
                fld->internal.xstart = si->xstart;
                fld->internal.xstop = si->xstop;
            } else {
                return ORC_ERR_FIELD_STOP;
            }
            if (bcs[1] != ORC_BC_PERIODIC) {
                fld->internal.ystart = si->ystart;
                fld->internal.ystop = si->ystop;
            } else {
                return ORC_ERR_FIELD_STOP;
            }
        } else {
            return ORC_ERR_FIELD_STOP;
        }
        break;
    case ORC_F_POINTS: /* cf_field_init :1004-1023 */
        fld->defined_on = ORC_F_POINTS;
        if (grid_offset == ORC_OFFSET_SW) { /* cf_sw_init :1027-1084 */
            if (bcs[0] == ORC_BC_PERIODIC) {
                fld->internal.xstart = si->xstart;
                fld->internal.xstop = si->xstop;
            } else {
                return ORC_ERR_FIELD_STOP; /* :1048 */
            }
            if (bcs[1] == ORC_BC_PERIODIC) {
                fld->internal.ystart = si->ystart;
                fld->internal.ystop = si->ystop;
            } else {
                return ORC_ERR_FIELD_STOP; /* :1059 */
            }
            init_periodic_bc_halos(bcs, fld);
        } else if (grid_offset == ORC_OFFSET_NE) { /* cf_ne_init :1088-1122 */
            if (bcs[0] != ORC_BC_PERIODIC) {
                fld->internal.xstart = si->xstart;
                fld->internal.xstop = si->xstop;
            } else {
                return ORC_ERR_FIELD_STOP;
            }
            if (bcs[1] != ORC_BC_PERIODIC) {
                fld->internal.ystart = si->ystart;
                fld->internal.ystop = si->ystop;
            } else {
                return ORC_ERR_FIELD_STOP;
            }
        } else {
            return ORC_ERR_FIELD_STOP;
        }
        break;
    case ORC_ALL_POINTS: /* field_init :628-648 */
        fld->defined_on = ORC_ALL_POINTS;
        fld->internal.xstart = 1;
        fld->internal.xstop = grid_nx;
        fld->internal.ystart = 1;
        fld->internal.ystop = grid_ny;
        fld->num_halos = 0;
        break;
    default:
        return ORC_ERR_POINTS; /* :591-593 */
    }

    /* :596-622 */
    fld->internal.nx = fld->internal.xstop - fld->internal.xstart + 1;
    fld->internal.ny = fld->internal.ystop - fld->internal.ystart + 1;
    fld->whole.xstart = fld->internal.xstart - NBOUNDARY;
    fld->whole.xstop = fld->internal.xstop + NBOUNDARY;
    fld->whole.ystart = fld->internal.ystart - NBOUNDARY;
    fld->whole.ystop = fld->internal.ystop + NBOUNDARY;
    fld->whole.nx = fld->whole.xstop - fld->whole.xstart + 1;
    fld->whole.ny = fld->whole.ystop - fld->whole.ystart + 1;
    return ORC_OK;
}

/* src/field_mod.f90:375-389 */
void orc_field_construct(const orc_subdomain *sub, const orc_field_bounds *fb, int nx, int ny,
                         const double *init_global_data, int gnx, double *data)
{
    size_t k, n = (size_t)nx * (size_t)ny;
    int ji, jj;
    for (k = 0; k < n; ++k) data[k] = 0.0;
    if (init_global_data != NULL) {
        int dx = sub->global.xstart - fb->internal.xstart;
        int dy = sub->global.ystart - fb->internal.ystart;
        for (jj = sub->internal.ystart; jj <= sub->internal.ystop; ++jj)
            for (ji = sub->internal.xstart; ji <= sub->internal.xstop; ++ji)
                data[IDX(nx, ji, jj)] = init_global_data[IDX(gnx, ji + dx, jj + dy)];
    }
}

/* src/field_mod.f90:1126-1133, 1152-1159 */
void orc_copy_field(const double *field_in, double *field_out, int nx, int ny)
{
    size_t k, n = (size_t)nx * (size_t)ny;
    for (k = 0; k < n; ++k) field_out[k] = field_in[k];
}

/* src/field_mod.f90:1140-1148, 1179-1187: array assignment, RHS evaluated first */
void orc_copy_patch(double *field, int nx, const orc_region *src, const orc_region *dest)
{
    int w = src->xstop - src->xstart + 1;
    int h = src->ystop - src->ystart + 1;
    int i, j;
    double *tmp;
    if (w <= 0 || h <= 0) return;
    tmp = (double *)malloc(sizeof(double) * (size_t)w * (size_t)h);
    for (j = 0; j < h; ++j)
        for (i = 0; i < w; ++i)
            tmp[(size_t)j * w + i] = field[IDX(nx, src->xstart + i, src->ystart + j)];
    for (j = 0; j < h; ++j)
        for (i = 0; i < w; ++i)
            field[IDX(nx, dest->xstart + i, dest->ystart + j)] = tmp[(size_t)j * w + i];
    free(tmp);
}

/* src/field_mod.f90:1191-1202 */
void orc_set_field(double *field, int nx, int ny, double val)
{
    size_t k, n = (size_t)nx * (size_t)ny;
    for (k = 0; k < n; ++k) field[k] = val;
}

/* src/field_mod.f90:1289-1302 (the global_sum of :1305 is the caller's business) */
double orc_array_checksum(const double *field, int nx, int xstart, int xstop, int ystart, int ystop)
{
    double val = 0.0;
    int i, j;
    for (j = ystart; j <= ystop; ++j)
        for (i = xstart; i <= xstop; ++i)
            val += fabs(field[IDX(nx, i, j)]);
    return val;
}

double orc_array_checksum_ld(const double *field, int nx, int xstart, int xstop, int ystart, int ystop)
{
    long double val = 0.0L;
    int i, j;
    for (j = ystart; j <= ystop; ++j)
        for (i = xstart; i <= xstop; ++i)
            val += fabsl((long double)field[IDX(nx, i, j)]);
    return (double)val;
}

/* SURVEY CS-6: for ihalo=1..num_halos: copy_field(fld, halo%source, halo%dest) */
void orc_apply_periodic_halos(double *field, int nx, const orc_field_bounds *fb)
{
    int ihalo;
    for (ihalo = 0; ihalo < fb->num_halos; ++ihalo)
        orc_copy_patch(field, nx, &fb->halo[ihalo].source, &fb->halo[ihalo].dest);
}

/* tests/device_computation/test_device_io.f90:70-90 (write) */
void orc_write_to_device(const double *host, double *device, int pitch,
                         int startx, int starty, int nx, int ny)
{
    size_t next_offset = ((size_t)(starty - 1) * pitch + (size_t)(startx - 1));
    int i, k;
    for (i = starty; i <= starty + ny - 1; ++i) {
        for (k = 0; k < nx; ++k)
            device[next_offset + k] = host[IDX(pitch, startx + k, i)];
        next_offset += pitch;
    }
}

/* tests/device_computation/test_device_io.f90:50-68 (read) */
void orc_read_from_device(const double *device, double *host, int pitch,
                          int startx, int starty, int nx, int ny)
{
    size_t next_offset = ((size_t)(starty - 1) * pitch + (size_t)(startx - 1));
    int i, k;
    for (i = starty; i <= starty + ny - 1; ++i) {
        for (k = 0; k < nx; ++k)
            host[IDX(pitch, startx + k, i)] = device[next_offset + k];
        next_offset += pitch;
    }
}

/* src/field_mod.f90:1313-1390 */
int orc_gather_inner_data(const orc_decomp *decomp, int nranks, const double *const *fields,
                          const int *pitches, const orc_field_bounds *bounds,
                          double *global_data)
{
    int ji, jj, rank, n, i;
    int gnx = decomp->global_nx;
    double *recv_buffer;

    if (nranks == 1) { /* :1333-1343 */
        const orc_field_bounds *fb = &bounds[0];
        int dx = fb->internal.xstart - 1;
        int dy = fb->internal.ystart - 1;
        for (jj = fb->internal.ystart; jj <= fb->internal.ystop; ++jj)
            for (ji = fb->internal.xstart; ji <= fb->internal.xstop; ++ji)
                global_data[IDX(gnx, ji - dx, jj - dy)] = fields[0][IDX(pitches[0], ji, jj)];
        return ORC_OK;
    }
    {
        /* :1348-1351 (every rank computes the same n) */
        int halo_x = bounds[0].internal.xstart - 1;
        int halo_y = bounds[0].internal.ystart - 1;
        n = (decomp->max_width - 2 * halo_x) * (decomp->max_height - 2 * halo_y);
    }
    recv_buffer = (double *)calloc((size_t)n * nranks, sizeof(double));
    if (!recv_buffer) return ORC_ERR_TOO_MANY;
    /* :1361-1371 every rank packs its send buffer; MPI_Gather concatenates by rank */
    for (rank = 1; rank <= nranks; ++rank) {
        const orc_field_bounds *fb = &bounds[rank - 1];
        i = (rank - 1) * n;
        for (jj = fb->internal.ystart; jj <= fb->internal.ystop; ++jj)
            for (ji = fb->internal.xstart; ji <= fb->internal.xstop; ++ji) {
                recv_buffer[i] = fields[rank - 1][IDX(pitches[rank - 1], ji, jj)];
                i = i + 1;
            }
    }
    /* :1373-1388 on the master */
    for (rank = 1; rank <= nranks; ++rank) {
        const orc_region *g = &decomp->subdomains[rank - 1].global;
        i = (rank - 1) * n;
        for (jj = g->ystart; jj <= g->ystop; ++jj)
            for (ji = g->xstart; ji <= g->xstop; ++ji) {
                global_data[IDX(gnx, ji, jj)] = recv_buffer[i];
                i = i + 1;
            }
    }
    free(recv_buffer);
    return ORC_OK;
}

/* ======================================================================== */
/* parallel_comms_mod                                                       */
/* ======================================================================== */

#define MAX_HALO_DEPTH 1 /* src/parallel_comms_mod.f90:48 */

/* src/parallel_comms_mod.f90:119-123 */
static const int west_[9]  = {0, 1, 0, 0, 0, 1, 0, 1, 0};
static const int east_[9]  = {0, 0, 1, 0, 0, 0, 1, 0, 1};
static const int south_[9] = {0, 0, 0, 1, 0, 1, 0, 0, 1};
static const int north_[9] = {0, 0, 0, 0, 1, 0, 1, 1, 0};

/* src/parallel_comms_mod.f90:1365-1398 */
int orc_iprocmap(const orc_decomp *decomp, int nprocp, int ia, int ja)
{
    int iproc;
    for (iproc = 1; iproc <= nprocp; ++iproc) {
        const orc_region *g = &decomp->subdomains[iproc - 1].global;
        if (g->xstart <= ia && ia <= g->xstop && g->ystart <= ja && ja <= g->ystop)
            return iproc;
    }
    return 0;
}

/* src/parallel_comms_mod.f90:1174-1267 */
static int addsend(orc_comms *c, int dir, int proc, const int *isrc, const int *jsrc,
                   const int *ides, const int *jdes, const int *nx, const int *ny)
{
    int icomm;
    if (proc < 0) return 0;
    c->nsend = c->nsend + 1;
    icomm = c->nsend;
    if (icomm > ORC_MAXCOMM) return ORC_ERR_MAXCOMM;
    c->dirsend[icomm - 1] = dir;
    c->destination[icomm - 1] = proc;
    c->isrcsend[icomm - 1] = isrc[0];
    c->jsrcsend[icomm - 1] = jsrc[0];
    c->idessend[icomm - 1] = ides[0];
    c->jdessend[icomm - 1] = jdes[0];
    c->nxsend[icomm - 1] = nx[0];
    c->nysend[icomm - 1] = ny[0];
    c->nsendp[icomm - 1] = nx[0] * ny[0];
    if (nx[0] < 1 || ny[0] < 1) return ORC_ERR_ARGS;
    return 0;
}

/* src/parallel_comms_mod.f90:1269-1363 */
static int addrecv(orc_comms *c, int dir, int proc, const int *isrc, const int *jsrc,
                   const int *ides, const int *jdes, const int *nx, const int *ny)
{
    int icomm;
    if (proc < 0) return 0;
    c->nrecv = c->nrecv + 1;
    icomm = c->nrecv;
    if (icomm > ORC_MAXCOMM) return ORC_ERR_MAXCOMM;
    c->dirrecv[icomm - 1] = dir;
    c->source[icomm - 1] = proc;
    c->isrcrecv[icomm - 1] = isrc[0];
    c->jsrcrecv[icomm - 1] = jsrc[0];
    c->idesrecv[icomm - 1] = ides[0];
    c->jdesrecv[icomm - 1] = jdes[0];
    c->nxrecv[icomm - 1] = nx[0];
    c->nyrecv[icomm - 1] = ny[0];
    c->nrecvp[icomm - 1] = nx[0] * ny[0];
    if (nx[0] < 1 || ny[0] < 1) return ORC_ERR_ARGS;
    return 0;
}

/* src/parallel_comms_mod.f90:178-1172 */
int orc_map_comms(const orc_decomp *decomp, int irank, int nprocp, int pbc,
                  const int halo_depths[2], orc_comms *c)
{
    int idirn, i1, i2, ihalo, iproc, iprocx, iprocy, j1, j2, nadd, naddmaxr, naddmaxs;
    int iinside, ioutside, ldiff0, ldiff1, imax_, imin_;
    int ielb_iproc, ieub_iproc, jelb_iproc, jeub_iproc;
    int idesr[MAX_HALO_DEPTH], jdesr[MAX_HALO_DEPTH], idess[MAX_HALO_DEPTH], jdess[MAX_HALO_DEPTH];
    int isrcr[MAX_HALO_DEPTH], jsrcr[MAX_HALO_DEPTH], isrcs[MAX_HALO_DEPTH], jsrcs[MAX_HALO_DEPTH];
    int nxr[MAX_HALO_DEPTH], nyr[MAX_HALO_DEPTH], nxs[MAX_HALO_DEPTH], nys[MAX_HALO_DEPTH];
    int addcorner, ierr, k;
    int halo_depthx, halo_depthy;
    int jelb, jeub, ielb, ieub;
    const orc_subdomain *subdomain;
#define SUBD(p) (decomp->subdomains[(p) - 1])

    halo_depthx = halo_depths[0];
    halo_depthy = halo_depths[1];
    if (halo_depthx > MAX_HALO_DEPTH || halo_depthy > MAX_HALO_DEPTH) return ORC_ERR_HALO_DEPTH;
    if (pbc) return ORC_ERR_PBC_DISTMEM;

    /* :246-261 */
    for (k = 0; k < ORC_MAXCOMM; ++k) {
        c->dirsend[k] = c->destination[k] = c->isrcsend[k] = c->jsrcsend[k] = -999;
        c->idessend[k] = c->jdessend[k] = c->nxsend[k] = c->nysend[k] = -999;
        c->dirrecv[k] = c->source[k] = c->idesrecv[k] = c->jdesrecv[k] = -999;
        c->nxrecv[k] = c->nyrecv[k] = -999;
        /* isrcrecv, jsrcrecv, nsendp, nrecvp are not reset by the reference; use the
         * same marker so that unused slots compare deterministically */
        c->isrcrecv[k] = c->jsrcrecv[k] = c->nsendp[k] = c->nrecvp[k] = -999;
    }
    c->nsend = 0;
    c->nrecv = 0;

    /* :277-283 */
    subdomain = &SUBD(irank);
    jelb = subdomain->global.ystart;
    jeub = subdomain->global.ystop;
    ielb = subdomain->global.xstart;
    ieub = subdomain->global.xstop;

    /* ---- :285-471 send in minus I (Iplus), receive what was sent in plus I (Iminus) */
    j1 = jelb;
    while (j1 <= jeub) {
        iproc = orc_iprocmap(decomp, nprocp, ielb - 1, j1);
        if (iproc > 0) {
            j2 = imin(jeub, SUBD(iproc).global.ystop);
            for (k = 0; k < MAX_HALO_DEPTH; ++k) isrcs[k] = subdomain->internal.xstart;
            for (ihalo = 1; ihalo <= halo_depthx; ++ihalo) {
                isrcr[ihalo - 1] = SUBD(iproc).internal.xstop - ihalo + 1;
                idesr[ihalo - 1] = ihalo;
                nxr[ihalo - 1] = ihalo;
                nxs[ihalo - 1] = ihalo;
            }
            for (k = 0; k < MAX_HALO_DEPTH; ++k) {
                idess[k] = SUBD(iproc).internal.xstop + 1;
                jsrcs[k] = j1 - subdomain->global.ystart + subdomain->internal.ystart;
                jdess[k] = j1 - SUBD(iproc).global.ystart + SUBD(iproc).internal.ystart;
                jdesr[k] = jsrcs[k];
                jsrcr[k] = jdess[k];
                nyr[k] = j2 - j1 + 1;
                nys[k] = nyr[k];
            }
            /* corners at the start :350-397 */
            naddmaxr = 0;
            naddmaxs = 0;
            for (ihalo = 1; ihalo <= halo_depthx; ++ihalo) {
                if (j1 - ihalo >= jelb && orc_iprocmap(decomp, nprocp, ielb - ihalo, j1) > 0)
                    naddmaxs = ihalo;
                if (j1 == jelb && orc_iprocmap(decomp, nprocp, ielb - ihalo, j1 - ihalo) == iproc)
                    naddmaxr = ihalo;
            }
            for (ihalo = 1; ihalo <= halo_depthx; ++ihalo) {
                nadd = imin(ihalo, naddmaxs);
                jdess[ihalo - 1] -= nadd;
                jsrcs[ihalo - 1] -= nadd;
                nys[ihalo - 1] += nadd;
                nadd = imin(ihalo, naddmaxr);
                jdesr[ihalo - 1] -= nadd;
                jsrcr[ihalo - 1] -= nadd;
                nyr[ihalo - 1] += nadd;
            }
            /* corners at the end :399-441 */
            naddmaxr = 0;
            naddmaxs = 0;
            for (ihalo = 1; ihalo <= halo_depthx; ++ihalo) {
                if (j2 + ihalo <= jeub && orc_iprocmap(decomp, nprocp, ielb - ihalo, j2) > 0)
                    naddmaxs = ihalo;
                if (j2 == jeub && orc_iprocmap(decomp, nprocp, ielb - ihalo, j2 + ihalo) == iproc)
                    naddmaxr = ihalo;
            }
            for (ihalo = 1; ihalo <= halo_depthx; ++ihalo) {
                nadd = imin(ihalo, naddmaxs);
                nys[ihalo - 1] += nadd;
                nadd = imin(ihalo, naddmaxr);
                nyr[ihalo - 1] += nadd;
            }
            /* :445-455 */
            ierr = addsend(c, ORC_IPLUS, iproc - 1, isrcs, jsrcs, idess, jdess, nxs, nys);
            if (ierr != 0) return ierr;
            ierr = addrecv(c, ORC_IMINUS, iproc - 1, isrcr, jsrcr, idesr, jdesr, nxr, nyr);
            if (ierr != 0) return ierr;
            j1 = j2 + 1;
        } else {
            j1 = j1 + 1;
        }
    }

    /* ---- :473-636 send in plus I (Iminus), receive what was sent in minus I (Iplus) */
    j1 = jelb;
    while (j1 <= jeub) {
        iproc = orc_iprocmap(decomp, nprocp, ieub + 1, j1);
        if (iproc > 0) {
            j2 = imin(jeub, SUBD(iproc).global.ystop);
            for (ihalo = 1; ihalo <= halo_depthx; ++ihalo) {
                isrcr[ihalo - 1] = SUBD(iproc).internal.xstart + ihalo;
                isrcs[ihalo - 1] = subdomain->internal.xstop - ihalo + 1;
                idess[ihalo - 1] = ihalo;
                nxr[ihalo - 1] = ihalo;
                nxs[ihalo - 1] = ihalo;
                idesr[ihalo - 1] = subdomain->internal.xstop + ihalo;
            }
            for (k = 0; k < MAX_HALO_DEPTH; ++k) {
                jdess[k] = j1 - SUBD(iproc).global.ystart + SUBD(iproc).internal.ystart;
                jsrcs[k] = j1 - subdomain->global.ystart + subdomain->internal.ystart;
                jdesr[k] = jsrcs[k];
                jsrcr[k] = jdess[k];
                nyr[k] = j2 - j1 + 1;
                nys[k] = nyr[k];
            }
            /* START corners :540-572 */
            naddmaxr = 0;
            naddmaxs = 0;
            for (ihalo = 1; ihalo <= halo_depthx; ++ihalo) {
                if (j1 - ihalo >= jelb && orc_iprocmap(decomp, nprocp, ieub + ihalo, j1) > 0)
                    naddmaxs = ihalo;
                if (j1 == jelb && orc_iprocmap(decomp, nprocp, ieub + ihalo, j1 - ihalo) == iproc)
                    naddmaxr = ihalo;
            }
            for (ihalo = 1; ihalo <= halo_depthx; ++ihalo) {
                nadd = imin(ihalo, naddmaxs);
                jdess[ihalo - 1] -= nadd;
                jsrcs[ihalo - 1] -= nadd;
                nys[ihalo - 1] += nadd;
                nadd = imin(ihalo, naddmaxr);
                jdesr[ihalo - 1] -= nadd;
                jsrcr[ihalo - 1] -= nadd;
                nyr[ihalo - 1] += nadd;
            }
            /* end corners :574-615 */
            naddmaxr = 0;
            naddmaxs = 0;
            for (ihalo = 1; ihalo <= halo_depthx; ++ihalo) {
                if (j2 + ihalo <= jeub && orc_iprocmap(decomp, nprocp, ieub + ihalo, j2) > 0)
                    naddmaxs = ihalo;
                if (j2 == jeub && orc_iprocmap(decomp, nprocp, ieub + ihalo, j2 + ihalo) == iproc)
                    naddmaxr = ihalo;
            }
            for (ihalo = 1; ihalo <= halo_depthx; ++ihalo) {
                nadd = imin(ihalo, naddmaxs);
                nys[ihalo - 1] += nadd;
                nadd = imin(ihalo, naddmaxr);
                nyr[ihalo - 1] += nadd;
            }
            /* :619-627 */
            ierr = addsend(c, ORC_IMINUS, iproc - 1, isrcs, jsrcs, idess, jdess, nxs, nys);
            if (ierr != 0) return ierr;
            ierr = addrecv(c, ORC_IPLUS, iproc - 1, isrcr, jsrcr, idesr, jdesr, nxr, nyr);
            if (ierr != 0) return ierr;
            j1 = j2 + 1;
        } else {
            j1 = j1 + 1;
        }
    }

    /* ---- :638-818 send in minus J (Jplus), receive what was sent in plus J (Jminus) */
    imin_ = ielb;
    imax_ = ieub;
    i1 = imin_;
    while (i1 <= imax_) {
        iproc = orc_iprocmap(decomp, nprocp, i1, jelb - 1);
        if (iproc > 0) {
            ieub_iproc = SUBD(iproc).global.xstop;
            i2 = imin(imax_, ieub_iproc);
            for (k = 0; k < MAX_HALO_DEPTH; ++k) {
                isrcs[k] = i1 - subdomain->global.xstart + subdomain->internal.xstart;
                idess[k] = i1 - SUBD(iproc).global.xstart + SUBD(iproc).internal.xstart;
                idesr[k] = isrcs[k];
                isrcr[k] = idess[k];
                nxr[k] = i2 - i1 + 1;
                nxs[k] = nxr[k];
                jsrcs[k] = subdomain->internal.ystart;
            }
            for (ihalo = 1; ihalo <= halo_depthy; ++ihalo) {
                jsrcr[ihalo - 1] = SUBD(iproc).internal.ystop - ihalo + 1;
                jdesr[ihalo - 1] = ihalo;
                nyr[ihalo - 1] = ihalo;
                nys[ihalo - 1] = ihalo;
            }
            for (k = 0; k < MAX_HALO_DEPTH; ++k) jdess[k] = SUBD(iproc).internal.ystop + 1;
            /* START corners :707-751 */
            naddmaxr = 0;
            naddmaxs = 0;
            for (ihalo = 1; ihalo <= halo_depthy; ++ihalo) {
                if (i1 - ihalo >= imin_ &&
                    orc_iprocmap(decomp, nprocp, i1 - ihalo, jelb - ihalo) > 0)
                    naddmaxs = ihalo;
                if (i1 == imin_ &&
                    orc_iprocmap(decomp, nprocp, i1 - ihalo, jelb - ihalo) == iproc)
                    naddmaxr = ihalo;
            }
            for (ihalo = 1; ihalo <= halo_depthy; ++ihalo) {
                nadd = imin(ihalo, naddmaxs);
                idess[ihalo - 1] -= nadd;
                isrcs[ihalo - 1] -= nadd;
                nxs[ihalo - 1] += nadd;
                nadd = imin(ihalo, naddmaxr);
                idesr[ihalo - 1] -= nadd;
                isrcr[ihalo - 1] -= nadd;
                nxr[ihalo - 1] += nadd;
            }
            /* end corners :753-795 */
            naddmaxr = 0;
            naddmaxs = 0;
            for (ihalo = 1; ihalo <= halo_depthy; ++ihalo) {
                if (i2 + ihalo <= imax_ && orc_iprocmap(decomp, nprocp, i2, jelb - ihalo) > 0)
                    naddmaxs = ihalo;
                if (i2 == imax_ &&
                    orc_iprocmap(decomp, nprocp, i2 + ihalo, jelb - ihalo) == iproc)
                    naddmaxr = ihalo;
            }
            for (ihalo = 1; ihalo <= halo_depthy; ++ihalo) {
                nadd = imin(ihalo, naddmaxs);
                nxs[ihalo - 1] += nadd;
                nadd = imin(ihalo, naddmaxr);
                nxr[ihalo - 1] += nadd;
            }
            /* :800-808 */
            ierr = addsend(c, ORC_JPLUS, iproc - 1, isrcs, jsrcs, idess, jdess, nxs, nys);
            if (ierr != 0) return ierr;
            ierr = addrecv(c, ORC_JMINUS, iproc - 1, isrcr, jsrcr, idesr, jdesr, nxr, nyr);
            if (ierr != 0) return ierr;
            i1 = i2 + 1;
        } else {
            i1 = i1 + 1;
        }
    }

    /* ---- :820-986 send in plus J (Jminus), receive what was sent in minus J (Jplus) */
    imin_ = ielb;
    imax_ = ieub;
    i1 = imin_;
    while (i1 <= imax_) {
        iproc = orc_iprocmap(decomp, nprocp, i1, jeub + 1);
        if (iproc > 0) {
            ieub_iproc = SUBD(iproc).global.xstop;
            i2 = imin(imax_, ieub_iproc);
            for (k = 0; k < MAX_HALO_DEPTH; ++k) {
                isrcs[k] = i1 - subdomain->global.xstart + subdomain->internal.xstart;
                idess[k] = i1 - SUBD(iproc).global.xstart + SUBD(iproc).internal.xstart;
                idesr[k] = isrcs[k];
                isrcr[k] = idess[k];
                nxr[k] = i2 - i1 + 1;
                nxs[k] = nxr[k];
                jsrcr[k] = SUBD(iproc).internal.ystart;
            }
            for (ihalo = 1; ihalo <= halo_depthy; ++ihalo) {
                jsrcs[ihalo - 1] = subdomain->internal.ystop - ihalo + 1;
                jdess[ihalo - 1] = ihalo;
                nyr[ihalo - 1] = ihalo;
                nys[ihalo - 1] = ihalo;
            }
            for (k = 0; k < MAX_HALO_DEPTH; ++k) jdesr[k] = subdomain->internal.ystop + 1;
            /* START corners :875-920 */
            naddmaxr = 0;
            naddmaxs = 0;
            for (ihalo = 1; ihalo <= halo_depthy; ++ihalo) {
                if (i1 - ihalo >= imin_ && orc_iprocmap(decomp, nprocp, i1, jeub + ihalo) > 0)
                    naddmaxs = ihalo;
                if (i1 == imin_ &&
                    orc_iprocmap(decomp, nprocp, i1 - ihalo, jeub + ihalo) == iproc)
                    naddmaxr = ihalo;
            }
            for (ihalo = 1; ihalo <= halo_depthy; ++ihalo) {
                nadd = imin(ihalo, naddmaxs);
                idess[ihalo - 1] -= nadd;
                isrcs[ihalo - 1] -= nadd;
                nxs[ihalo - 1] += nadd;
                nadd = imin(ihalo, naddmaxr);
                idesr[ihalo - 1] -= nadd;
                isrcr[ihalo - 1] -= nadd;
                nxr[ihalo - 1] += nadd;
            }
            /* end corners :922-964 */
            naddmaxr = 0;
            naddmaxs = 0;
            for (ihalo = 1; ihalo <= halo_depthy; ++ihalo) {
                if (i2 + ihalo <= imax_ && orc_iprocmap(decomp, nprocp, i2, jeub + ihalo) > 0)
                    naddmaxs = ihalo;
                if (i2 == imax_ &&
                    orc_iprocmap(decomp, nprocp, i2 + ihalo, jeub + ihalo) == iproc)
                    naddmaxr = ihalo;
            }
            for (ihalo = 1; ihalo <= halo_depthy; ++ihalo) {
                nadd = imin(ihalo, naddmaxs);
                nxs[ihalo - 1] += nadd;
                nadd = imin(ihalo, naddmaxr);
                nxr[ihalo - 1] += nadd;
            }
            /* :968-976 */
            ierr = addsend(c, ORC_JMINUS, iproc - 1, isrcs, jsrcs, idess, jdess, nxs, nys);
            if (ierr != 0) return ierr;
            ierr = addrecv(c, ORC_JPLUS, iproc - 1, isrcr, jsrcr, idesr, jdesr, nxr, nyr);
            if (ierr != 0) return ierr;
            i1 = i2 + 1;
        } else {
            i1 = i1 + 1;
        }
    }

    /* ---- :988-1170 diagonal communications */
    for (idirn = 5; idirn <= 8; ++idirn) {
        addcorner = 0;
        ioutside = west_[idirn] * (subdomain->global.xstart - 1) +
                   east_[idirn] * (subdomain->global.xstop + 1);
        iinside = west_[idirn] * subdomain->global.xstart +
                  east_[idirn] * subdomain->global.xstop;
        iprocx = orc_iprocmap(decomp, nprocp, ioutside,
                              south_[idirn] * jelb + north_[idirn] * jeub);
        iprocy = orc_iprocmap(decomp, nprocp, iinside,
                              south_[idirn] * (jelb - 1) + north_[idirn] * (jeub + 1));
        iproc = 0;
        for (ihalo = halo_depthx; ihalo >= 1; --ihalo) {
            iproc = orc_iprocmap(decomp, nprocp,
                                 ioutside - west_[idirn] * (ihalo - 1) + east_[idirn] * (ihalo - 1),
                                 south_[idirn] * (jelb - ihalo) + north_[idirn] * (jeub + ihalo));
            if (iproc > 0 && iprocx > 0 && iprocy > 0 && iproc != iprocx && iproc != iprocy) {
                ielb_iproc = SUBD(iproc).global.xstart;
                ieub_iproc = SUBD(iproc).global.xstop;
                jelb_iproc = SUBD(iproc).global.ystart;
                jeub_iproc = SUBD(iproc).global.ystop;
                if (ihalo == halo_depthx) addcorner = 1;

                ldiff0 = ielb_iproc - ieub;
                ldiff1 = ielb - ieub_iproc;
                nxs[ihalo - 1] = ihalo - east_[idirn] * (ldiff0 - 1) - west_[idirn] * (ldiff1 - 1);
                ldiff0 = jelb_iproc - jeub;
                ldiff1 = jelb - jeub_iproc;
                nys[ihalo - 1] = ihalo - north_[idirn] * (ldiff0 - 1) - south_[idirn] * (ldiff1 - 1);

                isrcs[ihalo - 1] = east_[idirn] * (subdomain->internal.xstop - ihalo + 1) +
                                   west_[idirn] * (subdomain->internal.xstart + ihalo - 1);
                jsrcs[ihalo - 1] = north_[idirn] * (subdomain->internal.ystop - ihalo + 1) +
                                   south_[idirn] * (subdomain->internal.ystart + ihalo - 1);
                idess[ihalo - 1] = west_[idirn] * (SUBD(iproc).internal.xstop + ihalo) +
                                   east_[idirn] * (SUBD(iproc).internal.xstart - ihalo);
                jdess[ihalo - 1] = south_[idirn] * (SUBD(iproc).internal.ystop + ihalo) +
                                   north_[idirn] * (SUBD(iproc).internal.ystart - ihalo);
                isrcr[ihalo - 1] = west_[idirn] * (SUBD(iproc).internal.xstop - ihalo + 1) +
                                   east_[idirn] * (SUBD(iproc).internal.xstart + ihalo - 1);
                jsrcr[ihalo - 1] = south_[idirn] * (SUBD(iproc).internal.ystop - ihalo + 1) +
                                   north_[idirn] * (SUBD(iproc).internal.ystart + ihalo - 1);
                idesr[ihalo - 1] = east_[idirn] * (subdomain->internal.xstop + ihalo) +
                                   west_[idirn] * (subdomain->internal.xstart - ihalo);
                jdesr[ihalo - 1] = north_[idirn] * (subdomain->internal.ystop + ihalo) +
                                   south_[idirn] * (subdomain->internal.ystart - ihalo);
            } else {
                isrcs[ihalo - 1] = 0; jsrcs[ihalo - 1] = 0;
                idess[ihalo - 1] = 0; jdess[ihalo - 1] = 0;
                isrcr[ihalo - 1] = 0; jsrcr[ihalo - 1] = 0;
                idesr[ihalo - 1] = 0; jdesr[ihalo - 1] = 0;
                nxs[ihalo - 1] = 0; nys[ihalo - 1] = 0;
            }
        }
        for (k = 0; k < MAX_HALO_DEPTH; ++k) {
            nxr[k] = nxs[k];
            nyr[k] = nys[k];
        }
        if (addcorner) {
            static const int opp_dirn[9] = {0, ORC_IMINUS, ORC_IPLUS, ORC_JMINUS, ORC_JPLUS,
                                            ORC_IMINUSJMINUS, ORC_IPLUSJPLUS,
                                            ORC_IMINUSJPLUS, ORC_IPLUSJMINUS};
            ierr = addsend(c, idirn, iproc - 1, isrcs, jsrcs, idess, jdess, nxs, nys);
            if (ierr != 0) return ierr;
            ierr = addrecv(c, opp_dirn[idirn], iproc - 1, isrcr, jsrcr, idesr, jdesr, nxr, nyr);
            if (ierr != 0) return ierr;
        }
    }
#undef SUBD
    return ORC_OK;
}

/* src/parallel_comms_mod.f90:1501-1855 for all ranks at once.  MPI semantics restated:
 * a send (rank r, tag = tag_orig + dirsend) matches, on the destination, the earliest
 * posted receive with source == r-1 and the same tag (dirrecv) -- non-overtaking order. */
int orc_exchange_generic_all(int nranks, double *const *b2, const int *pitches,
                             const orc_comms *comms, int comm1, int comm2, int comm3, int comm4)
{
    int enabled[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    int r, isend, irecv, i, j, ic;
    double **sendBuff;
    int *consumed; /* per destination rank, per recv slot */
    int rc = ORC_OK;

    /* :1557-1571 */
    if (comm1 > 0) enabled[comm1] = 1;
    if (comm2 > 0) enabled[comm2] = 1;
    if (comm3 > 0) enabled[comm3] = 1;
    if (comm4 > 0) enabled[comm4] = 1;
    enabled[ORC_IPLUSJPLUS] = enabled[ORC_IPLUS] && enabled[ORC_JPLUS];
    enabled[ORC_IMINUSJMINUS] = enabled[ORC_IMINUS] && enabled[ORC_JMINUS];
    enabled[ORC_IPLUSJMINUS] = enabled[ORC_IPLUS] && enabled[ORC_JMINUS];
    enabled[ORC_IMINUSJPLUS] = enabled[ORC_IMINUS] && enabled[ORC_JPLUS];

    sendBuff = (double **)calloc((size_t)nranks * ORC_MAXCOMM, sizeof(double *));
    consumed = (int *)calloc((size_t)nranks * ORC_MAXCOMM, sizeof(int));

    /* pack every send first (all ranks), :1637-1691; j outer, i inner */
    for (r = 0; r < nranks; ++r) {
        const orc_comms *c = &comms[r];
        for (isend = 0; isend < c->nsend; ++isend) {
            if (enabled[c->dirsend[isend]] && c->destination[isend] >= 0 && c->nxsend[isend] > 0) {
                int istart = c->isrcsend[isend], iend = istart + c->nxsend[isend] - 1;
                int jstart = c->jsrcsend[isend], jend = jstart + c->nysend[isend] - 1;
                double *buf = (double *)malloc(sizeof(double) * (size_t)c->nsendp[isend]);
                ic = 0;
                for (j = jstart; j <= jend; ++j)
                    for (i = istart; i <= iend; ++i)
                        buf[ic++] = b2[r][IDX(pitches[r], i, j)];
                sendBuff[(size_t)r * ORC_MAXCOMM + isend] = buf;
            }
        }
    }
    /* deliver + unpack, :1769-1837 */
    for (r = 0; r < nranks; ++r) {
        const orc_comms *c = &comms[r];
        for (isend = 0; isend < c->nsend; ++isend) {
            double *buf = sendBuff[(size_t)r * ORC_MAXCOMM + isend];
            int dest, found = -1;
            const orc_comms *d;
            if (!buf) continue;
            dest = c->destination[isend];
            d = &comms[dest];
            for (irecv = 0; irecv < d->nrecv; ++irecv) {
                if (consumed[(size_t)dest * ORC_MAXCOMM + irecv]) continue;
                if (!(enabled[d->dirrecv[irecv]] && d->source[irecv] >= 0 && d->nxrecv[irecv] > 0))
                    continue;
                if (d->source[irecv] == r && d->dirrecv[irecv] == c->dirsend[isend]) {
                    found = irecv;
                    break;
                }
            }
            if (found < 0 || d->nrecvp[found] != c->nsendp[isend]) {
                rc = ORC_ERR_ARGS; /* an MPI run would hang or truncate here */
                continue;
            }
            consumed[(size_t)dest * ORC_MAXCOMM + found] = 1;
            {
                int jstart = d->jdesrecv[found], jend = jstart + d->nyrecv[found] - 1;
                int istart = d->idesrecv[found], iend = istart + d->nxrecv[found] - 1;
                ic = 0;
                for (j = jstart; j <= jend; ++j)
                    for (i = istart; i <= iend; ++i)
                        b2[dest][IDX(pitches[dest], i, j)] = buf[ic++];
            }
        }
    }
    /* every enabled receive must have been matched */
    for (r = 0; r < nranks; ++r) {
        const orc_comms *c = &comms[r];
        for (irecv = 0; irecv < c->nrecv; ++irecv)
            if (enabled[c->dirrecv[irecv]] && c->source[irecv] >= 0 && c->nxrecv[irecv] > 0 &&
                !consumed[(size_t)r * ORC_MAXCOMM + irecv])
                rc = ORC_ERR_ARGS;
    }
    for (r = 0; r < nranks * ORC_MAXCOMM; ++r) free(sendBuff[r]);
    free(sendBuff);
    free(consumed);
    return rc;
}
