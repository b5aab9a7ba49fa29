/*
 * oracle/healpix_oracle.c -- TEST INFRASTRUCTURE ONLY (CPU oracle).
 *
 * Plain-C restatement of the part of the healpy / healpix_cxx surface that the
 * reference's Shell runners call (Runners/HealpixRunner.py:303,306,310,312,
 * 333,334,337,429,432,433; utils/io.py:350; utils/Pixel.py:479).
 *
 * healpy (unpinned in the reference's setup.py:21 / environment.yaml:14) is a
 * third-party dependency whose source is NOT under /root/reference and which is
 * not installed in this image.  What follows restates the published HEALPix
 * RING-scheme algorithms (Gorski et al. 2005, ApJ 622, 759; healpix_cxx
 * "Healpix_Base" semantics: ring_above, ring2z, get_ring_info_small,
 * get_ring_info2, pix2loc, loc2pix, query_disc (non-inclusive), get_interpol)
 * and the thin numpy wrappers healpy puts in front of them (ang2vec, vec2ang,
 * lonlat handling).  PARITY UNPINNED against the real healpy: the reference
 * ships no golden vectors for it.  It is
 * pinned instead by hand-derivable known answers and brute-force properties in
 * tests/test_oracle_healpix.py.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * load this file.  The product (baryonification_b200/) never does.
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -shared -fPIC).
 * -ffp-contract=off keeps every a*b+c as two roundings, which is what the
 * CUDA side reproduces with __dmul_rn/__dadd_rn in its index arithmetic.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define HPX_PI 3.141592653589793238462643383279502884197
#define HPX_TWOPI 6.283185307179586476925286766559005768394
#define HPX_HALFPI 1.570796326794896619231321691639751442099
#define HPX_INV_TWOPI (1.0 / HPX_TWOPI)
#define HPX_INV_HALFPI 0.6366197723675813430755350534900574
#define HPX_TWOTHIRD (2.0 / 3.0)
#define HPX_DEG2RAD (HPX_PI / 180.0)
#define HPX_RAD2DEG (180.0 / HPX_PI)

typedef int64_t i64;

typedef struct {
    i64 nside, npix, ncap;
    int order;           /* log2(nside) or -1 */
    double fact1, fact2; /* fact2 = 4/npix ; fact1 = 2*nside*fact2 */
} hpx_t;

static hpx_t hpx_make(i64 nside) {
    hpx_t h;
    h.nside = nside;
    h.npix = 12 * nside * nside;
    h.ncap = 2 * nside * (nside - 1);
    h.fact2 = 4.0 / (double)h.npix;
    h.fact1 = (double)(nside << 1) * h.fact2;
    h.order = -1;
    if ((nside & (nside - 1)) == 0) {
        int o = 0;
        while (((i64)1 << o) < nside) ++o;
        h.order = o;
    }
    return h;
}

static i64 hpx_isqrt(i64 v) {
    i64 r = (i64)sqrt((double)v + 0.5);
    /* exact correction for large arguments */
    while (r * r > v) --r;
    while ((r + 1) * (r + 1) <= v) ++r;
    return r;
}

static double hpx_fmodulo(double v1, double v2) {
    if (v1 >= 0) return (v1 < v2) ? v1 : fmod(v1, v2);
    double tmp = fmod(v1, v2) + v2;
    return (tmp == v2) ? 0.0 : tmp;
}

/* ring_above: index of the ring just above (north of) z, 0..4*nside-1 */
static i64 hpx_ring_above(const hpx_t *h, double z) {
    double az = fabs(z);
    if (az <= HPX_TWOTHIRD) return (i64)((double)h->nside * (2.0 - 1.5 * z));
    i64 iring = (i64)((double)h->nside * sqrt(3.0 * (1.0 - az)));
    return (z > 0) ? iring : 4 * h->nside - iring - 1;
}

static double hpx_ring2z(const hpx_t *h, i64 ring) {
    if (ring < h->nside) return 1.0 - (double)(ring * ring) * h->fact2;
    if (ring <= 3 * h->nside) return (double)(2 * h->nside - ring) * h->fact1;
    ring = 4 * h->nside - ring;
    return (double)(ring * ring) * h->fact2 - 1.0;
}

static void hpx_ring_info_small(const hpx_t *h, i64 ring, i64 *startpix,
                                i64 *ringpix, int *shifted) {
    if (ring < h->nside) {
        *shifted = 1;
        *ringpix = 4 * ring;
        *startpix = 2 * ring * (ring - 1);
    } else if (ring < 3 * h->nside) {
        *shifted = ((ring - h->nside) & 1) == 0;
        *ringpix = 4 * h->nside;
        *startpix = h->ncap + (ring - h->nside) * (*ringpix);
    } else {
        i64 nr = 4 * h->nside - ring;
        *shifted = 1;
        *ringpix = 4 * nr;
        *startpix = h->npix - 2 * nr * (nr + 1);
    }
}

static void hpx_ring_info2(const hpx_t *h, i64 ring, i64 *startpix, i64 *ringpix,
                           double *theta, int *shifted) {
    i64 northring = (ring > 2 * h->nside) ? 4 * h->nside - ring : ring;
    if (northring < h->nside) {
        double tmp = (double)(northring * northring) * h->fact2;
        double costheta = 1.0 - tmp;
        double sintheta = sqrt(tmp * (2.0 - tmp));
        *theta = atan2(sintheta, costheta);
        *ringpix = 4 * northring;
        *shifted = 1;
        *startpix = 2 * northring * (northring - 1);
    } else {
        *theta = acos((double)(2 * h->nside - northring) * h->fact1);
        *ringpix = 4 * h->nside;
        *shifted = ((northring - h->nside) & 1) == 0;
        *startpix = h->ncap + (northring - h->nside) * (*ringpix);
    }
    if (northring != ring) { /* southern hemisphere */
        *theta = HPX_PI - *theta;
        *startpix = h->npix - *startpix - *ringpix;
    }
}

/* pix -> (z, phi [, sin(theta)]) in the RING scheme */
static void hpx_pix2loc(const hpx_t *h, i64 pix, double *z, double *phi,
                        double *sth, int *have_sth) {
    *have_sth = 0;
    if (pix < h->ncap) { /* north polar cap */
        i64 iring = (1 + hpx_isqrt(1 + 2 * pix)) >> 1;
        i64 iphi = (pix + 1) - 2 * iring * (iring - 1);
        double tmp = (double)(iring * iring) * h->fact2;
        *z = 1.0 - tmp;
        if (*z > 0.99) {
            *sth = sqrt(tmp * (2.0 - tmp));
            *have_sth = 1;
        }
        *phi = ((double)iphi - 0.5) * HPX_HALFPI / (double)iring;
    } else if (pix < (h->npix - h->ncap)) { /* equatorial belt */
        i64 nl4 = 4 * h->nside;
        i64 ip = pix - h->ncap;
        i64 tmp = (h->order >= 0) ? ip >> (h->order + 2) : ip / nl4;
        i64 iring = tmp + h->nside, iphi = ip - nl4 * tmp + 1;
        double fodd = ((iring + h->nside) & 1) ? 1.0 : 0.5;
        *z = (double)(2 * h->nside - iring) * h->fact1;
        *phi = ((double)iphi - fodd) * HPX_PI * 0.75 * h->fact1;
    } else { /* south polar cap */
        i64 ip = h->npix - pix;
        i64 iring = (1 + hpx_isqrt(2 * ip - 1)) >> 1;
        i64 iphi = 4 * iring + 1 - (ip - 2 * iring * (iring - 1));
        double tmp = (double)(iring * iring) * h->fact2;
        *z = tmp - 1.0;
        if (*z < -0.99) {
            *sth = sqrt(tmp * (2.0 - tmp));
            *have_sth = 1;
        }
        *phi = ((double)iphi - 0.5) * HPX_HALFPI / (double)iring;
    }
}

static i64 hpx_loc2pix(const hpx_t *h, double z, double phi, double sth,
                       int have_sth) {
    double za = fabs(z);
    double tt = hpx_fmodulo(phi * HPX_INV_HALFPI, 4.0); /* in [0,4) */
    if (za <= HPX_TWOTHIRD) {
        i64 nl4 = 4 * h->nside;
        double temp1 = (double)h->nside * (0.5 + tt);
        double temp2 = (double)h->nside * z * 0.75;
        i64 jp = (i64)(temp1 - temp2); /* ascending edge line */
        i64 jm = (i64)(temp1 + temp2); /* descending edge line */
        i64 ir = h->nside + 1 + jp - jm; /* ring counted from z=2/3, 1..2n+1 */
        i64 kshift = 1 - (ir & 1);
        i64 t1 = jp + jm - h->nside + kshift + 1 + nl4 + nl4;
        i64 ip = (h->order > 0) ? ((t1 >> 1) & (nl4 - 1)) : ((t1 >> 1) % nl4);
        return h->ncap + (ir - 1) * nl4 + ip;
    }
    double tp = tt - (double)(i64)tt;
    double tmp = ((za < 0.99) || (!have_sth))
                     ? (double)h->nside * sqrt(3.0 * (1.0 - za))
                     : (double)h->nside * sth / sqrt((1.0 + za) / 3.0);
    i64 jp = (i64)(tp * tmp);
    i64 jm = (i64)((1.0 - tp) * tmp);
    i64 ir = jp + jm + 1; /* ring counted from the closest pole */
    i64 ip = (i64)(tt * (double)ir);
    if (ip >= 4 * ir) ip = 4 * ir - 1; /* guards the "must not happen" assert */
    return (z > 0) ? 2 * ir * (ir - 1) + ip : h->npix - 2 * ir * (ir + 1) + ip;
}

/* ------------------------------------------------------------------ */
/* exported, vectorised entry points (ctypes)                          */
/* ------------------------------------------------------------------ */

i64 hpxo_npix2nside(i64 npix) {
    i64 res = hpx_isqrt(npix / 12);
    return (npix == res * res * 12) ? res : -1;
}

double hpxo_nside2resol(i64 nside) {
    return sqrt(4.0 * HPX_PI / (double)(12 * nside * nside));
}

/* healpy.ang2vec(lon, lat, lonlat=True) */
void hpxo_ang2vec_lonlat(i64 n, const double *lon, const double *lat, double *vec) {
    for (i64 i = 0; i < n; ++i) {
        double theta = HPX_PI / 2.0 - lat[i] * HPX_DEG2RAD;
        double phi = lon[i] * HPX_DEG2RAD;
        double st = sin(theta);
        vec[3 * i + 0] = st * cos(phi);
        vec[3 * i + 1] = st * sin(phi);
        vec[3 * i + 2] = cos(theta);
    }
}

/* healpy.vec2ang(v, lonlat=True): theta = arccos(z/|v|), phi = arctan2(y,x)
 * (+2pi if negative); returns (degrees(phi), 90 - degrees(theta)). */
void hpxo_vec2ang_lonlat(i64 n, const double *vec, double *lon, double *lat) {
    for (i64 i = 0; i < n; ++i) {
        double x = vec[3 * i], y = vec[3 * i + 1], z = vec[3 * i + 2];
        double dnorm = sqrt(x * x + y * y + z * z);
        double theta = acos(z / dnorm);
        double phi = atan2(y, x);
        if (phi < 0) phi += HPX_TWOPI;
        lon[i] = phi * HPX_RAD2DEG;
        lat[i] = 90.0 - theta * HPX_RAD2DEG;
    }
}

/* healpy.pix2vec(nside, ipix) RING */
void hpxo_pix2vec_ring(i64 nside, i64 n, const i64 *pix, double *vec) {
    hpx_t h = hpx_make(nside);
    for (i64 i = 0; i < n; ++i) {
        double z, phi, sth;
        int have;
        hpx_pix2loc(&h, pix[i], &z, &phi, &sth, &have);
        if (!have) sth = sqrt((1.0 - z) * (1.0 + z));
        vec[3 * i + 0] = sth * cos(phi);
        vec[3 * i + 1] = sth * sin(phi);
        vec[3 * i + 2] = z;
    }
}

/* healpy.pix2ang(nside, ipix) RING -> theta, phi (radians) */
void hpxo_pix2ang_ring(i64 nside, i64 n, const i64 *pix, double *theta, double *phi) {
    hpx_t h = hpx_make(nside);
    for (i64 i = 0; i < n; ++i) {
        double z, sth;
        int have;
        hpx_pix2loc(&h, pix[i], &z, &phi[i], &sth, &have);
        theta[i] = have ? atan2(sth, z) : acos(z);
    }
}

/* healpy.ang2pix(nside, theta, phi) RING (theta, phi in radians) */
void hpxo_ang2pix_ring(i64 nside, i64 n, const double *theta, const double *phi,
                       i64 *pix) {
    hpx_t h = hpx_make(nside);
    for (i64 i = 0; i < n; ++i) {
        double th = theta[i];
        if ((th < 0.01) || (th > 3.14159 - 0.01))
            pix[i] = hpx_loc2pix(&h, cos(th), phi[i], sin(th), 1);
        else
            pix[i] = hpx_loc2pix(&h, cos(th), phi[i], 0.0, 0);
    }
}

/* pointing(vec3): theta = atan2(sqrt(x^2+y^2), z), phi = atan2(y,x) in [0,2pi) */
static void hpx_vec2pointing(const double *v, double *theta, double *phi) {
    *theta = atan2(sqrt(v[0] * v[0] + v[1] * v[1]), v[2]);
    double p = (v[0] == 0.0 && v[1] == 0.0) ? 0.0 : atan2(v[1], v[0]);
    if (p < 0.0) p += HPX_TWOPI;
    *phi = hpx_fmodulo(p, HPX_TWOPI);
}

/*
 * healpy.query_disc(nside, vec, radius, inclusive=False, nest=False).
 * Emits ascending half-open pixel ranges [lo,hi) into `ranges` (2 per entry);
 * returns the number of ranges, or the negated required count when `cap` is
 * too small.  Range merging follows rangeset::append (a new range that starts
 * at or before the end of the previous one extends it).
 */
static i64 rs_append(i64 *ranges, i64 nr, i64 cap, i64 v1, i64 v2) {
    if (v2 <= v1) return nr;
    if (nr > 0 && nr <= cap && v1 <= ranges[2 * (nr - 1) + 1]) {
        if (v2 > ranges[2 * (nr - 1) + 1]) ranges[2 * (nr - 1) + 1] = v2;
        return nr;
    }
    if (nr < cap) {
        ranges[2 * nr] = v1;
        ranges[2 * nr + 1] = v2;
    }
    return nr + 1;
}

i64 hpxo_query_disc_ranges(i64 nside, const double *vec, double radius, i64 cap,
                           i64 *ranges) {
    hpx_t h = hpx_make(nside);
    double theta, phi;
    hpx_vec2pointing(vec, &theta, &phi);
    i64 nr_out = 0;
    double rsmall = radius, rbig = radius;
    if (rsmall >= HPX_PI) return rs_append(ranges, 0, cap, 0, h.npix);
    if (rbig > HPX_PI) rbig = HPX_PI;
    double cosrbig = cos(rbig);
    double z0 = cos(theta);
    double xa = 1.0 / sqrt((1.0 - z0) * (1.0 + z0));

    double rlat1 = theta - rsmall;
    double zmax = cos(rlat1);
    i64 irmin = hpx_ring_above(&h, zmax) + 1;
    if ((rlat1 <= 0) && (irmin > 1)) { /* north pole inside the disc */
        i64 sp, rp;
        int dummy;
        hpx_ring_info_small(&h, irmin - 1, &sp, &rp, &dummy);
        nr_out = rs_append(ranges, nr_out, cap, 0, sp + rp);
    }
    double rlat2 = theta + rsmall;
    double zmin = cos(rlat2);
    i64 irmax = hpx_ring_above(&h, zmin);

    for (i64 iz = irmin; iz <= irmax; ++iz) {
        double z = hpx_ring2z(&h, iz);
        double x = (cosrbig - z * z0) * xa;
        double ysq = 1.0 - z * z - x * x;
        double dphi = (ysq <= 0) ? 0.0 : atan2(sqrt(ysq), x);
        if (dphi > 0) {
            i64 nr, ipix1;
            int shifted;
            hpx_ring_info_small(&h, iz, &ipix1, &nr, &shifted);
            double shift = shifted ? 0.5 : 0.0;
            i64 ipix2 = ipix1 + nr - 1;
            i64 ip_lo = (i64)floor((double)nr * HPX_INV_TWOPI * (phi - dphi) - shift) + 1;
            i64 ip_hi = (i64)floor((double)nr * HPX_INV_TWOPI * (phi + dphi) - shift);
            if (ip_hi >= nr) {
                ip_lo -= nr;
                ip_hi -= nr;
            }
            if (ip_lo < 0) {
                nr_out = rs_append(ranges, nr_out, cap, ipix1, ipix1 + ip_hi + 1);
                nr_out = rs_append(ranges, nr_out, cap, ipix1 + ip_lo + nr, ipix2 + 1);
            } else {
                nr_out = rs_append(ranges, nr_out, cap, ipix1 + ip_lo, ipix1 + ip_hi + 1);
            }
        }
    }
    if ((rlat2 >= HPX_PI) && (irmax + 1 < 4 * h.nside)) { /* south pole inside */
        i64 sp, rp;
        int dummy;
        hpx_ring_info_small(&h, irmax + 1, &sp, &rp, &dummy);
        nr_out = rs_append(ranges, nr_out, cap, sp, h.npix);
    }
    return (nr_out > cap) ? -nr_out : nr_out;
}

/*
 * healpy.get_interp_weights(nside, lon, lat, lonlat=True): 4 neighbours and
 * bilinear weights (sum == 1).  pix/wgt are [n][4].
 */
static void hpx_get_interpol(const hpx_t *h, double theta, double phi, i64 *pix,
                             double *wgt) {
    double z = cos(theta);
    i64 ir1 = hpx_ring_above(h, z);
    i64 ir2 = ir1 + 1;
    double theta1 = 0, theta2 = 0, w1, tmp, dphi;
    i64 sp, nr, i1, i2;
    int shift;
    if (ir1 > 0) {
        hpx_ring_info2(h, ir1, &sp, &nr, &theta1, &shift);
        dphi = HPX_TWOPI / (double)nr;
        tmp = (phi / dphi - 0.5 * shift);
        i1 = (tmp < 0) ? (i64)tmp - 1 : (i64)tmp;
        w1 = (phi - ((double)i1 + 0.5 * shift) * dphi) / dphi;
        i2 = i1 + 1;
        if (i1 < 0) i1 += nr;
        if (i2 >= nr) i2 -= nr;
        pix[0] = sp + i1;
        pix[1] = sp + i2;
        wgt[0] = 1 - w1;
        wgt[1] = w1;
    }
    if (ir2 < 4 * h->nside) {
        hpx_ring_info2(h, ir2, &sp, &nr, &theta2, &shift);
        dphi = HPX_TWOPI / (double)nr;
        tmp = (phi / dphi - 0.5 * shift);
        i1 = (tmp < 0) ? (i64)tmp - 1 : (i64)tmp;
        w1 = (phi - ((double)i1 + 0.5 * shift) * dphi) / dphi;
        i2 = i1 + 1;
        if (i1 < 0) i1 += nr;
        if (i2 >= nr) i2 -= nr;
        pix[2] = sp + i1;
        pix[3] = sp + i2;
        wgt[2] = 1 - w1;
        wgt[3] = w1;
    }
    if (ir1 == 0) {
        double wtheta = theta / theta2;
        wgt[2] *= wtheta;
        wgt[3] *= wtheta;
        double fac = (1 - wtheta) * 0.25;
        wgt[0] = fac;
        wgt[1] = fac;
        wgt[2] += fac;
        wgt[3] += fac;
        pix[0] = (pix[2] + 2) & 3;
        pix[1] = (pix[3] + 2) & 3;
    } else if (ir2 == 4 * h->nside) {
        double wtheta = (theta - theta1) / (HPX_PI - theta1);
        wgt[0] *= (1 - wtheta);
        wgt[1] *= (1 - wtheta);
        double fac = wtheta * 0.25;
        wgt[0] += fac;
        wgt[1] += fac;
        wgt[2] = fac;
        wgt[3] = fac;
        pix[2] = ((pix[0] + 2) & 3) + h->npix - 4;
        pix[3] = ((pix[1] + 2) & 3) + h->npix - 4;
    } else {
        double wtheta = (theta - theta1) / (theta2 - theta1);
        wgt[0] *= (1 - wtheta);
        wgt[1] *= (1 - wtheta);
        wgt[2] *= wtheta;
        wgt[3] *= wtheta;
    }
}

void hpxo_get_interp_weights_lonlat(i64 nside, i64 n, const double *lon,
                                    const double *lat, i64 *pix, double *wgt) {
    hpx_t h = hpx_make(nside);
    for (i64 i = 0; i < n; ++i) {
        double theta = HPX_PI / 2.0 - lat[i] * HPX_DEG2RAD;
        double phi = hpx_fmodulo(lon[i] * HPX_DEG2RAD, HPX_TWOPI); /* pointing::normalize */
        hpx_get_interpol(&h, theta, phi, pix + 4 * i, wgt + 4 * i);
    }
}

void hpxo_get_interp_weights(i64 nside, i64 n, const double *theta,
                             const double *phi, i64 *pix, double *wgt) {
    hpx_t h = hpx_make(nside);
    for (i64 i = 0; i < n; ++i)
        hpx_get_interpol(&h, theta[i], hpx_fmodulo(phi[i], HPX_TWOPI), pix + 4 * i,
                         wgt + 4 * i);
}
