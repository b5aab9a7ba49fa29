// bf_healpix.cuh -- HEALPix RING-scheme geometry on the device (Gorski et al. 2005; the
// healpix_cxx Healpix_Base semantics healpy exposes).  Replaces the healpy calls of
// Runners/HealpixRunner.py:303 (ang2vec), :306/:432 (query_disc, inclusive=False, nest=False),
// :310/:337 (get_interp_weights), :312/:333/:433 (pix2vec), :334 (vec2ang).
// Compiled with -fmad=false so that index arithmetic rounds like the CPU code it must agree with.
#pragma once
#include "bf_common.cuh"

#define BFH_PI 3.141592653589793238462643383279502884197
#define BFH_TWOPI 6.283185307179586476925286766559005768394
#define BFH_HALFPI 1.570796326794896619231321691639751442099
#define BFH_INV_TWOPI (1.0 / BFH_TWOPI)
#define BFH_INV_HALFPI 0.6366197723675813430755350534900574
#define BFH_TWOTHIRD (2.0 / 3.0)
#define BFH_DEG2RAD (BFH_PI / 180.0)
#define BFH_RAD2DEG (180.0 / BFH_PI)

struct bf_hpx {
    i64 nside, npix, ncap;
    int order;
    double fact1, fact2;
};

static inline bf_hpx bf_hpx_make(i64 nside) {
    bf_hpx h;
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

__device__ __forceinline__ i64 bfh_isqrt(i64 v) {
    i64 r = (i64)sqrt((double)v + 0.5);
    while (r * r > v) --r;
    while ((r + 1) * (r + 1) <= v) ++r;
    return r;
}

__device__ __forceinline__ double bfh_fmodulo(double v1, double v2) {
    if (v1 >= 0) return (v1 < v2) ? v1 : fmod(v1, v2);
    double tmp = fmod(v1, v2) + v2;
    return (tmp == v2) ? 0.0 : tmp;
}

__device__ __forceinline__ i64 bfh_ring_above(const bf_hpx &h, double z) {
    double az = fabs(z);
    if (az <= BFH_TWOTHIRD) return (i64)((double)h.nside * (2.0 - 1.5 * z));
    i64 iring = (i64)((double)h.nside * sqrt(3.0 * (1.0 - az)));
    return (z > 0) ? iring : 4 * h.nside - iring - 1;
}

__device__ __forceinline__ double bfh_ring2z(const bf_hpx &h, i64 ring) {
    if (ring < h.nside) return 1.0 - (double)(ring * ring) * h.fact2;
    if (ring <= 3 * h.nside) return (double)(2 * h.nside - ring) * h.fact1;
    ring = 4 * h.nside - ring;
    return (double)(ring * ring) * h.fact2 - 1.0;
}

__device__ __forceinline__ void bfh_ring_info_small(const bf_hpx &h, i64 ring, i64 &startpix, i64 &ringpix,
                                                    bool &shifted) {
    if (ring < h.nside) {
        shifted = true;
        ringpix = 4 * ring;
        startpix = 2 * ring * (ring - 1);
    } else if (ring < 3 * h.nside) {
        shifted = ((ring - h.nside) & 1) == 0;
        ringpix = 4 * h.nside;
        startpix = h.ncap + (ring - h.nside) * ringpix;
    } else {
        i64 nr = 4 * h.nside - ring;
        shifted = true;
        ringpix = 4 * nr;
        startpix = h.npix - 2 * nr * (nr + 1);
    }
}

__device__ __forceinline__ void bfh_ring_info2(const bf_hpx &h, i64 ring, i64 &startpix, i64 &ringpix,
                                               double &theta, bool &shifted) {
    i64 northring = (ring > 2 * h.nside) ? 4 * h.nside - ring : ring;
    if (northring < h.nside) {
        double tmp = (double)(northring * northring) * h.fact2;
        double costheta = 1.0 - tmp;
        double sintheta = sqrt(tmp * (2.0 - tmp));
        theta = atan2(sintheta, costheta);
        ringpix = 4 * northring;
        shifted = true;
        startpix = 2 * northring * (northring - 1);
    } else {
        theta = acos((double)(2 * h.nside - northring) * h.fact1);
        ringpix = 4 * h.nside;
        shifted = ((northring - h.nside) & 1) == 0;
        startpix = h.ncap + (northring - h.nside) * ringpix;
    }
    if (northring != ring) {
        theta = BFH_PI - theta;
        startpix = h.npix - startpix - ringpix;
    }
}

// ring number (1..4nside-1) and 1-based index in ring of a RING pixel
__device__ __forceinline__ void bfh_pix2ring(const bf_hpx &h, i64 pix, i64 &iring, i64 &iphi) {
    if (pix < h.ncap) {
        iring = (1 + bfh_isqrt(1 + 2 * pix)) >> 1;
        iphi = (pix + 1) - 2 * iring * (iring - 1);
    } else if (pix < h.npix - h.ncap) {
        i64 nl4 = 4 * h.nside, ip = pix - h.ncap;
        i64 tmp = (h.order >= 0) ? ip >> (h.order + 2) : ip / nl4;
        iring = tmp + h.nside;
        iphi = ip - nl4 * tmp + 1;
    } else {
        i64 ip = h.npix - pix;
        i64 ir = (1 + bfh_isqrt(2 * ip - 1)) >> 1;
        iphi = 4 * ir + 1 - (ip - 2 * ir * (ir - 1));
        iring = 4 * h.nside - ir;
    }
}

// z, phi (and sin(theta) near the poles) of pixel `iphi` (1-based) of ring `iring`
__device__ __forceinline__ void bfh_ringpix2loc(const bf_hpx &h, i64 iring, i64 iphi, double &z, double &phi,
                                                double &sth, bool &have_sth) {
    have_sth = false;
    if (iring < h.nside) {
        double tmp = (double)(iring * iring) * h.fact2;
        z = 1.0 - tmp;
        if (z > 0.99) { sth = sqrt(tmp * (2.0 - tmp)); have_sth = true; }
        phi = ((double)iphi - 0.5) * BFH_HALFPI / (double)iring;
    } else if (iring <= 3 * h.nside) {
        double fodd = ((iring + h.nside) & 1) ? 1.0 : 0.5;
        z = (double)(2 * h.nside - iring) * h.fact1;
        phi = ((double)iphi - fodd) * BFH_PI * 0.75 * h.fact1;
    } else {
        i64 ir = 4 * h.nside - iring;
        double tmp = (double)(ir * ir) * h.fact2;
        z = tmp - 1.0;
        if (z < -0.99) { sth = sqrt(tmp * (2.0 - tmp)); have_sth = true; }
        phi = ((double)iphi - 0.5) * BFH_HALFPI / (double)ir;
    }
}

__device__ __forceinline__ void bfh_pix2loc(const bf_hpx &h, i64 pix, double &z, double &phi, double &sth,
                                            bool &have_sth) {
    i64 iring, iphi;
    bfh_pix2ring(h, pix, iring, iphi);
    bfh_ringpix2loc(h, iring, iphi, z, phi, sth, have_sth);
}

__device__ __forceinline__ void bfh_loc2vec(double z, double phi, double sth, bool have_sth, double v[3]) {
    if (!have_sth) sth = sqrt((1.0 - z) * (1.0 + z));
    double s, c;
    sincos(phi, &s, &c);
    v[0] = sth * c;
    v[1] = sth * s;
    v[2] = z;
}

__device__ __forceinline__ void bfh_pix2vec(const bf_hpx &h, i64 pix, double v[3]) {
    double z, phi, sth;
    bool have;
    bfh_pix2loc(h, pix, z, phi, sth, have);
    bfh_loc2vec(z, phi, sth, have, v);
}

__device__ __forceinline__ i64 bfh_loc2pix(const bf_hpx &h, double z, double phi, double sth, bool have_sth) {
    double za = fabs(z);
    double tt = bfh_fmodulo(phi * BFH_INV_HALFPI, 4.0);
    if (za <= BFH_TWOTHIRD) {
        i64 nl4 = 4 * h.nside;
        double temp1 = (double)h.nside * (0.5 + tt);
        double temp2 = (double)h.nside * z * 0.75;
        i64 jp = (i64)(temp1 - temp2);
        i64 jm = (i64)(temp1 + temp2);
        i64 ir = h.nside + 1 + jp - jm;
        i64 kshift = 1 - (ir & 1);
        i64 t1 = jp + jm - h.nside + kshift + 1 + nl4 + nl4;
        i64 ip = (h.order > 0) ? ((t1 >> 1) & (nl4 - 1)) : ((t1 >> 1) % nl4);
        return h.ncap + (ir - 1) * nl4 + ip;
    }
    double tp = tt - (double)(i64)tt;
    double tmp = ((za < 0.99) || (!have_sth)) ? (double)h.nside * sqrt(3.0 * (1.0 - za))
                                              : (double)h.nside * sth / sqrt((1.0 + za) / 3.0);
    i64 jp = (i64)(tp * tmp);
    i64 jm = (i64)((1.0 - tp) * tmp);
    i64 ir = jp + jm + 1;
    i64 ip = (i64)(tt * (double)ir);
    if (ip >= 4 * ir) ip = 4 * ir - 1;
    return (z > 0) ? 2 * ir * (ir - 1) + ip : h.npix - 2 * ir * (ir + 1) + ip;
}

__device__ __forceinline__ i64 bfh_ang2pix(const bf_hpx &h, double theta, double phi) {
    if ((theta < 0.01) || (theta > 3.14159 - 0.01)) return bfh_loc2pix(h, cos(theta), phi, sin(theta), true);
    return bfh_loc2pix(h, cos(theta), phi, 0.0, false);
}

// healpy.ang2vec(lon, lat, lonlat=True)
__device__ __forceinline__ void bfh_ang2vec_lonlat(double lon, double lat, double v[3]) {
    double theta = BFH_PI / 2.0 - lat * BFH_DEG2RAD;
    double phi = lon * BFH_DEG2RAD;
    double st = sin(theta);
    v[0] = st * cos(phi);
    v[1] = st * sin(phi);
    v[2] = cos(theta);
}

// get_interpol(pointing(theta, phi)) with phi already normalised to [0, 2pi)
// (ir1 = ring above the point, 0 at the north pole: pix[0..1] lie on ring max(ir1, 1), pix[2..3] on min(ir1 + 1, 4N - 1))
// `theta_tab` (optional): co-latitudes of rings 0 .. 4 nside as bfh_ring_info2 computes them (filled by calling it, so
// the values are the same bits); the two acos / atan2 per point become two loads.  The integer ring data of
// bfh_ring_info_small equal bfh_ring_info2's.
__device__ __forceinline__ void bfh_ring_info2_tab(const bf_hpx &h, i64 ring, i64 &startpix, i64 &ringpix, double &theta,
                                                   bool &shifted, const double *__restrict__ theta_tab) {
    if (theta_tab) {
        bfh_ring_info_small(h, ring, startpix, ringpix, shifted);
        theta = __ldg(theta_tab + ring);
    } else {
        bfh_ring_info2(h, ring, startpix, ringpix, theta, shifted);
    }
}

__device__ __forceinline__ void bfh_get_interpol_r(const bf_hpx &h, double theta, double phi, i64 pix[4],
                                                   double wgt[4], i64 &ir1, const double *__restrict__ theta_tab = nullptr) {
    double z = cos(theta);
    ir1 = bfh_ring_above(h, z);
    i64 ir2 = ir1 + 1;
    double theta1 = 0, theta2 = 0, w1, tmp, dphi;
    i64 sp, nr, i1, i2;
    bool shift;
    if (ir1 > 0) {
        bfh_ring_info2_tab(h, ir1, sp, nr, theta1, shift, theta_tab);
        dphi = BFH_TWOPI / (double)nr;
        tmp = (phi / dphi - 0.5 * (shift ? 1.0 : 0.0));
        i1 = (tmp < 0) ? (i64)tmp - 1 : (i64)tmp;
        w1 = (phi - ((double)i1 + 0.5 * (shift ? 1.0 : 0.0)) * dphi) / dphi;
        i2 = i1 + 1;
        if (i1 < 0) i1 += nr;
        if (i2 >= nr) i2 -= nr;
        pix[0] = sp + i1; pix[1] = sp + i2;
        wgt[0] = 1 - w1;  wgt[1] = w1;
    }
    if (ir2 < 4 * h.nside) {
        bfh_ring_info2_tab(h, ir2, sp, nr, theta2, shift, theta_tab);
        dphi = BFH_TWOPI / (double)nr;
        tmp = (phi / dphi - 0.5 * (shift ? 1.0 : 0.0));
        i1 = (tmp < 0) ? (i64)tmp - 1 : (i64)tmp;
        w1 = (phi - ((double)i1 + 0.5 * (shift ? 1.0 : 0.0)) * dphi) / dphi;
        i2 = i1 + 1;
        if (i1 < 0) i1 += nr;
        if (i2 >= nr) i2 -= nr;
        pix[2] = sp + i1; pix[3] = sp + i2;
        wgt[2] = 1 - w1;  wgt[3] = w1;
    }
    if (ir1 == 0) {
        double wtheta = theta / theta2;
        wgt[2] *= wtheta; wgt[3] *= wtheta;
        double fac = (1 - wtheta) * 0.25;
        wgt[0] = fac; wgt[1] = fac; wgt[2] += fac; wgt[3] += fac;
        pix[0] = (pix[2] + 2) & 3;
        pix[1] = (pix[3] + 2) & 3;
    } else if (ir2 == 4 * h.nside) {
        double wtheta = (theta - theta1) / (BFH_PI - theta1);
        wgt[0] *= (1 - wtheta); wgt[1] *= (1 - wtheta);
        double fac = wtheta * 0.25;
        wgt[0] += fac; wgt[1] += fac; wgt[2] = fac; wgt[3] = fac;
        pix[2] = ((pix[0] + 2) & 3) + h.npix - 4;
        pix[3] = ((pix[1] + 2) & 3) + h.npix - 4;
    } else {
        double wtheta = (theta - theta1) / (theta2 - theta1);
        wgt[0] *= (1 - wtheta); wgt[1] *= (1 - wtheta);
        wgt[2] *= wtheta; wgt[3] *= wtheta;
    }
}

__device__ __forceinline__ void bfh_get_interpol(const bf_hpx &h, double theta, double phi, i64 pix[4],
                                                 double wgt[4]) {
    i64 ir1;
    bfh_get_interpol_r(h, theta, phi, pix, wgt, ir1);
}

__device__ __forceinline__ void bfh_interp_lonlat_r(const bf_hpx &h, double lon, double lat, i64 pix[4],
                                                    double wgt[4], i64 &ir1, const double *__restrict__ theta_tab = nullptr) {
    double theta = BFH_PI / 2.0 - lat * BFH_DEG2RAD;
    double phi = bfh_fmodulo(lon * BFH_DEG2RAD, BFH_TWOPI);
    bfh_get_interpol_r(h, theta, phi, pix, wgt, ir1, theta_tab);
}

__device__ __forceinline__ void bfh_interp_lonlat(const bf_hpx &h, double lon, double lat, i64 pix[4],
                                                  double wgt[4]) {
    i64 ir1;
    bfh_interp_lonlat_r(h, lon, lat, pix, wgt, ir1);
}

// The 4 deposits of one displaced pixel (HealpixRunner.py:333-341): new_vec = pix2vec(p) + offsets[p] ->
// vec2ang(lonlat=True) in DEGREES -> get_interp_weights(lonlat=True).  One definition for every re-binning variant,
// so that their deposits are the same bits.
__device__ __forceinline__ void bfh_displaced_deposits_r(const bf_hpx &h, i64 p, const double off[3], i64 cp[4],
                                                         double w[4], i64 &ir1,
                                                         const double *__restrict__ theta_tab = nullptr) {
    double v[3];
    bfh_pix2vec(h, p, v);
    double x = v[0] + off[0], y = v[1] + off[1], z = v[2] + off[2];
    double dnorm = sqrt(x * x + y * y + z * z);
    double theta = acos(z / dnorm);
    double phi = atan2(y, x);
    if (phi < 0) phi += BFH_TWOPI;
    bfh_interp_lonlat_r(h, phi * BFH_RAD2DEG, 90.0 - theta * BFH_RAD2DEG, cp, w, ir1, theta_tab);
}

__device__ __forceinline__ void bfh_displaced_deposits(const bf_hpx &h, i64 p, const double off[3], i64 cp[4],
                                                       double w[4]) {
    i64 ir1;
    bfh_displaced_deposits_r(h, p, off, cp, w, ir1);
}

// ------------------------------------------------------------------------------------------
// query_disc(inclusive=False, RING), decomposed for parallel use.
// bf_disc holds the per-disc constants; bfh_disc_ring() gives the pixel segment(s) of one ring.
// ------------------------------------------------------------------------------------------
struct bf_disc {
    double phi, z0, xa, cosr;
    i64 irmin, irmax;      // rings walked: irmin..irmax inclusive
    i64 north_end;         // pixels [0, north_end) included whole (north pole inside disc), else 0
    i64 south_begin;       // pixels [south_begin, npix) included whole, else npix
    bool all;              // radius >= pi
};

__device__ __forceinline__ bf_disc bfh_disc_setup(const bf_hpx &h, const double v[3], double radius) {
    bf_disc d;
    // pointing(vec3): theta = atan2(sqrt(x^2+y^2), z); phi = atan2(y, x) in [0, 2pi)
    double theta = atan2(sqrt(v[0] * v[0] + v[1] * v[1]), v[2]);
    double p = (v[0] == 0.0 && v[1] == 0.0) ? 0.0 : atan2(v[1], v[0]);
    if (p < 0.0) p += BFH_TWOPI;
    d.phi = bfh_fmodulo(p, BFH_TWOPI);
    d.all = radius >= BFH_PI;
    d.north_end = 0;
    d.south_begin = h.npix;
    d.irmin = 1;
    d.irmax = 0;
    d.z0 = d.xa = d.cosr = 0;
    if (d.all) { d.north_end = h.npix; return d; }
    d.cosr = cos(radius);
    d.z0 = cos(theta);
    d.xa = 1.0 / sqrt((1.0 - d.z0) * (1.0 + d.z0));
    double rlat1 = theta - radius;
    double zmax = cos(rlat1);
    d.irmin = bfh_ring_above(h, zmax) + 1;
    if ((rlat1 <= 0) && (d.irmin > 1)) {
        i64 sp, rp; bool dummy;
        bfh_ring_info_small(h, d.irmin - 1, sp, rp, dummy);
        d.north_end = sp + rp;
    }
    double rlat2 = theta + radius;
    double zmin = cos(rlat2);
    d.irmax = bfh_ring_above(h, zmin);
    if ((rlat2 >= BFH_PI) && (d.irmax + 1 < 4 * h.nside)) {
        i64 sp, rp; bool dummy;
        bfh_ring_info_small(h, d.irmax + 1, sp, rp, dummy);
        d.south_begin = sp;
    }
    return d;
}

// Pixel set of ring iz: [a0, a1) U [b0, b1) (absolute RING indices, a-range first, ascending,
// possibly empty).  Returns the pixel count.  Overlapping wrap ranges are merged like rangeset::append.
__device__ __forceinline__ i64 bfh_disc_ring(const bf_hpx &h, const bf_disc &d, i64 iz, i64 &a0, i64 &a1,
                                             i64 &b0, i64 &b1) {
    a0 = a1 = b0 = b1 = 0;
    double z = bfh_ring2z(h, iz);
    double x = (d.cosr - z * d.z0) * d.xa;
    double ysq = 1.0 - z * z - x * x;
    if (ysq <= 0) return 0;
    double dphi = atan2(sqrt(ysq), x);
    if (!(dphi > 0)) return 0;
    i64 nr, ipix1;
    bool shifted;
    bfh_ring_info_small(h, iz, ipix1, nr, shifted);
    double shift = shifted ? 0.5 : 0.0;
    i64 ip_lo = (i64)floor((double)nr * BFH_INV_TWOPI * (d.phi - dphi) - shift) + 1;
    i64 ip_hi = (i64)floor((double)nr * BFH_INV_TWOPI * (d.phi + dphi) - shift);
    if (ip_hi >= nr) { ip_lo -= nr; ip_hi -= nr; }
    if (ip_lo < 0) {
        a0 = ipix1; a1 = ipix1 + ip_hi + 1;
        b0 = ipix1 + ip_lo + nr; b1 = ipix1 + nr;
        if (a1 < a0) a1 = a0;
        if (b0 <= a1) { a1 = (b1 > a1) ? b1 : a1; b0 = b1 = 0; }  // merged
        else if (b1 < b0) b1 = b0;
    } else {
        a0 = ipix1 + ip_lo; a1 = ipix1 + ip_hi + 1;
        if (a1 < a0) a1 = a0;
    }
    return (a1 - a0) + (b1 - b0);
}
