/* orbx_sincos.h — the ONE definition of cos/sin used by computeOrbDescriptor's rotation
 * (reference: src/ORBextractor.cc:112-113, `(float)cos(angle)`, `(float)sin(angle)` through host libm).
 *
 * Host libm's cosf/sinf are not reproducible across glibc versions and are not what CUDA's
 * cosf/sinf compute, so "bit-exact descriptors" needs a definition both sides can evaluate
 * identically (determinisation 2).  This header evaluates sin and cos in
 * double precision with a Cody-Waite pi/2 reduction and the classic fdlibm minimax kernels,
 * using ONLY separately rounded IEEE multiplies and adds (no fma, no contraction), then rounds
 * once to float.  Compiled for host (oracle, -ffp-contract=off) and device (explicit *_rn
 * intrinsics) it yields the same bits.  Valid for |x| <= 8 (angles are in [0, 2*pi)).
 */
#ifndef ORBX_SINCOS_H
#define ORBX_SINCOS_H

#if defined(__CUDA_ARCH__)
#define ORBX_HD __host__ __device__ __forceinline__
#define ORBX_DMUL(a, b) __dmul_rn((a), (b))
#define ORBX_DADD(a, b) __dadd_rn((a), (b))
#define ORBX_RINT(a) rint(a)
#else
#if defined(__CUDACC__)
#define ORBX_HD __host__ __device__ inline
#else
#define ORBX_HD static inline
#endif
#include <math.h>
#define ORBX_DMUL(a, b) ((a) * (b))
#define ORBX_DADD(a, b) ((a) + (b))
#define ORBX_RINT(a) nearbyint(a)
#endif

ORBX_HD void orbx_sincosf(float xf, float* s_out, float* c_out)
{
    const double x = (double)xf;
    /* k = nearest multiple of pi/2 */
    const double k = ORBX_RINT(ORBX_DMUL(x, 6.36619772367581382433e-01));
    /* r = x - k*pio2, pio2 split so that k*hi is exact for small k */
    double r = ORBX_DADD(x, -ORBX_DMUL(k, 1.57079632673412561417e+00));
    r = ORBX_DADD(r, -ORBX_DMUL(k, 6.07710050650619224932e-11));
    const double z = ORBX_DMUL(r, r);
    /* sin kernel */
    double ps = 1.58969099521155010221e-10;
    ps = ORBX_DADD(ORBX_DMUL(ps, z), -2.50507602534068634195e-08);
    ps = ORBX_DADD(ORBX_DMUL(ps, z), 2.75573137070700676789e-06);
    ps = ORBX_DADD(ORBX_DMUL(ps, z), -1.98412698298579493134e-04);
    ps = ORBX_DADD(ORBX_DMUL(ps, z), 8.33333333332248946124e-03);
    ps = ORBX_DADD(ORBX_DMUL(ps, z), -1.66666666666666324348e-01);
    const double sr = ORBX_DADD(r, ORBX_DMUL(ORBX_DMUL(r, z), ps));
    /* cos kernel */
    double pc = -1.13596475577881948265e-11;
    pc = ORBX_DADD(ORBX_DMUL(pc, z), 2.08757232129817482790e-09);
    pc = ORBX_DADD(ORBX_DMUL(pc, z), -2.75573143513906633035e-07);
    pc = ORBX_DADD(ORBX_DMUL(pc, z), 2.48015872894767294178e-05);
    pc = ORBX_DADD(ORBX_DMUL(pc, z), -1.38888888888741095749e-03);
    pc = ORBX_DADD(ORBX_DMUL(pc, z), 4.16666666666666019037e-02);
    const double cr = ORBX_DADD(ORBX_DADD(1.0, -ORBX_DMUL(0.5, z)), ORBX_DMUL(ORBX_DMUL(z, z), pc));
    const int q = ((int)k) & 3;
    double s, c;
    if (q == 0)      { s = sr;  c = cr;  }
    else if (q == 1) { s = cr;  c = -sr; }
    else if (q == 2) { s = -sr; c = -cr; }
    else             { s = -cr; c = sr;  }
    *s_out = (float)s;
    *c_out = (float)c;
}

#endif /* ORBX_SINCOS_H */
