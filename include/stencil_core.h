/* stencil_core.h — the one place where the per-node arithmetic is written down.
 *
 * Scheme: report.pdf p.5 §2.1-2.2 (7-point Laplacian; Taylor first step;
 * leapfrog).  Every implementation in this repository (C++ sequential / OpenMP /
 * rank-threaded, CUDA generic, CUDA TMA z-marching) evaluates a node with exactly
 * this sequence of fused multiply-adds, so their fields agree bit for bit and
 * the 1/2/4/8-GPU traces are digit-identical like the reference's 1- and 2-GPU
 * traces (report.pdf p.15-16 §4.3).
 *
 * The order is chosen for the z-marching kernel: everything except the z+1
 * neighbour is known when plane z arrives (`w3d_partial`), the last term is added
 * when plane z+1 arrives (`w3d_finish`).
 */
#ifndef WAVE3D_STENCIL_CORE_H
#define WAVE3D_STENCIL_CORE_H

#if defined(__CUDACC__)
#define W3D_HD __host__ __device__ __forceinline__
#else
#define W3D_HD static inline
#endif

#include <math.h>

W3D_HD double w3d_partial(double c, double p, double xm, double xp, double ym, double yp,
                          double zm, double a, double b, double cx, double cy, double cz) {
  double dx = fma(-2.0, c, xm + xp);
  double dy = fma(-2.0, c, ym + yp);
  double dz = fma(-2.0, c, zm);
  double acc = a * c;
  acc = fma(b, p, acc);
  acc = fma(cx, dx, acc);
  acc = fma(cy, dy, acc);
  acc = fma(cz, dz, acc);
  return acc;
}

W3D_HD double w3d_finish(double acc, double zp, double cz) { return fma(cz, zp, acc); }

W3D_HD double w3d_node(double c, double p, double xm, double xp, double ym, double yp, double zm,
                       double zp, double a, double b, double cx, double cy, double cz) {
  return w3d_finish(w3d_partial(c, p, xm, xp, ym, yp, zm, a, b, cx, cy, cz), zp, cz);
}

/* |u - sxy * szt| with sxy = sx[i]*sy[j], szt = sz[k]*cos(a_t t_n). */
W3D_HD double w3d_abs_err(double u, double sxy, double szt) { return fabs(fma(-sxy, szt, u)); }

#endif
