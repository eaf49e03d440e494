/* wave3d_cpu.h — C ABI of the native CPU solvers (libwave3d_cpu.so).
 *
 * Native equivalents of the reference's CPU programs, which are described in its
 * report but not shipped as files:
 *   wave.cpp (sequential)            report.pdf p.20 §5.1, p.11-12 §4.2.1
 *   openmpwave.cpp (OpenMP)          report.pdf p.21 §5.2
 *   mpi.cpp / onlyMPI.cpp (MPI)      report.pdf p.21-24 §5.3
 *   mpiomp.cpp (MPI+OpenMP)          report.pdf p.24-26 §5.4
 *
 * Block layout shared with the CUDA library: a local block of (nx,ny,nz)
 * unknowns is stored with one ghost layer on every side as
 *   u[((k*(ny+2) + j) * pitch) + xoff + i],   i in [0,nx+1], j in [0,ny+2), k in [0,nz+2)
 * so i=0 / i=nx+1 are ghosts (Dirichlet boundary nodes, or the neighbour's edge).
 */
#ifndef WAVE3D_CPU_H
#define WAVE3D_CPU_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct W3dBlock {
  int32_t n[3];     /* unknowns per axis: x, y, z */
  int32_t xoff;     /* element offset of the low x ghost inside a row */
  int64_t pitch;    /* row pitch in elements (>= xoff + n[0] + 2) */
} W3dBlock;

/* number of doubles a block needs */
int64_t w3d_block_elems(const W3dBlock* b);

/* u = sx[i]*sy[j]*sz[k] on every node incl. ghosts; tables have n+2 entries. */
void w3d_cpu_init(double* u, const W3dBlock* b, const double* sx, const double* sy,
                  const double* sz, int threads);

/* In-place step on the sub-box lo[d] <= idx < hi[d] (1-based local unknown indices,
 * i.e. 1..n[d] inclusive-exclusive as [lo,hi)):
 *   prev <- a*cur + b*prev + sum_d c[d]*(cur_{+1} - 2 cur + cur_{-1})
 * (Taylor first step: a=1,b=0,c=tau^2/(2h^2); leapfrog: a=2,b=-1,c=tau^2/h^2;
 * report.pdf p.5 §2.1-2.2).  The operation order is fixed (see stencil_core.h) so
 * that CPU and CUDA results are bit-identical. */
void w3d_cpu_step(double* prev_next, const double* cur, const W3dBlock* b, double a, double bb,
                  const double c[3], const int32_t lo[3], const int32_t hi[3], int threads);

/* max |u - sx*sy*(sz*ct)| and sum of squares over all unknowns of the block. */
void w3d_cpu_error(const double* u, const W3dBlock* b, const double* sx, const double* sy,
                   const double* sz, double ct, double* max_err, double* sum_sq, int threads);

/* Copy low/high faces of axis d into the opposite ghosts (periodic self-wrap). */
void w3d_cpu_wrap(double* u, const W3dBlock* b, int axis);

/* pack / unpack one face (axis d, side 0=low 1=high). ghost=0: the edge layer of
 * unknowns (what a neighbour needs); ghost=1: the ghost layer. Buffer is
 * contiguous, fastest index first, over the *unknowns* of the two other axes. */
int64_t w3d_cpu_face_elems(const W3dBlock* b, int axis);
void w3d_cpu_pack(const double* u, const W3dBlock* b, int axis, int side, int ghost, double* buf);
void w3d_cpu_unpack(double* u, const W3dBlock* b, int axis, int side, int ghost, const double* buf);

/* Whole-run drivers ------------------------------------------------------- */
typedef struct W3dCpuRun {
  int32_t N[3];        /* intervals per axis */
  double  L[3];
  double  tau;
  int32_t K;
  int32_t periodic[3]; /* 0 Dirichlet, 1 periodic */
  int32_t threads;     /* OpenMP threads per rank (1 = sequential) */
  int32_t dims[3];     /* rank grid (1,1,1 = single block) */
  int32_t every;       /* report every n-th step */
  int32_t verbose;     /* print "Step ..." lines to stdout like the reference */
} W3dCpuRun;

/* phases: [total, init, compute, boundary, exchange] seconds (max over ranks),
 * same columns as report.pdf p.16 §4.4.  trace: K+1 rows of (max, l2), row n =
 * time level n (rows not divisible by `every` hold -1).  Returns 0 on success. */
int w3d_cpu_solve(const W3dCpuRun* run, double* trace /* 2*(K+1) */, double* phases /* 5 */);

const char* w3d_cpu_version(void);
int w3d_cpu_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
