// session.cu — time stepper and streaming session on top of the C ABI in wave3d.h.
//
// The reference's MPI+CUDA program drives its GPU from a host loop that copies halos in and out
// every step and prints the error as the run goes (report.pdf p.15-16 §4.3, p.16 §4.4 "H2D/D2H"
// column).  What is left of that host loop here — choose the arrays of the next launch, feed the
// step's time factor from pinned memory, read the step's error pair back — lives in native code so
// that a step-by-step user is bound by the GPU and not by an interpreter: at 512^3 on eight GPUs a
// pass is 105 us.
//
// Nothing here touches plan internals; it is written against the public entry points only.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>
#include <vector>

#include "wave3d.h"

namespace {
inline void coef_of(int n, double* a, double* b, double* cs) {
  // Taylor first step (report.pdf p.5 §2.2 item 2), leapfrog afterwards (item 3)
  *a = n == 1 ? 1.0 : 2.0;
  *b = n == 1 ? 0.0 : -1.0;
  *cs = n == 1 ? 0.5 : 1.0;
}
}  // namespace

struct W3dSession {
  W3dSessionDesc d;
  W3dStepper st;
  bool pairs;
  std::vector<cudaEvent_t> events;     // one per step index; a pair launch records the later one
  std::vector<int> event_of;           // step -> index into events, -1 = not launched
};

extern "C" {

void w3d_stepper_reset(W3dStepper* st) {
  if (!st) return;
  st->level = 0; st->cur = 0; st->prev = 1; st->reserved = 0;
}

int w3d_advance_pair(W3dPlan* plan, W3dStepper* st, const double* ct, double* err1, double* err2, double* err_fix,
                     void* stream) {
  if (!plan || !st || !ct) return W3D_EINVAL;
  const int n = st->level, cur = st->cur, prev = st->prev;
  // array roles rotate over the four arrays: (0,1)->(2,3), (3,2)->(0,1), (1,0)->(2,3), ...
  const int o1 = cur <= 1 ? 2 : 0, o2 = cur <= 1 ? 3 : 1;
  double a1, b1, c1, a2, b2, c2;
  coef_of(n + 1, &a1, &b1, &c1);
  coef_of(n + 2, &a2, &b2, &c2);
  int rc = w3d_step_pair(plan, n, cur, prev, o1, o2, a1, b1, c1, ct[n + 1], a2, b2, c2, ct[n + 2], err1, err2, stream);
  if (rc != W3D_OK) return rc;
  if (w3d_plan_pair_mode(plan) == W3D_PAIR_FIXUP) {
    rc = w3d_step_fixup(plan, n + 1, o1, cur, o2, a2, b2, c2, ct[n + 2], err_fix, stream);
    if (rc != W3D_OK) return rc;
  }
  st->level = n + 2; st->cur = o2; st->prev = o1;
  return W3D_OK;
}

int w3d_advance_single(W3dPlan* plan, W3dStepper* st, const double* ct, double* err, void* stream) {
  if (!plan || !st || !ct) return W3D_EINVAL;
  const int n = st->level + 1;
  double a, b, cs;
  coef_of(n, &a, &b, &cs);
  const int rc = w3d_plan_pairs_enabled(plan)
                     ? w3d_step_fused_on(plan, n - 1, st->cur, st->prev, a, b, cs, ct[n], err, stream)
                     : w3d_step_fused(plan, n - 1, a, b, cs, ct[n], err, stream);
  if (rc != W3D_OK) return rc;
  std::swap(st->cur, st->prev);
  st->level = n;
  return W3D_OK;
}

int w3d_session_create(const W3dSessionDesc* desc, W3dSession** out) {
  if (!desc || !out || !desc->plan || !desc->ct_host || !desc->ct_dev || !desc->err_dev || !desc->err_host ||
      desc->K < 1 || desc->slots < 1)
    return W3D_EINVAL;
  W3dSession* s = new (std::nothrow) W3dSession();
  if (!s) return W3D_ENOMEM;
  s->d = *desc;
  w3d_stepper_reset(&s->st);
  s->pairs = w3d_plan_pairs_enabled(desc->plan) != 0;
  s->events.assign((size_t)desc->K + 1, nullptr);
  s->event_of.assign((size_t)desc->K + 1, -1);
  for (auto& e : s->events)
    if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
      w3d_session_destroy(s);
      return W3D_ECUDA;
    }
  // the kernels read the time factor of step n from ct_dev[n], which the session refreshes every step
  const int rc = w3d_plan_set_ct_table(desc->plan, desc->ct_dev);
  if (rc != W3D_OK) { w3d_session_destroy(s); return rc; }
  *out = s;
  return W3D_OK;
}

void w3d_session_destroy(W3dSession* s) {
  if (!s) return;
  for (auto e : s->events)
    if (e) cudaEventDestroy(e);
  if (s->d.plan) w3d_plan_set_ct_table(s->d.plan, nullptr);
  delete s;
}

int w3d_session_begin(W3dSession* s, const W3dStepper* st) {
  if (!s) return W3D_EINVAL;
  if (st) s->st = *st; else w3d_stepper_reset(&s->st);
  std::fill(s->event_of.begin(), s->event_of.end(), -1);
  return W3D_OK;
}

int w3d_session_stepper(W3dSession* s, W3dStepper* out) {
  if (!s || !out) return W3D_EINVAL;
  *out = s->st;
  return W3D_OK;
}

int w3d_session_step_async(W3dSession* s, int n, void* stream) {
  if (!s || n < 1 || n > s->d.K) return W3D_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const W3dSessionDesc& d = s->d;
  // this step's input: 8 bytes from pinned host memory
  if (cudaMemcpyAsync(d.ct_dev + n, d.ct_host + n, sizeof(double), cudaMemcpyHostToDevice, st) != cudaSuccess) return W3D_ECUDA;
  const size_t row = (size_t)d.slots * 2;                      // doubles per step in the error tables
  int first = n, rc = W3D_OK;
  if (s->pairs && n == s->st.level + 2) {
    // second step of a pair: one launch produces both levels and both error pairs
    first = n - 1;
    double* fix = d.slots > 1 ? d.err_dev + n * row + 2 : nullptr;
    rc = w3d_advance_pair(d.plan, &s->st, d.ct_host, d.err_dev + (n - 1) * row, d.err_dev + n * row, fix, st);
  } else if (n == s->st.level + 1 && (!s->pairs || n == d.K)) {
    rc = w3d_advance_single(d.plan, &s->st, d.ct_host, d.err_dev + n * row, st);      // also: last step of an odd-length pair run
  } else if (s->pairs && n == s->st.level + 1) {
    return W3D_OK;                                             // first step of a pair: its input is uploaded, the launch follows
  } else {
    return W3D_EINVAL;                                         // steps must be fed in order
  }
  if (rc != W3D_OK) return rc;
  // this launch's results: back to pinned host memory, then an event the consumer waits on
  if (cudaMemcpyAsync(d.err_host + first * row, d.err_dev + first * row, (size_t)(n - first + 1) * row * sizeof(double),
                      cudaMemcpyDeviceToHost, st) != cudaSuccess)
    return W3D_ECUDA;
  if (cudaEventRecord(s->events[n], st) != cudaSuccess) return W3D_ECUDA;
  for (int q = first; q <= n; ++q) s->event_of[q] = n;
  return W3D_OK;
}

int w3d_session_result(W3dSession* s, int n, double out[2]) {
  if (!s || !out || n < 1 || n > s->d.K || s->event_of[n] < 0) return W3D_EINVAL;
  if (cudaEventSynchronize(s->events[s->event_of[n]]) != cudaSuccess) return W3D_ECUDA;
  const double* e = s->d.err_host + (size_t)n * s->d.slots * 2;
  double mx = 0.0, sum = 0.0;
  for (int q = 0; q < s->d.slots; ++q) { mx = std::max(mx, e[2 * q]); sum += e[2 * q + 1]; }
  out[0] = mx; out[1] = sum;
  return W3D_OK;
}

}  // extern "C"
