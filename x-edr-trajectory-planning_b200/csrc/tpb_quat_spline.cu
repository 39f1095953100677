// Quaternion B-spline evaluation: BSplineQ::EvalCurve (splines/bsplineq.cc:223-244) and
// BSplineQ::EvalCurveAndDerivative (:246-307) for many splines and parameters at once.
// FP64, -fmad=false.
//
//  k_quat_spline_eval<P, kDer>  one thread per (spline, parameter).  The spline's knots
//                               and control quaternions are staged in shared memory once
//                               per CTA; the cumulative basis (:309-333) comes from
//                               UpdateBasisAndDerivatives (bspline_base.cc:268-348, A2.3)
//                               unrolled in registers; every thread stores 32 (value) or
//                               64 (value + derivative) contiguous bytes.
//
// The arithmetic per evaluation is transcendental (P x (log, atan2, sin, cos, exp) for the
// value, three times that with the derivative), so the kernel is bound by the FP64 pipe,
// not by HBM: 32 / 64 bytes leave per evaluation.
#include <cuda_runtime.h>

#include <string>

#include "../../include/tpb.h"
#include "tpb_device.cuh"
#include "tpb_nvtx.h"
#include "tpb_quat.cuh"

namespace tpb {
namespace {

constexpr int kQuatThreads = 128;

struct QuatEvalArgs {
  int n;          // control points
  int S;          // splines
  long long E;    // parameters per spline
  const double *knots, *quats, *u, *u0, *du;
  double *out, *dout;
  int32_t* status;
};

__device__ __forceinline__ void store_quat4(double* p, const Quat& q) {
  // 32-byte aligned: p = base + 4 * index, base from cudaMalloc or a caller's array of rows
  double2* v = reinterpret_cast<double2*>(p);
  v[0] = make_double2(q.x, q.y);
  v[1] = make_double2(q.z, q.w);
}

template <int P, bool kDer>
__global__ void __launch_bounds__(kQuatThreads)
k_quat_spline_eval(QuatEvalArgs g) {
  extern __shared__ double smem[];
  const int s = path_of_block();
  if (s >= g.S) return;  // uniform per CTA
  const int nk = g.n + P + 1;
  double* s_knots = smem;          // [n + P + 1]
  double* s_q = smem + nk;         // [n][4]
  {
    const double* gk = g.knots + static_cast<size_t>(s) * nk;
    const double* gq = g.quats + static_cast<size_t>(s) * g.n * 4;
    for (int i = threadIdx.x; i < nk; i += blockDim.x) s_knots[i] = gk[i];
    for (int i = threadIdx.x; i < g.n * 4; i += blockDim.x) s_q[i] = gq[i];
  }
  __syncthreads();
  const long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= g.E) return;
  const size_t at = static_cast<size_t>(s) * g.E + e;
  const double u = g.u ? g.u[at] : g.u0[s] + static_cast<double>(e) * g.du[s];
  // range check of EvalCurve / EvalCurveAndDerivative (:229-232, :255-258); NaN fails it
  const bool in_range = (u >= s_knots[0]) && (u <= s_knots[nk - 1]);
  if (g.status) g.status[at] = in_range ? 0 : 11;
  if (!in_range) return;
  const int span = knot_span(s_knots, nk, P, u);
  double ders[(kDer ? 1 : 0) + 1][P + 1];
  basis_ders<P, kDer ? 1 : 0>(s_knots, span, u, ders);
  // cumulative basis functions that are not identically 0 or 1 (:309-333)
  double cum[P], dcum[P];
  cum[P - 1] = ders[0][P];
  if (kDer) dcum[P - 1] = ders[kDer ? 1 : 0][P];
#pragma unroll
  for (int i = P - 2; i >= 0; --i) {
    cum[i] = cum[i + 1] + ders[0][i + 1];
    if (kDer) dcum[i] = dcum[i + 1] + ders[kDer ? 1 : 0][i + 1];
  }
  const double* pts = s_q + static_cast<size_t>(span - P) * 4;
  if (!kDer) {
    Quat q = load_quat(pts);
#pragma unroll
    for (int i = 0; i < P; ++i) {
      const Quat rel = qmul(qinv(load_quat(pts + 4 * i)), load_quat(pts + 4 * (i + 1)));
      q = qmul(q, qpow(rel, cum[i]));
    }
    normalize_positive_real(q);
    store_quat4(g.out + at * 4, q);
    return;
  }
  // products of the relative control quaternions' powers (:263-272); pw[0] = identity
  Quat pw[P + 1], rel[P];
  pw[0] = Quat{0.0, 0.0, 0.0, 1.0};
#pragma unroll
  for (int k = 0; k < P; ++k) {
    rel[k] = qmul(qinv(load_quat(pts + 4 * k)), load_quat(pts + 4 * (k + 1)));
    pw[k + 1] = qmul(pw[k], qpow(rel[k], cum[k]));
  }
  const Quat p0 = load_quat(pts);
  Quat q = qmul(p0, pw[P]);
  // sum over the derivative of the product w.r.t. each factor (:277-294)
  Quat d{0.0, 0.0, 0.0, 0.0};
#pragma unroll
  for (int k = 0; k < P; ++k) {
    Quat dk = qpow_derivative(rel[k], cum[k]);
    dk.x *= dcum[k];
    dk.y *= dcum[k];
    dk.z *= dcum[k];
    dk.w *= dcum[k];
    const Quat t = qmul(qmul(qmul(pw[k], dk), qinv(pw[k + 1])), pw[P]);
    d.x += t.x;
    d.y += t.y;
    d.z += t.z;
    d.w += t.w;
  }
  d = qmul(p0, d);
  if (q.w < 0) {  // consistent signs for value and derivative (:299-304)
    q.x *= -1.0;
    q.y *= -1.0;
    q.z *= -1.0;
    q.w *= -1.0;
    const double z = sq_norm(q);
    if (z > 0) {
      const double n = sqrt(z);
      q.x /= n;
      q.y /= n;
      q.z /= n;
      q.w /= n;
    }
    d.x *= -1.0;
    d.y *= -1.0;
    d.z *= -1.0;
    d.w *= -1.0;
  }
  store_quat4(g.out + at * 4, q);
  store_quat4(g.dout + at * 4, d);
}

template <int P, bool kDer>
cudaError_t launch_one(const QuatEvalArgs& g, size_t smem, cudaStream_t st) {
  if (smem > 48 * 1024) {
    cudaError_t err = cudaFuncSetAttribute(k_quat_spline_eval<P, kDer>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           static_cast<int>(smem));
    if (err != cudaSuccess) return err;
  }
  const unsigned bx = static_cast<unsigned>((g.E + kQuatThreads - 1) / kQuatThreads);
  k_quat_spline_eval<P, kDer><<<path_grid(bx, g.S), kQuatThreads, smem, st>>>(g);
  return cudaGetLastError();
}

template <int P>
cudaError_t launch_degree(const QuatEvalArgs& g, size_t smem, cudaStream_t st) {
  return g.dout ? launch_one<P, true>(g, smem, st) : launch_one<P, false>(g, smem, st);
}

int32_t check(const tpb_quat_spline_eval_desc* d) {
  if (d == nullptr) return TPB_INVALID_ARGUMENT;
  if (d->degree < 1 || d->degree > 5) return TPB_OUT_OF_RANGE;
  if (d->num_points < d->degree + 1 || d->num_splines < 1 || d->evals_per_spline < 1)
    return TPB_INVALID_ARGUMENT;
  if (d->knots == nullptr || d->quats == nullptr || d->out == nullptr) return TPB_INVALID_ARGUMENT;
  if (d->u == nullptr && (d->u0 == nullptr || d->du == nullptr)) return TPB_INVALID_ARGUMENT;
  const size_t smem = (static_cast<size_t>(d->num_points) * 5 + d->degree + 1) * sizeof(double);
  if (smem > 200 * 1024) return TPB_OUT_OF_RANGE;
  return TPB_OK;
}

}  // namespace
}  // namespace tpb

extern "C" {

int32_t tpb_quat_spline_eval_device(const tpb_quat_spline_eval_desc* d, void* cuda_stream) {
  TPB_NVTX("tpb_quat_spline_eval");
  const int32_t rc = tpb::check(d);
  if (rc != TPB_OK) return rc;
  tpb::QuatEvalArgs g;
  g.n = d->num_points;
  g.S = d->num_splines;
  g.E = d->evals_per_spline;
  g.knots = d->knots;
  g.quats = d->quats;
  g.u = d->u;
  g.u0 = d->u0;
  g.du = d->du;
  g.out = d->out;
  g.dout = d->dout;
  g.status = d->status;
  const size_t smem = (static_cast<size_t>(g.n) * 5 + d->degree + 1) * sizeof(double);
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  cudaError_t err = cudaSuccess;
  switch (d->degree) {
    case 1: err = tpb::launch_degree<1>(g, smem, st); break;
    case 2: err = tpb::launch_degree<2>(g, smem, st); break;
    case 3: err = tpb::launch_degree<3>(g, smem, st); break;
    case 4: err = tpb::launch_degree<4>(g, smem, st); break;
    default: err = tpb::launch_degree<5>(g, smem, st); break;
  }
  return err == cudaSuccess ? TPB_OK : TPB_INTERNAL;
}

int32_t tpb_quat_spline_eval_host(const tpb_quat_spline_eval_desc* d, int32_t device) {
  const int32_t rc = tpb::check(d);
  if (rc != TPB_OK) return rc;
  if (cudaSetDevice(device) != cudaSuccess) return TPB_INTERNAL;
  const size_t S = d->num_splines, E = static_cast<size_t>(d->evals_per_spline);
  const size_t nk = d->num_points + d->degree + 1, nq = static_cast<size_t>(d->num_points) * 4;
  tpb_quat_spline_eval_desc dd = *d;
  void* bufs[8] = {nullptr};
  int nb = 0;
  bool ok = true;
  auto up = [&](const void* src, size_t bytes) -> void* {
    if (src == nullptr || !ok) return nullptr;
    void* p = nullptr;
    ok = cudaMalloc(&p, bytes) == cudaSuccess;
    if (!ok) return nullptr;
    bufs[nb++] = p;
    ok = cudaMemcpy(p, src, bytes, cudaMemcpyHostToDevice) == cudaSuccess;
    return p;
  };
  auto dev = [&](bool wanted, size_t bytes) -> void* {
    if (!wanted || !ok) return nullptr;
    void* p = nullptr;
    ok = cudaMalloc(&p, bytes) == cudaSuccess;
    if (ok) bufs[nb++] = p;
    return p;
  };
  dd.knots = static_cast<const double*>(up(d->knots, S * nk * sizeof(double)));
  dd.quats = static_cast<const double*>(up(d->quats, S * nq * sizeof(double)));
  dd.u = static_cast<const double*>(up(d->u, S * E * sizeof(double)));
  dd.u0 = static_cast<const double*>(up(d->u0, S * sizeof(double)));
  dd.du = static_cast<const double*>(up(d->du, S * sizeof(double)));
  // rows of out-of-range parameters stay as the caller left them: upload, then download
  dd.out = static_cast<double*>(up(d->out, S * E * 4 * sizeof(double)));
  dd.dout = static_cast<double*>(up(d->dout, S * E * 4 * sizeof(double)));
  dd.status = static_cast<int32_t*>(dev(d->status != nullptr, S * E * sizeof(int32_t)));
  int32_t out_rc = ok ? tpb_quat_spline_eval_device(&dd, nullptr) : TPB_INTERNAL;
  if (out_rc == TPB_OK) {
    ok = cudaMemcpy(d->out, dd.out, S * E * 4 * sizeof(double), cudaMemcpyDeviceToHost) == cudaSuccess;
    if (ok && d->dout)
      ok = cudaMemcpy(d->dout, dd.dout, S * E * 4 * sizeof(double), cudaMemcpyDeviceToHost) ==
           cudaSuccess;
    if (ok && d->status)
      ok = cudaMemcpy(d->status, dd.status, S * E * sizeof(int32_t), cudaMemcpyDeviceToHost) ==
           cudaSuccess;
    if (!ok) out_rc = TPB_INTERNAL;
  }
  for (int i = 0; i < nb; ++i) cudaFree(bufs[i]);
  return out_rc;
}

}  // extern "C"
