// Queries on solved profiles, batched on the device:
// TimeOptimalPathProfile::SampleIndexFromTime, GetPathParameterAndDerivatives,
// GetPreviousIndex, GetPreviousDiscreteValues (time_optimal_path_timing.cc:1497-1668)
// for Q query times per path.
//
// One CTA per (path, tile of 128 queries); thread 0 finds the path's low_idx_ /
// high_idx_ (:468-477) once, every thread then runs the reference's binary search for
// its own time and evaluates the closed form inside the segment in the reference's
// operation order (compiled with -fmad=false), so results equal the reference's bit
// for bit.  Per query: 8 B read + up to 7 x 8 + 8 B written; the per-path arrays
// (time, s, sd, sd2, sdd) are read through L1/L2.
#include "tpb_device.cuh"
#include "tpb_engine.h"

namespace tpb {
namespace {

constexpr int kQueryThreads = 128;

__global__ void __launch_bounds__(kQueryThreads)
k_profile_query(Dims dm, DeviceArrays a, ProfileQueryArgs p) {
  __shared__ int s_lo, s_hi;
  const int b = blockIdx.y;
  const int N = dm.N;
  const size_t o = static_cast<size_t>(b) * N;
  const double* time = a.time + o;
  const bool solved = a.status[b] == TPB_PATH_OK;
  if (threadIdx.x == 0 && solved) {
    int low = 0;
    while ((time[low] == time[low + 1]) && (low < N - 2)) low++;
    int high = N - 1;
    while ((time[high] == time[high - 1]) && (high >= low)) high--;
    s_lo = low;
    s_hi = high;
  }
  __syncthreads();
  const int qi = blockIdx.x * kQueryThreads + threadIdx.x;
  if (qi >= p.Q) return;
  const size_t out = static_cast<size_t>(b) * p.Q + qi;
  const double nan = quiet_nan();
  double rs = nan, rsd = nan, rsdd = nan, sk = nan, sdk = nan, sddk = nan, tk = nan;
  int prev = -1;
  int st = 3;  // bit 0: GetPathParameterAndDerivatives false, bit 1: GetPreviousIndex < 0
  if (solved) {
    const double* s = a.s + o;
    const double* sd = a.sd + o;
    const double* sd2 = a.sd2 + o;
    const double t = p.time[out];
    // SetSetupDone (:535-552): the profile's parameter range
    const double s_start = p.use_s_range ? a.s_start[b] : a.path_start[b];
    const double s_end = p.use_s_range ? a.s_end[b] : s_start + a.delta[b] * (N - 1);
    const double ds_ = (s_end - s_start) / (N - 1);
    const double inv_ds = 1.0 / ds_;
    // SampleIndexFromTime (:1497-1524)
    int k;
    if (t <= time[0]) {
      k = 0;
    } else if (t >= time[N - 1]) {
      k = N - 2;
    } else {
      int low = s_lo, high = s_hi;
      int mid = (low + high) / 2;
      while (t < time[mid] || t >= time[mid + 1]) {
        if (t < time[mid]) {
          high = mid;
        } else {
          low = mid;
        }
        mid = (low + high) / 2;
      }
      while ((time[mid] == time[mid + 1]) && (mid < N - 1)) mid++;
      k = mid;
    }
    // GetPathParameterAndDerivatives (:1550-1626)
    bool ok = true;
    if (t <= time[0]) {
      rs = s_start;
      rsd = sd[0];
      rsdd = 0.5 * inv_ds * (sd2[1] - sd2[0]);
    } else if (t >= time[N - 1]) {
      rs = s_end;
      rsd = sd[N - 1];
      rsdd = 0.0;
    } else if (time[k] == time[k + 1]) {
      rs = s[k + 1];
      rsd = sd[k + 1];
      rsdd = 0.5 * inv_ds * (sd2[k + 1] - sd2[k]);
    } else if (k > N - 2) {
      ok = false;
    } else {
      const double dt = t - time[k];
      const double sda = sd[k], sdb = sd[k + 1], sd2a = sd2[k], sd2b = sd2[k + 1];
      if (sda > 0 || sdb > 0) {
        double ds = sda * dt + dt * dt * 0.25 * inv_ds * (sd2b - sd2a);
        if (ds > ds_) ds = ds_;
        if (dt < 0 || ds < 0) {
          ok = false;
        } else {
          const double sn = s[k] + ds;
          rs = (s[k + 1] < sn) ? s[k + 1] : sn;  // std::min(s_[k] + ds, s_[k + 1])
          rsd = sqrt(sd2a + ds * inv_ds * (sd2b - sd2a));
          rsdd = 0.5 * inv_ds * (sd2b - sd2a);
        }
      } else {
        rs = s[k] + (s[k + 1] - s[k]) * dt / (time[k + 1] - time[k]);
        rsd = 0.0;
        rsdd = 0.0;
      }
    }
    st = ok ? 0 : 1;
    // GetPreviousIndex / GetPreviousDiscreteValues (:1628-1668)
    if (t < time[0]) {
      st |= 2;
    } else {
      prev = (t > time[N - 1]) ? N - 1 : k;
      sk = s[prev];
      sdk = sd[prev];
      sddk = a.sdd[o + prev];
      tk = time[prev];
    }
  }
  if (p.s) p.s[out] = rs;
  if (p.sd) p.sd[out] = rsd;
  if (p.sdd) p.sdd[out] = rsdd;
  if (p.prev_index) p.prev_index[out] = prev;
  if (p.sk) p.sk[out] = sk;
  if (p.sdk) p.sdk[out] = sdk;
  if (p.sddk) p.sddk[out] = sddk;
  if (p.tk) p.tk[out] = tk;
  if (p.status) p.status[out] = st;
}

// TimeOptimalPathProfile::SolutionSatisfiesConstraints (:492-518) for every path: the
// number of (sample, row) pairs whose value A sdd + B sd^2 lies outside [lower, upper] by
// more than kTiny.  One warp per path, lanes over samples; joint paths synthesise their
// 2 D rows from q', q'' as ConstraintSetup does (timeable_path_joint_spline.cc:320-343),
// everything else reads the stored rows.
__global__ void __launch_bounds__(128)
k_count_violations(Dims dm, DeviceArrays a, bool joint, double safety, int32_t* count) {
  const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (b >= dm.B) return;  // uniform per warp
  const int N = dm.N, D = dm.D, C = dm.C;
  const size_t o = static_cast<size_t>(b) * N;
  int bad = 0;
  if (a.status[b] == TPB_PATH_OK) {
    for (int i = lane; i < N; i += 32) {
      const double sddv = a.sdd[o + i], sd2v = a.sd2[o + i];
      if (joint) {
        for (int d = 0; d < D; ++d) {
          const double q1 = a.dq[(o + i) * D + d], q2 = a.ddq[(o + i) * D + d];
          const double am = a.amax[static_cast<size_t>(b) * D + d] * safety;
          const double vs = a.vmax[static_cast<size_t>(b) * D + d] * safety;
          const double v = q1 * sddv + q2 * sd2v;
          bad += ((v + kTiny < -am) || (v - kTiny > am)) ? 1 : 0;
          const double w = 0.0 * sddv + (q1 * q1) * sd2v;
          bad += ((w + kTiny < 0.0) || (w - kTiny > vs * vs)) ? 1 : 0;
        }
      } else {
        const double* c = a.cons + (o + i) * C * 4;
        for (int r = 0; r < C; ++r) {
          const double v = c[r * 4 + 0] * sddv + c[r * 4 + 1] * sd2v;
          bad += ((v + kTiny < c[r * 4 + 2]) || (v - kTiny > c[r * 4 + 3])) ? 1 : 0;
        }
      }
    }
  } else {
    bad = -1;  // "No valid solution." (kFailedPrecondition in the reference)
  }
  for (int s = 16; s > 0; s >>= 1) bad += __shfl_xor_sync(kFullMask, bad, s);
  if (lane == 0) count[b] = (a.status[b] == TPB_PATH_OK) ? bad : -1;
}

}  // namespace

int launch_count_violations(const Dims& d, const DeviceArrays& a, bool joint, double safety,
                            int32_t* count, cudaStream_t st) {
  k_count_violations<<<(d.B + 3) / 4, 128, 0, st>>>(d, a, joint, safety, count);
  return 1;
}

int launch_profile_query(const Dims& d, const DeviceArrays& a, const ProfileQueryArgs& p,
                         cudaStream_t st) {
  dim3 grid((p.Q + kQueryThreads - 1) / kQueryThreads, d.B);
  k_profile_query<<<grid, kQueryThreads, 0, st>>>(d, a, p);
  return 1;
}

}  // namespace tpb
