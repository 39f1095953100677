// Batched, device-resident TrajectoryBuffers (trajectory_buffer.cc:40-393,
// trajectory_buffer.h:30-184): the consumer format of the resampled
// trajectories.  B buffers of `cap` row slots each,
//
//   t [B][cap]      p, v, a [B][cap][D]      head, count, seq [B]
//
// buffer b holding rows head[b] .. head[b]+count[b]-1.  DiscardSegmentBefore moves
// head (no data moves); InsertSegment writes behind the kept rows and moves them to
// slot 0 only when the slots behind them do not suffice.
//
// Kernels:
//   k_tbuf_insert   1 CTA / buffer: thread 0 finds the insertion point (the
//                   reference's upper_bound with `<=` + timestep tolerance), the CTA
//                   streams the segment's rows in (contiguous on both sides: a pure
//                   HBM copy of (3D+1)*8 B per row read and written)
//   k_tbuf_sample   1 warp / 32 queries: bracket by a two-level search (pivots of the
//                   time stamps in shared memory), linear interpolation, coalesced writes
//   k_tbuf_discard, k_tbuf_append, k_tbuf_upto, k_tbuf_offset, k_tbuf_state,
//   k_tbuf_clear    1 thread / buffer (or row): bookkeeping
//   k_tbuf_stop_prepare -> k_rescale_to_stop (tpb_stop.cu, in place on rows
//   0..index of every selected buffer, results in scratch arrays of the same layout)
//   -> k_tbuf_stop_commit (1 CTA / buffer: velocity match test, InsertSegment of the
//   stopping segment whose positions are rows of the buffer itself)
//
// Results are bit-identical to the reference's for everything but
// eigenmath::InterpolateLinear (un-vendored, see tpb.h); -fmad=false keeps the
// interpolation un-contracted like the oracle's.
#include <cstdlib>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <string>
#include <vector>

#include "tpb_device.cuh"
#include "tpb_engine.h"

struct tpb_tbuf {
  int device = 0;
  int B = 0, D = 0, cap = 0;
  double tol = 1e-6;
  cudaStream_t stream = nullptr;
  double *t = nullptr, *p = nullptr, *v = nullptr, *a = nullptr;
  int32_t *head = nullptr, *count = nullptr, *seq = nullptr;
  // stop scratch (allocated by the first stop call)
  double *s_t = nullptr, *s_v = nullptr, *s_a = nullptr;
  int64_t* s_begin = nullptr;
  int32_t *s_len = nullptr, *s_idx = nullptr, *s_pending = nullptr, *s_count = nullptr,
          *s_switch = nullptr, *s_rstatus = nullptr;
  int64_t launches = 0;
  std::string error;
};

namespace tpb {
namespace {

constexpr int kInsertThreads = 256;
constexpr int32_t kPending = -1;
constexpr int32_t kSkipIndex = INT32_MIN;

struct TbufDev {
  int B, D, cap;
  double tol;
  double *t, *p, *v, *a;
  int32_t *head, *count, *seq;
};

// first i in [0, n) with x <= t[i] (c_upper_bound with the `a <= b` comparator,
// trajectory_buffer.cc:104-106, 161-162); n if none
__device__ __forceinline__ int first_ge(const double* __restrict__ t, int n, double x) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = lo + ((hi - lo) >> 1);
    if (x <= t[mid]) hi = mid; else lo = mid + 1;
  }
  return lo;
}
// first i with x < t[i] (c_upper_bound, :223, 241)
__device__ __forceinline__ int first_gt(const double* __restrict__ t, int n, double x) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = lo + ((hi - lo) >> 1);
    if (x < t[mid]) hi = mid; else lo = mid + 1;
  }
  return lo;
}
// first i with t[i] >= x (c_lower_bound, :376)
__device__ __forceinline__ int lower_bound(const double* __restrict__ t, int n, double x) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = lo + ((hi - lo) >> 1);
    if (t[mid] < x) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// eigenmath::InterpolateLinear(x, x0, x1, y0, y1) -- un-vendored, same statement as
// the oracle's
__device__ __forceinline__ double interpolate(double x, double x0, double x1, double y0,
                                              double y1) {
  if (x1 == x0) return y0;
  const double w = (x - x0) / (x1 - x0);
  return y0 + w * (y1 - y0);
}

// GetOffsetBracket (:232-251) on rows T[0..n)
__device__ __forceinline__ int bracket(const double* __restrict__ T, int n, double x, int* i0,
                                       int* i1) {
  if (n == 0) return TPB_FAILED_PRECONDITION;
  if (x < T[0] || x > T[n - 1]) return TPB_OUT_OF_RANGE;
  const int upper = first_gt(T, n, x);
  if (upper == n) {
    *i0 = *i1 = n - 1;
  } else {
    *i0 = upper - 1;
    *i1 = upper;
  }
  return TPB_OK;
}

// Moves n doubles from src to dst inside one array (any overlap), by the whole CTA.
__device__ void block_move(double* dst, const double* src, long n) {
  if (dst == src || n <= 0) return;
  const int T = blockDim.x;
  if (dst < src) {
    for (long base = 0; base < n; base += T) {
      const long i = base + threadIdx.x;
      double x = 0.0;
      if (i < n) x = src[i];
      __syncthreads();
      if (i < n) dst[i] = x;
      __syncthreads();
    }
  } else {
    for (long base = n; base > 0; base -= T) {
      const long i = base - 1 - threadIdx.x;
      double x = 0.0;
      if (i >= 0) x = src[i];
      __syncthreads();
      if (i >= 0) dst[i] = x;
      __syncthreads();
    }
  }
}

// Streams n doubles from src to dst (distinct arrays), four loads in flight per thread.
__device__ __forceinline__ void block_copy(double* __restrict__ dst,
                                           const double* __restrict__ src, long n) {
  const int T = blockDim.x;
  long i = threadIdx.x;
  for (; i + 3L * T < n; i += 4L * T) {
    const double x0 = src[i], x1 = src[i + T], x2 = src[i + 2L * T], x3 = src[i + 3L * T];
    dst[i] = x0;
    dst[i + T] = x1;
    dst[i + 2L * T] = x2;
    dst[i + 3L * T] = x3;
  }
  for (; i < n; i += T) dst[i] = src[i];
}

__global__ void k_tbuf_clear(TbufDev d, const int32_t* __restrict__ select) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= d.B) return;
  if (select != nullptr && select[b] == 0) return;
  d.head[b] = 0;
  d.count[b] = 0;
  d.seq[b] = 0;
}

// InsertSegment (:80-132)
__global__ void __launch_bounds__(kInsertThreads)
k_tbuf_insert(TbufDev d, const int64_t* __restrict__ begin, const int32_t* __restrict__ length,
              const double* __restrict__ times, const double* __restrict__ pos,
              const double* __restrict__ vel, const double* __restrict__ acc,
              int32_t* __restrict__ status) {
  const int b = blockIdx.x;
  const int n = length[b];
  __shared__ int sh_keep, sh_head_old, sh_head_new, sh_go;
  if (threadIdx.x == 0) {
    int st = TPB_OK, go = 0, keep = 0, head_old = d.head[b], head_new = head_old;
    if (n >= 0) {
      int seq = d.seq[b] + 1;  // :97
      if (n > 0) {
        const int cnt = d.count[b];
        const double* T = d.t + static_cast<size_t>(b) * d.cap + head_old;
        const double x = times[begin[b]];
        int upper = first_ge(T, cnt, x);
        bool reset = false;
        if (cnt == 0 || upper == 0) {  // :107-114
          reset = true;
        } else if (x - T[upper - 1] < d.tol) {  // :118-120
          upper--;
        }
        keep = reset ? 0 : upper;
        if (keep + n > d.cap) {
          st = TPB_RESOURCE_EXHAUSTED;
          seq -= 1;  // buffer unchanged
        } else {
          go = 1;
          if (keep == 0 || head_old + keep + n > d.cap) head_new = 0;
          if (reset) seq = 0;
          d.count[b] = keep + n;
          d.head[b] = head_new;
        }
      }
      d.seq[b] = seq;
    }
    if (status != nullptr) status[b] = st;
    sh_keep = keep;
    sh_head_old = head_old;
    sh_head_new = head_new;
    sh_go = go;
  }
  __syncthreads();
  if (!sh_go) return;
  const int keep = sh_keep, D = d.D;
  const size_t base = static_cast<size_t>(b) * d.cap;
  if (sh_head_new != sh_head_old && keep > 0) {  // make room: kept rows to slot 0
    const size_t o = base + sh_head_old, w = base + sh_head_new;
    block_move(d.t + w, d.t + o, keep);
    block_move(d.p + w * D, d.p + o * D, static_cast<long>(keep) * D);
    block_move(d.v + w * D, d.v + o * D, static_cast<long>(keep) * D);
    block_move(d.a + w * D, d.a + o * D, static_cast<long>(keep) * D);
  }
  const size_t dst = base + sh_head_new + keep, src = static_cast<size_t>(begin[b]);
  block_copy(d.t + dst, times + src, n);
  block_copy(d.p + dst * D, pos + src * D, static_cast<long>(n) * D);
  block_copy(d.v + dst * D, vel + src * D, static_cast<long>(n) * D);
  block_copy(d.a + dst * D, acc + src * D, static_cast<long>(n) * D);
}

// AppendSample (:134-148)
__global__ void k_tbuf_append(TbufDev d, const int32_t* __restrict__ select,
                              const double* __restrict__ time, const double* __restrict__ pos,
                              const double* __restrict__ vel, const double* __restrict__ acc,
                              int32_t* __restrict__ status) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= d.B) return;
  int st = TPB_OK;
  if (select == nullptr || select[b] != 0) {
    const int cnt = d.count[b], D = d.D;
    int head = d.head[b];
    const size_t base = static_cast<size_t>(b) * d.cap;
    const double x = time[b];
    if (cnt > 0 && d.t[base + head + cnt - 1] >= x) {
      st = TPB_INVALID_ARGUMENT;
    } else if (cnt + 1 > d.cap) {
      st = TPB_RESOURCE_EXHAUSTED;
    } else {
      if (cnt == 0) head = 0;
      if (head + cnt + 1 > d.cap) {  // rows to slot 0 (rare; this thread alone)
        for (int i = 0; i < cnt; ++i) {
          d.t[base + i] = d.t[base + head + i];
          for (int j = 0; j < D; ++j) {
            d.p[(base + i) * D + j] = d.p[(base + head + i) * D + j];
            d.v[(base + i) * D + j] = d.v[(base + head + i) * D + j];
            d.a[(base + i) * D + j] = d.a[(base + head + i) * D + j];
          }
        }
        head = 0;
      }
      const size_t r = base + head + cnt;
      d.t[r] = x;
      for (int j = 0; j < D; ++j) {
        d.p[r * D + j] = pos[static_cast<size_t>(b) * D + j];
        d.v[r * D + j] = vel[static_cast<size_t>(b) * D + j];
        d.a[r * D + j] = acc[static_cast<size_t>(b) * D + j];
      }
      d.head[b] = head;
      d.count[b] = cnt + 1;
    }
  }
  if (status != nullptr) status[b] = st;
}

// DiscardSegmentBefore(double) (:154-207)
__global__ void k_tbuf_discard(TbufDev d, const double* __restrict__ time_sec) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= d.B) return;
  const double x = time_sec[b];
  if (x != x) return;  // skip
  const int n = d.count[b];
  if (n == 0) return;
  const int head = d.head[b], D = d.D;
  const size_t base = static_cast<size_t>(b) * d.cap + head;
  double* T = d.t + base;
  if (x <= T[0]) return;
  if (x > T[n - 1]) {  // Clear()
    d.head[b] = 0;
    d.count[b] = 0;
    d.seq[b] = 0;
    return;
  }
  int offset = first_ge(T, n, x);
  if (offset <= 0) return;
  const bool close_to_existing = x - T[offset - 1] <= d.tol;
  const bool create_by_interpolation = fabs(T[offset] - x) > d.tol;
  if (close_to_existing || create_by_interpolation) --offset;
  if (create_by_interpolation) {
    // the bracket of x is (offset, offset + 1) here: T[offset] < x < T[offset + 1]
    const double x0 = T[offset], x1 = T[offset + 1];
    double* P = d.p + (base + offset) * D;
    double* V = d.v + (base + offset) * D;
    double* A = d.a + (base + offset) * D;
    for (int j = 0; j < D; ++j) {
      P[j] = interpolate(x, x0, x1, P[j], P[D + j]);
      V[j] = interpolate(x, x0, x1, V[j], V[D + j]);
      A[j] = interpolate(x, x0, x1, A[j], A[D + j]);
    }
    T[offset] = x;
  }
  d.head[b] = head + offset;
  d.count[b] = n - offset;
}

// GetPositionsUpToTime (:209-226)
__global__ void k_tbuf_upto(TbufDev d, const double* __restrict__ time_sec,
                            int32_t* __restrict__ num) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= d.B) return;
  const int n = d.count[b];
  const double* T = d.t + static_cast<size_t>(b) * d.cap + d.head[b];
  const double x = time_sec[b];
  int r = 0;
  if (n > 0 && !(x < T[0] || x > T[n - 1]) && x == x) r = first_gt(T, n, x) - 1;
  num[b] = r;
}

// Get{Position,Velocity,Acceleration}AtTime (:228-293).  A warp takes 32 consecutive
// queries of one buffer: lane l brackets query l, then the warp walks the 32*D output
// values of each requested array in order (coalesced writes; the bracket and weight of
// value e come from lane e / D by shuffle).
//
// The bracket is c_upper_bound in the buffer's time stamps, which are non-decreasing, so
// any search finds the same index.  The CTA (all its warps serve the same buffer) first
// stages every S-th time stamp in shared memory, S the smallest power of two that leaves
// at most kSamplePivots of them; a query is then located among the pivots without leaving
// the SM and needs log2(S) dependent global loads instead of log2(n).  The gather keeps
// the loads of kSampleBatch output values per lane in flight before the first store.
// (60 % -> 87 % of the measured HBM peak on the bench's 8.4 M queries.)
constexpr int kSamplePivots = 1024;
template <int kSampleBatch, int kSampleWarps, int kMinBlocks>
__global__ void __launch_bounds__(kSampleWarps * 32, kMinBlocks)
k_tbuf_sample(TbufDev d, int Q, const double* __restrict__ time_sec,
              double* __restrict__ out_p, double* __restrict__ out_v,
              double* __restrict__ out_a, int32_t* __restrict__ status) {
  __shared__ double s_pivot[kSamplePivots];
  const int D = d.D;
  const int b = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int q0 = (blockIdx.x * kSampleWarps + (threadIdx.x >> 5)) * 32;
  const int n = d.count[b];
  const size_t base = static_cast<size_t>(b) * d.cap + d.head[b];
  const double* __restrict__ T = d.t + base;
  int shift = 0;
  while (((n + (1 << shift) - 1) >> shift) > kSamplePivots) ++shift;
  const int npiv = (n + (1 << shift) - 1) >> shift;
  for (int k = threadIdx.x; k < npiv; k += kSampleWarps * 32) s_pivot[k] = T[k << shift];
  __syncthreads();
  if (q0 >= Q) return;
  const int nq = min(32, Q - q0);
  const size_t qrow = static_cast<size_t>(b) * Q + q0;
  int st = TPB_OUT_OF_RANGE, i0 = 0, i1 = 0;
  double w = 0.0;
  if (lane < nq) {
    const double x = time_sec[qrow + lane];
    if (x == x) {
      // GetOffsetBracket (:232-251)
      if (n == 0) {
        st = TPB_FAILED_PRECONDITION;
      } else if (!(x < s_pivot[0] || x > T[n - 1])) {
        st = TPB_OK;
        // first pivot above x (pivot 0 is not: x >= T[0]) ...
        int lo = 1, hi = npiv;
        while (lo < hi) {
          const int mid = lo + ((hi - lo) >> 1);
          if (x < s_pivot[mid]) hi = mid; else lo = mid + 1;
        }
        // ... then T[(lo-1) << shift] <= x and the first time stamp above x is in
        // ((lo-1) << shift, lo << shift] (or n)
        int glo = ((lo - 1) << shift) + 1, ghi = min(lo << shift, n);
        while (glo < ghi) {
          const int mid = glo + ((ghi - glo) >> 1);
          if (x < T[mid]) ghi = mid; else glo = mid + 1;
        }
        const int upper = glo;
        if (upper == n) {
          i0 = i1 = n - 1;
        } else {
          i0 = upper - 1;
          i1 = upper;
        }
        const double x0 = T[i0], x1 = T[i1];
        if (x1 == x0) {
          i1 = i0;  // InterpolateLinear returns y0
        } else {
          w = (x - x0) / (x1 - x0);
        }
      }
    }
    if (status != nullptr) status[qrow + lane] = st;
  }
  const double* __restrict__ src[3] = {d.p, d.v, d.a};
  double* __restrict__ dst[3] = {out_p, out_v, out_a};
  const int total = nq * D;
  for (int e0 = 0; e0 < total; e0 += 32 * kSampleBatch) {
    double y0[kSampleBatch][3], y1[kSampleBatch][3], wq[kSampleBatch];
    bool act[kSampleBatch], okq[kSampleBatch], same[kSampleBatch];
#pragma unroll
    for (int u = 0; u < kSampleBatch; ++u) {
      const int e = e0 + 32 * u + lane;
      const bool active = e < total;
      const int qq = active ? e / D : 0;
      const int j = e - qq * D;
      const int st_q = __shfl_sync(kFullMask, st, qq);
      const int i0_q = __shfl_sync(kFullMask, i0, qq);
      const int i1_q = __shfl_sync(kFullMask, i1, qq);
      wq[u] = __shfl_sync(kFullMask, w, qq);
      act[u] = active;
      okq[u] = active && st_q == TPB_OK;
      same[u] = i1_q == i0_q;
      const size_t r0 = (base + i0_q) * D + j, r1 = (base + i1_q) * D + j;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        y0[u][k] = quiet_nan();
        y1[u][k] = quiet_nan();
        if (okq[u] && dst[k] != nullptr) {
          y0[u][k] = src[k][r0];
          y1[u][k] = src[k][r1];  // r1 == r0 when the bracket is one sample
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kSampleBatch; ++u) {
      if (!act[u]) continue;
      const size_t o = qrow * D + e0 + 32 * u + lane;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        if (dst[k] == nullptr) continue;
        const double a0 = y0[u][k], a1 = y1[u][k];
        // NaN rows for queries outside the buffer; y0 itself for a one-sample bracket
        dst[k][o] = (okq[u] && !same[u]) ? a0 + wq[u] * (a1 - a0) : a0;
      }
    }
  }
}

// AddOffsetToTimestamps (:386-389)
__global__ void k_tbuf_offset(TbufDev d, const double* __restrict__ offset) {
  const int b = blockIdx.y;
  const int n = d.count[b];
  double* T = d.t + static_cast<size_t>(b) * d.cap + d.head[b];
  const double off = offset[b];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    T[i] = T[i] + off;
}

__global__ void k_tbuf_state(TbufDev d, int32_t* __restrict__ num, int32_t* __restrict__ seq,
                             double* __restrict__ front, double* __restrict__ back) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= d.B) return;
  const int n = d.count[b];
  const double* T = d.t + static_cast<size_t>(b) * d.cap + d.head[b];
  if (num != nullptr) num[b] = n;
  if (seq != nullptr) seq[b] = d.seq[b];
  if (front != nullptr) front[b] = n > 0 ? T[0] : quiet_nan();
  if (back != nullptr) back[b] = n > 0 ? T[n - 1] : quiet_nan();
}

// StopAtIndex (:295-331) / StopBeforeTime (:366-381) up to the call of
// RescaleTrajectoryBackwardToStop: argument checks, the "already standing at the
// last sample" case, and the (begin, length) of the rows to rescale.
__global__ void k_tbuf_stop_prepare(TbufDev d, const int32_t* __restrict__ index,
                                    const double* __restrict__ time_sec,
                                    const double* __restrict__ max_acc, double time_step,
                                    int64_t* __restrict__ begin, int32_t* __restrict__ length,
                                    int32_t* __restrict__ idx_out,
                                    int32_t* __restrict__ pending) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= d.B) return;
  const int n = d.count[b], D = d.D;
  const size_t base = static_cast<size_t>(b) * d.cap + d.head[b];
  const double* T = d.t + base;
  int st = kPending, idx = 0;
  if (index != nullptr) {
    idx = index[b];
    if (idx == kSkipIndex) st = TPB_OK;
  } else {
    const double x = time_sec[b];
    if (x != x || n == 0) {
      st = TPB_OK;  // skipped / nothing to stop (:369-371)
    } else if (x < T[0]) {
      st = TPB_OUT_OF_RANGE;
    } else {
      const int lower = lower_bound(T, n, x);
      idx = min(lower + 1, n - 1);
    }
  }
  if (st == kPending) {
    // (`index > GetNumSamples() - 1` is an unsigned comparison in the reference: an
    // empty buffer only rejects index <= 0 and then reads out of bounds)
    if (idx <= 0 || (n > 0 && idx > n - 1)) {
      st = TPB_OUT_OF_RANGE;
    } else if (n == 0) {
      st = TPB_INTERNAL;
    } else {
      double amin = max_acc[static_cast<size_t>(b) * D];
      for (int j = 1; j < D; ++j) {
        const double x = max_acc[static_cast<size_t>(b) * D + j];
        amin = x < amin ? x : amin;
      }
      if (amin <= 0.0 || time_step <= 0.0) {
        st = TPB_INVALID_ARGUMENT;
      } else if (idx == n - 1) {
        double vinf = 0.0;
        for (int j = 0; j < D; ++j) {
          const double x = fabs(d.v[(base + idx) * D + j]);
          vinf = x > vinf ? x : vinf;
        }
        if (vinf < 1e-4) {  // kVerySmall (:315-321)
          for (int j = 0; j < D; ++j) {
            d.v[(base + idx) * D + j] = 0.0;
            d.a[(base + idx) * D + j] = 0.0;
          }
          st = TPB_OK;
        }
      }
    }
  }
  begin[b] = static_cast<int64_t>(base);
  length[b] = st == kPending ? idx + 1 : 0;
  idx_out[b] = idx;
  pending[b] = st;
}

// StopAtIndex after RescaleTrajectoryBackwardToStop (:333-357): the velocity match
// test and InsertSegment of the stopping segment (scratch rows sw .. sw+m-1 of the
// buffer; its positions are the buffer's own rows sw .. sw+m-1).
__global__ void __launch_bounds__(kInsertThreads)
k_tbuf_stop_commit(TbufDev d, const int32_t* __restrict__ pending,
                   const int32_t* __restrict__ idx_in, const int32_t* __restrict__ rcount,
                   const int32_t* __restrict__ rswitch, const int32_t* __restrict__ rstatus,
                   const double* __restrict__ s_t, const double* __restrict__ s_v,
                   const double* __restrict__ s_a, int32_t* __restrict__ status) {
  const int b = blockIdx.x;
  const int D = d.D;
  __shared__ int sh_go, sh_keep, sh_head_old, sh_head_new;
  const int m = rcount[b], sw = rswitch[b];
  if (threadIdx.x == 0) {
    int st = pending[b], go = 0, keep = 0;
    const int head_old = d.head[b];
    int head_new = head_old;
    if (st == kPending) {
      const int n = d.count[b], idx = idx_in[b];
      const size_t base = static_cast<size_t>(b) * d.cap + head_old;
      const double* T = d.t + base;
      st = TPB_OK;
      if (rstatus[b] != TPB_PATH_OK) {
        st = TPB_INVALID_ARGUMENT;  // AreInputsValidForSampledTrajectory
      } else if (m == 0) {
        st = TPB_INTERNAL;  // CHECK(!rescaled_stop.times.empty()) in the reference
      } else {
        const double x = s_t[base + sw];
        if (m == idx) {  // every sample used: the initial velocity must match (:342-355)
          int i0 = 0, i1 = 0;
          st = bracket(T, n, x, &i0, &i1);
          if (st == TPB_OK) {
            double err = 0.0;
            for (int j = 0; j < D; ++j) {
              const double vj = interpolate(x, T[i0], T[i1], d.v[(base + i0) * D + j],
                                            d.v[(base + i1) * D + j]);
              const double e = fabs(vj - s_v[(base + sw) * D + j]);
              err = e > err ? e : err;
            }
            if (err > 1e-2) st = TPB_NOT_FOUND;  // kAcceptableMatchError
          }
        }
        if (st == TPB_OK) {  // InsertSegment (:97-131)
          int seq = d.seq[b] + 1;
          int upper = first_ge(T, n, x);
          bool reset = false;
          if (upper == 0) {
            reset = true;
          } else if (x - T[upper - 1] < d.tol) {
            upper--;
          }
          keep = reset ? 0 : upper;
          if (keep + m > d.cap || (head_old + keep + m > d.cap && head_old == 0)) {
            st = TPB_RESOURCE_EXHAUSTED;
          } else {
            go = 1;
            if (head_old + keep + m > d.cap) head_new = 0;
            if (reset) seq = 0;
            d.seq[b] = seq;
            d.count[b] = keep + m;
            d.head[b] = head_new;
          }
        }
      }
    }
    if (status != nullptr) status[b] = st;
    sh_go = go;
    sh_keep = keep;
    sh_head_old = head_old;
    sh_head_new = head_new;
  }
  __syncthreads();
  if (!sh_go) return;
  const int keep = sh_keep;
  const size_t slot0 = static_cast<size_t>(b) * d.cap;
  const size_t o = slot0 + sh_head_old, w = slot0 + sh_head_new;
  if (w != o && keep > 0) {  // kept rows to slot 0; they end before row head_old + sw
    block_move(d.t + w, d.t + o, keep);
    block_move(d.p + w * D, d.p + o * D, static_cast<long>(keep) * D);
    block_move(d.v + w * D, d.v + o * D, static_cast<long>(keep) * D);
    block_move(d.a + w * D, d.a + o * D, static_cast<long>(keep) * D);
  }
  block_move(d.p + (w + keep) * D, d.p + (o + sw) * D, static_cast<long>(m) * D);
  block_copy(d.t + w + keep, s_t + o + sw, m);
  block_copy(d.v + (w + keep) * D, s_v + (o + sw) * D, static_cast<long>(m) * D);
  block_copy(d.a + (w + keep) * D, s_a + (o + sw) * D, static_cast<long>(m) * D);
}

// Reserve: every buffer's rows to slot 0 of arrays with `new_cap` slots per buffer.
__global__ void __launch_bounds__(kInsertThreads)
k_tbuf_relocate(TbufDev d, int new_cap, double* __restrict__ nt, double* __restrict__ np_,
                double* __restrict__ nv, double* __restrict__ na) {
  const int b = blockIdx.x, D = d.D;
  const int n = d.count[b];
  const size_t src = static_cast<size_t>(b) * d.cap + d.head[b];
  const size_t dst = static_cast<size_t>(b) * new_cap;
  block_copy(nt + dst, d.t + src, n);
  block_copy(np_ + dst * D, d.p + src * D, static_cast<long>(n) * D);
  block_copy(nv + dst * D, d.v + src * D, static_cast<long>(n) * D);
  block_copy(na + dst * D, d.a + src * D, static_cast<long>(n) * D);
}

TbufDev dev_of(const tpb_tbuf* t) {
  return TbufDev{t->B, t->D, t->cap, t->tol, t->t, t->p, t->v, t->a, t->head, t->count, t->seq};
}

// Per-call staging of host arguments: inputs are copied to scratch device memory on
// the buffer's stream, outputs are copied back and the stream is synchronised when
// the call ends.  With on_device the pointers pass through untouched.
struct Stage {
  cudaStream_t st;
  bool dev;
  cudaError_t err = cudaSuccess;
  std::vector<void*> tmp;
  struct Back { void* host; void* device; size_t bytes; };
  std::vector<Back> backs;
  Stage(cudaStream_t s, bool on_device) : st(s), dev(on_device) {}
  void* scratch(size_t bytes) {
    void* q = nullptr;
    if (err == cudaSuccess) err = cudaMalloc(&q, bytes ? bytes : 1);
    if (err == cudaSuccess) tmp.push_back(q); else q = nullptr;
    return q;
  }
  template <typename T>
  const T* in(const T* h, size_t n) {
    if (h == nullptr || dev) return h;
    void* q = scratch(n * sizeof(T));
    if (q != nullptr && n > 0) err = cudaMemcpyAsync(q, h, n * sizeof(T), cudaMemcpyHostToDevice, st);
    return static_cast<const T*>(q);
  }
  template <typename T>
  T* out(T* h, size_t n) {
    if (h == nullptr || dev) return h;
    void* q = scratch(n * sizeof(T));
    if (q != nullptr) backs.push_back({h, q, n * sizeof(T)});
    return static_cast<T*>(q);
  }
  cudaError_t finish() {
    if (err == cudaSuccess) err = cudaGetLastError();
    for (const Back& k : backs)
      if (err == cudaSuccess && k.bytes > 0)
        err = cudaMemcpyAsync(k.host, k.device, k.bytes, cudaMemcpyDeviceToHost, st);
    if (!dev || !tmp.empty()) {
      const cudaError_t e = cudaStreamSynchronize(st);
      if (err == cudaSuccess) err = e;
    }
    for (void* q : tmp) cudaFree(q);
    tmp.clear();
    return err;
  }
};

int32_t fail(tpb_tbuf* t, int32_t code, const std::string& msg) {
  t->error = msg;
  return code;
}
int32_t done(tpb_tbuf* t, Stage& s, const char* what) {
  const cudaError_t e = s.finish();
  if (e != cudaSuccess) return fail(t, TPB_INTERNAL, std::string(what) + ": " + cudaGetErrorString(e));
  return TPB_OK;
}
inline int blocks_for(int n, int threads) { return (n + threads - 1) / threads; }

template <typename T>
cudaError_t dev_alloc(T** q, size_t n) {
  return cudaMalloc(reinterpret_cast<void**>(q), (n ? n : 1) * sizeof(T));
}

int32_t stop_common(tpb_tbuf* t, const int32_t* index, const double* time_sec,
                    const double* max_acc, double time_step, int32_t* status, int on_device) {
  if (t == nullptr || max_acc == nullptr) return TPB_INVALID_ARGUMENT;
  if (t->D > kMaxDofs) return fail(t, TPB_UNIMPLEMENTED, "stop operations need num_dofs <= 16");
  if (cudaSetDevice(t->device) != cudaSuccess) return fail(t, TPB_INTERNAL, "cudaSetDevice");
  const size_t B = t->B, D = t->D, R = B * static_cast<size_t>(t->cap);
  if (t->s_t == nullptr) {
    cudaError_t e = dev_alloc(&t->s_t, R);
    if (e == cudaSuccess) e = dev_alloc(&t->s_v, R * D);
    if (e == cudaSuccess) e = dev_alloc(&t->s_a, R * D);
    if (e != cudaSuccess) return fail(t, TPB_INTERNAL, std::string("stop scratch: ") + cudaGetErrorString(e));
  }
  if (t->s_begin == nullptr) {
    cudaError_t e = dev_alloc(&t->s_begin, B);
    if (e == cudaSuccess) e = dev_alloc(&t->s_len, B);
    if (e == cudaSuccess) e = dev_alloc(&t->s_idx, B);
    if (e == cudaSuccess) e = dev_alloc(&t->s_pending, B);
    if (e == cudaSuccess) e = dev_alloc(&t->s_count, B);
    if (e == cudaSuccess) e = dev_alloc(&t->s_switch, B);
    if (e == cudaSuccess) e = dev_alloc(&t->s_rstatus, B);
    if (e != cudaSuccess) return fail(t, TPB_INTERNAL, std::string("stop scratch: ") + cudaGetErrorString(e));
  }
  Stage s(t->stream, on_device != 0);
  const int32_t* d_index = s.in(index, B);
  const double* d_time = s.in(time_sec, B);
  const double* d_amax = s.in(max_acc, B * D);
  int32_t* d_status = s.out(status, B);
  if (s.err == cudaSuccess) {
    const TbufDev d = dev_of(t);
    k_tbuf_stop_prepare<<<blocks_for(t->B, 128), 128, 0, t->stream>>>(
        d, d_index, d_time, d_amax, time_step, t->s_begin, t->s_len, t->s_idx, t->s_pending);
    StopArgs a;
    a.num_traj = t->B;
    a.D = t->D;
    a.begin = t->s_begin;
    a.length = t->s_len;
    a.times = t->t;
    a.vel = t->v;
    a.acc = t->a;
    a.max_acc = d_amax;
    a.count = t->s_count;
    a.switch_index = t->s_switch;
    a.status = t->s_rstatus;
    a.out_times = t->s_t;
    a.out_vel = t->s_v;
    a.out_acc = t->s_a;
    launch_rescale_to_stop(a, t->stream);
    k_tbuf_stop_commit<<<t->B, kInsertThreads, 0, t->stream>>>(
        d, t->s_pending, t->s_idx, t->s_count, t->s_switch, t->s_rstatus, t->s_t, t->s_v, t->s_a,
        d_status);
    t->launches += 3;
  }
  return done(t, s, "tpb_tbuf_stop");
}

}  // namespace
}  // namespace tpb

using namespace tpb;

extern "C" {

int32_t tpb_tbuf_create(const tpb_tbuf_config* cfg, tpb_tbuf** out) {
  if (cfg == nullptr || out == nullptr) return TPB_INVALID_ARGUMENT;
  *out = nullptr;
  if (cfg->num_buffers < 1 || cfg->num_buffers > 65535 || cfg->num_dofs < 1 ||
      cfg->capacity_rows < 2)
    return TPB_INVALID_ARGUMENT;  // (one grid row per buffer)
  if (!(cfg->timestep_tolerance > 0)) return TPB_FAILED_PRECONDITION;  // :42-45
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return TPB_FAILED_PRECONDITION;  // no CPU fallback
  }
  if (cfg->device < 0 || cfg->device >= ndev) return TPB_INVALID_ARGUMENT;
  if (cudaSetDevice(cfg->device) != cudaSuccess) return TPB_INTERNAL;
  tpb_tbuf* t = new tpb_tbuf();
  t->device = cfg->device;
  t->B = cfg->num_buffers;
  t->D = cfg->num_dofs;
  t->cap = cfg->capacity_rows;
  t->tol = cfg->timestep_tolerance;
  const size_t B = t->B, D = t->D, R = B * static_cast<size_t>(t->cap);
  cudaError_t e = dev_alloc(&t->t, R);
  if (e == cudaSuccess) e = dev_alloc(&t->p, R * D);
  if (e == cudaSuccess) e = dev_alloc(&t->v, R * D);
  if (e == cudaSuccess) e = dev_alloc(&t->a, R * D);
  if (e == cudaSuccess) e = dev_alloc(&t->head, B);
  if (e == cudaSuccess) e = dev_alloc(&t->count, B);
  if (e == cudaSuccess) e = dev_alloc(&t->seq, B);
  if (e == cudaSuccess) e = cudaMemset(t->head, 0, B * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMemset(t->count, 0, B * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMemset(t->seq, 0, B * sizeof(int32_t));
  if (e != cudaSuccess) {
    fprintf(stderr, "tpb_tbuf_create: %s\n", cudaGetErrorString(e));
    tpb_tbuf_destroy(t);
    return TPB_INTERNAL;
  }
  *out = t;
  return TPB_OK;
}

int32_t tpb_tbuf_destroy(tpb_tbuf* t) {
  if (t == nullptr) return TPB_OK;
  cudaSetDevice(t->device);
  if (t->stream != nullptr) cudaStreamSynchronize(t->stream); else cudaDeviceSynchronize();
  void* all[] = {t->t, t->p, t->v, t->a, t->head, t->count, t->seq, t->s_t, t->s_v, t->s_a,
                 t->s_begin, t->s_len, t->s_idx, t->s_pending, t->s_count, t->s_switch,
                 t->s_rstatus};
  for (void* q : all)
    if (q != nullptr) cudaFree(q);
  delete t;
  return TPB_OK;
}

const char* tpb_tbuf_last_error(const tpb_tbuf* t) { return t ? t->error.c_str() : ""; }

int32_t tpb_tbuf_set_stream(tpb_tbuf* t, void* cuda_stream) {
  if (t == nullptr) return TPB_INVALID_ARGUMENT;
  t->stream = static_cast<cudaStream_t>(cuda_stream);
  return TPB_OK;
}

int32_t tpb_tbuf_sync(tpb_tbuf* t) {
  if (t == nullptr) return TPB_INVALID_ARGUMENT;
  cudaSetDevice(t->device);
  const cudaError_t e = cudaStreamSynchronize(t->stream);
  if (e != cudaSuccess) return fail(t, TPB_INTERNAL, cudaGetErrorString(e));
  return TPB_OK;
}

int64_t tpb_tbuf_kernel_launches(const tpb_tbuf* t) { return t ? t->launches : 0; }

int32_t tpb_tbuf_reserve(tpb_tbuf* t, int32_t capacity_rows) {
  if (t == nullptr || capacity_rows < 2) return TPB_INVALID_ARGUMENT;
  if (capacity_rows <= t->cap) return TPB_OK;
  cudaSetDevice(t->device);
  const size_t B = t->B, D = t->D, R = B * static_cast<size_t>(capacity_rows);
  double *nt = nullptr, *np_ = nullptr, *nv = nullptr, *na = nullptr;
  cudaError_t e = dev_alloc(&nt, R);
  if (e == cudaSuccess) e = dev_alloc(&np_, R * D);
  if (e == cudaSuccess) e = dev_alloc(&nv, R * D);
  if (e == cudaSuccess) e = dev_alloc(&na, R * D);
  if (e == cudaSuccess) {
    k_tbuf_relocate<<<t->B, kInsertThreads, 0, t->stream>>>(dev_of(t), capacity_rows, nt, np_, nv,
                                                            na);
    t->launches += 1;
    e = cudaMemsetAsync(t->head, 0, B * sizeof(int32_t), t->stream);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
  if (e != cudaSuccess) {
    for (void* q : {static_cast<void*>(nt), static_cast<void*>(np_), static_cast<void*>(nv),
                    static_cast<void*>(na)})
      if (q != nullptr) cudaFree(q);
    return fail(t, TPB_INTERNAL, std::string("tpb_tbuf_reserve: ") + cudaGetErrorString(e));
  }
  for (void* q : {static_cast<void*>(t->t), static_cast<void*>(t->p), static_cast<void*>(t->v),
                  static_cast<void*>(t->a), static_cast<void*>(t->s_t), static_cast<void*>(t->s_v),
                  static_cast<void*>(t->s_a)})
    if (q != nullptr) cudaFree(q);
  t->t = nt;
  t->p = np_;
  t->v = nv;
  t->a = na;
  t->s_t = t->s_v = t->s_a = nullptr;  // stop scratch follows the capacity
  t->cap = capacity_rows;
  return TPB_OK;
}

int32_t tpb_tbuf_clear(tpb_tbuf* t, const int32_t* select, int32_t on_device) {
  if (t == nullptr) return TPB_INVALID_ARGUMENT;
  cudaSetDevice(t->device);
  Stage s(t->stream, on_device != 0);
  const int32_t* d_sel = s.in(select, t->B);
  if (s.err == cudaSuccess) {
    k_tbuf_clear<<<blocks_for(t->B, 128), 128, 0, t->stream>>>(dev_of(t), d_sel);
    t->launches += 1;
  }
  return done(t, s, "tpb_tbuf_clear");
}

int32_t tpb_tbuf_insert_segments(tpb_tbuf* t, const int64_t* begin, const int32_t* length,
                                 int64_t total_rows, const double* times,
                                 const double* positions, const double* velocities,
                                 const double* accelerations, int32_t* status,
                                 int32_t on_device) {
  if (t == nullptr || begin == nullptr || length == nullptr || total_rows < 0)
    return TPB_INVALID_ARGUMENT;
  if (total_rows > 0 && (times == nullptr || positions == nullptr || velocities == nullptr ||
                         accelerations == nullptr))
    return fail(t, TPB_INVALID_ARGUMENT, "segment arrays missing");
  cudaSetDevice(t->device);
  const size_t B = t->B, D = t->D, R = static_cast<size_t>(total_rows);
  if (!on_device) {  // every segment must lie inside the packed arrays (:84-95)
    for (size_t b = 0; b < B; ++b)
      if (length[b] > 0 && (begin[b] < 0 || begin[b] + length[b] > total_rows))
        return fail(t, TPB_INVALID_ARGUMENT, "segment outside the packed arrays");
  }
  Stage s(t->stream, on_device != 0);
  const int64_t* d_begin = s.in(begin, B);
  const int32_t* d_len = s.in(length, B);
  const double* d_t = s.in(times, R);
  const double* d_p = s.in(positions, R * D);
  const double* d_v = s.in(velocities, R * D);
  const double* d_a = s.in(accelerations, R * D);
  int32_t* d_status = s.out(status, B);
  if (s.err == cudaSuccess) {
    k_tbuf_insert<<<t->B, kInsertThreads, 0, t->stream>>>(dev_of(t), d_begin, d_len, d_t, d_p,
                                                          d_v, d_a, d_status);
    t->launches += 1;
  }
  return done(t, s, "tpb_tbuf_insert_segments");
}

int32_t tpb_tbuf_append_samples(tpb_tbuf* t, const int32_t* select, const double* time,
                                const double* positions, const double* velocities,
                                const double* accelerations, int32_t* status,
                                int32_t on_device) {
  if (t == nullptr || time == nullptr || positions == nullptr || velocities == nullptr ||
      accelerations == nullptr)
    return TPB_INVALID_ARGUMENT;
  cudaSetDevice(t->device);
  const size_t B = t->B, D = t->D;
  Stage s(t->stream, on_device != 0);
  const int32_t* d_sel = s.in(select, B);
  const double* d_t = s.in(time, B);
  const double* d_p = s.in(positions, B * D);
  const double* d_v = s.in(velocities, B * D);
  const double* d_a = s.in(accelerations, B * D);
  int32_t* d_status = s.out(status, B);
  if (s.err == cudaSuccess) {
    k_tbuf_append<<<blocks_for(t->B, 128), 128, 0, t->stream>>>(dev_of(t), d_sel, d_t, d_p, d_v,
                                                                d_a, d_status);
    t->launches += 1;
  }
  return done(t, s, "tpb_tbuf_append_samples");
}

int32_t tpb_tbuf_discard_before(tpb_tbuf* t, const double* time_sec, int32_t on_device) {
  if (t == nullptr || time_sec == nullptr) return TPB_INVALID_ARGUMENT;
  cudaSetDevice(t->device);
  Stage s(t->stream, on_device != 0);
  const double* d_time = s.in(time_sec, t->B);
  if (s.err == cudaSuccess) {
    k_tbuf_discard<<<blocks_for(t->B, 128), 128, 0, t->stream>>>(dev_of(t), d_time);
    t->launches += 1;
  }
  return done(t, s, "tpb_tbuf_discard_before");
}

int32_t tpb_tbuf_positions_up_to_time(tpb_tbuf* t, const double* time_sec, int32_t* num,
                                      int32_t on_device) {
  if (t == nullptr || time_sec == nullptr || num == nullptr) return TPB_INVALID_ARGUMENT;
  cudaSetDevice(t->device);
  Stage s(t->stream, on_device != 0);
  const double* d_time = s.in(time_sec, t->B);
  int32_t* d_num = s.out(num, t->B);
  if (s.err == cudaSuccess) {
    k_tbuf_upto<<<blocks_for(t->B, 128), 128, 0, t->stream>>>(dev_of(t), d_time, d_num);
    t->launches += 1;
  }
  return done(t, s, "tpb_tbuf_positions_up_to_time");
}

int32_t tpb_tbuf_sample(tpb_tbuf* t, int32_t queries_per_buffer, const double* time_sec,
                        double* positions, double* velocities, double* accelerations,
                        int32_t* status, int32_t on_device) {
  if (t == nullptr || queries_per_buffer < 0 || (queries_per_buffer > 0 && time_sec == nullptr))
    return TPB_INVALID_ARGUMENT;
  if (queries_per_buffer == 0) return TPB_OK;
  cudaSetDevice(t->device);
  const size_t B = t->B, D = t->D, Q = queries_per_buffer;
  Stage s(t->stream, on_device != 0);
  const double* d_time = s.in(time_sec, B * Q);
  double* d_p = s.out(positions, B * Q * D);
  double* d_v = s.out(velocities, B * Q * D);
  double* d_a = s.out(accelerations, B * Q * D);
  int32_t* d_status = s.out(status, B * Q);
  if (s.err == cudaSuccess) {
    // measured on B200 (profiles/r1_summary.md v9): 2 values per lane in flight, 4 warps
    // per CTA, 8 CTAs/SM (64 registers) -- 0.77 ms for 8.4 M queries; larger batches or
    // CTAs lose more in occupancy than they gain in loads in flight
    constexpr int kBatch = 2, kWarps = 4, kMinBlocks = 8;
    const unsigned gx = static_cast<unsigned>((Q + kWarps * 32 - 1) / (kWarps * 32));
    k_tbuf_sample<kBatch, kWarps, kMinBlocks><<<dim3(gx, t->B), kWarps * 32, 0, t->stream>>>(
        dev_of(t), queries_per_buffer, d_time, d_p, d_v, d_a, d_status);
    t->launches += 1;
  }
  return done(t, s, "tpb_tbuf_sample");
}

int32_t tpb_tbuf_stop_at_index(tpb_tbuf* t, const int32_t* index, const double* max_acceleration,
                               double time_step, int32_t* status, int32_t on_device) {
  if (index == nullptr) return TPB_INVALID_ARGUMENT;
  return stop_common(t, index, nullptr, max_acceleration, time_step, status, on_device);
}

int32_t tpb_tbuf_stop_before_time(tpb_tbuf* t, const double* time_sec,
                                  const double* max_acceleration, double time_step,
                                  int32_t* status, int32_t on_device) {
  if (time_sec == nullptr) return TPB_INVALID_ARGUMENT;
  return stop_common(t, nullptr, time_sec, max_acceleration, time_step, status, on_device);
}

int32_t tpb_tbuf_add_offset(tpb_tbuf* t, const double* offset, int32_t on_device) {
  if (t == nullptr || offset == nullptr) return TPB_INVALID_ARGUMENT;
  cudaSetDevice(t->device);
  Stage s(t->stream, on_device != 0);
  const double* d_off = s.in(offset, t->B);
  if (s.err == cudaSuccess) {
    const unsigned gx = static_cast<unsigned>(std::min(blocks_for(t->cap, 256), 64));
    k_tbuf_offset<<<dim3(gx, t->B), 256, 0, t->stream>>>(dev_of(t), d_off);
    t->launches += 1;
  }
  return done(t, s, "tpb_tbuf_add_offset");
}

int32_t tpb_tbuf_get_state(tpb_tbuf* t, int32_t* num_samples, int32_t* sequence_number,
                           double* front_time, double* back_time, int32_t on_device) {
  if (t == nullptr) return TPB_INVALID_ARGUMENT;
  cudaSetDevice(t->device);
  Stage s(t->stream, on_device != 0);
  int32_t* d_num = s.out(num_samples, t->B);
  int32_t* d_seq = s.out(sequence_number, t->B);
  double* d_front = s.out(front_time, t->B);
  double* d_back = s.out(back_time, t->B);
  if (s.err == cudaSuccess) {
    k_tbuf_state<<<blocks_for(t->B, 128), 128, 0, t->stream>>>(dev_of(t), d_num, d_seq, d_front,
                                                               d_back);
    t->launches += 1;
  }
  return done(t, s, "tpb_tbuf_get_state");
}

int32_t tpb_tbuf_get_rows(tpb_tbuf* t, int32_t buffer, int32_t first_row, int32_t num_rows,
                          double* times, double* positions, double* velocities,
                          double* accelerations, int32_t on_device) {
  if (t == nullptr || buffer < 0 || buffer >= t->B || first_row < 0 || num_rows < 0)
    return TPB_INVALID_ARGUMENT;
  cudaSetDevice(t->device);
  int32_t hc[2] = {0, 0};
  cudaError_t e = cudaMemcpyAsync(&hc[0], t->head + buffer, sizeof(int32_t),
                                  cudaMemcpyDeviceToHost, t->stream);
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(&hc[1], t->count + buffer, sizeof(int32_t), cudaMemcpyDeviceToHost,
                        t->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
  if (e != cudaSuccess) return fail(t, TPB_INTERNAL, cudaGetErrorString(e));
  if (first_row + num_rows > hc[1]) return fail(t, TPB_OUT_OF_RANGE, "rows outside the buffer");
  const size_t D = t->D, r = static_cast<size_t>(buffer) * t->cap + hc[0] + first_row;
  const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
  const size_t n = static_cast<size_t>(num_rows);
  if (n > 0) {
    if (times != nullptr && e == cudaSuccess)
      e = cudaMemcpyAsync(times, t->t + r, n * sizeof(double), kind, t->stream);
    if (positions != nullptr && e == cudaSuccess)
      e = cudaMemcpyAsync(positions, t->p + r * D, n * D * sizeof(double), kind, t->stream);
    if (velocities != nullptr && e == cudaSuccess)
      e = cudaMemcpyAsync(velocities, t->v + r * D, n * D * sizeof(double), kind, t->stream);
    if (accelerations != nullptr && e == cudaSuccess)
      e = cudaMemcpyAsync(accelerations, t->a + r * D, n * D * sizeof(double), kind, t->stream);
    if (!on_device && e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
  }
  if (e != cudaSuccess) return fail(t, TPB_INTERNAL, cudaGetErrorString(e));
  return TPB_OK;
}

int32_t tpb_tbuf_devptr(tpb_tbuf* t, tpb_tbuf_array what, void** ptr, size_t* bytes) {
  if (t == nullptr || ptr == nullptr) return TPB_INVALID_ARGUMENT;
  const size_t B = t->B, D = t->D, R = B * static_cast<size_t>(t->cap);
  size_t n = 0;
  switch (what) {
    case TPB_TBUF_TIMES: *ptr = t->t; n = R * sizeof(double); break;
    case TPB_TBUF_POSITIONS: *ptr = t->p; n = R * D * sizeof(double); break;
    case TPB_TBUF_VELOCITIES: *ptr = t->v; n = R * D * sizeof(double); break;
    case TPB_TBUF_ACCELERATIONS: *ptr = t->a; n = R * D * sizeof(double); break;
    case TPB_TBUF_HEAD: *ptr = t->head; n = B * sizeof(int32_t); break;
    case TPB_TBUF_COUNT: *ptr = t->count; n = B * sizeof(int32_t); break;
    case TPB_TBUF_SEQUENCE: *ptr = t->seq; n = B * sizeof(int32_t); break;
    default: return TPB_INVALID_ARGUMENT;
  }
  if (bytes != nullptr) *bytes = n;
  return TPB_OK;
}

}  // extern "C"
