/* TEST INFRASTRUCTURE ONLY -- not part of the product.
 *
 * CPU restatement of the reference's TrajectoryBuffer
 * (/root/reference/trajectory_planning/trajectory_buffer.cc:40-393, declared in
 * trajectory_buffer.h:30-184) and of time.h:23-30 (TimeFromSec / TimeToSec), as
 * plain C over one growable buffer of [n][D] row-major rows.  Pinned bit-for-bit
 * against the reference's own object code (oracle/_ref: trajectory_buffer.cc
 * compiled verbatim against oracle/shim) on random operation sequences and on
 * the scenarios of trajectory_buffer_test.cc by tests/test_oracle_tbuf.py.
 *
 * UNPINNED: eigenmath::InterpolateLinear(x, x0, x1, y0, y1) is un-vendored
 * (x_edr_eigenmath v1.0.0); restated as y0 + ((x-x0)/(x1-x0))*(y1-y0), y0 when
 * x1 == x0 -- the same statement as oracle/shim/eigenmath/interpolation.h, so
 * the pin above covers everything but that one formula.
 *
 * Return values are absl::StatusCode numbers (0 ok, 3 invalid argument, 5 not
 * found, 9 failed precondition, 11 out of range).  Where the reference CHECK-
 * aborts (StopAtIndex on a sample that already stands still,
 * trajectory_buffer.cc:338) this returns 13 (internal) and leaves the buffer
 * unchanged.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "tp_oracle.h"

struct orc_tbuf {
  double tol;
  int seq;
  int n, cap, D;
  double *t, *p, *v, *a;
};

/* time.h:23-30: orc_time_from_sec / orc_time_to_sec live in tp_oracle.c */

int orc_tbuf_create(double timestep_tolerance, int D, orc_tbuf** out) {
  *out = NULL;
  if (timestep_tolerance <= 0) return 9; /* trajectory_buffer.cc:42-45 */
  orc_tbuf* b = (orc_tbuf*)calloc(1, sizeof(orc_tbuf));
  b->tol = timestep_tolerance;
  b->D = D;
  *out = b;
  return 0;
}
void orc_tbuf_destroy(orc_tbuf* b) {
  if (b == NULL) return;
  free(b->t);
  free(b->p);
  free(b->v);
  free(b->a);
  free(b);
}
static void reserve(orc_tbuf* b, int n) {
  if (n <= b->cap) return;
  int cap = b->cap > 0 ? b->cap : 64;
  while (cap < n) cap *= 2;
  b->t = (double*)realloc(b->t, sizeof(double) * (size_t)cap);
  b->p = (double*)realloc(b->p, sizeof(double) * (size_t)cap * b->D);
  b->v = (double*)realloc(b->v, sizeof(double) * (size_t)cap * b->D);
  b->a = (double*)realloc(b->a, sizeof(double) * (size_t)cap * b->D);
  b->cap = cap;
}
/* :64-70 */
void orc_tbuf_clear(orc_tbuf* b) {
  b->seq = 0;
  b->n = 0;
}
/* first index whose time is >= x (c_upper_bound with the `a <= b` comparator,
 * :104-106, 161-162) */
static int first_ge(const orc_tbuf* b, double x) {
  int lo = 0, hi = b->n;
  while (lo < hi) {
    const int mid = lo + (hi - lo) / 2;
    if (x <= b->t[mid]) hi = mid; else lo = mid + 1;
  }
  return lo;
}
/* first index whose time is > x (c_upper_bound, :223, 241) */
static int first_gt(const orc_tbuf* b, double x) {
  int lo = 0, hi = b->n;
  while (lo < hi) {
    const int mid = lo + (hi - lo) / 2;
    if (x < b->t[mid]) hi = mid; else lo = mid + 1;
  }
  return lo;
}
static void put_rows(orc_tbuf* b, int at, int n, const double* times, const double* pos,
                     const double* vel, const double* acc) {
  reserve(b, at + n);
  memcpy(b->t + at, times, sizeof(double) * (size_t)n);
  memcpy(b->p + (size_t)at * b->D, pos, sizeof(double) * (size_t)n * b->D);
  memcpy(b->v + (size_t)at * b->D, vel, sizeof(double) * (size_t)n * b->D);
  memcpy(b->a + (size_t)at * b->D, acc, sizeof(double) * (size_t)n * b->D);
  b->n = at + n;
}

/* InsertSegment (:80-132); the size checks (:84-95) are implied by the layout */
int orc_tbuf_insert_segment(orc_tbuf* b, int n, const double* times, const double* pos,
                            const double* vel, const double* acc) {
  b->seq++;
  if (n == 0) return 0;
  int upper = first_ge(b, times[0]);
  if (b->n == 0 || upper == 0) {
    put_rows(b, 0, n, times, pos, vel, acc);
    b->seq = 0;
    return 0;
  }
  if (times[0] - b->t[upper - 1] < b->tol) upper--;
  put_rows(b, upper, n, times, pos, vel, acc);
  return 0;
}

/* AppendSample (:134-148) */
int orc_tbuf_append_sample(orc_tbuf* b, double time, const double* pos, const double* vel,
                           const double* acc) {
  if (b->n > 0 && b->t[b->n - 1] >= time) return 3;
  put_rows(b, b->n, 1, &time, pos, vel, acc);
  return 0;
}

/* GetOffsetBracket (:232-251) */
static int bracket(const orc_tbuf* b, double time_sec, int* first, int* second) {
  if (b->n == 0) return 9;
  if (time_sec < b->t[0] || time_sec > b->t[b->n - 1]) return 11;
  const int upper = first_gt(b, time_sec);
  if (upper == b->n) {
    *first = *second = b->n - 1;
    return 0;
  }
  *first = upper - 1;
  *second = upper;
  return 0;
}
/* eigenmath::InterpolateLinear -- UNPINNED, see the header */
static void interpolate(double x, double x0, double x1, const double* y0, const double* y1,
                        int D, double* out) {
  if (x1 == x0) {
    for (int j = 0; j < D; ++j) out[j] = y0[j];
    return;
  }
  const double t = (x - x0) / (x1 - x0);
  for (int j = 0; j < D; ++j) out[j] = y0[j] + t * (y1[j] - y0[j]);
}
/* Get{Position,Velocity,Acceleration}AtTime (:253-293); which: 0, 1, 2 */
int orc_tbuf_get_at_time(const orc_tbuf* b, int which, double time_sec, double* out) {
  int i0, i1;
  const int st = bracket(b, time_sec, &i0, &i1);
  if (st != 0) return st;
  const double* rows = which == 0 ? b->p : which == 1 ? b->v : b->a;
  interpolate(time_sec, b->t[i0], b->t[i1], rows + (size_t)i0 * b->D, rows + (size_t)i1 * b->D,
              b->D, out);
  return 0;
}

/* DiscardSegmentBefore(double) (:154-207) */
void orc_tbuf_discard_before_sec(orc_tbuf* b, double time_sec) {
  if (b->n == 0) return;
  if (time_sec <= b->t[0]) return;
  if (time_sec > b->t[b->n - 1]) {
    orc_tbuf_clear(b);
    return;
  }
  int offset = first_ge(b, time_sec);
  if (offset <= 0) return;
  const int close_to_existing = time_sec - b->t[offset - 1] <= b->tol;
  const int create_by_interpolation = fabs(b->t[offset] - time_sec) > b->tol;
  if (close_to_existing || create_by_interpolation) --offset;
  if (create_by_interpolation) {
    const int D = b->D;
    double* tmp = (double*)malloc(sizeof(double) * 3 * (size_t)D);
    orc_tbuf_get_at_time(b, 0, time_sec, tmp);
    orc_tbuf_get_at_time(b, 1, time_sec, tmp + D);
    orc_tbuf_get_at_time(b, 2, time_sec, tmp + 2 * D);
    b->t[offset] = time_sec;
    memcpy(b->p + (size_t)offset * D, tmp, sizeof(double) * (size_t)D);
    memcpy(b->v + (size_t)offset * D, tmp + D, sizeof(double) * (size_t)D);
    memcpy(b->a + (size_t)offset * D, tmp + 2 * D, sizeof(double) * (size_t)D);
    free(tmp);
  }
  const int keep = b->n - offset;
  memmove(b->t, b->t + offset, sizeof(double) * (size_t)keep);
  memmove(b->p, b->p + (size_t)offset * b->D, sizeof(double) * (size_t)keep * b->D);
  memmove(b->v, b->v + (size_t)offset * b->D, sizeof(double) * (size_t)keep * b->D);
  memmove(b->a, b->a + (size_t)offset * b->D, sizeof(double) * (size_t)keep * b->D);
  b->n = keep;
}
/* DiscardSegmentBefore(absl::Time) (:150-152) */
void orc_tbuf_discard_before_ns(orc_tbuf* b, int64_t ns) {
  orc_tbuf_discard_before_sec(b, orc_time_to_sec(ns));
}

/* GetPositionsUpToTime (:209-226): the number of leading samples returned */
int orc_tbuf_positions_up_to_time_ns(const orc_tbuf* b, int64_t ns) {
  if (b->n == 0) return 0;
  const double time_sec = orc_time_to_sec(ns);
  if (time_sec < b->t[0] || time_sec > b->t[b->n - 1]) return 0;
  return first_gt(b, time_sec) - 1;
}

/* StopAtIndex (:295-358) */
int orc_tbuf_stop_at_index(orc_tbuf* b, int index, const double* max_acc, double time_step) {
  const int D = b->D;
  /* `index > GetNumSamples() - 1` compares int with size_t: for an empty buffer
   * the right side wraps to SIZE_MAX and only `index <= 0` can trigger. */
  if (index <= 0 || (b->n > 0 && index > b->n - 1)) return 11;
  if (b->n == 0) return 13; /* (the reference reads out of bounds here) */
  double amin = max_acc[0];
  for (int j = 1; j < D; ++j) amin = max_acc[j] < amin ? max_acc[j] : amin;
  if (amin <= 0.0) return 3;
  if (time_step <= 0.0) return 3;
  const double kVerySmall = 1e-4;
  if (index == b->n - 1) {
    double vinf = 0.0;
    for (int j = 0; j < D; ++j) {
      const double x = fabs(b->v[(size_t)index * D + j]);
      if (x > vinf) vinf = x;
    }
    if (vinf < kVerySmall) {
      for (int j = 0; j < D; ++j) {
        b->v[(size_t)index * D + j] = 0.0;
        b->a[(size_t)index * D + j] = 0.0;
      }
      return 0;
    }
  }
  const int ns = index + 1;
  double* ot = (double*)malloc(sizeof(double) * (size_t)ns);
  double* ov = (double*)malloc(sizeof(double) * (size_t)ns * D);
  double* oa = (double*)malloc(sizeof(double) * (size_t)ns * D);
  double* op = (double*)malloc(sizeof(double) * (size_t)ns * D);
  int count = 0, sw = 0, ret = 0;
  const int st = orc_rescale_to_stop(ns, D, max_acc, b->t, b->v, b->a, &count, &sw, ot, ov, oa);
  if (st != ORC_OK) {
    ret = 3; /* AreInputsValidForSampledTrajectory -> kInvalidArgument */
  } else if (count == 0) {
    ret = 13; /* CHECK(!rescaled_stop.times.empty()) in the reference */
  } else {
    if (count == index) { /* :342-355 */
      double vel0[64];
      double* vs = D <= 64 ? vel0 : (double*)malloc(sizeof(double) * (size_t)D);
      const int gs = orc_tbuf_get_at_time(b, 1, ot[0], vs);
      if (gs != 0) {
        ret = gs;
      } else {
        double err = 0.0;
        for (int j = 0; j < D; ++j) {
          const double x = fabs(vs[j] - ov[j]);
          if (x > err) err = x;
        }
        if (err > 1e-2) ret = 5;
      }
      if (vs != vel0) free(vs);
    }
    if (ret == 0) {
      memcpy(op, b->p + (size_t)sw * D, sizeof(double) * (size_t)count * D);
      ret = orc_tbuf_insert_segment(b, count, ot, op, ov, oa);
    }
  }
  free(ot);
  free(ov);
  free(oa);
  free(op);
  return ret;
}

/* StopBeforeTime(double) (:366-381) */
int orc_tbuf_stop_before_time(orc_tbuf* b, double time_sec, const double* max_acc,
                              double time_step) {
  if (b->n == 0) return 0;
  if (time_sec < b->t[0]) return 11;
  int lower = 0, hi = b->n; /* c_lower_bound: first time >= time_sec */
  while (lower < hi) {
    const int mid = lower + (hi - lower) / 2;
    if (b->t[mid] < time_sec) lower = mid + 1; else hi = mid;
  }
  int index = lower + 1;
  if (index > b->n - 1) index = b->n - 1;
  return orc_tbuf_stop_at_index(b, index, max_acc, time_step);
}

/* AddOffsetToTimestamps (:383-390) */
void orc_tbuf_add_offset(orc_tbuf* b, double offset) {
  for (int i = 0; i < b->n; ++i) b->t[i] = b->t[i] + offset;
}

int orc_tbuf_num_samples(const orc_tbuf* b) { return b->n; }
int orc_tbuf_sequence_number(const orc_tbuf* b) { return b->seq; }
/* GetStartTime / GetEndTime (:50-62) */
int64_t orc_tbuf_start_ns(const orc_tbuf* b) {
  return b->n == 0 ? orc_time_from_sec(0) : orc_time_from_sec(b->t[0]);
}
int64_t orc_tbuf_end_ns(const orc_tbuf* b) {
  return b->n == 0 ? 0 : orc_time_from_sec(b->t[b->n - 1]);
}
void orc_tbuf_rows(const orc_tbuf* b, double* times, double* pos, double* vel, double* acc) {
  memcpy(times, b->t, sizeof(double) * (size_t)b->n);
  memcpy(pos, b->p, sizeof(double) * (size_t)b->n * b->D);
  memcpy(vel, b->v, sizeof(double) * (size_t)b->n * b->D);
  memcpy(acc, b->a, sizeof(double) * (size_t)b->n * b->D);
}
