// Multi-GPU front-end of the batched path-timing engine (host code only).
//
// The unit of work of the reference is one PathTimingTrajectory::Plan of one path
// (path_timing_trajectory.cc:579-684); paths are independent, so a batch shards over
// the GPUs of a box without any exchange step: shard g owns the
// contiguous block [g * ceil(B / G), (g + 1) * ceil(B / G)) of the batch, has its own
// engine (one device, one stream) and is driven by its own host thread.  There is no
// data-path collective; the only "gather" is each shard writing its block of the
// caller's host arrays (durations, status, packed rows at the global row offset).
//
// A device may appear more than once in the device list: logical shards of one GPU
// (used by the tests on a single-GPU box, and to overlap one shard's device-to-host
// copies with another shard's kernels).
#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/tpb.h"

struct tpb_multi {
  tpb_config cfg{};
  std::vector<tpb_engine*> engine;
  std::vector<int32_t> device;
  std::vector<int32_t> first, count;  // of the last batch
  std::vector<int64_t> rows, row_base;
  int32_t num_paths = 0;
  std::string error;
};

namespace {

int32_t fail(tpb_multi* m, int32_t code, const std::string& msg) {
  if (m != nullptr) m->error = msg;
  return code;
}

// Runs fn(g) for every shard on its own thread; returns the first non-zero code.
template <typename Fn>
int32_t for_each_shard(tpb_multi* m, Fn fn) {
  const int G = static_cast<int>(m->engine.size());
  std::vector<int32_t> rc(G, TPB_OK);
  m->error.clear();
  if (G == 1) {
    rc[0] = fn(0);
  } else {
    std::vector<std::thread> pool;
    pool.reserve(G);
    for (int g = 0; g < G; ++g) pool.emplace_back([&, g]() { rc[g] = fn(g); });
    for (auto& t : pool) t.join();
  }
  for (int g = 0; g < G; ++g) {
    if (rc[g] != TPB_OK) {
      if (m->error.empty()) {  // (a message of the front-end itself stays)
        const char* msg = tpb_last_error(m->engine[g]);
        m->error = "shard " + std::to_string(g) + " (device " + std::to_string(m->device[g]) +
                   "): " + (msg ? msg : "");
      }
      return rc[g];
    }
  }
  return TPB_OK;
}

}  // namespace

extern "C" {

// Page-locked host memory for callers without the CUDA headers (the host C++ classes):
// device-to-host copies into it run at the full PCIe rate and asynchronously.
int32_t tpb_host_alloc(size_t bytes, void** ptr) {
  if (ptr == nullptr) return TPB_INVALID_ARGUMENT;
  *ptr = nullptr;
  if (bytes == 0) return TPB_OK;
  return cudaMallocHost(ptr, bytes) == cudaSuccess ? TPB_OK : TPB_INTERNAL;
}
int32_t tpb_host_free(void* ptr) {
  if (ptr == nullptr) return TPB_OK;
  return cudaFreeHost(ptr) == cudaSuccess ? TPB_OK : TPB_INTERNAL;
}

int32_t tpb_shard_range(int32_t num_paths, int32_t num_shards, int32_t shard, int32_t* first,
                        int32_t* count) {
  if (num_paths < 0 || num_shards < 1 || shard < 0 || shard >= num_shards || first == nullptr ||
      count == nullptr)
    return TPB_INVALID_ARGUMENT;
  const int64_t per = (static_cast<int64_t>(num_paths) + num_shards - 1) / num_shards;
  const int64_t lo = std::min<int64_t>(num_paths, per * shard);
  *first = static_cast<int32_t>(lo);
  *count = static_cast<int32_t>(std::min<int64_t>(per, num_paths - lo));
  return TPB_OK;
}

int32_t tpb_multi_create(const tpb_config* cfg, const int32_t* devices, int32_t num_shards,
                         tpb_multi** out) {
  if (cfg == nullptr || out == nullptr || num_shards < 1 || num_shards > 64)
    return TPB_INVALID_ARGUMENT;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return TPB_FAILED_PRECONDITION;
  auto* m = new tpb_multi;
  m->cfg = *cfg;
  const int G = num_shards;
  int32_t first = 0, per = 0;
  tpb_shard_range(cfg->max_paths, G, 0, &first, &per);
  for (int g = 0; g < G; ++g) {
    // devices == NULL: shard g on device g modulo the devices present
    const int32_t dev = devices ? devices[g] : g % ndev;
    tpb_config c = *cfg;
    c.device = dev;
    c.max_paths = std::max(per, 1);
    c.max_time_samples_total = (cfg->max_time_samples_total + G - 1) / G;
    tpb_engine* e = nullptr;
    const int32_t rc = tpb_create(&c, &e);
    if (rc != TPB_OK) {
      for (tpb_engine* x : m->engine) tpb_destroy(x);
      delete m;
      return rc;
    }
    m->engine.push_back(e);
    m->device.push_back(dev);
  }
  m->first.assign(G, 0);
  m->count.assign(G, 0);
  m->rows.assign(G, 0);
  m->row_base.assign(G, 0);
  *out = m;
  return TPB_OK;
}

int32_t tpb_multi_destroy(tpb_multi* m) {
  if (m == nullptr) return TPB_INVALID_ARGUMENT;
  for (tpb_engine* e : m->engine) tpb_destroy(e);
  delete m;
  return TPB_OK;
}

const char* tpb_multi_last_error(const tpb_multi* m) { return m ? m->error.c_str() : ""; }

int32_t tpb_multi_num_shards(const tpb_multi* m) {
  return m ? static_cast<int32_t>(m->engine.size()) : 0;
}

int32_t tpb_multi_shard(const tpb_multi* m, int32_t shard, int32_t* device, int32_t* first_path,
                        int32_t* num_paths, int64_t* first_row, int64_t* num_rows,
                        tpb_engine** engine) {
  if (m == nullptr || shard < 0 || shard >= static_cast<int32_t>(m->engine.size()))
    return TPB_INVALID_ARGUMENT;
  if (device) *device = m->device[shard];
  if (first_path) *first_path = m->first[shard];
  if (num_paths) *num_paths = m->count[shard];
  if (first_row) *first_row = m->row_base[shard];
  if (num_rows) *num_rows = m->rows[shard];
  if (engine) *engine = m->engine[shard];
  return TPB_OK;
}

int64_t tpb_multi_kernel_launches(const tpb_multi* m) {
  int64_t n = 0;
  if (m)
    for (tpb_engine* e : m->engine) n += tpb_kernel_launches(e);
  return n;
}

// Partition of a batch of num_paths over the shards (contiguous blocks).
static int32_t partition(tpb_multi* m, int32_t num_paths) {
  const int G = static_cast<int>(m->engine.size());
  if (num_paths < 1 || num_paths > m->cfg.max_paths)
    return fail(m, TPB_OUT_OF_RANGE, "num_paths outside [1, max_paths]");
  m->num_paths = num_paths;
  for (int g = 0; g < G; ++g) tpb_shard_range(num_paths, G, g, &m->first[g], &m->count[g]);
  std::fill(m->rows.begin(), m->rows.end(), 0);
  std::fill(m->row_base.begin(), m->row_base.end(), 0);
  return TPB_OK;
}

static int32_t upload_shard(tpb_multi* m, int g, const int32_t* num_waypoints,
                            const double* waypoints, const double* max_velocity,
                            const double* max_acceleration) {
  const size_t D = m->cfg.num_dofs, W = m->cfg.max_waypoints;
  const size_t f = static_cast<size_t>(m->first[g]);
  tpb_engine* e = m->engine[g];
  int32_t rc = tpb_set_joint_waypoints(e, m->count[g], num_waypoints ? num_waypoints + f : nullptr,
                                       waypoints + f * W * D, 0);
  if (rc != TPB_OK) return rc;
  return tpb_set_joint_limits(e, max_velocity + f * D, max_acceleration + f * D, 1, 0);
}

// Waypoints [B][max_waypoints][D] and limits [B][D] (host) -> the shards' GPUs.
int32_t tpb_multi_set_joint_paths(tpb_multi* m, int32_t num_paths, const int32_t* num_waypoints,
                                  const double* waypoints, const double* max_velocity,
                                  const double* max_acceleration) {
  if (m == nullptr || waypoints == nullptr || max_velocity == nullptr ||
      max_acceleration == nullptr)
    return TPB_INVALID_ARGUMENT;
  if (int32_t rc = partition(m, num_paths)) return rc;
  return for_each_shard(m, [&](int g) -> int32_t {
    if (m->count[g] == 0) return TPB_OK;
    return upload_shard(m, g, num_waypoints, waypoints, max_velocity, max_acceleration);
  });
}

// tpb_plan_joint_device on every shard at once; the results stay on their GPUs
// (tpb_multi_shard hands out the engines for device-resident consumers).
int32_t tpb_multi_plan_joint_device(tpb_multi* m, double start_sec, double time_step_sec,
                                    int64_t* total_samples) {
  if (m == nullptr) return TPB_INVALID_ARGUMENT;
  if (m->num_paths == 0) return fail(m, TPB_FAILED_PRECONDITION, "No path set");
  const bool want_rows = time_step_sec > 0;
  const int32_t rc = for_each_shard(m, [&](int g) -> int32_t {
    if (m->count[g] == 0) return TPB_OK;
    int64_t total = 0;
    const int32_t r = tpb_plan_joint_device(m->engine[g], start_sec, time_step_sec, 1,
                                            want_rows ? &total : nullptr);
    m->rows[g] = total;
    return r ? r : tpb_sync(m->engine[g]);
  });
  int64_t run = 0;
  for (size_t g = 0; g < m->engine.size(); ++g) {
    m->row_base[g] = run;
    run += m->rows[g];
  }
  if (total_samples) *total_samples = run;
  return rc;
}

// The optional final gather: a per-path array of every shard (TPB_ARR_DURATION,
// TPB_ARR_STATUS, TPB_ARR_LAST_EXTREMAL, or a [B][N] / [B][N][D] array) into one host
// array in path order.  bytes_per_path = the array's bytes per path.
int32_t tpb_multi_gather(tpb_multi* m, tpb_array what, void* dst, size_t bytes_per_path) {
  if (m == nullptr || dst == nullptr) return TPB_INVALID_ARGUMENT;
  if (m->num_paths == 0) return fail(m, TPB_FAILED_PRECONDITION, "No path set");
  return for_each_shard(m, [&](int g) -> int32_t {
    if (m->count[g] == 0) return TPB_OK;
    char* p = static_cast<char*>(dst) + static_cast<size_t>(m->first[g]) * bytes_per_path;
    return tpb_get(m->engine[g], what, p, static_cast<size_t>(m->count[g]) * bytes_per_path, 0);
  });
}

// tpb_plan_joint_host over all shards: the same inputs and outputs as for one engine,
// with num_paths up to the front-end's max_paths.  Every shard uploads its block,
// runs the device pipeline and publishes its row count; as soon as the shards before
// it have published theirs it knows its global row base and copies its block of every
// requested output straight into the caller's arrays.
int32_t tpb_multi_plan_joint_host(tpb_multi* m, tpb_plan_io* io) {
  if (m == nullptr || io == nullptr) return TPB_INVALID_ARGUMENT;
  if (int32_t rc = partition(m, io->num_paths)) return rc;
  const int G = static_cast<int>(m->engine.size());
  const size_t D = m->cfg.num_dofs, N = m->cfg.num_samples;
  const bool want_rows = io->time_step_sec > 0 && m->cfg.max_time_samples_total > 0;
  io->r_total = 0;
  if (G == 1) {
    // one shard: the engine's own pipelined call (the copies of finished chunks of the
    // batch overlap the kernels of the next ones)
    m->error.clear();
    const int32_t rc1 = tpb_plan_joint_host(m->engine[0], io);
    m->rows[0] = io->r_total;
    if (rc1 != TPB_OK) {
      const char* msg = tpb_last_error(m->engine[0]);
      m->error = msg ? msg : "";
    }
    return rc1;
  }
  std::mutex mu;
  std::condition_variable cv;
  std::vector<char> published(G, 0);
  bool failed = false;
  auto publish = [&](int g, int64_t rows, bool ok) {
    std::lock_guard<std::mutex> lock(mu);
    m->rows[g] = rows;
    published[g] = 1;
    if (!ok) failed = true;
    cv.notify_all();
  };
  const int32_t rc = for_each_shard(m, [&](int g) -> int32_t {
    tpb_engine* e = m->engine[g];
    const size_t f = static_cast<size_t>(m->first[g]), n = static_cast<size_t>(m->count[g]);
    if (n == 0) {
      publish(g, 0, true);
      return TPB_OK;
    }
    int64_t total = 0;
    int32_t r = upload_shard(m, g, io->num_waypoints, io->waypoints, io->max_velocity,
                             io->max_acceleration);
    if (r == TPB_OK)
      r = tpb_plan_joint_device(e, io->start_sec, io->time_step_sec, 1,
                                want_rows ? &total : nullptr);
    if (r == TPB_OK) r = tpb_sync(e);
    publish(g, r == TPB_OK ? total : 0, r == TPB_OK);
    if (r != TPB_OK) return r;
    // outputs that do not depend on the other shards
    if (io->duration) r = r ? r : tpb_get(e, TPB_ARR_DURATION, io->duration + f, n * 8, 0);
    if (io->status) r = r ? r : tpb_get(e, TPB_ARR_STATUS, io->status + f, n * 4, 0);
    const size_t pb = n * N * sizeof(double);
    if (io->profile_time) r = r ? r : tpb_get(e, TPB_ARR_TIME, io->profile_time + f * N, pb, 0);
    if (io->profile_s) r = r ? r : tpb_get(e, TPB_ARR_S, io->profile_s + f * N, pb, 0);
    if (io->profile_sd) r = r ? r : tpb_get(e, TPB_ARR_SD, io->profile_sd + f * N, pb, 0);
    if (io->profile_sdd) r = r ? r : tpb_get(e, TPB_ARR_SDD, io->profile_sdd + f * N, pb, 0);
    if (r != TPB_OK || !want_rows) return r;
    // the global row base: rows of all shards before this one
    int64_t base = 0;
    {
      std::unique_lock<std::mutex> lock(mu);
      cv.wait(lock, [&]() {
        for (int k = 0; k < g; ++k)
          if (!published[k]) return false;
        return true;
      });
      if (failed) return TPB_OK;  // the failing shard reports
      for (int k = 0; k < g; ++k) base += m->rows[k];
    }
    m->row_base[g] = base;
    const bool any_out = io->r_pos || io->r_vel || io->r_acc || io->r_s || io->r_sd ||
                         io->r_sdd || io->r_time;
    if (any_out && base + total > io->r_capacity) {
      std::lock_guard<std::mutex> lock(mu);
      fail(m, TPB_OUT_OF_RANGE, "resampled rows exceed r_capacity");
      return TPB_OUT_OF_RANGE;
    }
    if (io->r_offset) {
      std::vector<int64_t> local(n + 1);
      r = tpb_get(e, TPB_ARR_R_OFFSET, local.data(), (n + 1) * sizeof(int64_t), 0);
      if (r != TPB_OK) return r;
      // shard g writes entries [first, first + count); the last shard with paths also
      // writes the closing entry
      for (size_t i = 0; i < n; ++i) io->r_offset[f + i] = local[i] + base;
      if (f + n == static_cast<size_t>(m->num_paths)) io->r_offset[f + n] = local[n] + base;
    }
    const size_t b = static_cast<size_t>(base), t = static_cast<size_t>(total);
    auto rows_out = [&](double* dst, tpb_array what, size_t width) {
      if (dst == nullptr || t == 0 || r != TPB_OK) return;
      r = tpb_get(e, what, dst + b * width, t * width * sizeof(double), 0);
    };
    rows_out(io->r_pos, TPB_ARR_R_POS, D);
    rows_out(io->r_vel, TPB_ARR_R_VEL, D);
    rows_out(io->r_acc, TPB_ARR_R_ACC, D);
    rows_out(io->r_s, TPB_ARR_R_S, 1);
    rows_out(io->r_sd, TPB_ARR_R_SD, 1);
    rows_out(io->r_sdd, TPB_ARR_R_SDD, 1);
    rows_out(io->r_time, TPB_ARR_R_TIME, 1);
    return r;
  });
  for (int g = 0; g < G; ++g) io->r_total += m->rows[g];
  if (rc == TPB_OUT_OF_RANGE && m->error.empty())
    m->error = "resampled rows exceed a capacity";
  return rc;
}

}  // extern "C"
