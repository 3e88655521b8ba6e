// rtx_multi.cu — N GPUs of one process behind one handle (rtx_multi_*, include/rtx_cuda.h).
//
// ≙ Scene::render (raytracing/scene.rs:144-170): one blocking call that returns the whole image.  The reference
// spawns a thread per row and joins them in row order; here the image's rows are dealt to the GPUs in strips of 4
// (RTX_FLAG_SHARD_ROWS), one worker thread per GPU — alive for the lifetime of the handle — drives its device through
// rtx_scene_render_rows_to_host, and every GPU's copy engine puts its rows straight into the caller's buffer over its
// own PCIe link.  No process group, no NCCL, no Python: this is what a Rust `cuda::Scene::render` binds to get the
// whole box.  Pixels are a pure function of (scene, width, height, seed), so the result equals the one-GPU image.
#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <thread>

#include "rtx_internal.h"

struct rtx_multi {
  struct Worker {
    rtx_scene* scene = nullptr;
    int device = 0;
    cudaStream_t stream = nullptr;
    unsigned long long* d_counters = nullptr;
    std::thread thread;
    // the job of the current frame and its outcome
    rtx_status status = RTX_OK;
    std::string error;
    unsigned long long counters[8] = {};
    float device_ms = 0.f;
    uint32_t launches = 0;
  };
  std::vector<Worker> w;
  std::mutex render_mu;          // one frame at a time: held for the whole of rtx_multi_render
  std::mutex mu;                 // guards the fields below
  std::condition_variable cv_job, cv_done;
  uint64_t generation = 0;       // bumped per frame
  uint32_t pending = 0;          // workers still busy with the current frame
  bool stop = false;
  // frame arguments
  uint32_t width = 0, height = 0, flags = 0;
  uint64_t seed = 0;
  double* out_rgb = nullptr;
  uint32_t* out_ids = nullptr;
  mutable std::string last_error;
};

namespace rtx {

static void multi_worker(rtx_multi* m, uint32_t i) {
  rtx_multi::Worker& me = m->w[i];
  cudaSetDevice(me.device);
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  uint64_t seen = 0;
  for (;;) {
    uint32_t width, height, flags;
    uint64_t seed;
    double* rgb;
    uint32_t* ids;
    {
      std::unique_lock<std::mutex> lk(m->mu);
      m->cv_job.wait(lk, [&] { return m->stop || m->generation != seen; });
      if (m->stop) break;
      seen = m->generation;
      width = m->width, height = m->height, flags = m->flags, seed = m->seed, rgb = m->out_rgb, ids = m->out_ids;
    }
    const rtx_shard sh{i, (uint32_t)m->w.size()};
    rtx_status st = RTX_OK;
    cudaError_t e = cudaMemsetAsync(me.d_counters, 0, 8 * sizeof(unsigned long long), me.stream);
    if (e == cudaSuccess) e = cudaEventRecord(e0, me.stream);
    if (e == cudaSuccess) {
      st = rtx_scene_render_rows_to_host(me.scene, width, height, seed, flags, &sh, rgb, ids, me.d_counters, me.stream);
      if (st != RTX_OK) me.error = rtx_last_error(me.scene);
    }
    if (e == cudaSuccess && st == RTX_OK) e = cudaEventRecord(e1, me.stream);
    if (e == cudaSuccess && st == RTX_OK)
      e = cudaMemcpyAsync(me.counters, me.d_counters, sizeof(me.counters), cudaMemcpyDeviceToHost, me.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(me.stream);
    if (e != cudaSuccess) {
      st = RTX_ERR_CUDA;
      me.error = std::string("CUDA error on device ") + std::to_string(me.device) + ": " + cudaGetErrorString(e);
      cudaGetLastError();
    }
    me.device_ms = 0.f;
    if (st == RTX_OK) cudaEventElapsedTime(&me.device_ms, e0, e1);
    me.launches = rtx_scene_last_launches(me.scene);
    me.status = st;
    {
      std::lock_guard<std::mutex> lk(m->mu);
      if (--m->pending == 0) m->cv_done.notify_all();
    }
  }
  cudaEventDestroy(e0), cudaEventDestroy(e1);
}

template <typename F>
static rtx_status multi_for_each(rtx_multi* m, F f) {
  if (!m) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null multi handle");
  for (auto& w : m->w) {
    const rtx_status st = f(w.scene);
    if (st != RTX_OK) {
      m->last_error = rtx_last_error(w.scene);
      return fail(nullptr, st, m->last_error);
    }
  }
  return RTX_OK;
}

}  // namespace rtx

using namespace rtx;

extern "C" {

rtx_status rtx_multi_create(const rtx_camera* camera, const rtx_config* config, const int32_t* devices, uint32_t n,
                            rtx_multi** out) {
  if (!camera || !config || !devices || !out || n == 0) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "bad rtx_multi_create arguments");
  *out = nullptr;
  for (uint32_t i = 0; i < n; ++i)
    for (uint32_t j = 0; j < i; ++j)
      if (devices[i] == devices[j]) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "a device is listed twice");
  rtx_multi* m = new (std::nothrow) rtx_multi();
  if (!m) return fail(nullptr, RTX_ERR_OUT_OF_MEMORY, "host allocation failed");
  m->w.resize(n);
  rtx_status st = RTX_OK;
  for (uint32_t i = 0; i < n && st == RTX_OK; ++i) {
    rtx_multi::Worker& w = m->w[i];
    w.device = devices[i];
    st = rtx_scene_create(camera, config, devices[i], &w.scene);
    if (st != RTX_OK) break;
    cudaError_t e = cudaSetDevice(w.device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc((void**)&w.d_counters, 8 * sizeof(unsigned long long));
    if (e != cudaSuccess) st = cuda_fail(nullptr, e, "rtx_multi_create");
  }
  if (st != RTX_OK) {
    const std::string why = rtx_last_error(nullptr);
    rtx_multi_destroy(m);
    return fail(nullptr, st, why);
  }
  for (uint32_t i = 0; i < n; ++i) m->w[i].thread = std::thread(multi_worker, m, i);
  *out = m;
  return RTX_OK;
}

rtx_status rtx_multi_destroy(rtx_multi* m) {
  if (!m) return RTX_OK;
  {
    std::lock_guard<std::mutex> lk(m->mu);
    m->stop = true;
  }
  m->cv_job.notify_all();
  for (auto& w : m->w)
    if (w.thread.joinable()) w.thread.join();
  for (auto& w : m->w) {
    if (w.scene) {
      cudaSetDevice(w.device);
      if (w.stream) cudaStreamSynchronize(w.stream), cudaStreamDestroy(w.stream);
      cudaFree(w.d_counters);
      rtx_scene_destroy(w.scene);
    }
  }
  delete m;
  return RTX_OK;
}

uint32_t rtx_multi_device_count(const rtx_multi* m) { return m ? (uint32_t)m->w.size() : 0u; }
rtx_scene* rtx_multi_scene(rtx_multi* m, uint32_t i) { return (m && i < m->w.size()) ? m->w[i].scene : nullptr; }

rtx_status rtx_multi_set_camera(rtx_multi* m, const rtx_camera* c) {
  return multi_for_each(m, [&](rtx_scene* s) { return rtx_scene_set_camera(s, c); });
}
rtx_status rtx_multi_set_config(rtx_multi* m, const rtx_config* c) {
  return multi_for_each(m, [&](rtx_scene* s) { return rtx_scene_set_config(s, c); });
}
rtx_status rtx_multi_add_spheres(rtx_multi* m, const rtx_sphere* shapes, const rtx_material* mats, uint64_t n) {
  return multi_for_each(m, [&](rtx_scene* s) { return rtx_scene_add_spheres(s, shapes, mats, n); });
}
rtx_status rtx_multi_add_planes(rtx_multi* m, const rtx_plane* shapes, const rtx_material* mats, uint64_t n) {
  return multi_for_each(m, [&](rtx_scene* s) { return rtx_scene_add_planes(s, shapes, mats, n); });
}
rtx_status rtx_multi_add_triangles(rtx_multi* m, const rtx_triangle* shapes, const rtx_material* mats, uint64_t n) {
  return multi_for_each(m, [&](rtx_scene* s) { return rtx_scene_add_triangles(s, shapes, mats, n); });
}
rtx_status rtx_multi_clear_objects(rtx_multi* m) {
  return multi_for_each(m, [&](rtx_scene* s) { return rtx_scene_clear_objects(s); });
}

rtx_status rtx_multi_render(rtx_multi* m, uint32_t width, uint32_t height, uint64_t seed, uint32_t flags, double* out_rgb,
                            uint32_t* out_ids, rtx_stats* out_stats) {
  if (!m) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "null multi handle");
  if (!out_rgb) return fail(nullptr, RTX_ERR_INVALID_ARGUMENT, "out_rgb is null");
  const auto t0 = std::chrono::steady_clock::now();
  std::lock_guard<std::mutex> frame(m->render_mu);  // a concurrent caller's frame finishes first, results and all
  std::unique_lock<std::mutex> lk(m->mu);
  m->width = width, m->height = height, m->seed = seed, m->flags = flags, m->out_rgb = out_rgb, m->out_ids = out_ids;
  m->pending = (uint32_t)m->w.size();
  ++m->generation;
  m->cv_job.notify_all();
  m->cv_done.wait(lk, [&] { return m->pending == 0; });
  rtx_status st = RTX_OK;
  unsigned long long c[8] = {};
  float ms = 0.f;
  uint32_t launches = 0;
  for (auto& w : m->w) {
    if (w.status != RTX_OK && st == RTX_OK) st = w.status, m->last_error = w.error;
    for (int k = 0; k < 8; ++k) c[k] += w.counters[k];
    ms = std::max(ms, w.device_ms);
    launches += w.launches;
  }
  if (st != RTX_OK) return fail(nullptr, st, m->last_error);
  if (out_stats) {
    std::memset(out_stats, 0, sizeof(*out_stats));
    out_stats->rays = c[0], out_stats->paths = c[1], out_stats->triangle_panics = c[2], out_stats->hits = c[3];
    out_stats->nodes_visited = c[4], out_stats->prims_tested = c[5];
    out_stats->pixels = (uint64_t)width * height;
    out_stats->device_ms = ms;  // the slowest device, render + delivery
    out_stats->kernel_launches = launches;
    out_stats->accel_used = choose_accel(m->w[0].scene, flags);
    out_stats->total_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
  if (c[2] != 0)
    return fail(nullptr, RTX_ERR_REFERENCE_PANIC,
                std::to_string(c[2]) + " ray/triangle tests hit an assert_eq! of the reference (triangle.rs:76-94)");
  return RTX_OK;
}

}  // extern "C"
