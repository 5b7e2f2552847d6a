// include/tcs/batch_segmenter.h - tcs::BatchSegmenter: many independent scenes over the GPUs of one box, in C++.
//
// The reference segments one scene per RandomWalkerSegmentation::segment call on one CPU thread
// (include/impl/random_walker_segmentation.hpp:138-224).  Independent scenes are the path's natural shard (SURVEY.md 8e
// row 1): scene i goes to GPU i mod G, every GPU runs the whole pipeline for its scenes, and there is NO data-path
// collective and no NCCL.  Per GPU the driver keeps `contexts_per_device` worker threads, each with its own tcs_ctx
// (= its own CUDA stream): while one worker's batch is in its solve, the other uploads the next batch's clouds and
// builds its graphs - the copies and the launch-bound graph stages overlap the bandwidth-bound solve without any
// explicit double-buffering code.  A worker takes the next batch of its GPU from a shared cursor, so the GPU's workers
// stay busy until its scenes run out.  Batches go through tcs_segment_batch (block-diagonal joint solve).
//
// Host memory: outputs are received in per-worker pinned staging buffers and copied into the results (a device-to-host
// copy into pageable memory is staged by the driver and serialises the workers of a process); inputs are read where
// the caller keeps them - allocate them with tcs_host_alloc (pinned) for full PCIe speed, pageable inputs work but cap
// a multi-GPU process at the driver's staging rate (measured: 1.8 k scenes/s on 8 GPUs instead of 4 k+).
//
// Host code only: plain C++11 threads above the C ABI (include/tcs_b200.h); no CPU fallback - without a CUDA device the
// constructor throws.
#pragma once

#include <atomic>
#include <chrono>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../tcs_b200.h"

namespace tcs {

// one input scene: the caller keeps the buffers alive during segment()
struct SceneInput {
  const float* xyz = nullptr; size_t xyz_stride = 12;        // as tcs_cloud_upload (12 packed, 16 / 32 for point structs)
  const uint32_t* rgba = nullptr; size_t rgba_stride = 4;    // nullptr: no colour
  size_t n = 0;
  const float* seed_xyz = nullptr; const uint32_t* seed_label = nullptr; size_t ns = 0;
};

struct SceneResult {
  std::vector<int32_t> point_cluster;   // n: cluster of every input point (0..K-1 by ascending seed label, K = unlabelled)
  std::vector<uint32_t> vertex_colour;  // V
  size_t V = 0, E = 0;
  int ncolours = 0;
  int device = -1;
  int status = TCS_OK;
};

class BatchSegmenter {
 public:
  // devices: CUDA device indices to use (empty = device 0 only); batch: scenes per joint solve
  explicit BatchSegmenter(const std::vector<int>& devices = std::vector<int>(), int batch = 32, int contexts_per_device = 2)
      : devices_(devices.empty() ? std::vector<int>(1, 0) : devices), batch_(batch < 1 ? 1 : batch),
        per_device_(contexts_per_device < 1 ? 1 : contexts_per_device), resolution_(0.006f), preset_(TCS_PRESET_CLASS),
        mode_(TCS_MODE_PARITY), want_point_clusters_(true), seconds_(0.0) {
    tcs_solver_opts_default(&opts_);
    for (size_t d = 0; d < devices_.size(); ++d)
      for (int k = 0; k < per_device_; ++k) {
        tcs_ctx* c = nullptr;
        if (tcs_ctx_create(devices_[d], &c) != TCS_OK) {
          const std::string msg = std::string("tcs::BatchSegmenter: cannot create a CUDA context on device ") +
                                  std::to_string(devices_[d]) + " (no CPU fallback): " + tcs_last_error(nullptr);
          destroy();
          throw std::runtime_error(msg);
        }
        ctx_.push_back(c);
      }
    pinned_.resize(ctx_.size());
  }
  ~BatchSegmenter() { destroy(); }
  BatchSegmenter(const BatchSegmenter&) = delete;
  BatchSegmenter& operator=(const BatchSegmenter&) = delete;

  void setVoxelResolution(float r) { resolution_ = r; }
  void setPreset(int preset) { preset_ = preset; }            // TCS_PRESET_CLASS (the class path) | TCS_PRESET_CLI
  void setArithmeticMode(int mode) { mode_ = mode; }          // TCS_MODE_PARITY (reference summation order) | TCS_MODE_FAST
  void setSolverOptions(const tcs_solver_opts& o) { opts_ = o; }
  void setWantPointClusters(bool on) { want_point_clusters_ = on; }
  size_t numDevices() const { return devices_.size(); }
  double lastSeconds() const { return seconds_; }             // wall time of the last segment() call
  const std::vector<int>& lastIterations() const { return iterations_; }   // PCG iterations of every batch

  // Segments every scene; results come back in scene order.  Throws std::runtime_error on the first failing batch
  // (bad input, CUDA error); a batch that only hit max_iterations is reported through SceneResult::status.
  std::vector<SceneResult> segment(const std::vector<SceneInput>& scenes) {
    std::vector<SceneResult> results;
    segment(scenes, results);
    return results;
  }
  // Same, into `results` (resized to scenes.size()): the vectors of a previous call are reused, so a steady-state
  // pipeline allocates nothing - 24 worker threads faulting in 2.5 GB of fresh result pages per job contend on the
  // process's address-space lock and capped an 8-GPU process at 1.9 k scenes/s.
  void segment(const std::vector<SceneInput>& scenes, std::vector<SceneResult>& results) {
    const size_t G = devices_.size();
    results.resize(scenes.size());
    // scene i -> GPU i mod G; per GPU its scenes in order, cut into batches
    std::vector<std::vector<size_t>> mine(G);
    for (size_t i = 0; i < scenes.size(); ++i) mine[i % G].push_back(i);
    std::vector<std::atomic<size_t>> cursor(G);
    for (size_t d = 0; d < G; ++d) cursor[d].store(0);
    std::vector<std::string> errors(ctx_.size());
    std::vector<std::vector<int>> its(ctx_.size());
    const std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> workers;
    for (size_t w = 0; w < ctx_.size(); ++w)
      workers.emplace_back([&, w]() {
        const size_t d = w / (size_t)per_device_;
        std::vector<tcs_scene> arr;
        Pinned& pin = pinned_[w];
        // batches of at most batch_ scenes, sized so that the GPU's workers all get the same number of batches
        const size_t nd = mine[d].size(), per_round = (size_t)per_device_ * (size_t)batch_;
        const size_t rounds = nd ? (nd + per_round - 1) / per_round : 1;
        const size_t step = nd ? (nd + (size_t)per_device_ * rounds - 1) / ((size_t)per_device_ * rounds) : 1;
        while (errors[w].empty()) {
          const size_t b0 = cursor[d].fetch_add(step);
          if (b0 >= nd) break;
          const size_t b1 = b0 + step < nd ? b0 + step : nd;
          arr.assign(b1 - b0, tcs_scene());
          size_t words = 0;   // pinned staging of this batch's outputs: [point clusters][vertex colours] per scene
          for (size_t k = b0; k < b1; ++k) words += scenes[mine[d][k]].n * (want_point_clusters_ ? 2 : 1);
          if (!pin.reserve(words * 4)) {
            errors[w] = "tcs::BatchSegmenter: cannot allocate pinned host memory";
            break;
          }
          size_t at = 0;
          for (size_t k = b0; k < b1; ++k) {
            const SceneInput& in = scenes[mine[d][k]];
            SceneResult& out = results[mine[d][k]];
            tcs_scene& s = arr[k - b0];
            std::memset(&s, 0, sizeof s);
            s.xyz = in.xyz; s.xyz_stride = in.xyz_stride; s.rgba = in.rgba; s.rgba_stride = in.rgba_stride; s.n = in.n;
            s.seed_xyz = in.seed_xyz; s.seed_label = in.seed_label; s.ns = in.ns;
            (void)out;
            if (want_point_clusters_) { s.point_cluster = (int32_t*)pin.p + at; at += in.n; }
            s.vertex_colour = (uint32_t*)pin.p + at;     // capacity n >= V
            at += in.n;
          }
          tcs_solve_info info;
          const int st = tcs_segment_batch(ctx_[w], arr.data(), (int)arr.size(), resolution_, preset_, mode_, &opts_, &info);
          if (st != TCS_OK && st != TCS_ERR_NOT_CONVERGED) {
            errors[w] = std::string("tcs::BatchSegmenter: device ") + std::to_string(devices_[d]) + ": " + tcs_last_error(ctx_[w]);
            break;
          }
          its[w].push_back(info.iterations);
          for (size_t k = b0; k < b1; ++k) {
            SceneResult& out = results[mine[d][k]];
            const tcs_scene& s = arr[k - b0];
            out.V = s.V; out.E = s.E; out.ncolours = s.ncolours; out.device = devices_[d]; out.status = st;
            if (s.point_cluster) out.point_cluster.assign(s.point_cluster, s.point_cluster + s.n);
            out.vertex_colour.assign(s.vertex_colour, s.vertex_colour + s.V);
          }
        }
      });
    for (size_t w = 0; w < workers.size(); ++w) workers[w].join();
    seconds_ = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    iterations_.clear();
    for (size_t w = 0; w < its.size(); ++w) iterations_.insert(iterations_.end(), its[w].begin(), its[w].end());
    for (size_t w = 0; w < errors.size(); ++w)
      if (!errors[w].empty()) throw std::runtime_error(errors[w]);
  }

 private:
  struct Pinned {   // grow-only pinned buffer of one worker
    void* p = nullptr;
    size_t cap = 0;
    bool reserve(size_t bytes) {
      if (bytes <= cap) return true;
      tcs_host_free(p);
      cap = bytes + bytes / 4;
      p = tcs_host_alloc(cap);
      if (!p) cap = 0;
      return p != nullptr;
    }
    ~Pinned() { tcs_host_free(p); }
    Pinned() {}
    Pinned(const Pinned&) = delete;
    Pinned& operator=(const Pinned&) = delete;
    Pinned(Pinned&& o) : p(o.p), cap(o.cap) { o.p = nullptr; o.cap = 0; }
  };
  void destroy() {
    for (size_t i = 0; i < ctx_.size(); ++i) tcs_ctx_destroy(ctx_[i]);
    ctx_.clear();
  }
  std::vector<int> devices_;
  int batch_, per_device_;
  float resolution_;
  int preset_, mode_;
  bool want_point_clusters_;
  tcs_solver_opts opts_;
  std::vector<tcs_ctx*> ctx_;
  std::vector<Pinned> pinned_;
  double seconds_;
  std::vector<int> iterations_;
};

}  // namespace tcs
