// tcs_b200/cpp/batch_bench.cpp - tcs::BatchSegmenter (include/tcs/batch_segmenter.h) over the GPUs of one box:
// the C5 shape (BASELINE.json configs[4]) driven from C++ host code, no Python, no NCCL.
//   tcs_batch_bench scenes.bin [--gpus G] [--batch B] [--contexts C] [--repeat R] [--total N] [--fast] [--labels out.bin]
// scenes.bin (written by scripts/write_scenes.py): u64 count, then per scene
//   u64 n, f32 xyz[3n], u32 rgba[n], u64 ns, f32 seed_xyz[3ns], u32 seed_label[ns]
// --total N cycles the file's scenes to a job of N scenes.  Prints one JSON line per run.
// --labels writes, for the file's scenes in order: u64 V, u32 vertex_colour[V].
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "tcs/batch_segmenter.h"

// scene buffers live in pinned host memory (tcs_host_alloc): the copies to the GPUs run at PCIe speed
template <typename T>
struct PinnedArray {
  T* p = nullptr;
  size_t n = 0;
  bool resize(size_t count) {
    tcs_host_free(p);
    n = count;
    p = count ? (T*)tcs_host_alloc(count * sizeof(T)) : nullptr;
    return count == 0 || p != nullptr;
  }
  T* data() const { return p; }
  size_t size() const { return n; }
  ~PinnedArray() { tcs_host_free(p); }
  PinnedArray() {}
  PinnedArray(const PinnedArray&) = delete;
  PinnedArray& operator=(const PinnedArray&) = delete;
};
struct Scene {
  PinnedArray<float> xyz;
  PinnedArray<uint32_t> rgba;
  std::vector<float> seed_xyz;
  std::vector<uint32_t> seed_label;
};

static bool read_exact(FILE* f, void* p, size_t bytes) { return bytes == 0 || std::fread(p, 1, bytes, f) == bytes; }

int main(int argc, char** argv) {
  if (argc < 2) {
    std::fprintf(stderr, "usage: tcs_batch_bench scenes.bin [--gpus G] [--batch B] [--contexts C] [--repeat R] [--total N] [--fast] [--labels out.bin]\n");
    return 2;
  }
  int gpus = 1, batch = 32, contexts = 2, repeat = 3, fast = 0, clusters = 1, replicate = 0;
  size_t total = 0;
  std::string labels_path;
  for (int i = 2; i < argc; ++i) {
    const std::string a = argv[i];
    if (a == "--gpus" && i + 1 < argc) gpus = std::atoi(argv[++i]);
    else if (a == "--batch" && i + 1 < argc) batch = std::atoi(argv[++i]);
    else if (a == "--contexts" && i + 1 < argc) contexts = std::atoi(argv[++i]);
    else if (a == "--repeat" && i + 1 < argc) repeat = std::atoi(argv[++i]);
    else if (a == "--total" && i + 1 < argc) total = (size_t)std::atoll(argv[++i]);
    else if (a == "--fast") fast = 1;
    else if (a == "--no-clusters") clusters = 0;     // vertex colours only (no per-point cluster array)
    else if (a == "--replicate") replicate = 1;      // every GPU reads its own pinned copy of the scene file
    else if (a == "--labels" && i + 1 < argc) labels_path = argv[++i];
    else { std::fprintf(stderr, "tcs_batch_bench: unknown option %s\n", a.c_str()); return 2; }
  }
  FILE* f = std::fopen(argv[1], "rb");
  if (!f) { std::fprintf(stderr, "tcs_batch_bench: cannot open %s\n", argv[1]); return 2; }
  uint64_t count = 0;
  if (!read_exact(f, &count, 8) || count == 0 || count > (1u << 20)) { std::fprintf(stderr, "tcs_batch_bench: bad scene file\n"); return 2; }
  std::vector<Scene> file_scenes((size_t)count);
  bool pinned_ok = true;
  for (size_t s = 0; s < file_scenes.size(); ++s) {
    uint64_t n = 0, ns = 0;
    Scene& sc = file_scenes[s];
    if (!read_exact(f, &n, 8) || n > (1ull << 31)) { std::fprintf(stderr, "tcs_batch_bench: bad scene %zu\n", s); return 2; }
    pinned_ok = sc.xyz.resize(3 * n) && sc.rgba.resize(n);
    if (!pinned_ok) { std::fprintf(stderr, "tcs_batch_bench: tcs::BatchSegmenter needs a CUDA device (no CPU fallback): pinned host allocation failed\n"); return 3; }
    if (!read_exact(f, sc.xyz.data(), 12 * n) || !read_exact(f, sc.rgba.data(), 4 * n) || !read_exact(f, &ns, 8) || ns > (1u << 20)) {
      std::fprintf(stderr, "tcs_batch_bench: short scene %zu\n", s); return 2;
    }
    sc.seed_xyz.resize(3 * ns); sc.seed_label.resize(ns);
    if (!read_exact(f, sc.seed_xyz.data(), 12 * ns) || !read_exact(f, sc.seed_label.data(), 4 * ns)) {
      std::fprintf(stderr, "tcs_batch_bench: short scene %zu\n", s); return 2;
    }
  }
  std::fclose(f);
  if (total == 0) total = file_scenes.size();
  // --replicate: one pinned copy of the scenes per GPU (scene i of the job runs on GPU i mod G)
  std::vector<std::vector<Scene> > copies;
  if (replicate && gpus > 1) {
    copies.resize((size_t)gpus - 1);
    for (size_t c = 0; c < copies.size(); ++c) {
      copies[c] = std::vector<Scene>(file_scenes.size());
      for (size_t k = 0; k < file_scenes.size(); ++k) {
        Scene& d = copies[c][k];
        const Scene& o = file_scenes[k];
        if (!d.xyz.resize(o.xyz.size()) || !d.rgba.resize(o.rgba.size())) { std::fprintf(stderr, "tcs_batch_bench: pinned allocation failed\n"); return 3; }
        std::memcpy(d.xyz.data(), o.xyz.data(), o.xyz.size() * 4);
        std::memcpy(d.rgba.data(), o.rgba.data(), o.rgba.size() * 4);
        d.seed_xyz = o.seed_xyz; d.seed_label = o.seed_label;
      }
    }
  }
  std::vector<tcs::SceneInput> job(total);
  for (size_t i = 0; i < total; ++i) {
    const size_t dev = i % (size_t)gpus;
    const Scene& sc = (dev > 0 && !copies.empty()) ? copies[dev - 1][i % file_scenes.size()] : file_scenes[i % file_scenes.size()];
    tcs::SceneInput& in = job[i];
    in.xyz = sc.xyz.data(); in.xyz_stride = 12; in.rgba = sc.rgba.data(); in.rgba_stride = 4; in.n = sc.rgba.size();
    in.seed_xyz = sc.seed_xyz.data(); in.seed_label = sc.seed_label.data(); in.ns = sc.seed_label.size();
  }
  try {
    std::vector<int> devices;
    for (int d = 0; d < gpus; ++d) devices.push_back(d);
    tcs::BatchSegmenter seg(devices, batch, contexts);
    seg.setVoxelResolution(0.006f);
    seg.setArithmeticMode(fast ? TCS_MODE_FAST : TCS_MODE_PARITY);
    seg.setWantPointClusters(clusters != 0);
    std::vector<tcs::SceneResult> res;
    double best = 1e300;
    for (int r = 0; r < repeat + 1; ++r) {            // the first run warms allocators, module loading and the pools
      seg.segment(job, res);                         // reuses the result vectors of the previous run
      if (r > 0 && seg.lastSeconds() < best) best = seg.lastSeconds();
    }
    if (repeat < 1) best = seg.lastSeconds();
    size_t V = 0, bad = 0;
    for (size_t i = 0; i < res.size(); ++i) { V += res[i].V; bad += res[i].status != TCS_OK; }
    int it_min = 1 << 30, it_max = 0;
    for (size_t i = 0; i < seg.lastIterations().size(); ++i) {
      it_min = seg.lastIterations()[i] < it_min ? seg.lastIterations()[i] : it_min;
      it_max = seg.lastIterations()[i] > it_max ? seg.lastIterations()[i] : it_max;
    }
    std::printf("{\"tool\": \"tcs_batch_bench\", \"host\": \"C++ (tcs::BatchSegmenter, one thread + context per worker, no NCCL)\", "
                "\"n_gpus\": %d, \"contexts_per_gpu\": %d, \"batch\": %d, \"scenes\": %zu, \"mode\": \"%s\", \"seconds\": %.4f, "
                "\"scenes_per_s\": %.1f, \"vertices_per_s\": %.0f, \"vertices\": %zu, \"not_converged\": %zu, "
                "\"pcg_iterations_per_batch_min_max\": [%d, %d], \"includes\": \"host buffers in, labels back on the host (e2e)\"}\n",
                gpus, contexts, batch, total, fast ? "fast" : "parity", best, (double)total / best, (double)V / best, V, bad,
                it_max ? it_min : 0, it_max);
    if (!labels_path.empty()) {
      FILE* o = std::fopen(labels_path.c_str(), "wb");
      if (!o) { std::fprintf(stderr, "tcs_batch_bench: cannot write %s\n", labels_path.c_str()); return 2; }
      for (size_t i = 0; i < file_scenes.size() && i < res.size(); ++i) {
        const uint64_t v = res[i].V;
        std::fwrite(&v, 8, 1, o);
        std::fwrite(res[i].vertex_colour.data(), 4, v, o);
      }
      std::fclose(o);
    }
    return bad ? 4 : 0;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "tcs_batch_bench: %s\n", e.what());
    return 3;
  }
}
