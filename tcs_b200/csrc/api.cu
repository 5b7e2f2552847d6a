// tcs_b200/csrc/api.cu - context management, cloud upload, graph accessors of the C ABI (include/tcs_b200.h).
#include <cub/cub.cuh>

#include <cstdlib>
#include "common.cuh"

int tcs_fail(tcs_ctx* ctx, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf;
  return code;
}

static thread_local std::string g_create_err;

extern "C" const char* tcs_version(void) { return "tcs_b200 0.1 (sm_100a)"; }

extern "C" int tcs_ctx_create_on_stream(int device, void* cuda_stream, tcs_ctx** out) {
  if (!out) return TCS_ERR_INPUT;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0 || device < 0 || device >= count) {
    g_create_err = std::string("no usable CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "bad index");
    return TCS_ERR_CUDA;
  }
  tcs_ctx* ctx = new tcs_ctx;
  ctx->device = device;
  if (cudaSetDevice(device) != cudaSuccess) { delete ctx; return TCS_ERR_CUDA; }
  if (cuda_stream) {
    ctx->stream = (cudaStream_t)cuda_stream;
  } else {
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return TCS_ERR_CUDA; }
    ctx->own_stream = true;
  }
  cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
  // keep freed blocks in the stream-ordered pool: allocation inside the pipeline never hits the driver twice
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
    uint64_t thr = UINT64_MAX;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
  }
  if (cudaMallocHost(&ctx->h_pin, 4096) != cudaSuccess || cudaMalloc(&ctx->d_scal, 4096) != cudaSuccess) {
    delete ctx;
    return TCS_ERR_CUDA;
  }
  for (int i = 0; i < 32; ++i) cudaEventCreate(&ctx->ev[i]);
  if (const char* v = getenv("TCS_PDL")) ctx->pdl = atoi(v) != 0;
  if (const char* v = getenv("TCS_PDL_ROWS")) ctx->pdl_rows = atoll(v);
  *out = ctx;
  return TCS_OK;
}

extern "C" int tcs_ctx_create(int device, tcs_ctx** out) { return tcs_ctx_create_on_stream(device, nullptr, out); }

extern "C" void tcs_ctx_destroy(tcs_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->own_cloud && ctx->d_cloud) cudaFree(ctx->d_cloud);
  if (ctx->d_stage) cudaFree(ctx->d_stage);
  if (ctx->h_pin) cudaFreeHost(ctx->h_pin);
  if (ctx->d_scal) cudaFree(ctx->d_scal);
  for (int i = 0; i < 32; ++i) cudaEventDestroy(ctx->ev[i]);
  if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

// page-locked host memory usable from every device of the process: host buffers handed to the library are copied at
// full PCIe speed and asynchronously only when they are pinned (pageable memory goes through the driver's staging copy)
extern "C" void* tcs_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (bytes == 0 || cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}
extern "C" void tcs_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

extern "C" const char* tcs_last_error(const tcs_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }
extern "C" void* tcs_ctx_stream(tcs_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
extern "C" long long tcs_ctx_launch_count(const tcs_ctx* ctx) { return ctx ? ctx->launches : 0; }

extern "C" int tcs_ctx_synchronize(tcs_ctx* ctx) {
  if (!ctx) return TCS_ERR_INPUT;
  CK(cudaStreamSynchronize(ctx->stream));
  return TCS_OK;
}

// ---- helpers ------------------------------------------------------------------------------------------------
int tcs_impl_read_u32(tcs_ctx* ctx, const uint32_t* d_ptr, uint32_t* out) {
  CK(cudaMemcpyAsync(ctx->h_pin, d_ptr, 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  *out = *(uint32_t*)ctx->h_pin;
  return TCS_OK;
}

__global__ void k_set_last(const uint32_t* in, uint32_t* out, size_t n) {
  // out holds the exclusive sum of in[0..n); append the total as out[n]
  out[n] = n ? out[n - 1] + in[n - 1] : 0u;
}

int tcs_impl_scan_u32(tcs_ctx* ctx, const uint32_t* d_in, uint32_t* d_out, size_t n) {
  if (n > 0) {
    size_t bytes = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, bytes, d_in, d_out, (int)n, ctx->stream));
    DBuf<char> tmp;
    TRY(tmp.alloc(ctx, bytes));
    CK(cub::DeviceScan::ExclusiveSum(tmp.p, bytes, d_in, d_out, (int)n, ctx->stream));
    ctx->launches += 2;
  }
  LAUNCH(ctx, k_set_last, 1, 1, 0, d_in, d_out, n);
  return TCS_OK;
}

static int ensure_stage(tcs_ctx* ctx, size_t bytes) {
  if (ctx->stage_cap >= bytes) return TCS_OK;
  if (ctx->d_stage) { CK(cudaStreamSynchronize(ctx->stream)); CK(cudaFree(ctx->d_stage)); ctx->d_stage = nullptr; }
  CK(cudaMalloc(&ctx->d_stage, bytes));
  ctx->stage_cap = bytes;
  return TCS_OK;
}

// ---- cloud upload -------------------------------------------------------------------------------------------
// stage layout: xyz bytes (n * xyz_stride) then rgba bytes (n * rgba_stride); repack to 16-byte records
__global__ void k_pack_cloud(const unsigned char* __restrict__ xyz, size_t xyz_stride,
                             const unsigned char* __restrict__ rgba, size_t rgba_stride, size_t n,
                             float4* __restrict__ out) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* p = (const float*)(xyz + i * xyz_stride);
  float4 r;
  r.x = p[0]; r.y = p[1]; r.z = p[2];
  r.w = rgba ? __uint_as_float(*(const uint32_t*)(rgba + i * rgba_stride)) : 0.0f;
  out[i] = r;
}

extern "C" int tcs_cloud_upload(tcs_ctx* ctx, const float* xyz, size_t xyz_stride, const uint32_t* rgba,
                                size_t rgba_stride, size_t n) {
  if (!ctx) return TCS_ERR_INPUT;
  if (n && (!xyz || xyz_stride < 12 || (xyz_stride & 3) || (rgba && (rgba_stride < 4 || (rgba_stride & 3)))))
    return tcs_fail(ctx, TCS_ERR_INPUT, "tcs_cloud_upload: bad pointers/strides");
  if (n >= 0x7FFFFFFFu) return tcs_fail(ctx, TCS_ERR_INPUT, "tcs_cloud_upload: more than 2^31 points");
  CK(cudaSetDevice(ctx->device));
  if (!ctx->own_cloud || ctx->cloud_cap < n) {
    if (ctx->own_cloud && ctx->d_cloud) { CK(cudaStreamSynchronize(ctx->stream)); CK(cudaFree(ctx->d_cloud)); }
    ctx->d_cloud = nullptr;
    if (n) CK(cudaMalloc((void**)&ctx->d_cloud, n * sizeof(float4)));
    ctx->cloud_cap = n;
    ctx->own_cloud = true;
  }
  ctx->n_points = n;
  ctx->has_color = rgba != nullptr;
  if (n == 0) return TCS_OK;
  // interleaved AoS (pcl point structs): xyz and rgba live in the same records -> one copy
  const unsigned char* xb = (const unsigned char*)xyz;
  const unsigned char* cb = (const unsigned char*)rgba;
  bool same_records = rgba && rgba_stride == xyz_stride && cb > xb && (size_t)(cb - xb) + 4 <= xyz_stride;
  size_t xyz_bytes = (n - 1) * xyz_stride + 12;
  if (same_records) xyz_bytes = (n - 1) * xyz_stride + (size_t)(cb - xb) + 4;
  size_t xyz_pad = (xyz_bytes + 15) & ~(size_t)15;
  size_t rgba_bytes = (rgba && !same_records) ? (n - 1) * rgba_stride + 4 : 0;
  TRY(ensure_stage(ctx, xyz_pad + rgba_bytes));
  unsigned char* st = (unsigned char*)ctx->d_stage;
  CK(cudaMemcpyAsync(st, xyz, xyz_bytes, cudaMemcpyHostToDevice, ctx->stream));
  const unsigned char* d_rgba = nullptr;
  if (same_records) d_rgba = st + (cb - xb);
  else if (rgba) {
    CK(cudaMemcpyAsync(st + xyz_pad, rgba, rgba_bytes, cudaMemcpyHostToDevice, ctx->stream));
    d_rgba = st + xyz_pad;
  }
  LAUNCH(ctx, k_pack_cloud, div_up(n, 256), 256, 0, st, xyz_stride, d_rgba, rgba_stride, n, ctx->d_cloud);
  CK(cudaGetLastError());
  return TCS_OK;
}

extern "C" int tcs_cloud_set_device(tcs_ctx* ctx, const void* d_xyzc, size_t n, int has_color) {
  if (!ctx || (n && !d_xyzc)) return TCS_ERR_INPUT;
  if (n >= 0x7FFFFFFFu) return tcs_fail(ctx, TCS_ERR_INPUT, "more than 2^31 points");
  if (ctx->own_cloud && ctx->d_cloud) { CK(cudaStreamSynchronize(ctx->stream)); CK(cudaFree(ctx->d_cloud)); }
  ctx->d_cloud = (float4*)d_xyzc;
  ctx->own_cloud = false;
  ctx->cloud_cap = 0;
  ctx->n_points = n;
  ctx->has_color = has_color != 0;
  return TCS_OK;
}
