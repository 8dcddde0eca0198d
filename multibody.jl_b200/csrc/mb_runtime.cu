// libmultibody_b200 runtime: C ABI (include/mbb200.h), device management, sharding,
// host<->device staging (streamed: finished chunks of trajectories are copied back while the
// ensemble kernel is still running), ensemble-statistic reduction (NCCL across processes),
// FP64 peak probe.
// No CPU fallback: every compute entry fails with MBB200_ERR_NO_DEVICE without a GPU.
//
// Threading ("re-entrant per model handle"): every entry point takes the
// registry lock shared (init / shutdown take it exclusively); a solve owns the devices it uses
// for its duration (per-device mutex, always acquired in device order), so concurrent callers
// serialise per device instead of racing on the stream, the work counter and the workspace.
#include <dlfcn.h>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <thread>
#include <vector>

#pragma GCC visibility push(default)
#include "../../include/mbb200.h"
#pragma GCC visibility pop
#include "mb_registry.h"

// ---------------------------------------------------------------------------------------
// registry
// ---------------------------------------------------------------------------------------
static MbKernelEntry* g_models = nullptr;
void mb_register_model(MbKernelEntry* e) {
  e->next = g_models;
  g_models = e;
}
const MbKernelEntry* mb_find_model(const char* name) {
  for (const MbKernelEntry* e = g_models; e; e = e->next)
    if (std::strcmp(e->name, name) == 0) return e;
  return nullptr;
}

// ---------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
#define CU(call)                                                                                 \
  do {                                                                                           \
    cudaError_t e_ = (call);                                                                     \
    if (e_ != cudaSuccess)                                                                       \
      return fail(MBB200_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),       \
                  __FILE__, __LINE__);                                                           \
  } while (0)

// ---------------------------------------------------------------------------------------
// devices
// ---------------------------------------------------------------------------------------
struct Workspace {          // grow-only device scratch of one device
  void* ptr = nullptr;
  size_t bytes = 0;
};
struct Device {
  int ordinal = -1;
  int sm_count = 0;
  cudaStream_t stream = nullptr;    // uploads + kernels
  cudaStream_t cstream = nullptr;   // downloads of finished chunks, concurrent with the kernel
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  unsigned long long* queue = nullptr;
  unsigned* d_done = nullptr;       // per-chunk count of retired trajectories (device)
  int* h_flags = nullptr;           // per-chunk "complete" flags: pinned host memory, written by the kernel
  int* d_flags = nullptr;           // device alias of h_flags
  int flag_cap = 0;
  Workspace ws;
  Workspace ws2;                    // kernel scratch of models that ask for it (MbKernelEntry::scratch_bytes)
  double last_ms = 0.0;
  int last_launches = 0;
  std::mutex mu;                    // held by the call that owns stream / queue / ws
  unsigned char* static_scratch = nullptr;   // device address of mb_static_scratch on THIS device (mbb200_solve_device_static)
};
static std::vector<std::unique_ptr<Device>> g_dev;
static std::shared_mutex g_rw;
typedef std::shared_lock<std::shared_mutex> RLock;

namespace mbjit { struct Loaded; }
struct mbb200_model {
  const MbKernelEntry* entry;
  std::vector<double> tables;
  std::vector<double*> d_tables;   // per device
  // models compiled at load time from an MBIR2 image (mb_jit.inc): their own descriptor and the loaded modules
  MbKernelEntry own = {};
  std::string own_name;
  mbjit::Loaded* jit = nullptr;
};

extern "C" int mbb200_abi_version(void) { return MBB200_ABI_VERSION; }
extern "C" const char* mbb200_last_error(void) { return g_err; }
extern "C" int mbb200_device_count(void) {
  RLock lk(g_rw);
  return (int)g_dev.size();
}

static void destroy_device(Device& d) {
  cudaSetDevice(d.ordinal);
  if (d.stream) cudaStreamSynchronize(d.stream);
  if (d.cstream) cudaStreamSynchronize(d.cstream);
  if (d.ws.ptr) cudaFree(d.ws.ptr);
  if (d.ws2.ptr) cudaFree(d.ws2.ptr);
  if (d.queue) cudaFree(d.queue);
  if (d.d_done) cudaFree(d.d_done);
  if (d.h_flags) cudaFreeHost(d.h_flags);
  if (d.ev0) cudaEventDestroy(d.ev0);
  if (d.ev1) cudaEventDestroy(d.ev1);
  if (d.stream) cudaStreamDestroy(d.stream);
  if (d.cstream) cudaStreamDestroy(d.cstream);
}

static void comm_destroy_locked();

extern "C" int mbb200_init_devices(const int* ordinals, int n) {
  std::unique_lock<std::shared_mutex> lk(g_rw);
  if (!g_dev.empty()) {
    bool same = (int)g_dev.size() == n;
    for (int i = 0; same && i < n; i++) same = g_dev[i]->ordinal == ordinals[i];
    if (same) return MBB200_OK;
    return fail(MBB200_ERR_BAD_ARG, "already initialised with a different device set; call mbb200_shutdown first");
  }
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return fail(MBB200_ERR_NO_DEVICE, "no CUDA device: %s", e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  if (n < 1 || !ordinals) return fail(MBB200_ERR_BAD_ARG, "n_gpus must be >= 1");
  std::vector<std::unique_ptr<Device>> devs;
  auto bail = [&](int rc) {
    for (auto& d : devs) destroy_device(*d);
    return rc;
  };
  for (int i = 0; i < n; i++) {
    if (ordinals[i] < 0 || ordinals[i] >= count)
      return bail(fail(MBB200_ERR_NO_DEVICE, "device ordinal %d not present (%d visible)", ordinals[i], count));
    devs.emplace_back(new Device);
    Device& d = *devs.back();
    d.ordinal = ordinals[i];
    cudaDeviceProp prop;
    cudaError_t ce = cudaSetDevice(d.ordinal);
    if (ce == cudaSuccess) ce = cudaGetDeviceProperties(&prop, d.ordinal);
    if (ce != cudaSuccess) return bail(fail(MBB200_ERR_CUDA, "device %d: %s", d.ordinal, cudaGetErrorString(ce)));
    if (prop.major < 10)
      return bail(fail(MBB200_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", d.ordinal, prop.major, prop.minor));
    d.sm_count = prop.multiProcessorCount;
    d.flag_cap = 4096;
    ce = cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&d.cstream, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaEventCreate(&d.ev0);
    if (ce == cudaSuccess) ce = cudaEventCreate(&d.ev1);
    if (ce == cudaSuccess) ce = cudaMalloc(&d.queue, sizeof(unsigned long long));
    if (ce == cudaSuccess) ce = cudaMalloc(&d.d_done, sizeof(unsigned) * d.flag_cap);
    if (ce == cudaSuccess) ce = cudaHostAlloc(&d.h_flags, sizeof(int) * d.flag_cap, cudaHostAllocMapped | cudaHostAllocPortable);
    if (ce == cudaSuccess) ce = cudaHostGetDevicePointer((void**)&d.d_flags, d.h_flags, 0);
    if (ce != cudaSuccess) return bail(fail(MBB200_ERR_CUDA, "device %d setup: %s", d.ordinal, cudaGetErrorString(ce)));
  }
  g_dev.swap(devs);
  return MBB200_OK;
}

extern "C" int mbb200_init(int n_gpus) {
  if (n_gpus < 1) return fail(MBB200_ERR_BAD_ARG, "n_gpus must be >= 1");
  std::vector<int> ord(n_gpus);
  for (int i = 0; i < n_gpus; i++) ord[i] = i;
  return mbb200_init_devices(ord.data(), n_gpus);
}

extern "C" void mbb200_shutdown(void) {
  std::unique_lock<std::shared_mutex> lk(g_rw);
  comm_destroy_locked();
  for (auto& d : g_dev) {
    std::lock_guard<std::mutex> dl(d->mu);
    destroy_device(*d);
  }
  g_dev.clear();
}

static int ws_reserve(Device& d, size_t bytes) {     // caller holds d.mu
  if (d.ws.bytes >= bytes) return MBB200_OK;
  if (d.ws.ptr) CU(cudaFree(d.ws.ptr));
  d.ws.ptr = nullptr;
  d.ws.bytes = 0;
  CU(cudaMalloc(&d.ws.ptr, bytes));
  d.ws.bytes = bytes;
  return MBB200_OK;
}

// ---------------------------------------------------------------------------------------
// models
// ---------------------------------------------------------------------------------------
// image tables: [t0, h, nk] + per table nk (value, z) pairs -> device layout of mb_common.cuh: 8-double header +
// (nk + 1) interval rows x ntab x 4 polynomial coefficients
static int convert_tables(const double* img, long long n_table, std::vector<double>& out) {
  const double t0 = img[0], hh = img[1];
  const long long nk = (long long)img[2];
  const long long ntab = nk > 0 ? (n_table - 3) / (2 * nk) : 0;
  if (nk < 0 || (nk > 0 && (nk < 2 || !(hh > 0.0) || 3 + 2 * nk * ntab != n_table)))
    return fail(MBB200_ERR_BAD_ARG, "inconsistent table block (nk=%lld, n_table=%lld)", nk, n_table);
  out.assign(MB_TB_HEADER + (size_t)(nk + 1) * ntab * 4, 0.0);
  out[0] = t0; out[1] = hh > 0.0 ? hh : 1.0; out[2] = (double)nk; out[3] = (double)ntab;
  for (long long k = 0; k < ntab; k++) {
    const double* T = img + 3 + (size_t)k * 2 * nk;
    for (long long row = 0; row <= nk; row++) {
      double* c = out.data() + MB_TB_HEADER + ((size_t)row * ntab + k) * 4;
      if (row == 0) { c[0] = T[0]; continue; }
      if (row == nk) { c[0] = T[2 * (nk - 1)]; continue; }
      const long long i = row - 1;
      const double v0 = T[2 * i], z0 = T[2 * i + 1], v1 = T[2 * i + 2], z1 = T[2 * i + 3];
      c[0] = v0;
      c[1] = (v1 - v0) / hh - hh * (2.0 * z0 + z1) / 6.0;
      c[2] = 0.5 * z0;
      c[3] = (z1 - z0) / (6.0 * hh);
    }
  }
  return MBB200_OK;
}
static int upload_tables(mbb200_model* m) {
  m->d_tables.assign(g_dev.size(), nullptr);
  for (size_t i = 0; i < g_dev.size(); i++) {
    cudaError_t ce = cudaSetDevice(g_dev[i]->ordinal);
    if (ce == cudaSuccess) ce = cudaMalloc(&m->d_tables[i], m->tables.size() * sizeof(double));
    if (ce == cudaSuccess) ce = cudaMemcpy(m->d_tables[i], m->tables.data(), m->tables.size() * sizeof(double), cudaMemcpyHostToDevice);
    if (ce != cudaSuccess) return fail(MBB200_ERR_CUDA, "table upload failed: %s", cudaGetErrorString(ce));
  }
  return MBB200_OK;
}

#include "mb_jit.inc"
static int jit_model_load(const void* image, size_t len, mbb200_model** out);
static void jit_unload(mbb200_model* m);

static void model_free_locked(mbb200_model* m);
extern "C" int mbb200_model_load(const void* image, size_t len, mbb200_model** out) {
  if (!image || !out || len < sizeof(mbb200_image_header)) return fail(MBB200_ERR_BAD_ARG, "bad model image");
  RLock rl(g_rw);
  if (g_dev.empty()) return fail(MBB200_ERR_NOT_INIT, "mbb200_init has not been called");
  mbb200_image_header h;
  std::memcpy(&h, image, sizeof(h));
  if (std::memcmp(h.magic, "MBIR2", 5) == 0) return jit_model_load(image, len, out);      // the image carries the model
  if (std::memcmp(h.magic, "MBIR1", 5) != 0) return fail(MBB200_ERR_BAD_ARG, "bad image magic");
  h.name[sizeof(h.name) - 1] = 0;
  const MbKernelEntry* e = mb_find_model(h.name);
  if (!e) return fail(MBB200_ERR_UNKNOWN_MODEL, "no compiled kernel set for model '%s'", h.name);
  if (e->n != h.n || e->m != h.m)
    return fail(MBB200_ERR_BAD_ARG, "image of '%s' has n=%d m=%d, compiled model has n=%d m=%d", h.name, h.n, h.m, e->n, e->m);
  mbb200_model* m = new mbb200_model;
  if (h.n_table < 3 || len < sizeof(h) + (size_t)h.n_table * sizeof(double))
  {
    delete m;
    return fail(MBB200_ERR_BAD_ARG, "truncated table data");
  }
  m->entry = e;
  {
    std::vector<double> img(h.n_table);
    std::memcpy(img.data(), (const char*)image + sizeof(h), (size_t)h.n_table * sizeof(double));
    int trc = convert_tables(img.data(), (long long)h.n_table, m->tables);
    if (trc) { delete m; return trc; }
  }
  m->d_tables.assign(g_dev.size(), nullptr);
  for (size_t i = 0; i < g_dev.size(); i++) {
    cudaError_t ce = cudaSetDevice(g_dev[i]->ordinal);
    if (ce == cudaSuccess) ce = cudaMalloc(&m->d_tables[i], m->tables.size() * sizeof(double));
    if (ce == cudaSuccess) ce = cudaMemcpy(m->d_tables[i], m->tables.data(), m->tables.size() * sizeof(double), cudaMemcpyHostToDevice);
    if (ce != cudaSuccess) {
      int rc = fail(MBB200_ERR_CUDA, "table upload failed: %s", cudaGetErrorString(ce));
      model_free_locked(m);
      return rc;
    }
  }
  *out = m;
  return MBB200_OK;
}

static void model_free_locked(mbb200_model* m) {
  if (!m) return;
  jit_unload(m);
  for (size_t i = 0; i < m->d_tables.size() && i < g_dev.size(); i++)
    if (m->d_tables[i]) {
      cudaSetDevice(g_dev[i]->ordinal);
      cudaFree(m->d_tables[i]);
    }
  delete m;
}
extern "C" void mbb200_model_free(mbb200_model* m) {
  RLock rl(g_rw);
  model_free_locked(m);
}

extern "C" int mbb200_model_info(const mbb200_model* m, int32_t* n, int32_t* mm) {
  if (!m) return fail(MBB200_ERR_BAD_ARG, "null model");
  if (n) *n = m->entry->n;
  if (mm) *mm = m->entry->m;
  return MBB200_OK;
}

extern "C" int mbb200_model_flops(const mbb200_model* m, int64_t flops[5]) {
  if (!m || !flops) return fail(MBB200_ERR_BAD_ARG, "null argument");
  for (int i = 0; i < 5; i++) flops[i] = m->entry->flops[i];
  return MBB200_OK;
}

// ---------------------------------------------------------------------------------------
// models compiled at load time (MBIR2 images; mb_jit.inc)
// ---------------------------------------------------------------------------------------
static int jit_model_load(const void* image, size_t len, mbb200_model** out) {
  mbjit::Image im;
  int rc = mbjit::parse(image, len, im);
  if (rc) return rc;
  mbjit::Compiled cc;
  rc = mbjit::compile(im, cc);                 // cache hit: no NVRTC needed
  if (rc) return rc;
  rc = mbjit::bind_api();                      // the driver API is needed to load the cubin either way
  if (rc) return rc;
  mbb200_model* m = new mbb200_model;
  m->own_name = im.name;
  m->own = MbKernelEntry{};
  m->own.name = m->own_name.c_str();
  m->own.n = im.n; m->own.m = im.m; m->own.nlu = im.nlu; m->own.is_ode = im.is_ode; m->own.nobs = im.nobs;
  for (int k = 0; k < 5; k++) m->own.flops[k] = im.flops[k];
  m->entry = &m->own;
  rc = convert_tables(im.tables.data(), (long long)im.tables.size(), m->tables);
  if (rc) { delete m; return rc; }
  m->jit = new mbjit::Loaded;
  m->jit->mod.assign(g_dev.size(), nullptr);
  for (int k = 0; k < mbjit::NKERN; k++) m->jit->fn[k].assign(g_dev.size(), nullptr);
  rc = upload_tables(m);
  for (size_t i = 0; rc == MBB200_OK && i < g_dev.size(); i++) {
    cudaError_t ce = cudaSetDevice(g_dev[i]->ordinal);
    if (ce == cudaSuccess) ce = cudaFree(0);                                    // make the primary context current
    if (ce != cudaSuccess) { rc = fail(MBB200_ERR_CUDA, "device %d: %s", g_dev[i]->ordinal, cudaGetErrorString(ce)); break; }
    int dr = mbjit::g_api.ModuleLoadData(&m->jit->mod[i], cc.cubin.data());
    if (dr != 0) { rc = fail(MBB200_ERR_CUDA, "cuModuleLoadData failed (%d) for model '%s'", dr, im.name.c_str()); break; }
    for (int k = 0; k < mbjit::NKERN; k++)
      if (!cc.names[k].empty()) {
        dr = mbjit::g_api.ModuleGetFunction(&m->jit->fn[k][i], m->jit->mod[i], cc.names[k].c_str());
        if (dr != 0) { rc = fail(MBB200_ERR_CUDA, "kernel %s not found in the compiled image (%d)", cc.names[k].c_str(), dr); break; }
      }
  }
  if (rc) { model_free_locked(m); return rc; }
  *out = m;
  return MBB200_OK;
}

// Parse and compile an MBIR2 image into the cubin cache WITHOUT a device (warm the cache on a build machine; also what
// the CPU-side tests exercise: NVRTC needs no GPU).
extern "C" int mbb200_image_precompile(const void* image, size_t len) {
  if (!image || len < 8 || std::memcmp(image, "MBIR2", 5) != 0) return fail(MBB200_ERR_BAD_ARG, "not an MBIR2 image");
  mbjit::Image im;
  int rc = mbjit::parse(image, len, im);
  if (rc) return rc;
  mbjit::Compiled cc;
  return mbjit::compile(im, cc);
}

static void jit_unload(mbb200_model* m) {
  if (!m->jit) return;
  for (size_t i = 0; i < m->jit->mod.size() && i < g_dev.size(); i++)
    if (m->jit->mod[i]) {
      cudaSetDevice(g_dev[i]->ordinal);
      mbjit::g_api.ModuleUnload(m->jit->mod[i]);
    }
  delete m->jit;
  m->jit = nullptr;
}

// same launch geometry as MB_DEFINE_MODEL (csrc/mb_model_def.cuh): 128-thread CTAs, persistent grid
static cudaError_t jit_launch(const mbb200_model* m, int dev_index, int alg, const MbArgs& a, int sm_count, cudaStream_t st) {
  void* fn = (alg >= 1 && alg <= 3) ? m->jit->fn[alg - 1][dev_index] : (alg == 4 ? m->jit->fn[4][dev_index] : nullptr);
  if (!fn) return cudaErrorNotSupported;
  int bps = 1;
  if (mbjit::g_api.Occupancy(&bps, fn, 128, 0) != 0 || bps < 1) bps = 1;
  long long grid = (long long)sm_count * bps;
  const long long need = (a.N + 127) / 128;
  if (grid > need) grid = need;
  MbArgs args = a;
  void* params[1] = {&args};
  const int dr = mbjit::g_api.LaunchKernel(fn, (unsigned)grid, 1, 1, 128, 1, 1, 0, st, params, nullptr);
  return dr == 0 ? cudaSuccess : cudaErrorLaunchFailure;
}
static cudaError_t jit_launch_observed(const mbb200_model* m, int dev_index, const MbObsArgs& a, int sm_count, cudaStream_t st) {
  void* fn = m->jit->fn[3][dev_index];
  if (!fn) return cudaErrorNotSupported;
  long long blocks = (a.N * (long long)a.n_save + 255) / 256;
  if (blocks > (long long)sm_count * 8) blocks = (long long)sm_count * 8;
  if (blocks < 1) blocks = 1;
  MbObsArgs args = a;
  void* params[1] = {&args};
  const int dr = mbjit::g_api.LaunchKernel(fn, (unsigned)blocks, 1, 1, 256, 1, 1, 0, st, params, nullptr);
  return dr == 0 ? cudaSuccess : cudaErrorLaunchFailure;
}

// ---------------------------------------------------------------------------------------
// ensemble statistics: count, sum(x), sum(x^2) over successful trajectories
// (SciML `reduction`/`u_init` of an EnsembleProblem)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) mb_reduce_kernel(const double* __restrict__ out_u, const int* __restrict__ rc,
                                                        long long N, int rows, double* __restrict__ res) {
  const int row = blockIdx.x;   // rows = n_save*n value rows, plus one extra block for the count
  double s = 0.0, s2 = 0.0;
  if (row < rows) {
    const double* x = out_u + (size_t)row * N;
    for (long long i = threadIdx.x; i < N; i += blockDim.x)
      if (rc[i] == 1) { const double v = x[i]; s += v; s2 += v * v; }
  } else {
    for (long long i = threadIdx.x; i < N; i += blockDim.x) s += (rc[i] == 1) ? 1.0 : 0.0;
  }
  __shared__ double sh[2][8];
  for (int o = 16; o > 0; o >>= 1) { s += __shfl_down_sync(0xffffffffu, s, o); s2 += __shfl_down_sync(0xffffffffu, s2, o); }
  if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = s; sh[1][threadIdx.x >> 5] = s2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0;
    for (int w = 0; w < 8; w++) { a += sh[0][w]; b += sh[1][w]; }
    if (row < rows) { res[1 + row] = a; res[1 + rows + row] = b; }
    else res[0] = a;
  }
}

// ---------------------------------------------------------------------------------------
// cross-process reduction of the ensemble statistic ("one
// ncclAllReduce(sum, f64) over [count, sum x, sum x^2] ... only when the caller asks for an
// ensemble statistic").  NCCL is bound at run time (dlopen of libnccl.so.2: the copy already
// loaded in the process, MBB200_NCCL_LIB, or the system one), so the library itself links
// only cudart and single-process users need no NCCL at all.
// ---------------------------------------------------------------------------------------
namespace {
struct NcclId { char internal[128]; };
typedef void* NcclComm;
struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(NcclId*) = nullptr;
  int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;
NcclComm g_comm = nullptr;
int g_comm_rank = 0, g_comm_size = 1;
std::mutex g_comm_mu;
constexpr int kNcclFloat64 = 8, kNcclSum = 0;

int nccl_bind() {
  if (g_nccl.handle) return MBB200_OK;
  const char* cands[4] = {getenv("MBB200_NCCL_LIB"), "libnccl.so.2", "libnccl.so", nullptr};
  void* h = nullptr;
  for (int i = 0; i < 3 && !h; i++)
    if (cands[i] && cands[i][0]) h = dlopen(cands[i], RTLD_NOW | RTLD_GLOBAL);
  if (!h) return fail(MBB200_ERR_UNSUPPORTED, "NCCL not found (set MBB200_NCCL_LIB): %s", dlerror());
  g_nccl.GetUniqueId = (int (*)(NcclId*))dlsym(h, "ncclGetUniqueId");
  g_nccl.CommInitRank = (int (*)(NcclComm*, int, NcclId, int))dlsym(h, "ncclCommInitRank");
  g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t))dlsym(h, "ncclAllReduce");
  g_nccl.CommDestroy = (int (*)(NcclComm))dlsym(h, "ncclCommDestroy");
  g_nccl.GetErrorString = (const char* (*)(int))dlsym(h, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy) {
    dlclose(h);
    return fail(MBB200_ERR_UNSUPPORTED, "libnccl lacks a required entry point");
  }
  g_nccl.handle = h;
  return MBB200_OK;
}
const char* nccl_err(int rc) { return g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "nccl error"; }
}  // namespace

static void comm_destroy_locked() {
  std::lock_guard<std::mutex> lk(g_comm_mu);
  if (g_comm) {
    if (!g_dev.empty()) cudaSetDevice(g_dev[0]->ordinal);
    g_nccl.CommDestroy(g_comm);
  }
  g_comm = nullptr;
  g_comm_rank = 0;
  g_comm_size = 1;
}

extern "C" int mbb200_comm_unique_id(void* id_out, size_t len) {
  if (!id_out || len < sizeof(NcclId)) return fail(MBB200_ERR_BAD_ARG, "id buffer must hold %d bytes", (int)sizeof(NcclId));
  std::lock_guard<std::mutex> lk(g_comm_mu);
  int rc = nccl_bind();
  if (rc) return rc;
  NcclId id;
  int nrc = g_nccl.GetUniqueId(&id);
  if (nrc != 0) return fail(MBB200_ERR_CUDA, "ncclGetUniqueId: %s", nccl_err(nrc));
  std::memcpy(id_out, &id, sizeof(id));
  return MBB200_OK;
}

extern "C" int mbb200_comm_init(const void* id, size_t len, int rank, int nranks) {
  RLock rl(g_rw);
  if (g_dev.empty()) return fail(MBB200_ERR_NOT_INIT, "mbb200_init has not been called");
  if (!id || len < sizeof(NcclId) || nranks < 1 || rank < 0 || rank >= nranks) return fail(MBB200_ERR_BAD_ARG, "bad communicator arguments");
  std::lock_guard<std::mutex> lk(g_comm_mu);
  if (g_comm) return fail(MBB200_ERR_BAD_ARG, "communicator already initialised");
  int rc = nccl_bind();
  if (rc) return rc;
  NcclId nid;
  std::memcpy(&nid, id, sizeof(nid));
  CU(cudaSetDevice(g_dev[0]->ordinal));          // one process per GPU: the communicator lives on device 0 of this process
  NcclComm c = nullptr;
  int nrc = g_nccl.CommInitRank(&c, nranks, nid, rank);
  if (nrc != 0) return fail(MBB200_ERR_CUDA, "ncclCommInitRank: %s", nccl_err(nrc));
  g_comm = c;
  g_comm_rank = rank;
  g_comm_size = nranks;
  return MBB200_OK;
}

extern "C" int mbb200_comm_info(int* rank, int* nranks) {
  std::lock_guard<std::mutex> lk(g_comm_mu);
  if (rank) *rank = g_comm_rank;
  if (nranks) *nranks = g_comm_size;
  return g_comm ? 1 : 0;
}

// sum over ranks, in place, device buffer on device 0, ordered after everything already on `st`
static int comm_allreduce_dev(Device& d, double* d_buf, size_t count, cudaStream_t st) {
  std::lock_guard<std::mutex> lk(g_comm_mu);
  if (!g_comm || count == 0) return MBB200_OK;
  int nrc = g_nccl.AllReduce(d_buf, d_buf, count, kNcclFloat64, kNcclSum, g_comm, st);
  if (nrc != 0) return fail(MBB200_ERR_CUDA, "ncclAllReduce: %s", nccl_err(nrc));
  return MBB200_OK;
}

extern "C" int mbb200_comm_allreduce(double* host_buf, int64_t count) {
  RLock rl(g_rw);
  if (g_dev.empty()) return fail(MBB200_ERR_NOT_INIT, "mbb200_init has not been called");
  if (!host_buf || count < 0) return fail(MBB200_ERR_BAD_ARG, "bad buffer");
  {
    std::lock_guard<std::mutex> lk(g_comm_mu);
    if (!g_comm) return MBB200_OK;                // single process: the local sum is the global sum
  }
  Device& d = *g_dev[0];
  std::lock_guard<std::mutex> dl(d.mu);
  CU(cudaSetDevice(d.ordinal));
  int rc = ws_reserve(d, sizeof(double) * (size_t)count);
  if (rc) return rc;
  CU(cudaMemcpyAsync(d.ws.ptr, host_buf, sizeof(double) * count, cudaMemcpyHostToDevice, d.stream));
  rc = comm_allreduce_dev(d, (double*)d.ws.ptr, (size_t)count, d.stream);
  if (rc) return rc;
  CU(cudaMemcpyAsync(host_buf, d.ws.ptr, sizeof(double) * count, cudaMemcpyDeviceToHost, d.stream));
  CU(cudaStreamSynchronize(d.stream));
  return MBB200_OK;
}

extern "C" void mbb200_comm_destroy(void) {
  RLock rl(g_rw);
  comm_destroy_locked();
}

static int check_opts(const mbb200_model* m, const mbb200_opts* o, long long N, int n_save) {
  if (!m || !o) return fail(MBB200_ERR_BAD_ARG, "null model or options");
  if (g_dev.empty()) return fail(MBB200_ERR_NOT_INIT, "mbb200_init has not been called");
  if (o->alg != MBB200_ALG_TSIT5 && o->alg != MBB200_ALG_RODAS5P && o->alg != MBB200_ALG_RODAS4 && o->alg != MBB200_ALG_FBDF) return fail(MBB200_ERR_BAD_ARG, "unknown algorithm id %d", o->alg);
  if (o->alg == MBB200_ALG_TSIT5 && !m->entry->is_ode)
    return fail(MBB200_ERR_UNSUPPORTED, "Tsit5 needs an explicit ODE; model '%s' has a singular mass matrix (use Rodas5P or Rodas4)", m->entry->name);
  if (N < 0 || n_save < 0) return fail(MBB200_ERR_BAD_ARG, "negative size");
  if (!(o->abstol > 0.0) || !(o->reltol > 0.0)) return fail(MBB200_ERR_BAD_ARG, "abstol and reltol must be positive");
  if (!(o->t1 >= o->t0)) return fail(MBB200_ERR_BAD_ARG, "tspan must be increasing");
  if (o->pow_mode != 0 && o->pow_mode != 1) return fail(MBB200_ERR_BAD_ARG, "pow_mode must be 0 (fastpower) or 1 (exact)");
  if (o->qsteady_min > o->qsteady_max) return fail(MBB200_ERR_BAD_ARG, "qsteady_min > qsteady_max");
  return MBB200_OK;
}

// trajectories per completion chunk: a power of two, at least 8192, at most flag_cap chunks per shard
static int chunk_shift_for(const Device& d, long long Ns) {
  int sh = 13;
  while (((Ns + (1ll << sh) - 1) >> sh) > d.flag_cap) sh++;
  return sh;
}

// enqueue one shard on device d (everything already resident); timing events around the kernels.
// `stream_chunks`: the kernel reports finished chunks of 2^chunk_shift trajectories through d.h_flags.
static int enqueue_shard(Device& d, const mbb200_model* m, int dev_index, const mbb200_opts* o, long long Ns,
                         const double* d_u0, const double* d_p, const double* d_saveat, int n_save, double* d_out_u,
                         double* d_tfinal, int* d_rc, int* d_stats, double* d_reduce, bool stream_chunks, int chunk_shift) {
  MbArgs a;
  a.N = Ns; a.u0 = d_u0; a.p = d_p; a.tb = m->d_tables[dev_index]; a.saveat = d_saveat; a.n_save = n_save;
  a.out_u = d_out_u; a.out_tfinal = d_tfinal; a.retcode = d_rc; a.stats = d_stats; a.queue = d.queue;
  a.t0 = o->t0; a.t1 = o->t1; a.abstol = o->abstol; a.reltol = o->reltol; a.dt0 = o->dt0; a.dtmin = o->dtmin;
  a.dtmax = o->dtmax; a.maxiters = o->maxiters;
  a.qs_min = o->qsteady_min; a.qs_max = o->qsteady_max; a.pow_mode = o->pow_mode;
  if (a.qs_min == 0.0 && a.qs_max == 0.0) { a.qs_min = 1.0; a.qs_max = (o->alg == MBB200_ALG_TSIT5) ? 1.0 : 1.2; }
  a.done = nullptr; a.flags = nullptr; a.chunk_shift = 0;
  a.scratch = nullptr;
  if (m->entry->scratch_bytes && Ns > 0) {        // warp-per-trajectory kernels (long ropes): W, stage vectors, sweep scratch
    const size_t sb = m->entry->scratch_bytes(o->alg, Ns, d.sm_count);
    if (sb && d.ws2.bytes < sb) {
      CU(cudaStreamSynchronize(d.stream));
      if (d.ws2.ptr) CU(cudaFree(d.ws2.ptr));
      d.ws2.ptr = nullptr; d.ws2.bytes = 0;
      CU(cudaMalloc(&d.ws2.ptr, sb));
      d.ws2.bytes = sb;
    }
    a.scratch = sb ? (double*)d.ws2.ptr : nullptr;
    if (const char* pe = sb ? getenv("MBB200_L2_PERSIST") : nullptr) {   // tuning knob (off by default): pin the
      int maxp = 0, maxw = 0;                                           // kernel scratch in the L2 set-aside, value = MB (<= 1: all)
      cudaDeviceGetAttribute(&maxp, cudaDevAttrMaxPersistingL2CacheSize, d.ordinal);
      cudaDeviceGetAttribute(&maxw, cudaDevAttrMaxAccessPolicyWindowSize, d.ordinal);
      size_t want = (size_t)atoll(pe) << 20;
      if (want <= (1u << 20) || want > sb) want = sb;
      const size_t setaside = want < (size_t)maxp ? want : (size_t)maxp;
      CU(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, setaside));
      cudaStreamAttrValue v;
      memset(&v, 0, sizeof(v));
      v.accessPolicyWindow.base_ptr = d.ws2.ptr;
      v.accessPolicyWindow.num_bytes = setaside < (size_t)maxw ? setaside : (size_t)maxw;
      v.accessPolicyWindow.hitRatio = 1.0f;
      v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
      v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
      CU(cudaStreamSetAttribute(d.stream, cudaStreamAttributeAccessPolicyWindow, &v));
    }
  }
  d.last_launches = 0;
  CU(cudaMemsetAsync(d.queue, 0, sizeof(unsigned long long), d.stream));
  if (stream_chunks && Ns > 0) {
    const long long nch = (Ns + (1ll << chunk_shift) - 1) >> chunk_shift;
    CU(cudaMemsetAsync(d.d_done, 0, sizeof(unsigned) * (size_t)nch, d.stream));
    for (long long c = 0; c < nch; c++) d.h_flags[c] = 0;
    a.done = d.d_done; a.flags = d.d_flags; a.chunk_shift = chunk_shift;
  }
  CU(cudaEventRecord(d.ev0, d.stream));
  if (Ns > 0) {
    cudaError_t e = m->jit ? jit_launch(m, dev_index, o->alg, a, d.sm_count, d.stream) : m->entry->launch(o->alg, a, d.sm_count, d.stream);
    if (e != cudaSuccess && o->alg == MBB200_ALG_FBDF && (e == cudaErrorInvalidValue || e == cudaErrorNotSupported))
      return fail(MBB200_ERR_UNSUPPORTED, "model '%s' has no FBDF kernel", m->entry->name);
    if (e != cudaSuccess) return fail(MBB200_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
    d.last_launches++;
    if ((o->flags & MBB200_FLAG_REDUCE) && d_reduce && n_save > 0) {
      const int rows = n_save * m->entry->n;
      mb_reduce_kernel<<<rows + 1, 256, 0, d.stream>>>(d_out_u, d_rc, Ns, rows, d_reduce);
      CU(cudaGetLastError());
      d.last_launches++;
    }
  }
  CU(cudaEventRecord(d.ev1, d.stream));
  return MBB200_OK;
}

static int finish_shard(Device& d) {
  CU(cudaSetDevice(d.ordinal));
  CU(cudaStreamSynchronize(d.stream));
  CU(cudaStreamSynchronize(d.cstream));
  float ms = 0.f;
  CU(cudaEventElapsedTime(&ms, d.ev0, d.ev1));
  d.last_ms = ms;
  return MBB200_OK;
}

extern "C" int mbb200_solve_device(mbb200_model* m, const mbb200_opts* o, int dev_index, int64_t N, const double* d_u0,
                                   const double* d_p, const double* saveat, int32_t n_save, double* d_out_u,
                                   double* d_tfinal, int32_t* d_rc, int32_t* d_stats, double* d_reduce) {
  RLock rl(g_rw);
  int rc = check_opts(m, o, N, n_save);
  if (rc) return rc;
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index %d", dev_index);
  if (!d_u0 || !d_tfinal || !d_rc || !d_stats || (n_save > 0 && (!saveat || !d_out_u)) || (m->entry->m > 0 && !d_p))
    return fail(MBB200_ERR_BAD_ARG, "null buffer");
  Device& d = *g_dev[dev_index];
  std::lock_guard<std::mutex> dl(d.mu);
  CU(cudaSetDevice(d.ordinal));
  rc = ws_reserve(d, sizeof(double) * (size_t)(n_save > 0 ? n_save : 1));
  if (rc) return rc;
  if (n_save > 0) CU(cudaMemcpyAsync(d.ws.ptr, saveat, sizeof(double) * n_save, cudaMemcpyHostToDevice, d.stream));
  rc = enqueue_shard(d, m, dev_index, o, N, d_u0, d_p, (const double*)d.ws.ptr, n_save, d_out_u, d_tfinal, d_rc, d_stats,
                     d_reduce, false, 0);
  if (rc == MBB200_OK && (o->flags & MBB200_FLAG_REDUCE) && d_reduce && n_save > 0 && dev_index == 0) {
    if (N == 0) CU(cudaMemsetAsync(d_reduce, 0, sizeof(double) * (2 * (size_t)n_save * m->entry->n + 1), d.stream));
    rc = comm_allreduce_dev(d, d_reduce, 2 * (size_t)n_save * m->entry->n + 1, d.stream);   // no-op without a communicator
  }
  int rc2 = finish_shard(d);     // always drain what was enqueued before returning
  return rc ? rc : rc2;
}

// ---------------------------------------------------------------------------------------
// Allocation-free variant of mbb200_solve_device (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture,
// tight serving loops, harnesses that forbid it).
//
// The ensemble kernel needs two pieces of scratch memory: the trajectory work counter and a
// device copy of the save times.  mbb200_solve_device takes them from cudaMalloc'd memory (the
// counter at init, the save times from the grow-only workspace); here both live in ONE
// file-scope device array of fixed size, which every GPU that loads the library holds until the
// process exits:
//   bytes [0, 8)                       work counter (cleared at the start of every call)
//   bytes [256, 256 + 8 * MAX_SAVE)    save times of the current call
// In host code `mb_static_scratch` is only a host-side stand-in for the symbol, NOT a device
// address; the address on a given device comes from cudaGetSymbolAddress with that device
// current (done once per device, below) and that pointer is what the kernel receives.
// Every call shares the array and it keeps the previous call's contents, so the counter is
// cleared on the stream first; two calls that overlap on one GPU would race on it (inside this
// library they cannot: a call owns its device until its kernels have completed).
// ---------------------------------------------------------------------------------------
#define MB_STATIC_SAVE_OFFSET 256
static __device__ __align__(256) unsigned char mb_static_scratch[MB_STATIC_SAVE_OFFSET + sizeof(double) * MBB200_STATIC_MAX_SAVE];

extern "C" int mbb200_solve_device_static(mbb200_model* m, const mbb200_opts* o, int dev_index, int64_t N,
                                          const double* d_u0, const double* d_p, const double* saveat, int32_t n_save,
                                          double* d_out_u, double* d_tfinal, int32_t* d_rc, int32_t* d_stats,
                                          double* d_reduce) {
  RLock rl(g_rw);
  int rc = check_opts(m, o, N, n_save);
  if (rc) return rc;
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index %d", dev_index);
  if (n_save > MBB200_STATIC_MAX_SAVE)
    return fail(MBB200_ERR_BAD_ARG, "mbb200_solve_device_static: n_save = %d exceeds the static scratch limit MBB200_STATIC_MAX_SAVE = %d "
                                    "(use mbb200_solve_device for longer save grids)", n_save, MBB200_STATIC_MAX_SAVE);
  if (!d_u0 || !d_tfinal || !d_rc || !d_stats || (n_save > 0 && (!saveat || !d_out_u)) || (m->entry->m > 0 && !d_p))
    return fail(MBB200_ERR_BAD_ARG, "null buffer");
  if (m->entry->scratch_bytes)
    return fail(MBB200_ERR_UNSUPPORTED, "model '%s' needs per-warp device scratch that does not fit the static array; use mbb200_solve_device", m->entry->name);
  Device& d = *g_dev[dev_index];
  std::lock_guard<std::mutex> dl(d.mu);
  CU(cudaSetDevice(d.ordinal));
  if (!d.static_scratch) {                       // once per device, with that device current
    void* ptr = nullptr;
    CU(cudaGetSymbolAddress(&ptr, mb_static_scratch));
    if (!ptr) return fail(MBB200_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
    d.static_scratch = (unsigned char*)ptr;
  }
  unsigned long long* s_queue = (unsigned long long*)d.static_scratch;
  double* s_saveat = (double*)(d.static_scratch + MB_STATIC_SAVE_OFFSET);
  MbArgs a;
  a.N = N; a.u0 = d_u0; a.p = d_p; a.tb = m->d_tables[dev_index]; a.saveat = s_saveat; a.n_save = n_save;
  a.out_u = d_out_u; a.out_tfinal = d_tfinal; a.retcode = d_rc; a.stats = d_stats; a.queue = s_queue;
  a.t0 = o->t0; a.t1 = o->t1; a.abstol = o->abstol; a.reltol = o->reltol; a.dt0 = o->dt0; a.dtmin = o->dtmin;
  a.dtmax = o->dtmax; a.maxiters = o->maxiters;
  a.qs_min = o->qsteady_min; a.qs_max = o->qsteady_max; a.pow_mode = o->pow_mode;
  if (a.qs_min == 0.0 && a.qs_max == 0.0) { a.qs_min = 1.0; a.qs_max = (o->alg == MBB200_ALG_TSIT5) ? 1.0 : 1.2; }
  a.done = nullptr; a.flags = nullptr; a.chunk_shift = 0;
  a.scratch = nullptr;
  d.last_launches = 0;
  d.last_ms = 0.0;
  // the array keeps the previous call's contents: the counter must start at zero
  CU(cudaMemsetAsync(s_queue, 0, sizeof(unsigned long long), d.stream));
  if (n_save > 0) CU(cudaMemcpyAsync(s_saveat, saveat, sizeof(double) * n_save, cudaMemcpyHostToDevice, d.stream));
  CU(cudaEventRecord(d.ev0, d.stream));
  if (N > 0) {
    cudaError_t e = m->jit ? jit_launch(m, dev_index, o->alg, a, d.sm_count, d.stream) : m->entry->launch(o->alg, a, d.sm_count, d.stream);
    if (e != cudaSuccess) {
      cudaStreamSynchronize(d.stream);
      if (o->alg == MBB200_ALG_FBDF && (e == cudaErrorInvalidValue || e == cudaErrorNotSupported))
        return fail(MBB200_ERR_UNSUPPORTED, "model '%s' has no FBDF kernel", m->entry->name);
      return fail(MBB200_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
    }
    d.last_launches++;
    if ((o->flags & MBB200_FLAG_REDUCE) && d_reduce && n_save > 0) {   // statistic of THIS device only (no collective here)
      const int rows = n_save * m->entry->n;
      mb_reduce_kernel<<<rows + 1, 256, 0, d.stream>>>(d_out_u, d_rc, N, rows, d_reduce);
      cudaError_t e2 = cudaGetLastError();
      if (e2 != cudaSuccess) {
        cudaStreamSynchronize(d.stream);
        return fail(MBB200_ERR_CUDA, "statistics kernel launch failed: %s", cudaGetErrorString(e2));
      }
      d.last_launches++;
    }
  }
  CU(cudaEventRecord(d.ev1, d.stream));
  CU(cudaStreamSynchronize(d.stream));
  float ms = 0.f;
  CU(cudaEventElapsedTime(&ms, d.ev0, d.ev1));
  d.last_ms = ms;
  return MBB200_OK;
}

// Page-locks a caller buffer for the duration of a call when it is ordinary pageable memory
// (what a Julia `Array` is): copies from pageable memory are staged by the driver and are
// neither asynchronous nor fast.  Failure to register is not an error (the copy still works).
struct HostPin {
  void* p = nullptr;
  std::thread th;
  static bool wanted(const void* ptr, size_t bytes) {
    if (!ptr || bytes < (size_t(1) << 20) || getenv("MBB200_NO_HOST_REGISTER")) return false;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, ptr) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeUnregistered;
  }
  void pin(const void* ptr, size_t bytes) {                 // synchronous (inputs: needed before the first copy)
    if (!wanted(ptr, bytes)) return;
    if (cudaHostRegister(const_cast<void*>(ptr), bytes, cudaHostRegisterPortable) == cudaSuccess) p = const_cast<void*>(ptr);
    else cudaGetLastError();
  }
  // outputs: page-locking (which also faults the pages in) runs on a helper thread while the kernel
  // integrates; wait() is called before the first chunk is copied back
  void pin_async(const void* ptr, size_t bytes, int ordinal) {
    if (!wanted(ptr, bytes)) return;
    th = std::thread([this, ptr, bytes, ordinal]() {
      cudaSetDevice(ordinal);
      if (cudaHostRegister(const_cast<void*>(ptr), bytes, cudaHostRegisterPortable) == cudaSuccess) p = const_cast<void*>(ptr);
      else cudaGetLastError();
    });
  }
  void wait() { if (th.joinable()) th.join(); }
  ~HostPin() { wait(); if (p) cudaHostUnregister(p); }
};

extern "C" int mbb200_solve(mbb200_model* m, const mbb200_opts* o, int64_t N, const double* u0, const double* p,
                            const double* saveat, int32_t n_save, double* out_u, double* out_tfinal, int32_t* retcode,
                            int32_t* stats, double* reduce_out) {
  RLock rl(g_rw);
  int rc = check_opts(m, o, N, n_save);
  if (rc) return rc;
  const int n = m->entry->n, mp = m->entry->m;
  if (!u0 || !out_tfinal || !retcode || !stats || (n_save > 0 && (!saveat || !out_u)) || (mp > 0 && !p))
    return fail(MBB200_ERR_BAD_ARG, "null buffer");
  const int G = (int)g_dev.size();
  const bool want_red = (o->flags & MBB200_FLAG_REDUCE) && reduce_out && n_save > 0;
  const int rows = n_save * n;
  std::vector<std::unique_lock<std::mutex>> own;          // devices are owned in index order: no lock cycles
  for (int g = 0; g < G; g++) own.emplace_back(g_dev[g]->mu);
  HostPin pins[6];
  pins[0].pin(u0, sizeof(double) * (size_t)N * n);
  pins[1].pin(p, sizeof(double) * (size_t)N * mp);
  const int ord0 = g_dev[0]->ordinal;
  struct Shard { long long off = 0, Ns = 0, nch = 0, next = 0; int shift = 0; bool started = false;
                 double *u0 = 0, *p = 0, *sv = 0, *out = 0, *tf = 0, *red = 0; int *rc = 0, *st = 0; };
  std::vector<Shard> sh(G);
  // on any failure: drain every stream that has work of this call in flight, then report
  auto drain = [&](int code) {
    for (int g = 0; g < G; g++)
      if (sh[g].started) {
        cudaSetDevice(g_dev[g]->ordinal);
        cudaStreamSynchronize(g_dev[g]->stream);
        cudaStreamSynchronize(g_dev[g]->cstream);
      }
    return code;
  };
#define CUD(call)                                                                                         \
  do {                                                                                                    \
    cudaError_t e_ = (call);                                                                              \
    if (e_ != cudaSuccess)                                                                                \
      return drain(fail(MBB200_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__)); \
  } while (0)
  for (int g = 0; g < G; g++) {
    Device& d = *g_dev[g];
    Shard& s = sh[g];
    s.off = N * g / G;
    s.Ns = N * (g + 1) / G - s.off;
    d.last_ms = 0.0;
    d.last_launches = 0;
    if (s.Ns == 0) continue;                               // nothing to do on this device (N < number of devices)
    CUD(cudaSetDevice(d.ordinal));
    const size_t Np = ((size_t)s.Ns + 31) & ~(size_t)31;    // keep every sub-array 256-byte aligned
    size_t nd = Np * (size_t)(n + mp + rows + 1) + (size_t)((n_save + 31) & ~31) + (size_t)(2 * rows + 1 + 31);
    size_t bytes = nd * sizeof(double) + Np * 6 * sizeof(int) + 1024;
    rc = ws_reserve(d, bytes);
    if (rc) return drain(rc);
    double* w = (double*)d.ws.ptr;
    s.u0 = w; w += Np * n;
    s.p = w; w += Np * mp;
    s.out = w; w += Np * rows;
    s.tf = w; w += Np;
    s.sv = w; w += (n_save + 31) & ~31;
    s.red = w; w += (2 * rows + 1 + 31) & ~31;
    s.rc = (int*)w;
    s.st = s.rc + Np;
    s.shift = chunk_shift_for(d, s.Ns);
    s.nch = (s.Ns + (1ll << s.shift) - 1) >> s.shift;
    s.started = true;
    const size_t wb = (size_t)s.Ns * sizeof(double);
    CUD(cudaMemcpy2DAsync(s.u0, wb, u0 + s.off, (size_t)N * sizeof(double), wb, n, cudaMemcpyHostToDevice, d.stream));
    if (mp > 0) CUD(cudaMemcpy2DAsync(s.p, wb, p + s.off, (size_t)N * sizeof(double), wb, mp, cudaMemcpyHostToDevice, d.stream));
    if (n_save > 0) CUD(cudaMemcpyAsync(s.sv, saveat, sizeof(double) * n_save, cudaMemcpyHostToDevice, d.stream));
    rc = enqueue_shard(d, m, g, o, s.Ns, s.u0, s.p, s.sv, n_save, s.out, s.tf, s.rc, s.st, want_red ? s.red : nullptr, true, s.shift);
    if (rc) return drain(rc);
  }
  // Streamed read-back: a chunk of trajectories is copied to the caller's buffers as soon as the
  // kernel has retired all of it (per-chunk counters, flag in mapped host memory), on the copy
  // stream, while the kernel keeps integrating the rest.  When the kernel has finished, whatever
  // is left is copied at once.
  pins[2].pin_async(out_u, sizeof(double) * (size_t)N * rows, ord0);      // overlaps the kernels just enqueued
  pins[3].pin(out_tfinal, sizeof(double) * (size_t)N);
  pins[4].pin(retcode, sizeof(int) * (size_t)N);
  pins[5].pin(stats, sizeof(int) * (size_t)N * 5);
  pins[2].wait();
  int pending = 0;
  for (int g = 0; g < G; g++) pending += sh[g].started ? 1 : 0;
  while (pending > 0) {
    bool progressed = false;
    for (int g = 0; g < G; g++) {
      Shard& s = sh[g];
      if (!s.started || s.next >= s.nch) continue;
      Device& d = *g_dev[g];
      long long c1 = s.next;
      while (c1 < s.nch && *(volatile int*)&d.h_flags[c1]) c1++;
      if (c1 == s.next) {
        cudaError_t q = cudaEventQuery(d.ev1);
        if (q == cudaSuccess) c1 = s.nch;                  // kernel done: everything is ready
        else if (q != cudaErrorNotReady) return drain(fail(MBB200_ERR_CUDA, "ensemble kernel failed: %s", cudaGetErrorString(q)));
        else continue;
      }
      CUD(cudaSetDevice(d.ordinal));
      const long long a0 = s.next << s.shift;
      long long a1 = c1 << s.shift;
      if (a1 > s.Ns) a1 = s.Ns;
      const size_t cw = (size_t)(a1 - a0) * sizeof(double), ci = (size_t)(a1 - a0) * sizeof(int);
      const size_t wb = (size_t)s.Ns * sizeof(double), wi = (size_t)s.Ns * sizeof(int);
      if (n_save > 0)
        CUD(cudaMemcpy2DAsync(out_u + s.off + a0, (size_t)N * sizeof(double), s.out + a0, wb, cw, rows, cudaMemcpyDeviceToHost, d.cstream));
      CUD(cudaMemcpyAsync(out_tfinal + s.off + a0, s.tf + a0, cw, cudaMemcpyDeviceToHost, d.cstream));
      CUD(cudaMemcpyAsync(retcode + s.off + a0, s.rc + a0, ci, cudaMemcpyDeviceToHost, d.cstream));
      CUD(cudaMemcpy2DAsync(stats + s.off + a0, (size_t)N * sizeof(int), s.st + a0, wi, ci, 5, cudaMemcpyDeviceToHost, d.cstream));
      s.next = c1;
      progressed = true;
      if (s.next >= s.nch) pending--;
    }
    if (!progressed) std::this_thread::sleep_for(std::chrono::microseconds(50));
  }
#undef CUD
  if (want_red) std::memset(reduce_out, 0, sizeof(double) * (2 * rows + 1));
  std::vector<double> part(want_red ? 2 * rows + 1 : 0);
  int first_err = MBB200_OK;
  for (int g = 0; g < G; g++) {
    if (!sh[g].started) continue;
    rc = finish_shard(*g_dev[g]);
    if (rc) { if (!first_err) first_err = rc; continue; }
    if (want_red && !first_err) {
      cudaError_t ce = cudaMemcpy(part.data(), sh[g].red, sizeof(double) * part.size(), cudaMemcpyDeviceToHost);
      if (ce != cudaSuccess) { first_err = fail(MBB200_ERR_CUDA, "statistics read-back failed: %s", cudaGetErrorString(ce)); continue; }
      for (size_t i = 0; i < part.size(); i++) reduce_out[i] += part[i];
    }
  }
  if (first_err) return first_err;
  if (want_red) {
    bool have_comm;
    { std::lock_guard<std::mutex> lk(g_comm_mu); have_comm = g_comm != nullptr; }
    if (have_comm) {                                       // sum over the processes of the job (NCCL)
      Device& d = *g_dev[0];
      CU(cudaSetDevice(d.ordinal));
      rc = ws_reserve(d, sizeof(double) * part.size());
      if (rc) return rc;
      CU(cudaMemcpyAsync(d.ws.ptr, reduce_out, sizeof(double) * part.size(), cudaMemcpyHostToDevice, d.stream));
      rc = comm_allreduce_dev(d, (double*)d.ws.ptr, part.size(), d.stream);
      if (rc) { cudaStreamSynchronize(d.stream); return rc; }
      CU(cudaMemcpyAsync(reduce_out, d.ws.ptr, sizeof(double) * part.size(), cudaMemcpyDeviceToHost, d.stream));
      CU(cudaStreamSynchronize(d.stream));
    }
  }
  return MBB200_OK;
}

// Observed (eliminated) variables of saved states, host buffers: what `sol(t, idxs = expr)` does
// after the reference's solve (test/test_robot.jl:321-325).
extern "C" int mbb200_model_observed_count(const mbb200_model* m, int32_t* nobs) {
  if (!m || !nobs) return fail(MBB200_ERR_BAD_ARG, "null argument");
  *nobs = m->entry->nobs;
  return MBB200_OK;
}

extern "C" int mbb200_observed(mbb200_model* m, int64_t N, const double* u, const double* p, const double* saveat,
                               int32_t n_save, double* out_obs) {
  RLock rl(g_rw);
  if (!m) return fail(MBB200_ERR_BAD_ARG, "null model");
  if (g_dev.empty()) return fail(MBB200_ERR_NOT_INIT, "mbb200_init has not been called");
  const int n = m->entry->n, mp = m->entry->m, nobs = m->entry->nobs;
  if (nobs == 0) return fail(MBB200_ERR_UNSUPPORTED, "model '%s' has no observed variables", m->entry->name);
  if (N < 0 || n_save < 0) return fail(MBB200_ERR_BAD_ARG, "negative size");
  if (N == 0 || n_save == 0) return MBB200_OK;
  if (!u || !saveat || !out_obs || (mp > 0 && !p)) return fail(MBB200_ERR_BAD_ARG, "null buffer");
  Device& d = *g_dev[0];
  std::lock_guard<std::mutex> dl(d.mu);
  CU(cudaSetDevice(d.ordinal));
  // processed in slabs of save points so the workspace stays bounded
  const size_t per_save = (size_t)N * (n + nobs) * sizeof(double);
  int slab = (int)((size_t(1) << 30) / (per_save ? per_save : 1));
  if (slab < 1) slab = 1;
  if (slab > n_save) slab = n_save;
  const size_t bytes = (size_t)slab * per_save + (size_t)N * mp * sizeof(double) + (size_t)(n_save + 32) * sizeof(double);
  int rc = ws_reserve(d, bytes);
  if (rc) return rc;
  double* d_u = (double*)d.ws.ptr;
  double* d_o = d_u + (size_t)slab * N * n;
  double* d_p = d_o + (size_t)slab * N * nobs;
  double* d_sv = d_p + (size_t)N * mp;
  if (mp > 0) CU(cudaMemcpyAsync(d_p, p, (size_t)N * mp * sizeof(double), cudaMemcpyHostToDevice, d.stream));
  CU(cudaMemcpyAsync(d_sv, saveat, (size_t)n_save * sizeof(double), cudaMemcpyHostToDevice, d.stream));
  d.last_launches = 0;
  for (int s0 = 0; s0 < n_save; s0 += slab) {
    const int ns = (n_save - s0 < slab) ? n_save - s0 : slab;
    CU(cudaMemcpyAsync(d_u, u + (size_t)s0 * n * N, (size_t)ns * n * N * sizeof(double), cudaMemcpyHostToDevice, d.stream));
    MbObsArgs a;
    a.N = N; a.n_save = ns; a.u = d_u; a.p = d_p; a.tb = m->d_tables[0]; a.saveat = d_sv + s0; a.out = d_o;
    cudaError_t e = m->jit ? jit_launch_observed(m, 0, a, d.sm_count, d.stream) : m->entry->launch_observed(a, d.sm_count, d.stream);
    if (e != cudaSuccess) return fail(MBB200_ERR_CUDA, "observed kernel launch failed: %s", cudaGetErrorString(e));
    d.last_launches++;
    CU(cudaMemcpyAsync(out_obs + (size_t)s0 * nobs * N, d_o, (size_t)ns * nobs * N * sizeof(double), cudaMemcpyDeviceToHost, d.stream));
  }
  CU(cudaStreamSynchronize(d.stream));
  return MBB200_OK;
}

extern "C" int mbb200_last_timing(int dev_index, double* kernel_ms, int32_t* n_launches) {
  RLock rl(g_rw);
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index %d", dev_index);
  if (kernel_ms) *kernel_ms = g_dev[dev_index]->last_ms;
  if (n_launches) *n_launches = g_dev[dev_index]->last_launches;
  return MBB200_OK;
}

// ---------------------------------------------------------------------------------------
// FP64 roofline denominator: dependent-free DFMA chains (MEASURED_PEAKS.json holds HBM and
// BF16 only; BASELINE.md section 4 asks for the FP64 FMA rate measured on the box)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) mb_dfma_probe(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

extern "C" int mbb200_measure_fp64_peak(int dev_index, double* flops_per_s) {
  RLock rl(g_rw);
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index %d", dev_index);
  if (!flops_per_s) return fail(MBB200_ERR_BAD_ARG, "null output");
  Device& d = *g_dev[dev_index];
  std::lock_guard<std::mutex> dl(d.mu);
  CU(cudaSetDevice(d.ordinal));
  const int blocks = d.sm_count * 8, threads = 256, iters = 1 << 15;
  int rc = ws_reserve(d, sizeof(double) * (size_t)blocks * threads);
  if (rc) return rc;
  double best = 0.0;
  for (int rep = 0; rep < 5; rep++) {
    CU(cudaEventRecord(d.ev0, d.stream));
    mb_dfma_probe<<<blocks, threads, 0, d.stream>>>((double*)d.ws.ptr, iters, 0.999999, 1e-9);
    CU(cudaGetLastError());
    CU(cudaEventRecord(d.ev1, d.stream));
    CU(cudaStreamSynchronize(d.stream));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, d.ev0, d.ev1));
    const double fl = 2.0 * 8.0 * (double)iters * blocks * threads / (ms * 1e-3);
    if (rep > 0 && fl > best) best = fl;   // first repetition is warm-up
  }
  *flops_per_s = best;
  return MBB200_OK;
}

// ---------------------------------------------------------------------------------------
// memory helpers
// ---------------------------------------------------------------------------------------
extern "C" int mbb200_dev_alloc(int dev_index, size_t bytes, void** out) {
  RLock rl(g_rw);
  if (dev_index < 0 || dev_index >= (int)g_dev.size() || !out) return fail(MBB200_ERR_BAD_ARG, "bad argument");
  CU(cudaSetDevice(g_dev[dev_index]->ordinal));
  CU(cudaMalloc(out, bytes ? bytes : 1));
  return MBB200_OK;
}
extern "C" int mbb200_dev_free(int dev_index, void* ptr) {
  RLock rl(g_rw);
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index");
  CU(cudaSetDevice(g_dev[dev_index]->ordinal));
  CU(cudaFree(ptr));
  return MBB200_OK;
}
extern "C" int mbb200_dev_upload(int dev_index, void* dst, const void* src, size_t bytes) {
  RLock rl(g_rw);
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index");
  CU(cudaSetDevice(g_dev[dev_index]->ordinal));
  CU(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
  return MBB200_OK;
}
extern "C" int mbb200_dev_download(int dev_index, void* dst, const void* src, size_t bytes) {
  RLock rl(g_rw);
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index");
  CU(cudaSetDevice(g_dev[dev_index]->ordinal));
  CU(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
  return MBB200_OK;
}
extern "C" int mbb200_host_alloc_pinned(size_t bytes, void** out) {
  RLock rl(g_rw);
  if (!out) return fail(MBB200_ERR_BAD_ARG, "null output");
  if (g_dev.empty()) return fail(MBB200_ERR_NOT_INIT, "mbb200_init has not been called");
  CU(cudaMallocHost(out, bytes ? bytes : 1));
  return MBB200_OK;
}
extern "C" int mbb200_host_free_pinned(void* ptr) {
  CU(cudaFreeHost(ptr));
  return MBB200_OK;
}
