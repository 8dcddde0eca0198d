// libmultibody_b200 runtime: C ABI (include/mbb200.h), device management, sharding,
// host<->device staging, ensemble-statistic reduction, FP64 peak probe.
// No CPU fallback: every compute entry fails with MBB200_ERR_NO_DEVICE without a GPU.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
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
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  unsigned long long* queue = nullptr;
  Workspace ws;
  double last_ms = 0.0;
  int last_launches = 0;
};
static std::vector<Device> g_dev;
static std::mutex g_mu;

struct mbb200_model {
  const MbKernelEntry* entry;
  std::vector<double> tables;
  std::vector<double*> d_tables;   // per device
};

extern "C" int mbb200_abi_version(void) { return MBB200_ABI_VERSION; }
extern "C" const char* mbb200_last_error(void) { return g_err; }
extern "C" int mbb200_device_count(void) { return (int)g_dev.size(); }

extern "C" int mbb200_init_devices(const int* ordinals, int n) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_dev.empty()) {
    bool same = (int)g_dev.size() == n;
    for (int i = 0; same && i < n; i++) same = g_dev[i].ordinal == ordinals[i];
    if (same) return MBB200_OK;
    return fail(MBB200_ERR_BAD_ARG, "already initialised with a different device set; call mbb200_shutdown first");
  }
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return fail(MBB200_ERR_NO_DEVICE, "no CUDA device: %s", e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  if (n < 1) return fail(MBB200_ERR_BAD_ARG, "n_gpus must be >= 1");
  std::vector<Device> devs(n);
  for (int i = 0; i < n; i++) {
    if (ordinals[i] < 0 || ordinals[i] >= count)
      return fail(MBB200_ERR_NO_DEVICE, "device ordinal %d not present (%d visible)", ordinals[i], count);
    Device& d = devs[i];
    d.ordinal = ordinals[i];
    CU(cudaSetDevice(d.ordinal));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, d.ordinal));
    if (prop.major < 10)
      return fail(MBB200_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", d.ordinal, prop.major, prop.minor);
    d.sm_count = prop.multiProcessorCount;
    CU(cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&d.ev0));
    CU(cudaEventCreate(&d.ev1));
    CU(cudaMalloc(&d.queue, sizeof(unsigned long long)));
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
  std::lock_guard<std::mutex> lk(g_mu);
  for (Device& d : g_dev) {
    cudaSetDevice(d.ordinal);
    cudaStreamSynchronize(d.stream);
    if (d.ws.ptr) cudaFree(d.ws.ptr);
    if (d.queue) cudaFree(d.queue);
    cudaEventDestroy(d.ev0);
    cudaEventDestroy(d.ev1);
    cudaStreamDestroy(d.stream);
  }
  g_dev.clear();
}

static int ws_reserve(Device& d, size_t bytes) {
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
extern "C" int mbb200_model_load(const void* image, size_t len, mbb200_model** out) {
  if (!image || !out || len < sizeof(mbb200_image_header)) return fail(MBB200_ERR_BAD_ARG, "bad model image");
  if (g_dev.empty()) return fail(MBB200_ERR_NOT_INIT, "mbb200_init has not been called");
  mbb200_image_header h;
  std::memcpy(&h, image, sizeof(h));
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
    // image tables: [t0, h, nk] + per table nk (value, z) pairs -> device layout of
    // mb_common.cuh: 8-double header + (nk + 1) interval rows x ntab x 4 polynomial coefficients
    std::vector<double> img(h.n_table);
    std::memcpy(img.data(), (const char*)image + sizeof(h), (size_t)h.n_table * sizeof(double));
    const double t0 = img[0], hh = img[1];
    const long long nk = (long long)img[2];
    const long long ntab = nk > 0 ? ((long long)h.n_table - 3) / (2 * nk) : 0;
    if (nk < 0 || (nk > 0 && (nk < 2 || !(hh > 0.0) || 3 + 2 * nk * ntab != (long long)h.n_table))) {
      delete m;
      return fail(MBB200_ERR_BAD_ARG, "inconsistent table block (nk=%lld, n_table=%lld)", nk, (long long)h.n_table);
    }
    m->tables.assign(MB_TB_HEADER + (size_t)(nk + 1) * ntab * 4, 0.0);
    m->tables[0] = t0; m->tables[1] = hh > 0.0 ? hh : 1.0; m->tables[2] = (double)nk; m->tables[3] = (double)ntab;
    for (long long k = 0; k < ntab; k++) {
      const double* T = img.data() + 3 + (size_t)k * 2 * nk;
      for (long long row = 0; row <= nk; row++) {
        double* c = m->tables.data() + MB_TB_HEADER + ((size_t)row * ntab + k) * 4;
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
  }
  m->d_tables.assign(g_dev.size(), nullptr);
  for (size_t i = 0; i < g_dev.size(); i++) {
    cudaError_t ce = cudaSetDevice(g_dev[i].ordinal);
    if (ce == cudaSuccess) ce = cudaMalloc(&m->d_tables[i], m->tables.size() * sizeof(double));
    if (ce == cudaSuccess) ce = cudaMemcpy(m->d_tables[i], m->tables.data(), m->tables.size() * sizeof(double), cudaMemcpyHostToDevice);
    if (ce != cudaSuccess) {
      int rc = fail(MBB200_ERR_CUDA, "table upload failed: %s", cudaGetErrorString(ce));
      mbb200_model_free(m);
      return rc;
    }
  }
  *out = m;
  return MBB200_OK;
}

extern "C" void mbb200_model_free(mbb200_model* m) {
  if (!m) return;
  for (size_t i = 0; i < m->d_tables.size() && i < g_dev.size(); i++)
    if (m->d_tables[i]) {
      cudaSetDevice(g_dev[i].ordinal);
      cudaFree(m->d_tables[i]);
    }
  delete m;
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

static int check_opts(const mbb200_model* m, const mbb200_opts* o, long long N, int n_save) {
  if (!m || !o) return fail(MBB200_ERR_BAD_ARG, "null model or options");
  if (g_dev.empty()) return fail(MBB200_ERR_NOT_INIT, "mbb200_init has not been called");
  if (o->alg != MBB200_ALG_TSIT5 && o->alg != MBB200_ALG_RODAS5P && o->alg != MBB200_ALG_RODAS4) return fail(MBB200_ERR_BAD_ARG, "unknown algorithm id %d", o->alg);
  if (o->alg == MBB200_ALG_TSIT5 && !m->entry->is_ode)
    return fail(MBB200_ERR_UNSUPPORTED, "Tsit5 needs an explicit ODE; model '%s' has a singular mass matrix (use Rodas5P or Rodas4)", m->entry->name);
  if (N < 0 || n_save < 0) return fail(MBB200_ERR_BAD_ARG, "negative size");
  if (!(o->abstol > 0.0) || !(o->reltol > 0.0)) return fail(MBB200_ERR_BAD_ARG, "abstol and reltol must be positive");
  if (!(o->t1 >= o->t0)) return fail(MBB200_ERR_BAD_ARG, "tspan must be increasing");
  if (o->pow_mode != 0 && o->pow_mode != 1) return fail(MBB200_ERR_BAD_ARG, "pow_mode must be 0 (fastpower) or 1 (exact)");
  if (o->qsteady_min > o->qsteady_max) return fail(MBB200_ERR_BAD_ARG, "qsteady_min > qsteady_max");
  return MBB200_OK;
}

// enqueue one shard on device d (everything already resident); timing events around the kernels
static int enqueue_shard(Device& d, const mbb200_model* m, int dev_index, const mbb200_opts* o, long long Ns,
                         const double* d_u0, const double* d_p, const double* d_saveat, int n_save, double* d_out_u,
                         double* d_tfinal, int* d_rc, int* d_stats, double* d_reduce) {
  MbArgs a;
  a.N = Ns; a.u0 = d_u0; a.p = d_p; a.tb = m->d_tables[dev_index]; a.saveat = d_saveat; a.n_save = n_save;
  a.out_u = d_out_u; a.out_tfinal = d_tfinal; a.retcode = d_rc; a.stats = d_stats; a.queue = d.queue;
  a.t0 = o->t0; a.t1 = o->t1; a.abstol = o->abstol; a.reltol = o->reltol; a.dt0 = o->dt0; a.dtmin = o->dtmin;
  a.dtmax = o->dtmax; a.maxiters = o->maxiters;
  a.qs_min = o->qsteady_min; a.qs_max = o->qsteady_max; a.pow_mode = o->pow_mode;
  if (a.qs_min == 0.0 && a.qs_max == 0.0) { a.qs_min = 1.0; a.qs_max = (o->alg == MBB200_ALG_TSIT5) ? 1.0 : 1.2; }
  d.last_launches = 0;
  CU(cudaMemsetAsync(d.queue, 0, sizeof(unsigned long long), d.stream));
  CU(cudaEventRecord(d.ev0, d.stream));
  if (Ns > 0) {
    cudaError_t e = m->entry->launch(o->alg, a, d.sm_count, d.stream);
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
  float ms = 0.f;
  CU(cudaEventElapsedTime(&ms, d.ev0, d.ev1));
  d.last_ms = ms;
  return MBB200_OK;
}

extern "C" int mbb200_solve_device(mbb200_model* m, const mbb200_opts* o, int dev_index, int64_t N, const double* d_u0,
                                   const double* d_p, const double* saveat, int32_t n_save, double* d_out_u,
                                   double* d_tfinal, int32_t* d_rc, int32_t* d_stats, double* d_reduce) {
  int rc = check_opts(m, o, N, n_save);
  if (rc) return rc;
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index %d", dev_index);
  if (!d_u0 || !d_tfinal || !d_rc || !d_stats || (n_save > 0 && (!saveat || !d_out_u)) || (m->entry->m > 0 && !d_p))
    return fail(MBB200_ERR_BAD_ARG, "null buffer");
  Device& d = g_dev[dev_index];
  CU(cudaSetDevice(d.ordinal));
  rc = ws_reserve(d, sizeof(double) * (size_t)(n_save > 0 ? n_save : 1));
  if (rc) return rc;
  if (n_save > 0) CU(cudaMemcpyAsync(d.ws.ptr, saveat, sizeof(double) * n_save, cudaMemcpyHostToDevice, d.stream));
  rc = enqueue_shard(d, m, dev_index, o, N, d_u0, d_p, (const double*)d.ws.ptr, n_save, d_out_u, d_tfinal, d_rc, d_stats, d_reduce);
  if (rc) return rc;
  return finish_shard(d);
}

extern "C" int mbb200_solve(mbb200_model* m, const mbb200_opts* o, int64_t N, const double* u0, const double* p,
                            const double* saveat, int32_t n_save, double* out_u, double* out_tfinal, int32_t* retcode,
                            int32_t* stats, double* reduce_out) {
  int rc = check_opts(m, o, N, n_save);
  if (rc) return rc;
  const int n = m->entry->n, mp = m->entry->m;
  if (!u0 || !out_tfinal || !retcode || !stats || (n_save > 0 && (!saveat || !out_u)) || (mp > 0 && !p))
    return fail(MBB200_ERR_BAD_ARG, "null buffer");
  const int G = (int)g_dev.size();
  const bool want_red = (o->flags & MBB200_FLAG_REDUCE) && reduce_out && n_save > 0;
  const int rows = n_save * n;
  struct Shard { long long off, Ns; double *u0, *p, *sv, *out, *tf, *red; int *rc, *st; };
  std::vector<Shard> sh(G);
  for (int g = 0; g < G; g++) {
    Device& d = g_dev[g];
    Shard& s = sh[g];
    s.off = N * g / G;
    s.Ns = N * (g + 1) / G - s.off;
    CU(cudaSetDevice(d.ordinal));
    const size_t Np = ((size_t)s.Ns + 31) & ~(size_t)31;    // keep every sub-array 256-byte aligned
    size_t nd = Np * (size_t)(n + mp + rows + 1) + (size_t)((n_save + 31) & ~31) + (size_t)(2 * rows + 1 + 31);
    size_t bytes = nd * sizeof(double) + Np * 6 * sizeof(int) + 1024;
    rc = ws_reserve(d, bytes);
    if (rc) return rc;
    double* w = (double*)d.ws.ptr;
    s.u0 = w; w += Np * n;
    s.p = w; w += Np * mp;
    s.out = w; w += Np * rows;
    s.tf = w; w += Np;
    s.sv = w; w += (n_save + 31) & ~31;
    s.red = w; w += (2 * rows + 1 + 31) & ~31;
    s.rc = (int*)w;
    s.st = s.rc + Np;
    if (s.Ns == 0) continue;
    const size_t wb = (size_t)s.Ns * sizeof(double);
    CU(cudaMemcpy2DAsync(s.u0, wb, u0 + s.off, (size_t)N * sizeof(double), wb, n, cudaMemcpyHostToDevice, d.stream));
    if (mp > 0) CU(cudaMemcpy2DAsync(s.p, wb, p + s.off, (size_t)N * sizeof(double), wb, mp, cudaMemcpyHostToDevice, d.stream));
    if (n_save > 0) CU(cudaMemcpyAsync(s.sv, saveat, sizeof(double) * n_save, cudaMemcpyHostToDevice, d.stream));
    rc = enqueue_shard(d, m, g, o, s.Ns, s.u0, s.p, s.sv, n_save, s.out, s.tf, s.rc, s.st, want_red ? s.red : nullptr);
    if (rc) return rc;
    if (n_save > 0)
      CU(cudaMemcpy2DAsync(out_u + s.off, (size_t)N * sizeof(double), s.out, wb, wb, rows, cudaMemcpyDeviceToHost, d.stream));
    CU(cudaMemcpyAsync(out_tfinal + s.off, s.tf, wb, cudaMemcpyDeviceToHost, d.stream));
    CU(cudaMemcpyAsync(retcode + s.off, s.rc, (size_t)s.Ns * sizeof(int), cudaMemcpyDeviceToHost, d.stream));
    CU(cudaMemcpy2DAsync(stats + s.off, (size_t)N * sizeof(int), s.st, (size_t)s.Ns * sizeof(int), (size_t)s.Ns * sizeof(int), 5,
                         cudaMemcpyDeviceToHost, d.stream));
  }
  if (want_red) std::memset(reduce_out, 0, sizeof(double) * (2 * rows + 1));
  std::vector<double> part(want_red ? 2 * rows + 1 : 0);
  for (int g = 0; g < G; g++) {
    rc = finish_shard(g_dev[g]);
    if (rc) return rc;
    if (want_red && sh[g].Ns > 0) {
      CU(cudaMemcpy(part.data(), sh[g].red, sizeof(double) * part.size(), cudaMemcpyDeviceToHost));
      for (size_t i = 0; i < part.size(); i++) reduce_out[i] += part[i];
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
  if (!m) return fail(MBB200_ERR_BAD_ARG, "null model");
  if (g_dev.empty()) return fail(MBB200_ERR_NOT_INIT, "mbb200_init has not been called");
  const int n = m->entry->n, mp = m->entry->m, nobs = m->entry->nobs;
  if (nobs == 0) return fail(MBB200_ERR_UNSUPPORTED, "model '%s' has no observed variables", m->entry->name);
  if (N < 0 || n_save < 0) return fail(MBB200_ERR_BAD_ARG, "negative size");
  if (N == 0 || n_save == 0) return MBB200_OK;
  if (!u || !saveat || !out_obs || (mp > 0 && !p)) return fail(MBB200_ERR_BAD_ARG, "null buffer");
  Device& d = g_dev[0];
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
    cudaError_t e = m->entry->launch_observed(a, d.sm_count, d.stream);
    if (e != cudaSuccess) return fail(MBB200_ERR_CUDA, "observed kernel launch failed: %s", cudaGetErrorString(e));
    d.last_launches++;
    CU(cudaMemcpyAsync(out_obs + (size_t)s0 * nobs * N, d_o, (size_t)ns * nobs * N * sizeof(double), cudaMemcpyDeviceToHost, d.stream));
  }
  CU(cudaStreamSynchronize(d.stream));
  return MBB200_OK;
}

extern "C" int mbb200_last_timing(int dev_index, double* kernel_ms, int32_t* n_launches) {
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index %d", dev_index);
  if (kernel_ms) *kernel_ms = g_dev[dev_index].last_ms;
  if (n_launches) *n_launches = g_dev[dev_index].last_launches;
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
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index %d", dev_index);
  if (!flops_per_s) return fail(MBB200_ERR_BAD_ARG, "null output");
  Device& d = g_dev[dev_index];
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
  if (dev_index < 0 || dev_index >= (int)g_dev.size() || !out) return fail(MBB200_ERR_BAD_ARG, "bad argument");
  CU(cudaSetDevice(g_dev[dev_index].ordinal));
  CU(cudaMalloc(out, bytes ? bytes : 1));
  return MBB200_OK;
}
extern "C" int mbb200_dev_free(int dev_index, void* ptr) {
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index");
  CU(cudaSetDevice(g_dev[dev_index].ordinal));
  CU(cudaFree(ptr));
  return MBB200_OK;
}
extern "C" int mbb200_dev_upload(int dev_index, void* dst, const void* src, size_t bytes) {
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index");
  CU(cudaSetDevice(g_dev[dev_index].ordinal));
  CU(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
  return MBB200_OK;
}
extern "C" int mbb200_dev_download(int dev_index, void* dst, const void* src, size_t bytes) {
  if (dev_index < 0 || dev_index >= (int)g_dev.size()) return fail(MBB200_ERR_BAD_ARG, "bad device index");
  CU(cudaSetDevice(g_dev[dev_index].ordinal));
  CU(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
  return MBB200_OK;
}
extern "C" int mbb200_host_alloc_pinned(size_t bytes, void** out) {
  if (!out) return fail(MBB200_ERR_BAD_ARG, "null output");
  if (g_dev.empty()) return fail(MBB200_ERR_NOT_INIT, "mbb200_init has not been called");
  CU(cudaMallocHost(out, bytes ? bytes : 1));
  return MBB200_OK;
}
extern "C" int mbb200_host_free_pinned(void* ptr) {
  CU(cudaFreeHost(ptr));
  return MBB200_OK;
}
