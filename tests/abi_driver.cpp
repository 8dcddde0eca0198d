// C++ driver of the C ABI ("a C++ test driver and a Python ctypes driver exercise the same ABI").
// Loads libmultibody_b200.so with dlopen, loads a model IMAGE from a file (MBIR1: names a compiled kernel set; MBIR2: carries
// the model and is compiled at load time), solves an ensemble described by a small binary job file and writes the results.
//
//   abi_driver <libmultibody_b200.so> <image file> <job file> <result file>
//
// job file (little endian): int32 alg, n, m, n_save; int64 N; double t0, t1, abstol, reltol; double u0[n][N]; double p[m][N];
//                           double saveat[n_save]
// result file:              double out_u[n_save][n][N]; double t_final[N]; int32 retcode[N]; int32 stats[5][N]
#include <dlfcn.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iterator>
#include <vector>
#include "../include/mbb200.h"

#define BIND(name) auto p_##name = reinterpret_cast<decltype(&name)>(dlsym(h, #name)); if (!p_##name) { std::fprintf(stderr, "missing %s\n", #name); return 2; }

int main(int argc, char** argv) {
  if (argc != 5) { std::fprintf(stderr, "usage: abi_driver lib image job result\n"); return 2; }
  void* h = dlopen(argv[1], RTLD_NOW);
  if (!h) { std::fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
  BIND(mbb200_abi_version) BIND(mbb200_init) BIND(mbb200_model_load) BIND(mbb200_model_info) BIND(mbb200_solve)
  BIND(mbb200_last_error) BIND(mbb200_model_free) BIND(mbb200_shutdown)
  if (p_mbb200_abi_version() != MBB200_ABI_VERSION) { std::fprintf(stderr, "ABI version mismatch\n"); return 2; }
  std::ifstream fi(argv[2], std::ios::binary);
  std::vector<char> image((std::istreambuf_iterator<char>(fi)), std::istreambuf_iterator<char>());
  std::ifstream fj(argv[3], std::ios::binary);
  int32_t hd[4];
  int64_t N;
  double sc[4];
  fj.read((char*)hd, sizeof(hd)); fj.read((char*)&N, 8); fj.read((char*)sc, sizeof(sc));
  const int alg = hd[0], n = hd[1], m = hd[2], ns = hd[3];
  std::vector<double> u0((size_t)n * N), p((size_t)(m > 0 ? m : 1) * N), sv((size_t)ns);
  fj.read((char*)u0.data(), sizeof(double) * (size_t)n * N);
  fj.read((char*)p.data(), sizeof(double) * (size_t)m * N);
  fj.read((char*)sv.data(), sizeof(double) * (size_t)ns);
  if (!fj) { std::fprintf(stderr, "short job file\n"); return 2; }
  if (p_mbb200_init(1) != MBB200_OK) { std::fprintf(stderr, "init: %s\n", p_mbb200_last_error()); return 3; }
  mbb200_model* model = nullptr;
  if (p_mbb200_model_load(image.data(), image.size(), &model) != MBB200_OK) { std::fprintf(stderr, "model_load: %s\n", p_mbb200_last_error()); return 3; }
  int32_t mn = 0, mm = 0;
  p_mbb200_model_info(model, &mn, &mm);
  if (mn != n || mm != m) { std::fprintf(stderr, "model has n=%d m=%d, job has n=%d m=%d\n", mn, mm, n, m); return 3; }
  mbb200_opts o;
  std::memset(&o, 0, sizeof(o));
  o.alg = alg; o.t0 = sc[0]; o.t1 = sc[1]; o.abstol = sc[2]; o.reltol = sc[3];
  std::vector<double> out((size_t)ns * n * N), tf((size_t)N);
  std::vector<int32_t> rc((size_t)N), st((size_t)5 * N);
  const int r = p_mbb200_solve(model, &o, N, u0.data(), m > 0 ? p.data() : nullptr, sv.data(), ns, out.data(), tf.data(), rc.data(), st.data(), nullptr);
  if (r != MBB200_OK) { std::fprintf(stderr, "solve: %s\n", p_mbb200_last_error()); return 3; }
  std::ofstream fo(argv[4], std::ios::binary);
  fo.write((const char*)out.data(), sizeof(double) * out.size());
  fo.write((const char*)tf.data(), sizeof(double) * tf.size());
  fo.write((const char*)rc.data(), sizeof(int32_t) * rc.size());
  fo.write((const char*)st.data(), sizeof(int32_t) * st.size());
  long ok = 0;
  for (int64_t i = 0; i < N; i++) ok += rc[(size_t)i] == MBB200_RC_SUCCESS;
  std::printf("abi_driver: model n=%d m=%d, %lld trajectories, %ld successful\n", n, m, (long long)N, ok);
  p_mbb200_model_free(model);
  p_mbb200_shutdown();
  return ok == N ? 0 : 1;
}
