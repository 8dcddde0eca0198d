/* libmultibody_b200 -- C ABI of the B200 ensemble simulator for Multibody.jl models.
 *
 * This is the drop-in boundary of the hot path named by BASELINE.json `north_star`:
 *   ssys = multibody(model)                      /root/reference/src/Multibody.jl:166-188
 *   prob = ODEProblem(ssys, u0map, tspan)        /root/reference/test/test_robot.jl:272-287
 *   solve(EnsembleProblem(prob; prob_func), Rodas5P(), EnsembleB200(); trajectories, saveat,
 *         abstol, reltol)                        (the reference's single-trajectory form of this
 *                                                 call: test/test_robot.jl:288, docs/src/examples/
 *                                                 pendulum.md:68; the ensemble form is SciMLBase's
 *                                                 `__solve(::AbstractEnsembleProblem, alg,
 *                                                 ::EnsembleAlgorithm)`)
 * The Julia side (`EnsembleB200 <: SciMLBase.EnsembleAlgorithm`) binds exactly
 * these entry points with `ccall`.  Plain pointers and sizes only; the caller owns every host
 * pointer for the duration of a call, the library owns all device memory and handles.
 *
 * Return convention: 0 = ok, negative = library error (text via mbb200_last_error()).
 * Per-trajectory integrator outcomes are NOT call failures; they are reported in retcode[]
 * with SciMLBase.ReturnCode values (what `successful_retcode(sol)` inspects,
 * test/test_robot.jl:289).  There is no CPU fallback: without a CUDA device every compute
 * entry returns MBB200_ERR_NO_DEVICE.
 */
#ifndef MBB200_H
#define MBB200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define MBB200_ABI_VERSION 2

/* algorithm ids (OrdinaryDiffEq names: Tsit5(), Rodas5P(), Rodas4(), FBDF()).
 * FBDF (test/test_fourbar.jl:45,71) is served by a variable-order BDF in the published NDF formulation: the same
 * method family and tolerances, not the step sequence of OrdinaryDiffEqBDF. */
enum { MBB200_ALG_TSIT5 = 1, MBB200_ALG_RODAS5P = 2, MBB200_ALG_RODAS4 = 3, MBB200_ALG_FBDF = 4 };

/* SciMLBase.ReturnCode values written to retcode[] */
enum {
  MBB200_RC_DEFAULT = 0, MBB200_RC_SUCCESS = 1, MBB200_RC_DTNAN = 3, MBB200_RC_MAXITERS = 4,
  MBB200_RC_DTLESSTHANMIN = 5, MBB200_RC_UNSTABLE = 6,
  MBB200_RC_INITIALFAILURE = 7   /* DAE: no consistent algebraic states found for this trajectory */
};

/* library errors */
enum {
  MBB200_OK = 0, MBB200_ERR_NO_DEVICE = -1, MBB200_ERR_BAD_ARG = -2, MBB200_ERR_UNKNOWN_MODEL = -3,
  MBB200_ERR_CUDA = -4, MBB200_ERR_NOT_INIT = -5, MBB200_ERR_UNSUPPORTED = -6
};

/* solver options: the SciML `solve` keyword arguments of the reference call sites */
typedef struct mbb200_opts {
  int32_t alg;        /* MBB200_ALG_*                                                   */
  int32_t flags;      /* bit 0: also accumulate ensemble statistics into reduce_out     */
  double t0, t1;      /* tspan                                                          */
  double abstol;      /* default of the reference call: 1e-6                            */
  double reltol;      /* default of the reference call: 1e-3                            */
  double dt0;         /* initial step, 0 = automatic (Hairer)                           */
  double dtmin;       /* 0 = default                                                    */
  double dtmax;       /* 0 = t1 - t0                                                    */
  int64_t maxiters;   /* 0 = 100000                                                     */
  /* PI-controller details of OrdinaryDiffEq (ABI version 2):
   * dead band of `step_accept_controller!`: when qsteady_min <= q <= qsteady_max the step size
   * is kept.  0, 0 = the method's default: (1, 1) for Tsit5, (1, 6/5) for the Rosenbrock methods
   * (`qsteady_max_default(::OrdinaryDiffEqAdaptiveImplicitAlgorithm)`). */
  double qsteady_min, qsteady_max;
  int32_t pow_mode;   /* 0 = `fastpower` of the reference's controller (Float32 exp2 of an          */
                      /* approximate log2), 1 = exact double-precision pow                         */
  int32_t reserved;   /* must be 0                                                                 */
} mbb200_opts;

#define MBB200_FLAG_REDUCE 1

typedef struct mbb200_model mbb200_model;   /* opaque compiled-model handle */

/* Model image ("MBIR1"): what the Julia glue serialises from the reduced system.  In this
 * build the straight-line CUDA of each model is generated and compiled ahead of time, so
 * the image names the compiled kernel set and carries the data that is not code: the
 * shared interpolation tables (KinematicPTP, src/robot/path_planning.jl:191-193). */
typedef struct mbb200_image_header {
  char magic[8];        /* "MBIR1\0\0\0"          */
  char name[48];        /* model name, NUL padded */
  int32_t n, m;         /* states, parameters (validated against the compiled model) */
  int64_t n_table;      /* number of doubles of table data following the header       */
} mbb200_image_header;

/* Model image that CARRIES the model ("MBIR2"): the scheduled expression DAGs of rhs / dT / observed / factor / solve
 * of the reduced system plus its tables -- what the Julia glue serialises from `multibody(model)` + `ODEProblem`
 * (/root/reference/src/Multibody.jl:166-188) for a model that is not compiled into the library.  Byte layout:
 * multibody.jl_b200/codegen/mbir.py and INTEGRATION.md.  mbb200_model_load prints it as CUDA C, appends the library's
 * integrator templates, compiles for sm_100a with NVRTC (found at run time: libnvrtc.so.12 / $MBB200_NVRTC_LIB) and
 * caches the cubin by image hash in $MBB200_CACHE_DIR (default ~/.cache/mbb200; MBB200_NO_CACHE=1 disables).
 * mbb200_image_precompile does the parse + compile step alone, without a device (cache warm-up on a build machine). */
int mbb200_image_precompile(const void* image, size_t len);

int mbb200_abi_version(void);

/* Initialise `n_gpus` devices (ordinals 0..n_gpus-1 of this process).  Idempotent. */
int mbb200_init(int n_gpus);
/* Same with explicit device ordinals (one process per GPU: pass {LOCAL_RANK}, 1). */
int mbb200_init_devices(const int* ordinals, int n);
int mbb200_device_count(void);     /* devices initialised, 0 before init */

int mbb200_model_load(const void* image, size_t len, mbb200_model** out);
int mbb200_model_info(const mbb200_model* model, int32_t* n, int32_t* m);
/* algorithmic flop counts of the generated code: rhs, jac, dT, factor (jac+LU), solve */
int mbb200_model_flops(const mbb200_model* model, int64_t flops[5]);
void mbb200_model_free(mbb200_model* model);

/* Whole ensemble, HOST buffers (the call the Julia glue makes).  Layouts (struct of arrays):
 *   u0        [n][N]            initial states
 *   p         [m][N]            tunable parameters (may be NULL when m == 0)
 *   saveat    [n_save]          ascending save times inside [t0, t1]
 *   out_u     [n_save][n][N]    saved states (dense output of the bracketing step)
 *   out_tfinal[N], retcode [N], stats [5][N] = naccept, nreject, nf, njac, nsolve
 *   reduce_out[2*n_save*n + 1]  count, sum(x), sum(x^2) over successful trajectories
 *                               (nullable; filled when flags & MBB200_FLAG_REDUCE)
 * Trajectories are sharded in contiguous blocks over the initialised devices.  Host buffers may be
 * ordinary pageable memory: they are page-locked for the duration of the call.  Results are
 * copied back chunk by chunk while the kernel is still integrating the rest of the ensemble.
 * Thread-safe: concurrent calls serialise per device. */
int mbb200_solve(mbb200_model* model, const mbb200_opts* opts, int64_t N,
                 const double* u0, const double* p, const double* saveat, int32_t n_save,
                 double* out_u, double* out_tfinal, int32_t* retcode, int32_t* stats,
                 double* reduce_out);

/* Same, DEVICE buffers on device `dev_index` (index into the initialised devices): no
 * host<->device traffic; returns after the kernels have completed. `saveat` stays a host
 * pointer.  Used when the ensemble is generated or consumed on the GPU. */
int mbb200_solve_device(mbb200_model* model, const mbb200_opts* opts, int dev_index, int64_t N,
                        const double* d_u0, const double* d_p, const double* saveat, int32_t n_save,
                        double* d_out_u, double* d_out_tfinal, int32_t* d_retcode, int32_t* d_stats,
                        double* d_reduce_out);

/* Allocation-free variant of mbb200_solve_device.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same arguments, layouts and results as
 * mbb200_solve_device; the kernel's scratch memory (trajectory work counter + device copy of
 * `saveat`) lives in a static device array of 256 + 8 * MBB200_STATIC_MAX_SAVE bytes that every GPU
 * loading the library holds until exit, so the call performs no cudaMalloc / cudaFree and calls no
 * allocating library routine.  Limits and caveats: n_save <= MBB200_STATIC_MAX_SAVE (larger values
 * are rejected with MBB200_ERR_BAD_ARG); with MBB200_FLAG_REDUCE the statistic is that of this
 * device only (no cross-process sum); all calls on one GPU share the array, which keeps the last
 * call's contents (the counter is cleared at the start of every call) -- calls that overlap on the
 * same GPU would race on it (this library serialises its own calls per device). */
#define MBB200_STATIC_MAX_SAVE 4096
int mbb200_solve_device_static(mbb200_model* model, const mbb200_opts* opts, int dev_index, int64_t N,
                               const double* d_u0, const double* d_p, const double* saveat, int32_t n_save,
                               double* d_out_u, double* d_out_tfinal, int32_t* d_retcode, int32_t* d_stats,
                               double* d_reduce_out);

/* Cross-process ensemble statistics: one process per GPU, the only
 * collective of the path is the sum of reduce_out over the ranks.  Rank 0 obtains an id with
 * mbb200_comm_unique_id (128 bytes, a ncclUniqueId), the launcher distributes it (the Julia glue:
 * through MPI.jl / Distributed.jl / a file; bench.py: through its torchrun rendezvous), every
 * rank calls mbb200_comm_init after mbb200_init.  From then on mbb200_solve /
 * mbb200_solve_device (device index 0) with MBB200_FLAG_REDUCE return the statistic summed over
 * ALL ranks (ncclAllReduce, f64, on the solve stream); every rank must make the call.  NCCL is
 * bound at run time (libnccl.so.2 already in the process, $MBB200_NCCL_LIB, or the system copy). */
int mbb200_comm_unique_id(void* id_out, size_t len);
int mbb200_comm_init(const void* id, size_t len, int rank, int nranks);
int mbb200_comm_info(int* rank, int* nranks);            /* returns 1 when a communicator is active */
int mbb200_comm_allreduce(double* host_buf, int64_t count);   /* in-place sum over ranks (no-op without a communicator) */
void mbb200_comm_destroy(void);

/* Observed (eliminated) variables of saved states -- the `sol(tv, idxs = expr)` read-back every
 * reference assertion goes through (test/test_robot.jl:321-325).  `u` is the
 * out_u of a solve ([n_save][n][N], host), out_obs is [n_save][nobs][N]; the order of the
 * nobs variables is fixed by the model (generated header comment / ReducedSystem.observed_names). */
int mbb200_model_observed_count(const mbb200_model* model, int32_t* nobs);
int mbb200_observed(mbb200_model* model, int64_t N, const double* u, const double* p, const double* saveat,
                    int32_t n_save, double* out_obs);

/* Device time (ms, CUDA events on the launching stream) and number of kernel launches of
 * the most recent solve call on device `dev_index`. */
int mbb200_last_timing(int dev_index, double* kernel_ms, int32_t* n_launches);

/* Measured FP64 FMA throughput of device `dev_index` in FLOP/s (dependent-free DFMA loop;
 * the roofline denominator of the RHS/LU-bound kernels). */
int mbb200_measure_fp64_peak(int dev_index, double* flops_per_s);

/* Device memory helpers for callers without a CUDA binding of their own. */
int mbb200_dev_alloc(int dev_index, size_t bytes, void** out);
int mbb200_dev_free(int dev_index, void* ptr);
int mbb200_dev_upload(int dev_index, void* dst, const void* src, size_t bytes);
int mbb200_dev_download(int dev_index, void* dst, const void* src, size_t bytes);
int mbb200_host_alloc_pinned(size_t bytes, void** out);
int mbb200_host_free_pinned(void* ptr);

const char* mbb200_last_error(void);   /* thread-local, valid until the next call */
void mbb200_shutdown(void);

#ifdef __cplusplus
}
#endif
#endif /* MBB200_H */
