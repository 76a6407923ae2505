// C-ABI implementation (include/kolmo_b200.h): workspace, launch sequencing,
// CUDA-graph replay and chunked (L2-blocked) stepping.  sm_100a only.
#include "../../include/kolmo_b200.h"

#include <cuda_runtime.h>

#include <cmath>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "col_tma.cuh"
#include "resident.cuh"

namespace {

thread_local std::string g_err;

int fail(const std::string& msg) {
  g_err = msg;
  return 1;
}

}  // namespace

// shared with slab.cu
int kf_set_error(const char* msg) { return fail(msg); }

namespace {

#define KF_CUDA(expr)                                                                     \
  do {                                                                                    \
    cudaError_t e__ = (expr);                                                             \
    if (e__ != cudaSuccess)                                                               \
      return fail(std::string(#expr) + ": " + cudaGetErrorString(e__));                   \
  } while (0)

// reference coefficients, hydrogym/jax/solvers/base.py:13-15
const double kAlpha[6] = {0, 0.1496590219993, 0.3704009573644, 0.6222557631345, 0.9582821306748, 1};
const double kBeta[5] = {0, -0.4178904745, -1.192151694643, -1.697784692471, -1.514183444257};
const double kGamma[5] = {0.1496590219993, 0.3792103129999, 0.8229550293869, 0.6994504559488, 0.1530572479681};

struct Stage {  // type-erased per-stage scalars
  double beta, gdt, munu;
  int store_h;  // the RK register written by the last sub-stage is never read (the next step starts with beta = 0)
};

// Type-erased launch table, one instance per (dtype, N).
struct Ops {
  size_t elem;  // bytes per complex
  int tc, col_threads, row_groups, row_threads;
  size_t col_smem, row_smem;
  cudaError_t (*prepare)();
  void (*fill_tw)(double* re_im);  // transposed twiddle table of the plan (fft.cuh), N complex
  // TMA-fed persistent column kernel of the RK sub-stage: set-up (tensor maps, grid) and launch
  int (*tma_setup)(struct Ctx&);
  void (*col_tma)(const struct Ctx&, int nb, const Stage&, cudaStream_t);
  // all of these work on the chunk-local workspace; nb = environments in the chunk
  void (*col)(const struct Ctx&, int nb, const Stage&, int mode, cudaStream_t);
  void (*row)(const struct Ctx&, int nb, cudaStream_t);
  // obs_dev / acc already offset to the chunk's first environment
  void (*obs)(const struct Ctx&, int nb, double* acc, size_t acc_stride, void* obs_dev, double obs_w, cudaStream_t);
  void (*fields)(const struct Ctx&, int nb, void* out, int nfields, size_t fstride, cudaStream_t);
  void (*finalize)(const struct Ctx&, int nsnap, void* mean_tke, int* nonfinite, cudaStream_t);
  void (*import_)(const struct Ctx&, int nb, const void* user, size_t user_env, cudaStream_t);
  void (*export_)(const struct Ctx&, int nb, void* user, size_t user_env, cudaStream_t);
  void (*export_h)(const struct Ctx&, int nb, void* user, size_t user_env, cudaStream_t);  // the RK register h
  void (*forcing)(const struct Ctx&, int nb, const void* actions, cudaStream_t);
  void (*forcing_to)(const struct Ctx&, void* G, int nb, const void* actions, cudaStream_t);
  // on-chip-resident RK steps (resident.cuh): whole batch, user layout in place; nullptr where the grid does not fit
  cudaError_t (*resident_prepare)() = nullptr;
  cudaError_t (*resident)(const struct Ctx&, void* omega_user, const void* G, int B, int n_steps, cudaStream_t) = nullptr;
  int resident_cluster = 0;
  bool resident_default = false;  // taken unless KF_RESIDENT=0; otherwise only with KF_RESIDENT=1
};

struct Ctx {
  kf_config cfg;
  const Ops* ops;
  int N, MH, chunk;
  size_t plane;  // complex elements per env of a main/x-space array
  // device buffers (void*: complex<T>)
  // chunk-local workspace
  void *wmain = nullptr, *wnyq = nullptr, *hmain = nullptr, *hnyq = nullptr, *G = nullptr;
  void *fld = nullptr, *advx = nullptr, *tw = nullptr, *F = nullptr, *C = nullptr;
  double* acc = nullptr;  // [B][acc_snaps][4]
  int acc_snaps = 0;
  // device staging for kf_step_host (allocated on first use)
  void *d_omega = nullptr, *d_act = nullptr, *d_obs = nullptr, *d_tke = nullptr;
  int* d_nonf = nullptr;
  size_t bytes = 0;
  long long launches = 0;
  mutable cudaError_t launch_err = cudaSuccess;  // first failed cudaLaunchKernelEx of a stage kernel
  bool have_constants = false;
  // dense control spectrum (general control_field of RungeKuttaCrankNicolson.solve): [N][MH] + [N]; owner = lane 0
  void *ctrl_main = nullptr, *ctrl_nyq = nullptr;
  bool ctrl_on = false, ctrl_owner = false;
  // on-chip-resident path (small grids): forcing rows of the WHOLE batch
  bool resident_on = false;
  void* Gfull = nullptr;
  // TMA column kernel: tensor maps over advx / hmain / wmain and the persistent grid size
  CUtensorMap tm_adv, tm_h, tm_w;
  bool use_tma = false;
  int tma_ctas = 0;
  // programmatic dependent launch of the stage kernels (KF_PDL=1; see pdl_wait in kernels.cuh)
  bool pdl = false;
  // CUDA graphs of 4 RK steps, one per number of environments they were captured for (a full chunk, and the
  // balanced remainder of a batch that is not a multiple of lanes x chunk)
  static constexpr int kGraphs = 2;
  cudaGraphExec_t graph[kGraphs] = {nullptr, nullptr};
  int graph_nb[kGraphs] = {0, 0};
  // L2 policy hints (KF_L2_HINTS, default on when the chunk leaves the batch's persistent state no room in L2)
  int l2_hints = 0;
  // Second lane: an identical chunk-sized workspace driven on its own stream.  Consecutive chunks alternate
  // between the lanes, so the row kernel of one chunk overlaps the column kernel of the other (they saturate
  // different parts of the machine): measured 0.956 -> 0.882 ms per RK step at 256^2 x 256 fp32.
  Ctx* lane2 = nullptr;
  cudaStream_t s2 = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
};


// TC: columns per column-kernel CTA; GROUPS: FFT groups per row-kernel CTA;
// CB / RB: __launch_bounds__ min-blocks (register caps) of the column / row kernels.
template <typename T, class P, int TC, int GROUPS, int CB = 1, int RB = 1, int TB = 1, bool XB = false> struct Impl {
  using C = kf::Cx<T>;
  static constexpr int N = P::L, MH = N / 2;
  using C2 = kf::Cx<typename kf::PairOf<T>::type>;  // lane-paired complex (two transforms per element)
  static size_t col_smem() { return kf::col_exchange_bytes<T, N, TC>() + sizeof(C) * 2 * (size_t)N; }
  static size_t row_smem() { return sizeof(C2) * GROUPS * kf::row_exchange_elems<T, N>() + sizeof(C) * (size_t)N; }

  static cudaError_t prepare() {
    cudaError_t e;
    e = cudaFuncSetAttribute(kf::col_kernel<T, P, TC, kf::kColStage, CB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)col_smem());
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(kf::col_kernel<T, P, TC, kf::kColInit, CB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)col_smem());
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(kf::row_kernel<T, P, GROUPS, RB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)row_smem());
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(kf::row_fields_kernel<T, P, GROUPS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)row_smem());
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(kf::row_obs_kernel<T, P, GROUPS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)row_smem());
  }

  static kf::StageArgs<T> args(const Ctx& c) {
    kf::StageArgs<T> a;
    a.wmain = (C*)c.wmain; a.wnyq = (C*)c.wnyq; a.hmain = (C*)c.hmain; a.hnyq = (C*)c.hnyq;
    a.G = (const C*)c.G; a.advx = (C*)c.advx;
    a.ctrl_main = c.ctrl_on ? (const C*)c.ctrl_main : nullptr;
    a.ctrl_nyq = c.ctrl_on ? (const C*)c.ctrl_nyq : nullptr;
    a.fa = (C2*)c.fld; a.fb = a.fa + (size_t)c.chunk * c.plane;
    a.tw = (const C*)c.tw;
    a.beta = 0; a.gdt = 0; a.munu = 0;
    a.scale = (T)(1.0 / ((double)N * (double)N));
    a.dealias = c.cfg.dealias;
    a.store_h = 1;
    a.l2_hints = c.l2_hints;
    return a;
  }

  static void col(const Ctx& c, int nb, const Stage& st, int mode, cudaStream_t s) {
    kf::StageArgs<T> a = args(c);
    a.beta = (T)st.beta; a.gdt = (T)st.gdt; a.munu = (T)st.munu; a.store_h = st.store_h;
    const dim3 grid(MH / TC, nb);
    if (mode == kf::kColStage) kf::col_kernel<T, P, TC, kf::kColStage, CB><<<grid, P::TPF * TC, col_smem(), s>>>(a);
    else kf::col_kernel<T, P, TC, kf::kColInit, CB><<<grid, P::TPF * TC, col_smem(), s>>>(a);
  }
  // launch of a stage kernel, with the programmatic-serialization attribute when the handle asks for it
  template <class K, class... A>
  static void launch_stage(const Ctx& c, K kernel, dim3 grid, int threads, size_t smem, cudaStream_t s, A... kargs) {
    cudaLaunchConfig_t lc = {};
    lc.gridDim = grid; lc.blockDim = dim3(threads); lc.dynamicSmemBytes = smem; lc.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = c.pdl ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&lc, kernel, kargs...);
    if (e != cudaSuccess && c.launch_err == cudaSuccess) c.launch_err = e;  // reported by run_rk_steps / drive
  }
  static void row(const Ctx& c, int nb, cudaStream_t s) {
    kf::StageArgs<T> a = args(c);
    launch_stage(c, kf::row_kernel<T, P, GROUPS, RB>, dim3(N / 2 / GROUPS, nb), P::TPF * GROUPS, row_smem(), s, a);
  }
  static void obs(const Ctx& c, int nb, double* acc, size_t acc_stride, void* obs_dev, double obs_w, cudaStream_t s) {
    kf::ObsArgs<T> a;
    a.fa = (const C2*)c.fld; a.fb = a.fa + (size_t)c.chunk * c.plane; a.tw = (const C*)c.tw;
    a.acc = acc; a.acc_stride = acc_stride; a.obs = (T*)obs_dev;
    a.nobs = obs_dev ? c.cfg.obs_size : 0;
    a.sx = a.sy = N / c.cfg.obs_size;
    a.obs_w = (T)obs_w;
    kf::row_obs_kernel<T, P, GROUPS><<<dim3(N / GROUPS, nb), P::TPF * GROUPS, row_smem(), s>>>(a);
  }
  static void fields(const Ctx& c, int nb, void* out, int nfields, size_t fstride, cudaStream_t s) {
    kf::ObsArgs<T> a;
    a.fa = (const C2*)c.fld; a.fb = a.fa + (size_t)c.chunk * c.plane; a.tw = (const C*)c.tw;
    a.acc = nullptr; a.acc_stride = 0; a.obs = nullptr; a.nobs = 0; a.sx = a.sy = 1; a.obs_w = 0;
    kf::row_fields_kernel<T, P, GROUPS><<<dim3(N / GROUPS, nb), P::TPF * GROUPS, row_smem(), s>>>(a, (T*)out, nfields, fstride);
  }
  static void finalize(const Ctx& c, int nsnap, void* mean_tke, int* nonfinite, cudaStream_t s) {
    const int B = c.cfg.B;
    kf::tke_finalize_kernel<T><<<(B + 127) / 128, 128, 0, s>>>(c.acc, B, nsnap, N, N, (T*)mean_tke, nonfinite);
  }
  static void import_(const Ctx& c, int nb, const void* user, size_t user_env, cudaStream_t s) {
    kf::import_kernel<T><<<dim3(32, nb), 256, 0, s>>>((const C*)user, user_env, (C*)c.wmain, (C*)c.wnyq, N, MH);
  }
  static void export_(const Ctx& c, int nb, void* user, size_t user_env, cudaStream_t s) {
    kf::export_kernel<T><<<dim3(32, nb), 256, 0, s>>>((C*)user, user_env, (const C*)c.wmain, (const C*)c.wnyq, N, MH);
  }
  static void export_h(const Ctx& c, int nb, void* user, size_t user_env, cudaStream_t s) {
    kf::export_kernel<T><<<dim3(32, nb), 256, 0, s>>>((C*)user, user_env, (const C*)c.hmain, (const C*)c.hnyq, N, MH);
  }
  static void forcing(const Ctx& c, int nb, const void* actions, cudaStream_t s) {
    kf::forcing_row_kernel<T><<<dim3((MH + 1 + 127) / 128, nb), 128, 0, s>>>(
        (C*)c.G, (const C*)c.F, (const C*)c.C, (const T*)actions, (T)c.cfg.min_action, (T)c.cfg.max_action, MH + 1);
  }
  static void forcing_to(const Ctx& c, void* G, int nb, const void* actions, cudaStream_t s) {
    kf::forcing_row_kernel<T><<<dim3((MH + 1 + 127) / 128, nb), 128, 0, s>>>(
        (C*)G, (const C*)c.F, (const C*)c.C, (const T*)actions, (T)c.cfg.min_action, (T)c.cfg.max_action, MH + 1);
  }
  static int tma_setup(Ctx& c);
  static void col_tma(const Ctx& c, int nb, const Stage& st, cudaStream_t s) {
    kf::StageArgs<T> a = args(c);
    a.beta = (T)st.beta; a.gdt = (T)st.gdt; a.munu = (T)st.munu; a.store_h = st.store_h;
    const int ntiles = nb * (MH / TC);
    const int grid = ntiles < c.tma_ctas ? ntiles : c.tma_ctas;
    launch_stage(c, kf::col_tma_kernel<T, P, TC, TB, XB>, dim3(grid), P::TPF * TC, kf::ColTmaSmem<T, P, TC, XB>::bytes(),
                 s, a, c.tm_adv, c.tm_h, c.tm_w, ntiles);
  }
  static Ops make() {
    Ops o;
    o.elem = sizeof(C);
    o.tc = TC; o.col_threads = P::TPF * TC; o.row_groups = GROUPS; o.row_threads = P::TPF * GROUPS;
    o.col_smem = col_smem(); o.row_smem = row_smem();
    o.prepare = prepare; o.fill_tw = kf::fill_twiddles<P>; o.tma_setup = tma_setup; o.col_tma = col_tma; o.col = col; o.row = row; o.obs = obs; o.fields = fields; o.finalize = finalize;
    o.import_ = import_; o.export_ = export_; o.export_h = export_h; o.forcing = forcing; o.forcing_to = forcing_to;
    return o;
  }
};

// cuTensorMapEncodeTiled, fetched through the runtime so that the library has no link-time
// dependency on libcuda (it must load on GPU-less build hosts).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

template <typename T, class P, int TC, int GROUPS, int CB, int RB, int TB, bool XB>
int Impl<T, P, TC, GROUPS, CB, RB, TB, XB>::tma_setup(Ctx& c) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return fail("cuTensorMapEncodeTiled is not available from this driver");
  const size_t smem = kf::ColTmaSmem<T, P, TC, XB>::bytes();
  auto kern = kf::col_tma_kernel<T, P, TC, TB, XB>;
  KF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0, sms = 0;
  KF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, P::TPF * TC, smem));
  KF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c.cfg.device));
  if (per_sm < 1) return fail("col_tma_kernel does not fit on an SM");
  if (const char* e = std::getenv("KF_COL_PER_SM"))  // experiment: fewer resident column CTAs leave room for row CTAs
    if (std::atoi(e) >= 1 && std::atoi(e) < per_sm) per_sm = std::atoi(e);
  c.tma_ctas = per_sm * sms;
  const CUtensorMapDataType dt = sizeof(T) == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64;
  const cuuint64_t dims[2] = {(cuuint64_t)(2 * MH), (cuuint64_t)c.chunk * N};
  const cuuint64_t strides[1] = {(cuuint64_t)(2 * MH) * sizeof(T)};
  const cuuint32_t box[2] = {(cuuint32_t)(2 * TC), (cuuint32_t)(N < 256 ? N : 256)};
  const cuuint32_t estr[2] = {1, 1};
  void* bases[3] = {c.advx, c.hmain, c.wmain};
  CUtensorMap* maps[3] = {&c.tm_adv, &c.tm_h, &c.tm_w};
  for (int i = 0; i < 3; ++i) {
    CUresult r = enc(maps[i], dt, 2, bases[i], dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
  }
  c.use_tma = true;
  return 0;
}

// Tuning knobs of the headline configurations; overridable one by one (-DKF_F256_TC=8 ...) for experiments.
//   TC: columns per column-kernel CTA   GROUPS: FFT groups per row-kernel CTA
//   CB / RB / TB: min CTAs per SM (register caps) of col_kernel / row_kernel / col_tma_kernel
#ifndef KF_F256_TC
#define KF_F256_TC 8
#endif
#ifndef KF_F256_GROUPS
#define KF_F256_GROUPS 8
#endif
#ifndef KF_F256_CB
#define KF_F256_CB 2
#endif
#ifndef KF_F256_RB
#define KF_F256_RB 4
#endif
#ifndef KF_F256_TB
#define KF_F256_TB 2
#endif
#ifndef KF_F256_XB
#define KF_F256_XB false
#endif
#define KF_CFG_F256 KF_F256_TC, KF_F256_GROUPS, KF_F256_CB, KF_F256_RB, KF_F256_TB, KF_F256_XB
#ifndef KF_F256_TPF  // FFT plan of the 256^2 fp32 kernels: threads per transform and the radix sequence
#define KF_F256_TPF 16
#define KF_F256_R0 16
#define KF_F256_R1 16
#define KF_F256_R2 1
#endif
#define KF_PLAN_F256 256, KF_F256_TPF, KF_F256_R0, KF_F256_R1, KF_F256_R2
// fp64: 16 elements per thread (one exchange fewer per transform than 8) beats the lower register pressure of 8:
// 256^2 row kernel 242 -> 186 us, column kernel 300 -> 278 us per launch (256 envs)
#ifndef KF_D256_RB
#define KF_D256_RB 2
#endif
#ifndef KF_D256_GROUPS
#define KF_D256_GROUPS 8
#endif
#define KF_CFG_D256 8, KF_D256_GROUPS, 2, KF_D256_RB, 2
#ifdef KF_D256_E8
#define KF_PLAN_D256 256, 32, 8, 4, 8
#else
#define KF_PLAN_D256 256, 16, 16, 16
#endif
#ifdef KF_DOTHER_E8
#define KF_IMPL_D128 Impl<double, Plan<128, 16, 8, 2, 8>, 8, 8>
#define KF_IMPL_D512 Impl<double, Plan<512, 64, 8, 8, 8>, 4, 2>
#define KF_IMPL_D1024 Impl<double, Plan<1024, 128, 8, 2, 8, 8>, 4, 1>
#else
#define KF_IMPL_D128 Impl<double, Plan<128, 8, 16, 8>, 8, 16>
#define KF_IMPL_D512 Impl<double, Plan<512, 32, 8, 8, 8>, 4, 4>
#define KF_IMPL_D1024 Impl<double, Plan<1024, 64, 16, 4, 16>, 4, 2>
#endif
#define KF_CFG_F128 16, 16
#ifndef KF_F1024_GROUPS
#define KF_F1024_GROUPS 2
#endif
#ifndef KF_F1024_RB
#define KF_F1024_RB 3
#endif
#ifndef KF_F1024_TB
#define KF_F1024_TB 1
#endif
#ifndef KF_F1024_R1
#define KF_F1024_R1 4
#endif
#ifndef KF_F1024_R2
#define KF_F1024_R2 16
#endif
#ifndef KF_F1024_TC
#define KF_F1024_TC 8
#endif
#define KF_CFG_F1024 KF_F1024_TC, KF_F1024_GROUPS, 1, KF_F1024_RB, KF_F1024_TB
#ifdef KF_F1024_E32
#define KF_PLAN_F1024 1024, 32, 32, 32
#else
#define KF_PLAN_F1024 1024, 64, 16, KF_F1024_R1, KF_F1024_R2
#endif

// Launcher of resident_kernel<T, P, CS, MINB> (one cluster of CS CTAs per environment)
template <typename T, class P, int CS, int MINB> struct Resident {
  using C = kf::Cx<T>;
  static cudaError_t prepare() {
    cudaError_t e = cudaFuncSetAttribute(kf::resident_kernel<T, P, CS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kf::ResidentSmem<T, P, CS>::bytes());
    return e;
  }
  static cudaError_t launch(const Ctx& c, void* omega_user, const void* G, int B, int n_steps, cudaStream_t s) {
    kf::ResidentArgs<T> a;
    a.omega = (C*)omega_user;
    a.G = (const C*)G;
    a.tw = (const C*)c.tw;
    for (int k = 0; k < 5; ++k) {
      a.beta[k] = (T)kBeta[k];
      a.gdt[k] = (T)(kGamma[k] * c.cfg.dt);                                     // base.py:55
      a.munu[k] = (T)(0.5 * c.cfg.dt * (kAlpha[k + 1] - kAlpha[k]) * c.cfg.nu);  // base.py:54, kolmogorov.py:209-220
    }
    a.scale = (T)(1.0 / ((double)P::L * (double)P::L));
    a.dealias = c.cfg.dealias;
    a.n_steps = n_steps;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)B * CS);
    lc.blockDim = dim3(256);
    lc.dynamicSmemBytes = kf::ResidentSmem<T, P, CS>::bytes();
    lc.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    lc.attrs = at;
    lc.numAttrs = CS > 1 ? 1 : 0;
    return cudaLaunchKernelEx(&lc, kf::resident_kernel<T, P, CS, MINB>, a);
  }
  static void attach(Ops& o, bool by_default) {
    o.resident_prepare = prepare;
    o.resident = launch;
    o.resident_cluster = CS;
    o.resident_default = by_default;
  }
};

template <class R> Ops with_resident(Ops o, bool by_default) {
  R::attach(o, by_default);
  return o;
}

const Ops* find_ops(int dtype, int N) {
  using kf::Plan;
  // on-chip-resident kernels (resident.cuh): 64^2 one CTA per environment (default path); 128^2 / 256^2 fp32 as
  // clusters of 2 / 8 CTAs exchanging through distributed shared memory (opt-in: KF_RESIDENT=1)
  static const Ops f64_[] = {with_resident<Resident<float, Plan<64, 8, 8, 8>, 1, 2>>(Impl<float, Plan<64, 8, 8, 8>, 16, 16>::make(), true)};
  static const Ops f128[] = {with_resident<Resident<float, Plan<128, 8, 16, 8>, 2, 1>>(Impl<float, Plan<128, 8, 16, 8>, KF_CFG_F128>::make(), false)};
#if KF_F256_TPF == 16
  static const Ops f256[] = {with_resident<Resident<float, Plan<256, 16, 16, 16>, 8, 1>>(Impl<float, Plan<KF_PLAN_F256>, KF_CFG_F256>::make(), false)};
#else
  static const Ops f256[] = {Impl<float, Plan<KF_PLAN_F256>, KF_CFG_F256>::make()};
#endif
  static const Ops f512[] = {Impl<float, Plan<512, 32, 8, 8, 8>, 8, 4>::make()};
  static const Ops f1024[] = {Impl<float, Plan<KF_PLAN_F1024>, KF_CFG_F1024>::make()};
  static const Ops d64[] = {with_resident<Resident<double, Plan<64, 8, 8, 8>, 1, 1>>(Impl<double, Plan<64, 8, 8, 8>, 16, 16>::make(), true)};
  static const Ops d128[] = {KF_IMPL_D128::make()};
  static const Ops d256[] = {Impl<double, Plan<KF_PLAN_D256>, KF_CFG_D256>::make()};
  static const Ops d512[] = {KF_IMPL_D512::make()};
  static const Ops d1024[] = {KF_IMPL_D1024::make()};
  const bool f = dtype == KF_F32;
  switch (N) {
    case 64: return f ? f64_ : d64;
    case 128: return f ? f128 : d128;
    case 256: return f ? f256 : d256;
    case 512: return f ? f512 : d512;
    case 1024: return f ? f1024 : d1024;
    default: return nullptr;
  }
}

}  // namespace

namespace {

Stage stage_of(const Ctx& c, int s) {
  Stage st;
  st.beta = kBeta[s];
  st.gdt = kGamma[s] * c.cfg.dt;                                     // base.py:55
  st.munu = 0.5 * c.cfg.dt * (kAlpha[s + 1] - kAlpha[s]) * c.cfg.nu;  // base.py:54, kolmogorov.py:209-220
  st.store_h = (s != 4);
  return st;
}

// one RK solver step (5 sub-stages, 10 launches) on the chunk workspace;
// expects the x-space fields of the current omega_hat to be present.
void enqueue_rk_step(Ctx& c, int nb, cudaStream_t s) {
  for (int k = 0; k < 5; ++k) {
    c.ops->row(c, nb, s);
    if (c.use_tma) c.ops->col_tma(c, nb, stage_of(c, k), s);
    else c.ops->col(c, nb, stage_of(c, k), kf::kColStage, s);
  }
}

void drop_graphs(Ctx& c) {
  for (int i = 0; i < Ctx::kGraphs; ++i) {
    if (c.graph[i]) cudaGraphExecDestroy(c.graph[i]);
    c.graph[i] = nullptr;
    c.graph_nb[i] = 0;
  }
}

// graph of 4 RK steps on nb environments of the lane's workspace; *out = nullptr if no slot is free (direct launches)
int build_graph(Ctx& c, int nb, cudaGraphExec_t* out) {
  *out = nullptr;
  int slot = -1;
  for (int i = 0; i < Ctx::kGraphs; ++i) {
    if (c.graph[i] && c.graph_nb[i] == nb) { *out = c.graph[i]; return 0; }
    if (!c.graph[i] && slot < 0) slot = i;
  }
  if (slot < 0) return 0;
  const int steps = 4;
  cudaGraph_t g = nullptr;
  // capture on a private stream (the caller's may be the legacy default stream, which cannot capture)
  cudaStream_t cs = nullptr;
  KF_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
  cudaError_t ce = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
  if (ce == cudaSuccess) {
    for (int i = 0; i < steps; ++i) enqueue_rk_step(c, nb, cs);
    ce = cudaStreamEndCapture(cs, &g);
  }
  cudaStreamDestroy(cs);
  if (ce != cudaSuccess) return fail(std::string("graph capture: ") + cudaGetErrorString(ce));
  cudaError_t e = cudaGraphInstantiate(&c.graph[slot], g, 0);
  cudaGraphDestroy(g);
  if (e != cudaSuccess) return fail(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e));
  c.graph_nb[slot] = nb;
  *out = c.graph[slot];
  return 0;
}

// n RK steps on the chunk workspaces of up to two lanes, enqueued in lockstep (graph launch by graph launch) so
// that neither stream's queue runs dry while the host is still feeding the other one
// (fields must be current on entry, are current on exit)
int run_rk_steps(Ctx* const* L, const int* nb, const cudaStream_t* ls, int nl, int n) {
  int done = 0;
  cudaGraphExec_t gx[2] = {nullptr, nullptr};
  bool any_graph = false;
  for (int l = 0; l < nl; ++l) {
    if (nb[l] > 0 && L[l]->cfg.use_graph && n >= 8) {
      if (int rc = build_graph(*L[l], nb[l], &gx[l])) return rc;
      any_graph = any_graph || gx[l] != nullptr;
    }
  }
  const int gsteps = 4;
  while (any_graph && n - done >= gsteps) {
    for (int l = 0; l < nl; ++l) {
      if (nb[l] <= 0) continue;
      if (gx[l]) {
        KF_CUDA(cudaGraphLaunch(gx[l], ls[l]));
      } else {
        for (int i = 0; i < gsteps; ++i) enqueue_rk_step(*L[l], nb[l], ls[l]);
      }
      L[l]->launches += 10LL * gsteps;
    }
    done += gsteps;
  }
  for (; done < n; ++done) {
    for (int l = 0; l < nl; ++l) {
      if (nb[l] <= 0) continue;
      enqueue_rk_step(*L[l], nb[l], ls[l]);
      L[l]->launches += 10;
    }
  }
  KF_CUDA(cudaGetLastError());
  for (int l = 0; l < nl; ++l)
    if (L[l]->launch_err != cudaSuccess) {
      const cudaError_t e = L[l]->launch_err;
      L[l]->launch_err = cudaSuccess;
      return fail(std::string("stage kernel launch: ") + cudaGetErrorString(e));
    }
  return 0;
}

int ensure_acc(Ctx& c, int nsnap) {
  if (nsnap <= c.acc_snaps) return 0;
  if (c.acc) cudaFree(c.acc);
  c.acc = nullptr;
  KF_CUDA(cudaMalloc(&c.acc, sizeof(double) * 4 * (size_t)c.cfg.B * nsnap));
  c.acc_snaps = nsnap;
  return 0;
}

// Core driver shared by kf_rollout / kf_rk_steps / kf_step.
//   n_outer blocks of n_inner RK steps; after each block: snapshot to trajectory
//   (if given) and reward/obs reductions (if want_obs).  Chunks go to the lanes alternately; lane 0 runs on the
//   caller's stream, lane 1 on the handle's second stream, forked / joined with events.
int drive(Ctx& c, void* omega, const void* actions, void* traj, int n_outer, int n_inner, bool want_obs,
          void* obs_dev, void* tke_dev, int* nonf_dev, cudaStream_t s) {
  if (!c.have_constants) return fail("kf_set_constants has not been called");
  const size_t el = c.ops->elem, real = el / 2;
  const int B = c.cfg.B, N = c.N, MH1 = c.MH + 1;
  const size_t user_env = (size_t)N * MH1;  // complex elements per env in user layout
  const int nobs2 = c.cfg.obs_size * c.cfg.obs_size;
  const int nl = (c.lane2 && B > c.chunk) ? 2 : 1;
  Ctx* L[2] = {&c, c.lane2};
  cudaStream_t ls[2] = {s, c.s2};
  if (want_obs) {
    if (int rc = ensure_acc(c, n_outer)) return rc;
    KF_CUDA(cudaMemsetAsync(c.acc, 0, sizeof(double) * 4 * (size_t)B * n_outer, s));
    if (obs_dev) KF_CUDA(cudaMemsetAsync(obs_dev, 0, real * (size_t)B * nobs2, s));
  }
  if (c.resident_on && !c.ctrl_on) {
    // On-chip-resident path: the whole batch advances in place (user layout) inside ONE kernel per block of n_inner
    // RK steps; the chunk workspace only serves the reward / observation reductions of the snapshots.
    c.ops->forcing_to(c, c.Gfull, B, actions, s);
    c.launches += 1;
    for (int j = 0; j < n_outer; ++j) {
      KF_CUDA(c.ops->resident(c, omega, c.Gfull, B, n_inner, s));
      c.launches += 1;
      if (traj)
        KF_CUDA(cudaMemcpy2DAsync((char*)traj + el * user_env * j, el * user_env * n_outer, omega, el * user_env,
                                  el * user_env, B, cudaMemcpyDeviceToDevice, s));
      if (want_obs) {
        for (int b0 = 0; b0 < B; b0 += c.chunk) {
          const int nb = B - b0 < c.chunk ? B - b0 : c.chunk;
          c.ops->import_(c, nb, (const char*)omega + el * user_env * b0, user_env, s);
          c.ops->col(c, nb, stage_of(c, 0), kf::kColInit, s);
          c.ops->obs(c, nb, c.acc + 4 * ((size_t)b0 * n_outer + j), 4 * (size_t)n_outer,
                     obs_dev ? (char*)obs_dev + real * (size_t)b0 * nobs2 : nullptr, 1.0 / n_outer, s);
          c.launches += 3;
        }
      }
    }
    if (want_obs) {
      c.ops->finalize(c, n_outer, tke_dev, nonf_dev, s);
      c.launches += 1;
    }
    KF_CUDA(cudaGetLastError());
    return 0;
  }
  if (nl == 2) {
    KF_CUDA(cudaEventRecord(c.ev_fork, s));
    KF_CUDA(cudaStreamWaitEvent(c.s2, c.ev_fork, 0));
  }
  for (int g0 = 0; g0 < B; g0 += nl * c.chunk) {
    int b0[2], nb[2];
    // a last round with fewer than lanes x chunk environments is split evenly between the lanes
    const int left = B - g0;
    const int per = (nl == 2 && left < 2 * c.chunk) ? (left + 1) / 2 : c.chunk;
    for (int l = 0; l < nl; ++l) {
      b0[l] = g0 + l * per;
      nb[l] = B - b0[l] < per ? B - b0[l] : per;
      if (nb[l] <= 0) { nb[l] = 0; continue; }
      char* om = (char*)omega + el * user_env * b0[l];
      c.ops->import_(*L[l], nb[l], om, user_env, ls[l]);
      c.ops->forcing(*L[l], nb[l], actions ? (const char*)actions + real * 4 * b0[l] : nullptr, ls[l]);
      c.ops->col(*L[l], nb[l], stage_of(c, 0), kf::kColInit, ls[l]);  // fields of the initial state
      L[l]->launches += 3;
    }
    for (int j = 0; j < n_outer; ++j) {
      if (int rc = run_rk_steps(L, nb, ls, nl, n_inner)) return rc;
      for (int l = 0; l < nl; ++l) {
        if (nb[l] <= 0) continue;
        if (traj) {
          char* tj = (char*)traj + el * user_env * ((size_t)b0[l] * n_outer + j);
          c.ops->export_(*L[l], nb[l], tj, user_env * n_outer, ls[l]);
          L[l]->launches += 1;
        }
        if (want_obs) {
          c.ops->obs(*L[l], nb[l], c.acc + 4 * ((size_t)b0[l] * n_outer + j), 4 * (size_t)n_outer,
                     obs_dev ? (char*)obs_dev + real * (size_t)b0[l] * nobs2 : nullptr, 1.0 / n_outer, ls[l]);
          L[l]->launches += 1;
        }
      }
    }
    for (int l = 0; l < nl; ++l) {
      if (nb[l] <= 0) continue;
      c.ops->export_(*L[l], nb[l], (char*)omega + el * user_env * b0[l], user_env, ls[l]);
      L[l]->launches += 1;
    }
  }
  if (nl == 2) {
    KF_CUDA(cudaEventRecord(c.ev_join, c.s2));
    KF_CUDA(cudaStreamWaitEvent(s, c.ev_join, 0));
  }
  if (want_obs) {
    c.ops->finalize(c, n_outer, tke_dev, nonf_dev, s);
    c.launches += 1;
  }
  KF_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

struct kf_handle_s {
  Ctx c;
};

namespace {

int dev_alloc(Ctx& c, void** p, size_t bytes) {
  KF_CUDA(cudaMalloc(p, bytes));
  c.bytes += bytes;
  return 0;
}

void free_all(Ctx& c) {
  if (c.lane2) {
    free_all(*c.lane2);
    delete c.lane2;
    c.lane2 = nullptr;
  }
  if (c.s2) cudaStreamDestroy(c.s2);
  if (c.ev_fork) cudaEventDestroy(c.ev_fork);
  if (c.ev_join) cudaEventDestroy(c.ev_join);
  void* ptrs[] = {c.wmain, c.wnyq, c.hmain, c.hnyq, c.G, c.fld, c.advx, c.tw, c.F, c.C,
                  c.acc, c.d_omega, c.d_act, c.d_obs, c.d_tke, c.d_nonf, c.ctrl_owner ? c.ctrl_main : nullptr,
                  c.ctrl_owner ? c.ctrl_nyq : nullptr, c.Gfull};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  drop_graphs(c);
}

// one lane: chunk-sized workspace, twiddles, TMA descriptors
int setup_lane(Ctx& c, const kf_config* cfg, const Ops* ops, int chunk) {
  c.cfg = *cfg;
  c.ops = ops;
  c.N = cfg->N;
  c.MH = cfg->N / 2;
  c.plane = (size_t)c.N * c.MH;
  c.chunk = chunk;
  const size_t el = ops->elem;
  const size_t cp = (size_t)c.chunk * c.plane;
  if (int rc = dev_alloc(c, &c.wmain, el * cp)) return rc;
  if (int rc = dev_alloc(c, &c.hmain, el * cp)) return rc;
  if (int rc = dev_alloc(c, &c.wnyq, el * (size_t)c.chunk * c.N)) return rc;
  if (int rc = dev_alloc(c, &c.hnyq, el * (size_t)c.chunk * c.N)) return rc;
  if (int rc = dev_alloc(c, &c.G, el * (size_t)c.chunk * (c.MH + 1))) return rc;
  if (int rc = dev_alloc(c, &c.fld, el * 4 * cp)) return rc;
  if (int rc = dev_alloc(c, &c.advx, el * cp)) return rc;
  if (int rc = dev_alloc(c, &c.tw, el * c.N)) return rc;
  if (int rc = dev_alloc(c, &c.F, el * (c.MH + 1))) return rc;
  if (int rc = dev_alloc(c, &c.C, el * 4 * (c.MH + 1))) return rc;
  std::vector<double> t64(2 * (size_t)c.N);
  ops->fill_tw(t64.data());  // transposed twiddle table, computed in fp64 and rounded once
  if (cfg->dtype == KF_F32) {
    std::vector<float> t32(t64.begin(), t64.end());
    KF_CUDA(cudaMemcpy(c.tw, t32.data(), el * c.N, cudaMemcpyHostToDevice));
  } else {
    KF_CUDA(cudaMemcpy(c.tw, t64.data(), el * c.N, cudaMemcpyHostToDevice));
  }
  if (const char* e = std::getenv("KF_PDL")) c.pdl = std::atoi(e) != 0;
  // L2-blocked stepping (a chunk smaller than the batch lives in L2 by design): stream omega_hat / h with the
  // evict_first policy so the fields and advection rows keep the cache (+1 % at 256^2 x 256 fp32, profiles/r2_sweep_*)
  c.l2_hints = (chunk < cfg->B) ? 1 : 0;
  if (const char* e = std::getenv("KF_L2_HINTS")) c.l2_hints = std::atoi(e) & 1;
  if (!std::getenv("KF_NO_TMA")) return ops->tma_setup(c);
  return 0;
}

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    int cur = -1;
    cudaGetDevice(&cur);
    if (prev >= 0 && cur != prev) cudaSetDevice(prev);
  }
};

}  // namespace

extern "C" {

const char* kf_last_error(void) { return g_err.c_str(); }

int kf_create(kf_handle* out, const kf_config* cfg) {
  if (!out || !cfg) return fail("kf_create: null argument");
  *out = nullptr;
  if (cfg->N != cfg->M) return fail("kf_create: only square grids are supported (N == M)");
  if (cfg->dtype != KF_F32 && cfg->dtype != KF_F64) return fail("kf_create: dtype must be KF_F32 or KF_F64");
  const Ops* ops = find_ops(cfg->dtype, cfg->N);
  if (!ops) return fail("kf_create: grid size must be one of 64, 128, 256, 512, 1024");
  if (cfg->B < 1) return fail("kf_create: B must be >= 1");
  if (cfg->drag != 0.0) return fail("kf_create: drag must be 0 (the reference has no drag term)");
  if (cfg->n_inner < 1 || cfg->n_outer < 1) return fail("kf_create: n_inner and n_outer must be >= 1");
  if (cfg->obs_size < 1 || cfg->N % cfg->obs_size != 0) return fail("kf_create: obs_size must divide the grid size");
  if (!(cfg->dt > 0.0) || !(cfg->nu >= 0.0)) return fail("kf_create: need dt > 0 and nu >= 0");
  int ndev = 0;
  KF_CUDA(cudaGetDeviceCount(&ndev));
  if (cfg->device < 0 || cfg->device >= ndev) return fail("kf_create: no such CUDA device");
  DeviceGuard guard(cfg->device);
  cudaDeviceProp prop;
  KF_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major != 10) return fail("kf_create: this library is built for sm_100a (B200) only");

  // chunking: an explicit chunk is honoured; by default an fp32 batch on a grid of at most 512^2 is split into two
  // chunks, one per lane, when each lane's launches still have >= 256 column tiles.  Sustained measurements
  // (tools/gpu_sustained.py, % of the HBM roofline, two lanes vs one): 256^2 x 32: 77.6 vs 69.4; 256^2 x 16:
  // 49.9 vs 54.4 (128 tiles per lane: too few); 512^2 x 16: 58.8 vs 49.1; 512^2 x 32: 59.6 vs 56.4; 1024^2 x 16:
  // 53.0 vs 55.9 and 1024^2 x 64: 55.0 vs 60.3 (one lane wins); fp64 256^2 x 256: 2.62 vs 2.29 ms (one lane wins).
  // KF_LANES=1 / KF_LANES=2 force one / two lanes (experiments).
  const char* lanes_env = std::getenv("KF_LANES");
  const bool forced1 = lanes_env && lanes_env[0] == '1', forced2 = lanes_env && lanes_env[0] == '2';
  const bool two_lanes_wanted = forced2 || (!forced1 && cfg->dtype == KF_F32 && cfg->N <= 512);
  int chunk = (cfg->chunk > 0 && cfg->chunk < cfg->B) ? cfg->chunk : cfg->B;
  const bool enough_tiles = (cfg->B / 2) * (cfg->N / 2 / ops->tc) >= 256;
  if (cfg->chunk <= 0 && two_lanes_wanted && (forced2 ? cfg->B >= 2 : enough_tiles)) {
    chunk = (cfg->B + 1) / 2;
    // L2 time blocking: if both lanes' working sets (7 fields per environment) fit in L2 together, the 4 x-space
    // fields and the advection rows never reach HBM between the two kernels -- less DRAM power, so higher
    // sustained clocks under the 1 kW cap (256^2 x 256: 0.912 -> 0.869 ms per RK step; 128^2 x 1024: 0.905 ->
    // 0.836).  Largest power-of-two chunk with 2 * chunk * 7C <= 95 % of L2, provided a launch still has >= 512
    // column tiles (512^2 and 1024^2 do not: they keep B/2).
    int l2 = 0;
    cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, cfg->device);
    const double ws_env = 7.0 * (double)ops->elem * cfg->N * (cfg->N / 2 + 1);
    int c2 = 1;
    while (2.0 * (2 * c2) * ws_env <= 0.95 * (double)l2) c2 *= 2;
    if (l2 > 0 && c2 < chunk && c2 * (cfg->N / 2 / ops->tc) >= 512) chunk = c2;
  }
  const bool two_lanes = two_lanes_wanted && cfg->B > chunk;

  kf_handle h = new kf_handle_s();
  int rc = 0;
  cudaError_t pe = ops->prepare();
  if (pe != cudaSuccess) rc = fail(std::string("kernel attribute setup: ") + cudaGetErrorString(pe));
  if (!rc) rc = setup_lane(h->c, cfg, ops, chunk);
  if (!rc && ops->resident) {
    // on-chip-resident stepping is the default where a grid fits (64^2); KF_RESIDENT=0 selects the two-kernel path
    const char* e = std::getenv("KF_RESIDENT");
    if (e ? e[0] == '1' : ops->resident_default) {
      cudaError_t re = ops->resident_prepare();
      if (re != cudaSuccess) rc = fail(std::string("resident kernel set-up: ") + cudaGetErrorString(re));
      if (!rc) rc = dev_alloc(h->c, &h->c.Gfull, ops->elem * (size_t)cfg->B * (cfg->N / 2 + 1));
      if (!rc) h->c.resident_on = true;
    }
  }
  if (!rc && two_lanes && !h->c.resident_on) {
    Ctx& c = h->c;
    c.lane2 = new Ctx();
    rc = setup_lane(*c.lane2, cfg, ops, chunk);
    cudaError_t e = cudaSuccess;
    if (!rc) e = cudaStreamCreateWithFlags(&c.s2, cudaStreamNonBlocking);
    if (!rc && e == cudaSuccess) e = cudaEventCreateWithFlags(&c.ev_fork, cudaEventDisableTiming);
    if (!rc && e == cudaSuccess) e = cudaEventCreateWithFlags(&c.ev_join, cudaEventDisableTiming);
    if (!rc && e != cudaSuccess) rc = fail(std::string("second lane set-up: ") + cudaGetErrorString(e));
  }
  if (rc) {
    free_all(h->c);
    delete h;
    return rc;
  }
  *out = h;
  return 0;
}

int kf_destroy(kf_handle h) {
  if (!h) return 0;
  DeviceGuard guard(h->c.cfg.device);
  cudaDeviceSynchronize();
  free_all(h->c);
  delete h;
  return 0;
}

int kf_set_constants(kf_handle h, const void* forcing_row, const void* control_rows) {
  if (!h || !forcing_row || !control_rows) return fail("kf_set_constants: null argument");
  Ctx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  const size_t row = c.ops->elem * (c.MH + 1);
  for (Ctx* l : {&c, c.lane2}) {
    if (!l) continue;
    KF_CUDA(cudaMemcpy(l->F, forcing_row, row, cudaMemcpyHostToDevice));
    KF_CUDA(cudaMemcpy(l->C, control_rows, 4 * row, cudaMemcpyHostToDevice));
    l->have_constants = true;
  }
  return 0;
}

int kf_rollout(kf_handle h, void* omega_hat_dev, const void* actions_dev, void* trajectory_dev, void* stream) {
  if (!h || !omega_hat_dev) return fail("kf_rollout: null argument");
  Ctx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  return drive(c, omega_hat_dev, actions_dev, trajectory_dev, c.cfg.n_outer, c.cfg.n_inner, false, nullptr, nullptr,
               nullptr, (cudaStream_t)stream);
}

int kf_rk_steps(kf_handle h, void* omega_hat_dev, const void* actions_dev, int n_steps, void* stream) {
  if (!h || !omega_hat_dev) return fail("kf_rk_steps: null argument");
  if (n_steps < 0) return fail("kf_rk_steps: n_steps must be >= 0");
  Ctx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  return drive(c, omega_hat_dev, actions_dev, nullptr, 1, n_steps, false, nullptr, nullptr, nullptr,
               (cudaStream_t)stream);
}

int kf_step(kf_handle h, void* omega_hat_dev, const void* actions_dev, void* trajectory_dev, void* obs_dev,
            void* mean_tke_dev, int* nonfinite_dev, void* stream) {
  if (!h || !omega_hat_dev) return fail("kf_step: null argument");
  Ctx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  return drive(c, omega_hat_dev, actions_dev, trajectory_dev, c.cfg.n_outer, c.cfg.n_inner, true, obs_dev,
               mean_tke_dev, nonfinite_dev, (cudaStream_t)stream);
}

long long kf_launch_count(kf_handle h) {
  return h ? h->c.launches + (h->c.lane2 ? h->c.lane2->launches : 0) : 0;
}
size_t kf_workspace_bytes(kf_handle h) { return h ? h->c.bytes + (h->c.lane2 ? h->c.lane2->bytes : 0) : 0; }
int kf_chunk(kf_handle h) { return h ? h->c.chunk : 0; }
int kf_lanes(kf_handle h) { return h ? (h->c.lane2 ? 2 : 1) : 0; }

}  // extern "C"

extern "C" {

int kf_reward_obs(kf_handle h, const void* trajectory_dev, int n_snap, void* obs_dev, void* mean_tke_dev,
                  int* nonfinite_dev, void* stream) {
  if (!h || !trajectory_dev) return fail("kf_reward_obs: null argument");
  if (n_snap < 1) return fail("kf_reward_obs: n_snap must be >= 1");
  Ctx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t el = c.ops->elem, real = el / 2;
  const int B = c.cfg.B;
  const size_t user_env = (size_t)c.N * (c.MH + 1);
  const int nobs2 = c.cfg.obs_size * c.cfg.obs_size;
  if (int rc = ensure_acc(c, n_snap)) return rc;
  KF_CUDA(cudaMemsetAsync(c.acc, 0, sizeof(double) * 4 * (size_t)B * n_snap, s));
  if (obs_dev) KF_CUDA(cudaMemsetAsync(obs_dev, 0, real * (size_t)B * nobs2, s));
  for (int b0 = 0; b0 < B; b0 += c.chunk) {
    const int nb = (B - b0 < c.chunk) ? (B - b0) : c.chunk;
    for (int j = 0; j < n_snap; ++j) {
      const char* tj = (const char*)trajectory_dev + el * user_env * ((size_t)b0 * n_snap + j);
      c.ops->import_(c, nb, tj, user_env * n_snap, s);
      c.ops->col(c, nb, stage_of(c, 0), kf::kColInit, s);  // fields of the snapshot -> x-space
      c.ops->obs(c, nb, c.acc + 4 * ((size_t)b0 * n_snap + j), 4 * (size_t)n_snap,
                 obs_dev ? (char*)obs_dev + real * (size_t)b0 * nobs2 : nullptr, 1.0 / n_snap, s);
      c.launches += 3;
    }
  }
  c.ops->finalize(c, n_snap, mean_tke_dev, nonfinite_dev, s);
  c.launches += 1;
  KF_CUDA(cudaGetLastError());
  return 0;
}

int kf_physical_fields(kf_handle h, const void* omega_hat_dev, void* out_dev, int nfields, void* stream) {
  if (!h || !omega_hat_dev || !out_dev) return fail("kf_physical_fields: null argument");
  if (nfields < 1 || nfields > 4) return fail("kf_physical_fields: nfields must be 1..4 (u, v, d(omega)/dx, d(omega)/dy)");
  Ctx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t el = c.ops->elem, real = el / 2;
  const int B = c.cfg.B;
  const size_t user_env = (size_t)c.N * (c.MH + 1), phys_env = (size_t)c.N * c.N;
  for (int b0 = 0; b0 < B; b0 += c.chunk) {
    const int nb = (B - b0 < c.chunk) ? (B - b0) : c.chunk;
    c.ops->import_(c, nb, (const char*)omega_hat_dev + el * user_env * b0, user_env, s);
    c.ops->col(c, nb, stage_of(c, 0), kf::kColInit, s);
    c.ops->fields(c, nb, (char*)out_dev + real * phys_env * b0, nfields, phys_env * B, s);
    c.launches += 3;
  }
  KF_CUDA(cudaGetLastError());
  return 0;
}

int kf_step_host(kf_handle h, void* omega_hat_host, const void* actions_host, void* obs_host, void* mean_tke_host,
                 int* nonfinite_host) {
  if (!h || !omega_hat_host) return fail("kf_step_host: null argument");
  Ctx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  const size_t el = c.ops->elem, real = el / 2;
  const size_t B = c.cfg.B;
  const size_t om_bytes = el * B * c.N * (c.MH + 1);
  const size_t nobs2 = (size_t)c.cfg.obs_size * c.cfg.obs_size;
  if (!c.d_omega) {
    if (int rc = dev_alloc(c, &c.d_omega, om_bytes)) return rc;
    if (int rc = dev_alloc(c, &c.d_act, real * B * 4)) return rc;
    if (int rc = dev_alloc(c, &c.d_obs, real * B * nobs2)) return rc;
    if (int rc = dev_alloc(c, &c.d_tke, real * B)) return rc;
    if (int rc = dev_alloc(c, (void**)&c.d_nonf, sizeof(int) * B)) return rc;
  }
  cudaStream_t s = cudaStreamPerThread;
  KF_CUDA(cudaMemcpyAsync(c.d_omega, omega_hat_host, om_bytes, cudaMemcpyHostToDevice, s));
  if (actions_host) KF_CUDA(cudaMemcpyAsync(c.d_act, actions_host, real * B * 4, cudaMemcpyHostToDevice, s));
  if (int rc = drive(c, c.d_omega, actions_host ? c.d_act : nullptr, nullptr, c.cfg.n_outer, c.cfg.n_inner, true,
                     c.d_obs, c.d_tke, c.d_nonf, s))
    return rc;
  KF_CUDA(cudaMemcpyAsync(omega_hat_host, c.d_omega, om_bytes, cudaMemcpyDeviceToHost, s));
  if (obs_host) KF_CUDA(cudaMemcpyAsync(obs_host, c.d_obs, real * B * nobs2, cudaMemcpyDeviceToHost, s));
  if (mean_tke_host) KF_CUDA(cudaMemcpyAsync(mean_tke_host, c.d_tke, real * B, cudaMemcpyDeviceToHost, s));
  if (nonfinite_host) KF_CUDA(cudaMemcpyAsync(nonfinite_host, c.d_nonf, sizeof(int) * B, cudaMemcpyDeviceToHost, s));
  KF_CUDA(cudaStreamSynchronize(s));
  return 0;
}

}  // extern "C"

extern "C" {

// Measurement aid for bench.py: per-launch duration of the column kernel and of the row kernel measured INSIDE the
// real launch sequence of a sub-stage (row, column, row, column, ... on one lane's chunk workspace, one CUDA event
// between any two launches), so each kernel finds the caches in the state its predecessor leaves them in.  The
// workspace contents are advanced as in a real run (each launch is a genuine sub-stage): scratch state only.
int kf_profile_kernels(kf_handle h, int reps, float* ms_col, float* ms_row, void* stream) {
  if (!h || reps < 1 || reps > 4096) return fail("kf_profile_kernels: bad argument");
  Ctx& c = h->c;
  if (!c.have_constants) return fail("kf_set_constants has not been called");
  DeviceGuard guard(c.cfg.device);
  cudaStream_t s = (cudaStream_t)stream;
  std::vector<cudaEvent_t> ev(2 * (size_t)reps + 1);
  for (auto& e : ev) KF_CUDA(cudaEventCreate(&e));
  const int nb = c.chunk;
  auto col = [&](int k) {
    if (c.use_tma) c.ops->col_tma(c, nb, stage_of(c, k), s);
    else c.ops->col(c, nb, stage_of(c, k), kf::kColStage, s);
  };
  c.ops->forcing(c, nb, nullptr, s);
  for (int i = 0; i < 3; ++i) {  // warm-up
    c.ops->row(c, nb, s);
    col(1);
  }
  for (int i = 0; i < reps; ++i) {
    KF_CUDA(cudaEventRecord(ev[2 * i], s));
    c.ops->row(c, nb, s);
    KF_CUDA(cudaEventRecord(ev[2 * i + 1], s));
    col(1 + i % 3);  // sub-stages 1..3 move all 9C (stage 0 does not read h, stage 4 does not write it)
  }
  KF_CUDA(cudaEventRecord(ev[2 * reps], s));
  KF_CUDA(cudaEventSynchronize(ev[2 * reps]));
  double tr = 0, tc = 0;
  for (int i = 0; i < reps; ++i) {
    float a = 0, b = 0;
    KF_CUDA(cudaEventElapsedTime(&a, ev[2 * i], ev[2 * i + 1]));
    KF_CUDA(cudaEventElapsedTime(&b, ev[2 * i + 1], ev[2 * i + 2]));
    tr += a;
    tc += b;
  }
  if (ms_col) *ms_col = (float)(tc / reps);
  if (ms_row) *ms_row = (float)(tr / reps);
  c.launches += 2LL * reps + 7;
  for (auto& e : ev) cudaEventDestroy(e);
  KF_CUDA(cudaGetLastError());
  return 0;
}

// Dense control spectrum for the general control_field of RungeKuttaCrankNicolson.solve (base.py:104-105,
// kolmogorov.py:259-272): control_hat_host [N][M/2+1] complex in the handle's dtype, added to the nonlinear term of
// every environment in every sub-stage after the (optional) 2/3 mask; NULL removes it.  Synchronous.
int kf_set_control_field(kf_handle h, const void* control_hat_host) {
  if (!h) return fail("kf_set_control_field: null handle");
  Ctx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  KF_CUDA(cudaDeviceSynchronize());
  for (Ctx* l : {&c, c.lane2})  // captured graphs hold the old kernel arguments
    if (l) drop_graphs(*l);
  if (!control_hat_host) {
    for (Ctx* l : {&c, c.lane2})
      if (l) l->ctrl_on = false;
    return 0;
  }
  const size_t el = c.ops->elem;
  const int N = c.N, MH = c.MH;
  if (!c.ctrl_main) {
    if (int rc = dev_alloc(c, &c.ctrl_main, el * (size_t)N * MH)) return rc;
    if (int rc = dev_alloc(c, &c.ctrl_nyq, el * (size_t)N)) return rc;
    c.ctrl_owner = true;
  }
  // split the user layout [N][MH+1] into main [N][MH] + nyq [N] on the host side of the copy (strided 2-D copies)
  KF_CUDA(cudaMemcpy2D(c.ctrl_main, el * MH, control_hat_host, el * (MH + 1), el * MH, N, cudaMemcpyHostToDevice));
  KF_CUDA(cudaMemcpy2D(c.ctrl_nyq, el, (const char*)control_hat_host + el * MH, el * (MH + 1), el, N,
                       cudaMemcpyHostToDevice));
  for (Ctx* l : {&c, c.lane2}) {
    if (!l) continue;
    l->ctrl_main = c.ctrl_main;
    l->ctrl_nyq = c.ctrl_nyq;
    l->ctrl_on = true;
  }
  return 0;
}

// PseudoSpectralNavierStokes2D.nonlinear_terms (kolmogorov.py:222-257) of omega_hat_dev [B][N][M/2+1]:
// out_dev [B][N][M/2+1] = dealiasing(rfftn(-(grad_x*u + grad_y*v))) + forcing_term + control, the control being
// sum_m clip(a_m) C_m from actions_dev [B][4] (NULL = none) plus the dense spectrum of kf_set_control_field if one is
// set.  Runs one sub-stage's kernels (x-space fields, row kernel, forward column pass with beta = gamma = mu = 0, which
// leaves omega_hat untouched and the term in the RK register) and exports the register.  Scratches the workspace.
int kf_nonlinear_terms(kf_handle h, const void* omega_hat_dev, const void* actions_dev, void* out_dev, void* stream) {
  if (!h || !omega_hat_dev || !out_dev) return fail("kf_nonlinear_terms: null argument");
  Ctx& c = h->c;
  if (!c.have_constants) return fail("kf_set_constants has not been called");
  DeviceGuard guard(c.cfg.device);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t el = c.ops->elem, real = el / 2;
  const int B = c.cfg.B;
  const size_t user_env = (size_t)c.N * (c.MH + 1);
  Stage st;
  st.beta = 0.0; st.gdt = 0.0; st.munu = 0.0; st.store_h = 1;
  for (int b0 = 0; b0 < B; b0 += c.chunk) {
    const int nb = (B - b0 < c.chunk) ? (B - b0) : c.chunk;
    c.ops->import_(c, nb, (const char*)omega_hat_dev + el * user_env * b0, user_env, s);
    c.ops->forcing(c, nb, actions_dev ? (const char*)actions_dev + real * 4 * b0 : nullptr, s);
    c.ops->col(c, nb, st, kf::kColInit, s);
    c.ops->row(c, nb, s);
    if (c.use_tma) c.ops->col_tma(c, nb, st, s);
    else c.ops->col(c, nb, st, kf::kColStage, s);
    c.ops->export_h(c, nb, (char*)out_dev + el * user_env * b0, user_env, s);
    c.launches += 6;
  }
  KF_CUDA(cudaGetLastError());
  if (c.launch_err != cudaSuccess) {
    const cudaError_t e = c.launch_err;
    c.launch_err = cudaSuccess;
    return fail(std::string("stage kernel launch: ") + cudaGetErrorString(e));
  }
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant of kf_reward_obs (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops, harnesses that forbid
// it).  kf_reward_obs keeps its per-(environment, snapshot) partial sums -- 4 doubles each -- in a buffer it
// cudaMallocs on first use and grows on demand (ensure_acc); here that scratch is a file-scope device array of a
// fixed, modest size that every GPU loading the library holds until exit.
// ---------------------------------------------------------------------------------------------------------------
// KF_NOALLOC_MAX_ENV_SNAPSHOTS (include/kolmo_b200.h) = 16384: B * n_snap <= this; the array is 16384 * 4 * 8 B = 512 KiB per GPU

static __device__ __align__(256) double g_noalloc_acc[4 * KF_NOALLOC_MAX_ENV_SNAPSHOTS];

namespace {

// Device address of g_noalloc_acc on `device` (which must be current).  In HOST code the array's name is only a
// host-side stand-in for the symbol, not a device address: the kernels must be given what cudaGetSymbolAddress
// returns, looked up once per device.
int noalloc_scratch(int device, double** out) {
  constexpr int kMaxDevices = 64;
  static double* table[kMaxDevices] = {};
  if (device < 0 || device >= kMaxDevices) return fail("kf_reward_obs_noalloc: device ordinal out of range");
  if (!table[device]) {
    double* p = nullptr;
    KF_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_acc));
    if (!p) return fail("kf_reward_obs_noalloc: cudaGetSymbolAddress returned a null address");
    table[device] = p;
  }
  *out = table[device];
  return 0;
}

}  // namespace

extern "C" int kf_reward_obs_noalloc(kf_handle h, const void* trajectory_dev, int n_snap, void* obs_dev,
                                     void* mean_tke_dev, int* nonfinite_dev, void* stream) {
  if (!h || !trajectory_dev) return fail("kf_reward_obs_noalloc: null argument");
  if (n_snap < 1) return fail("kf_reward_obs_noalloc: n_snap must be >= 1");
  Ctx& c = h->c;
  const long long need = (long long)c.cfg.B * (long long)n_snap;
  if (need > KF_NOALLOC_MAX_ENV_SNAPSHOTS)
    return fail("kf_reward_obs_noalloc: B * n_snap = " + std::to_string(need) + " exceeds the static scratch of " +
                std::to_string(KF_NOALLOC_MAX_ENV_SNAPSHOTS) + " environment-snapshots (use kf_reward_obs)");
  DeviceGuard guard(c.cfg.device);
  double* acc = nullptr;
  if (int rc = noalloc_scratch(c.cfg.device, &acc)) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const size_t el = c.ops->elem, real = el / 2;
  const int B = c.cfg.B;
  const size_t user_env = (size_t)c.N * (c.MH + 1);
  const int nobs2 = c.cfg.obs_size * c.cfg.obs_size;
  // the array keeps the previous call's sums: the part this call accumulates into must start at zero
  KF_CUDA(cudaMemsetAsync(acc, 0, sizeof(double) * 4 * (size_t)B * n_snap, s));
  if (obs_dev) KF_CUDA(cudaMemsetAsync(obs_dev, 0, real * (size_t)B * nobs2, s));
  for (int b0 = 0; b0 < B; b0 += c.chunk) {
    const int nb = (B - b0 < c.chunk) ? (B - b0) : c.chunk;
    for (int j = 0; j < n_snap; ++j) {
      const char* tj = (const char*)trajectory_dev + el * user_env * ((size_t)b0 * n_snap + j);
      c.ops->import_(c, nb, tj, user_env * n_snap, s);
      c.ops->col(c, nb, stage_of(c, 0), kf::kColInit, s);  // fields of the snapshot -> x-space
      c.ops->obs(c, nb, acc + 4 * ((size_t)b0 * n_snap + j), 4 * (size_t)n_snap,
                 obs_dev ? (char*)obs_dev + real * (size_t)b0 * nobs2 : nullptr, 1.0 / n_snap, s);
      c.launches += 3;
    }
  }
  Ctx view = c;  // shallow copy (owns nothing): the finalize kernel reads its sums through Ctx::acc
  view.acc = acc;
  c.ops->finalize(view, n_snap, mean_tke_dev, nonfinite_dev, s);
  c.launches += 1;
  KF_CUDA(cudaGetLastError());
  return 0;
}
