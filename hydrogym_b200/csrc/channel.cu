// 3-D channel-flow environment step: C ABI (include/channel_b200.h) and the per-mode projection kernel.
// Reference: C = hydrogym/jax/envs/channel.py, S = hydrogym/jax/solvers/base.py.
#include "channel.cuh"

#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/channel_b200.h"
#include "../../include/kolmo_b200.h"

int kf_set_error(const char* msg);  // kolmo_b200.cu

namespace kc {

template <typename T> struct PArgs {
  Consts<T> c;
  int B, jets_on;
  const C2<T>* S;      // u* (S:148-152)
  C2<T>* S0;           // new base state
  const C2<T>* M0hat;
  const T* gain;       // [B] at t + dt
  T dt, target;
};

// Pressure projection C:311-372 + walls + constant-mass-flux shift S:166-183, per (kx, ky) mode along z.
// grid (37 * 37, ceil(B / kEnvsPerCta)): the CTA of (a, b) = (|kx|, |ky|) applies the shared inverse to the (up to 4)
// modes (+-a, +-b) of kEnvsPerCta environments at once; 4 x 72 threads: thread (i, slot) owns component i of one mode
// in every environment of the CTA.  The matrices are streamed through L1 (each element is used by the 4 slots), the
// operand vectors are broadcast from shared memory.
constexpr int kEnvsPerCta = 4;   // (1 when the batch is smaller)
template <typename T, int E> __global__ void __launch_bounds__(4 * N) project_kernel(const PArgs<T> a) {
  __shared__ C2<T> vec[E][4][N];
  __shared__ double red[E][N];
  const int ia = blockIdx.x / NAB, ib = blockIdx.x % NAB;
  const int tid = threadIdx.x, i = tid % N, slot = tid / N;
  const T* ainv = a.c.ainv_t + (size_t)blockIdx.x * NN;
  const T* dzt = a.c.dzt;
  const int sx = slot >> 1, sy = slot & 1;
  const bool live = !(sx && (ia == 0 || ia == N / 2)) && !(sy && (ib == 0 || ib == N / 2));
  const int kxi = sx ? N - ia : ia, kyi = sy ? N - ib : ib;
  const size_t fs = (size_t)a.B * V;
  const int b0 = blockIdx.y * E, nb = min(E, a.B - b0);
  const size_t g0 = (size_t)b0 * V + ((size_t)kxi * N + kyi) * N + i;
  // this CTA's inverse is needed by the second product only: on its way into L2 / L1 from the start
  for (int l = tid; l * 128 < NN * (int)sizeof(T); l += 4 * N) prefetch_l2(reinterpret_cast<const char*>(ainv) + l * 128);
  grid_dependency_wait();
  C2<T> u[E], v[E], w[E];
#pragma unroll
  for (int e = 0; e < E; ++e) {
    u[e] = mk<T>(0, 0); v[e] = u[e]; w[e] = u[e];
    if (live && e < nb) { u[e] = a.S[g0 + (size_t)e * V]; v[e] = a.S[fs + g0 + (size_t)e * V]; w[e] = a.S[2 * fs + g0 + (size_t)e * V]; }
  }
  auto put = [&](const C2<T> (&x)[E]) {
#pragma unroll
    for (int e = 0; e < E; ++e) vec[e][slot][i] = x[e];
  };
  auto matvec = [&](const T* mt, C2<T> (&out)[E]) {
#pragma unroll
    for (int e = 0; e < E; ++e) out[e] = mk<T>(0, 0);
#pragma unroll (E == 1 ? 72 : 24)
    for (int j = 0; j < N; ++j) {
      const T m = __ldg(mt + j * N + i);
#pragma unroll
      for (int e = 0; e < E; ++e) {
        const C2<T> x = vec[e][slot][j];
        out[e].x = fma(m, x.x, out[e].x);
        out[e].y = fma(m, x.y, out[e].y);
      }
    }
  };
  const T kxv = kxi == N / 2 ? T(0) : a.c.kx[kxi];   // `.real` after the derivative: Nyquist rows drop out
  const T kyv = kyi == N / 2 ? T(0) : a.c.ky[kyi];
  const T idt = T(1) / a.dt;
  const bool wall = i == 0 || i == N - 1;
  C2<T> t[E], p[E];
  put(w);
  __syncthreads();
  matvec(dzt, t);                                                          // dw/dz
#pragma unroll
  for (int e = 0; e < E; ++e) {
    C2<T> rhs = idt * (kxv * mul_i(u[e]) + kyv * mul_i(v[e]) + t[e]);     // C:316-322
    if (wall) rhs = idt * w[e];                                           // C:324-333 Neumann data
    if (ia == 0 && ib == 0 && i == N / 2) rhs = mk<T>(0, 0);              // C:335-337 mean-pressure pin
    t[e] = rhs;
  }
  __syncthreads();
  put(t);
  __syncthreads();
  matvec(ainv, p);                                                         // C:340
  __syncthreads();
  put(p);
  __syncthreads();
  matvec(dzt, t);                                                          // dp/dz
#pragma unroll
  for (int e = 0; e < E; ++e) {
    u[e] = u[e] - a.dt * (kxv * mul_i(p[e]));                              // C:343-344
    v[e] = v[e] - a.dt * (kyv * mul_i(p[e]));
    w[e] = w[e] - a.dt * t[e];                                             // C:346-349
    if (wall) { u[e] = mk<T>(0, 0); v[e] = u[e]; w[e] = u[e]; }            // walls at t + dt (C:351-357)
    if (a.jets_on && i >= 1 && i <= 5 && live && e < nb)
      w[e] = a.gain[b0 + e] * a.M0hat[(size_t)(b0 + e) * NN + kxi * N + kyi];
  }
  if (ia == 0 && ib == 0) {  // S:166-183: only the mean mode sees the bulk-velocity shift
    if (slot == 0) {
#pragma unroll
      for (int e = 0; e < E; ++e) red[e][i] = (double)u[e].x;
    }
    __syncthreads();
    if (slot == 0 && !wall) {
#pragma unroll
      for (int e = 0; e < E; ++e) {
        double s = 0.0;
        for (int j = 0; j < N; ++j) s += red[e][j];
        u[e].x += (a.target - (T)(s / ((double)NN * N))) * T(NN);
      }
    }
  }
#pragma unroll
  for (int e = 0; e < E; ++e)
    if (live && e < nb) { a.S0[g0 + (size_t)e * V] = u[e]; a.S0[fs + g0 + (size_t)e * V] = v[e]; a.S0[2 * fs + g0 + (size_t)e * V] = w[e]; }
}

template <typename T> struct JetArgs {
  Consts<T> c;
  int B;
  const T* actions;   // [B][24] raw
  T lo, hi;
  T* M0; T* M0dx; T* M0dy; C2<T>* M0hat;
};
// Jet pattern of each environment's clipped action (C:194-217 is linear in the amplitudes): grid (ceil(5184/256), B)
template <typename T> __global__ void jets_kernel(const JetArgs<T> a) {
  __shared__ T act[NACT];
  const int b = blockIdx.y;
  if (threadIdx.x < NACT) {
    T v = a.actions[b * NACT + threadIdx.x];
    act[threadIdx.x] = v < a.lo ? a.lo : (v > a.hi ? a.hi : v);   // C:493
  }
  __syncthreads();
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= NN) return;
  T m = 0, mx = 0, my = 0; C2<T> mh = mk<T>(0, 0);
  for (int k = 0; k < NACT; ++k) {
    const T v = act[k];
    m = fma(v, a.c.jets[k * NN + q], m);
    mx = fma(v, a.c.jets_dx[k * NN + q], mx);
    my = fma(v, a.c.jets_dy[k * NN + q], my);
    const C2<T> h = a.c.jets_hat[k * NN + q];
    mh.x = fma(v, h.x, mh.x); mh.y = fma(v, h.y, mh.y);
  }
  const size_t o = (size_t)b * NN + q;
  a.M0[o] = m; a.M0dx[o] = mx; a.M0dy[o] = my; a.M0hat[o] = mh;
}

template <typename T> struct ObsArgs {
  int B, k_det, wss_plane;
  const T* phys;   // [3][B][V]
  T* obs;          // [B][128]
  T* reward;       // [B]
  T nu_over_dz;    // nu / (z[k] - z[0])
  int probe[8];    // C:69-73
};
// Observation probes C:76-85 and wall shear stress C:88-92: grid (B), 256 threads
template <typename T> __global__ void obs_kernel(const ObsArgs<T> a) {
  __shared__ double red[256];
  const int b = blockIdx.x, tid = threadIdx.x;
  const T* U = a.phys + (size_t)b * V;
  const T* W = a.phys + ((size_t)2 * a.B + b) * V;
  if (tid < 128) {
    const int c = tid / 64, xi = a.probe[(tid % 64) / 8], yi = a.probe[tid % 8];
    a.obs[b * 128 + tid] = (c ? W : U)[((size_t)xi * N + yi) * N + a.k_det];
  }
  double s = 0.0;
  for (int q = tid; q < NN; q += 256) s += (double)(a.nu_over_dz * (U[(size_t)q * N + a.wss_plane] - U[(size_t)q * N]));
  red[tid] = s;
  __syncthreads();
  for (int h = 128; h > 0; h >>= 1) { if (tid < h) red[tid] += red[tid + h]; __syncthreads(); }
  if (tid == 0) a.reward[b] = (T)(-red[0] / NN);   // C:502
}

}  // namespace kc

// ------------------------------------------------------------------------------------------------ host side
#define KC_CUDA(x)                                                                                     \
  do {                                                                                                 \
    cudaError_t e_ = (x);                                                                              \
    if (e_ != cudaSuccess) return kf_set_error((std::string(#x) + ": " + cudaGetErrorString(e_)).c_str()); \
  } while (0)

struct kc_ctx {
  kc_config cfg;
  size_t w = 4;   // bytes per real
  std::vector<void*> allocs;
  size_t bytes = 0;
  long long launches = 0;
  bool have_ops = false, have_spectral = false;
  double z[KC_N];
  // constants
  void *tw = nullptr, *kx = nullptr, *ky = nullptr, *dzt = nullptr, *dzzt = nullptr, *ainv_t = nullptr;
  void *jets = nullptr, *jets_dx = nullptr, *jets_dy = nullptr, *jets_hat = nullptr;
  // per-environment arrays
  void *phys = nullptr, *S0 = nullptr, *S = nullptr, *ACC = nullptr, *A = nullptr, *AX = nullptr, *F = nullptr, *BN = nullptr;
  void *M0 = nullptr, *M0dx = nullptr, *M0dy = nullptr, *M0hat = nullptr, *gain = nullptr;
  int gain_steps = 0;
  std::vector<char> gain_host;
  // kc_step replays one CUDA graph of its 700-odd launches (transforms + nsteps solver steps) when the step count
  // and dt repeat; KC_GRAPH=0 launches directly
  bool use_graph = true;
  // programmatic dependent launch between the kernels of a step: measured slower in fp32 (268 -> 290 us per solver
  // step at B = 1: the early-resident waiting CTAs hold one of the two CTA slots of an SM), neutral in fp64: opt-in, KC_PDL=1
  bool pdl = false;
  cudaGraphExec_t graph = nullptr;
  cudaStream_t own = nullptr;          // the graph runs here, ordered after / before the caller's stream by events
  cudaEvent_t ev_in = nullptr, ev_out = nullptr;
  int graph_n = -1;
  double graph_dt = 0.0;
  long long graph_launches = 0;
};

namespace {

int dev_alloc(kc_ctx& c, void** p, size_t n) {
  KC_CUDA(cudaMalloc(p, n));
  KC_CUDA(cudaMemset(*p, 0, n));
  c.allocs.push_back(*p);
  c.bytes += n;
  return 0;
}

// fp64 host array -> working precision on the device
template <typename T> int upload(void* dst, const double* src, size_t n) {
  std::vector<T> tmp(n);
  for (size_t i = 0; i < n; ++i) tmp[i] = (T)src[i];
  KC_CUDA(cudaMemcpy(dst, tmp.data(), n * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}

template <typename T> kc::Consts<T> consts(const kc_ctx& c) {
  kc::Consts<T> k;
  k.tw = (const kc::C2<T>*)c.tw; k.kx = (const T*)c.kx; k.ky = (const T*)c.ky;
  k.dzt = (const T*)c.dzt; k.dzzt = (const T*)c.dzzt; k.ainv_t = (const T*)c.ainv_t;
  k.jets = (const T*)c.jets; k.jets_dx = (const T*)c.jets_dx; k.jets_dy = (const T*)c.jets_dy;
  k.jets_hat = (const kc::C2<T>*)c.jets_hat;
  return k;
}

// C:231-234 (the reference's ramp, including its negative branch beyond action_time); the kernels use -gain (C:237)
double jet_gain(double t) {
  double gain = 0.3;
  if (t < 10.0) gain = gain * (t * 0.1);
  if (t >= 50.0 - 10.0) gain = gain * ((50.0 - t) * 0.1);
  return gain;
}

template <typename T> struct Impl {
  using C = kc::C2<T>;
  static constexpr size_t kTile = (size_t)kc::NN * sizeof(C);
  static constexpr size_t kSmemX = 2 * kTile + kc::NN * sizeof(T) + kc::N * sizeof(C);
  static constexpr size_t kSmemY2 = kTile + kc::N * sizeof(C);
  
  static int init(kc_ctx&) {
    KC_CUDA(cudaFuncSetAttribute(kc::x_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemX));
    KC_CUDA(cudaFuncSetAttribute(kc::y1_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemX));
    KC_CUDA(cudaFuncSetAttribute(kc::y2_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemY2));
    return 0;
  }

  static dim3 tiles(const kc_ctx& c) { return dim3(3, kc::N, c.cfg.B); }

  // every kernel of a step waits (griddepcontrol.wait) before its first dependent access, so consecutive launches may
  // overlap the next kernel's prologue (constants, cp.async of the matrices) with the previous kernel's tail
  template <typename A> static void launch(kc_ctx& c, void (*kernel)(const A), dim3 grid, int threads, size_t smem,
                                           cudaStream_t s, const A& args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = c.pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, args);
    c.launches += 1;
  }

  static kc::XArgs<T> xargs(const kc_ctx& c, int mode, int stage, const T* gain, double dt) {
    kc::XArgs<T> a;
    a.c = consts<T>(c);
    a.mode = mode; a.stage = stage; a.B = c.cfg.B; a.jets_on = c.cfg.jets_on;
    a.BN = (const C*)c.BN; a.S0 = (C*)c.S0; a.S = (C*)c.S; a.ACC = (C*)c.ACC; a.A = (C*)c.A; a.AX = (C*)c.AX;
    a.RHS = nullptr; a.M0hat = (const C*)c.M0hat; a.gain = gain;
    a.nu = (T)c.cfg.nu; a.force = (T)(c.cfg.fx * kc::NN);
    const double as[4] = {0.5, 0.5, 1.0, 0.0}, ws[4] = {1.0, 2.0, 2.0, 1.0};
    a.a_dt = (T)(as[stage & 3] * dt); a.acc_w = (T)ws[stage & 3]; a.dt6 = (T)(dt / 6.0);
    return a;
  }

  static kc::YArgs<T> yargs(const kc_ctx& c, const T* gain, int bc, int mode) {
    kc::YArgs<T> a;
    a.c = consts<T>(c);
    a.B = c.cfg.B; a.bc = bc; a.jets_on = c.cfg.jets_on; a.mode = mode;
    a.A = (const C*)c.A; a.AX = (const C*)c.AX; a.F = (T*)c.F;
    a.M0 = (const T*)c.M0; a.M0dx = (const T*)c.M0dx; a.M0dy = (const T*)c.M0dy; a.gain = gain;
    a.phys = (const T*)c.phys; a.phys_out = (T*)c.phys; a.BN = (C*)c.BN;
    return a;
  }

  static int check(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return kf_set_error((std::string(what) + ": " + cudaGetErrorString(e)).c_str());
    return 0;
  }

  // physical state -> S0 (C:156-161)
  static int to_spectral(kc_ctx& c, cudaStream_t s) {
    launch(c, kc::y2_kernel<T>, tiles(c), kc::THREADS, kSmemY2, s, yargs(c, nullptr, 0, 1));
    launch(c, kc::x_kernel<T>, tiles(c), kc::THREADS, kSmemX, s, xargs(c, kc::kXForward, 0, nullptr, 0.0));
    c.have_spectral = true;
    return check("to_spectral");
  }

  // S0 -> physical state (C:163-168)
  static int to_physical(kc_ctx& c, cudaStream_t s) {
    launch(c, kc::x_kernel<T>, tiles(c), kc::THREADS, kSmemX, s, xargs(c, kc::kXFirst, 0, nullptr, 0.0));
    launch(c, kc::y1_kernel<T>, tiles(c), kc::THREADS, kSmemX, s, yargs(c, nullptr, 0, 0));
    return check("to_physical");
  }

  static int jets(kc_ctx& c, const void* actions, double lo, double hi, cudaStream_t s) {
    kc::JetArgs<T> a;
    a.c = consts<T>(c); a.B = c.cfg.B; a.actions = (const T*)actions; a.lo = (T)lo; a.hi = (T)hi;
    a.M0 = (T*)c.M0; a.M0dx = (T*)c.M0dx; a.M0dy = (T*)c.M0dy; a.M0hat = (C*)c.M0hat;
    kc::jets_kernel<T><<<dim3((kc::NN + 255) / 256, c.cfg.B), 256, 0, s>>>(a);
    c.launches += 1;
    return check("jets_kernel");
  }

  // -gain(t) of every (step, {t, t + dt/2, t + dt}, environment) of a call, one upload
  static int gains(kc_ctx& c, const double* time_host, int n, double dt, cudaStream_t s) {
    const int B = c.cfg.B;
    const size_t need = (size_t)n * 3 * B * sizeof(T);
    if (n > c.gain_steps) return kf_set_error("channel: too many steps in one call (max 4096)");
    KC_CUDA(cudaStreamSynchronize(s));  // the previous call's upload buffer may still be in flight
    c.gain_host.resize(need);
    T* g = (T*)c.gain_host.data();
    const double off[3] = {0.0, 0.5 * dt, dt};
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < 3; ++j)
        for (int b = 0; b < B; ++b) g[((size_t)i * 3 + j) * B + b] = (T)(-jet_gain(time_host[b] + i + off[j]));
    KC_CUDA(cudaMemcpyAsync(c.gain, g, need, cudaMemcpyHostToDevice, s));
    return 0;
  }

  // one right-hand-side evaluation's physical-space half: Y1 (walls at this stage's time) + Y2
  static void y_passes(kc_ctx& c, const T* gain, cudaStream_t s) {
    launch(c, kc::y1_kernel<T>, tiles(c), kc::THREADS, kSmemX, s, yargs(c, gain, 1, 0));
    launch(c, kc::y2_kernel<T>, tiles(c), kc::THREADS, kSmemY2, s, yargs(c, gain, 0, 0));
  }

  // S:126-185 x n on S0 (gains already uploaded)
  static int rk4_steps(kc_ctx& c, int n, double dt, cudaStream_t s) {
    const int B = c.cfg.B;
    for (int i = 0; i < n; ++i) {
      const T* g = (const T*)c.gain + (size_t)i * 3 * B;
      const T* gs[4] = {g, g + B, g + B, g + 2 * B};   // t, t + dt/2, t + dt/2, t + dt
      launch(c, kc::x_kernel<T>, tiles(c), kc::THREADS, kSmemX, s, xargs(c, kc::kXFirst, 0, gs[0], dt));
      for (int st = 0; st < 4; ++st) {
        y_passes(c, gs[st], s);
        launch(c, kc::x_kernel<T>, tiles(c), kc::THREADS, kSmemX, s, xargs(c, st == 3 ? kc::kXLast : kc::kXMid, st, gs[st], dt));
      }
      kc::PArgs<T> p;
      p.c = consts<T>(c); p.B = B; p.jets_on = c.cfg.jets_on; p.S = (const C*)c.S; p.S0 = (C*)c.S0;
      p.M0hat = (const C*)c.M0hat; p.gain = gs[3]; p.dt = (T)dt; p.target = (T)c.cfg.target_bulk_u;
      if (B >= kc::kEnvsPerCta)
        launch(c, kc::project_kernel<T, kc::kEnvsPerCta>, dim3(kc::NAB * kc::NAB, (B + kc::kEnvsPerCta - 1) / kc::kEnvsPerCta), 4 * kc::N, 0, s, p);
      else
        launch(c, kc::project_kernel<T, 1>, dim3(kc::NAB * kc::NAB, B), 4 * kc::N, 0, s, p);
    }
    return check("rk4_steps");
  }

  static int rhs(kc_ctx& c, void* out, cudaStream_t s) {
    const T* g = (const T*)c.gain;
    launch(c, kc::x_kernel<T>, tiles(c), kc::THREADS, kSmemX, s, xargs(c, kc::kXFirst, 0, g, 0.0));
    y_passes(c, g, s);
    kc::XArgs<T> a = xargs(c, kc::kXMid, 0, g, 0.0);
    a.RHS = (C*)out;
    launch(c, kc::x_kernel<T>, tiles(c), kc::THREADS, kSmemX, s, a);
    return check("rhs");
  }

  static int obs(kc_ctx& c, void* obs_dev, void* reward_dev, cudaStream_t s) {
    kc::ObsArgs<T> a;
    a.B = c.cfg.B; a.k_det = c.cfg.k_det; a.wss_plane = c.cfg.wss_plane;
    a.phys = (const T*)c.phys; a.obs = (T*)obs_dev; a.reward = (T*)reward_dev;
    a.nu_over_dz = (T)(c.cfg.nu / (c.z[c.cfg.wss_plane] - c.z[0]));
    for (int i = 0; i < 8; ++i) a.probe[i] = (int)((double)(kc::N - 1) * i / 7.0);   // linspace(0, 71, 8).astype(int), C:69-73
    kc::obs_kernel<T><<<c.cfg.B, 256, 0, s>>>(a);
    c.launches += 1;
    return check("obs_kernel");
  }
};

struct Ops {
  int (*init)(kc_ctx&);
  int (*to_spectral)(kc_ctx&, cudaStream_t);
  int (*to_physical)(kc_ctx&, cudaStream_t);
  int (*jets)(kc_ctx&, const void*, double, double, cudaStream_t);
  int (*gains)(kc_ctx&, const double*, int, double, cudaStream_t);
  int (*rk4_steps)(kc_ctx&, int, double, cudaStream_t);
  int (*rhs)(kc_ctx&, void*, cudaStream_t);
  int (*obs)(kc_ctx&, void*, void*, cudaStream_t);
};
template <typename T> Ops make_ops() {
  return {Impl<T>::init, Impl<T>::to_spectral, Impl<T>::to_physical, Impl<T>::jets, Impl<T>::gains, Impl<T>::rk4_steps,
          Impl<T>::rhs, Impl<T>::obs};
}
const Ops& ops(const kc_ctx& c) {
  static const Ops f32 = make_ops<float>(), f64 = make_ops<double>();
  return c.cfg.dtype == KF_F64 ? f64 : f32;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ C ABI
extern "C" {

int kc_create(kc_handle* out, const kc_config* cfg) {
  if (!out || !cfg) return kf_set_error("kc_create: null argument");
  if (cfg->B < 1 || cfg->B > 4096) return kf_set_error("kc_create: B must be 1..4096");
  if (cfg->dtype != KF_F32 && cfg->dtype != KF_F64) return kf_set_error("kc_create: dtype must be KF_F32 or KF_F64");
  if (cfg->k_det < 0 || cfg->k_det >= KC_N || cfg->wss_plane < 1 || cfg->wss_plane >= KC_N)
    return kf_set_error("kc_create: k_det / wss_plane out of range");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || cfg->device < 0 || cfg->device >= ndev)
    return kf_set_error("kc_create: no such CUDA device (hydrogym_b200 has no CPU fallback)");
  cudaDeviceProp prop;
  KC_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major != 10) return kf_set_error("kc_create: the kernels are built for sm_100a (B200) only");
  KC_CUDA(cudaSetDevice(cfg->device));
  kc_ctx* c = new kc_ctx();
  c->cfg = *cfg;
  c->w = cfg->dtype == KF_F64 ? 8 : 4;
  const size_t w = c->w, B = cfg->B, V = kc::V, NN = kc::NN;
  c->gain_steps = 4096;
  if (const char* e = getenv("KC_GRAPH")) c->use_graph = atoi(e) != 0;
  if (const char* e = getenv("KC_PDL")) c->pdl = atoi(e) != 0;
  if (cudaStreamCreateWithFlags(&c->own, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_in, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_out, cudaEventDisableTiming) != cudaSuccess) {
    kc_destroy(c);
    return kf_set_error("kc_create: stream / event creation failed");
  }
  struct { void** p; size_t n; } want[] = {
      {&c->tw, KC_N * 2 * w}, {&c->kx, KC_N * w}, {&c->ky, KC_N * w}, {&c->dzt, NN * w}, {&c->dzzt, NN * w},
      {&c->ainv_t, (size_t)KC_NAB * KC_NAB * NN * w}, {&c->jets, KC_ACTIONS * NN * w}, {&c->jets_dx, KC_ACTIONS * NN * w},
      {&c->jets_dy, KC_ACTIONS * NN * w}, {&c->jets_hat, KC_ACTIONS * NN * 2 * w},
      {&c->phys, 3 * B * V * w}, {&c->S0, 3 * B * V * 2 * w}, {&c->S, 3 * B * V * 2 * w}, {&c->ACC, 3 * B * V * 2 * w},
      {&c->A, 3 * B * V * 2 * w}, {&c->AX, 3 * B * V * 2 * w}, {&c->F, 12 * B * V * w}, {&c->BN, 3 * B * V * 2 * w},
      {&c->M0, B * NN * w}, {&c->M0dx, B * NN * w}, {&c->M0dy, B * NN * w}, {&c->M0hat, B * NN * 2 * w},
      {&c->gain, (size_t)c->gain_steps * 3 * B * w}};
  for (auto& q : want)
    if (int rc = dev_alloc(*c, q.p, q.n)) { kc_destroy(c); return rc; }
  if (int rc = ops(*c).init(*c)) { kc_destroy(c); return rc; }
  // wavenumbers as the reference builds them (C:114-118): fftfreq(n, d = L/n) * 2.0 * pi
  double kx[KC_N], ky[KC_N], tw[2 * KC_N];
  const double vx = 1.0 / (KC_N * (cfg->lx / KC_N)), vy = 1.0 / (KC_N * (cfg->ly / KC_N));
  for (int i = 0; i < KC_N; ++i) {
    const int p = i < KC_N / 2 ? i : i - KC_N;
    kx[i] = p * vx * 2.0 * M_PI;
    ky[i] = p * vy * 2.0 * M_PI;
    tw[2 * i] = std::cos(2.0 * M_PI * i / KC_N);
    tw[2 * i + 1] = -std::sin(2.0 * M_PI * i / KC_N);
  }
  int rc = 0;
  if (cfg->dtype == KF_F64) {
    rc = upload<double>(c->kx, kx, KC_N) || upload<double>(c->ky, ky, KC_N) || upload<double>(c->tw, tw, 2 * KC_N);
  } else {
    rc = upload<float>(c->kx, kx, KC_N) || upload<float>(c->ky, ky, KC_N) || upload<float>(c->tw, tw, 2 * KC_N);
  }
  if (rc) { kc_destroy(c); return rc; }
  *out = c;
  return 0;
}

int kc_destroy(kc_handle h) {
  if (!h) return 0;
  cudaSetDevice(h->cfg.device);
  cudaDeviceSynchronize();
  if (h->graph) cudaGraphExecDestroy(h->graph);
  if (h->own) cudaStreamDestroy(h->own);
  if (h->ev_in) cudaEventDestroy(h->ev_in);
  if (h->ev_out) cudaEventDestroy(h->ev_out);
  for (void* p : h->allocs) cudaFree(p);
  delete h;
  return 0;
}

int kc_set_operators(kc_handle h, const double* z, const double* dz, const double* dzz, const double* a_inv,
                     const double* jets, const double* jets_dx, const double* jets_dy, const double* jets_hat) {
  if (!h || !z || !dz || !dzz || !a_inv || !jets || !jets_dx || !jets_dy || !jets_hat)
    return kf_set_error("kc_set_operators: null argument");
  kc_ctx& c = *h;
  KC_CUDA(cudaSetDevice(c.cfg.device));
  std::memcpy(c.z, z, sizeof(c.z));
  const size_t NN = kc::NN;
  // matrices are stored transposed ([j][i]) so that threads with consecutive i read consecutive words
  std::vector<double> t(NN), ta((size_t)KC_NAB * KC_NAB * NN);
  auto transpose = [&](const double* m, double* o) {
    for (int i = 0; i < KC_N; ++i)
      for (int j = 0; j < KC_N; ++j) o[j * KC_N + i] = m[i * KC_N + j];
  };
  for (size_t m = 0; m < (size_t)KC_NAB * KC_NAB; ++m) transpose(a_inv + m * NN, ta.data() + m * NN);
  int rc = 0;
  auto up = [&](void* dst, const double* src, size_t n) {
    if (rc) return;
    rc = c.cfg.dtype == KF_F64 ? upload<double>(dst, src, n) : upload<float>(dst, src, n);
  };
  transpose(dz, t.data());
  up(c.dzt, t.data(), NN);
  transpose(dzz, t.data());
  up(c.dzzt, t.data(), NN);
  up(c.ainv_t, ta.data(), ta.size());
  up(c.jets, jets, KC_ACTIONS * NN);
  up(c.jets_dx, jets_dx, KC_ACTIONS * NN);
  up(c.jets_dy, jets_dy, KC_ACTIONS * NN);
  up(c.jets_hat, jets_hat, KC_ACTIONS * NN * 2);
  if (rc) return rc;
  c.have_ops = true;
  return 0;
}

int kc_set_state(kc_handle h, const void* U, const void* V, const void* W, void* stream) {
  if (!h || !U || !V || !W) return kf_set_error("kc_set_state: null argument");
  kc_ctx& c = *h;
  KC_CUDA(cudaSetDevice(c.cfg.device));
  const size_t n = (size_t)c.cfg.B * kc::V * c.w;
  const void* src[3] = {U, V, W};
  for (int f = 0; f < 3; ++f)
    KC_CUDA(cudaMemcpyAsync((char*)c.phys + f * n, src[f], n, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  c.have_spectral = false;
  return 0;
}

int kc_get_state(kc_handle h, void* U, void* V, void* W, void* stream) {
  if (!h || !U || !V || !W) return kf_set_error("kc_get_state: null argument");
  kc_ctx& c = *h;
  KC_CUDA(cudaSetDevice(c.cfg.device));
  const size_t n = (size_t)c.cfg.B * kc::V * c.w;
  void* dst[3] = {U, V, W};
  for (int f = 0; f < 3; ++f)
    KC_CUDA(cudaMemcpyAsync(dst[f], (char*)c.phys + f * n, n, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}

static int ready(kc_handle h, const char* who) {
  if (!h) return kf_set_error((std::string(who) + ": null handle").c_str());
  if (!h->have_ops) return kf_set_error((std::string(who) + ": call kc_set_operators first").c_str());
  cudaError_t e = cudaSetDevice(h->cfg.device);
  if (e != cudaSuccess) return kf_set_error(cudaGetErrorString(e));
  return 0;
}

int kc_to_spectral(kc_handle h, void* stream) {
  if (int rc = ready(h, "kc_to_spectral")) return rc;
  return ops(*h).to_spectral(*h, (cudaStream_t)stream);
}

int kc_to_physical(kc_handle h, void* stream) {
  if (int rc = ready(h, "kc_to_physical")) return rc;
  if (!h->have_spectral) return kf_set_error("kc_to_physical: no spectral state (call kc_to_spectral first)");
  return ops(*h).to_physical(*h, (cudaStream_t)stream);
}

int kc_rk4_steps(kc_handle h, const void* actions_dev, const double* time_host, int n, double dt, double lo, double hi,
                 void* stream) {
  if (int rc = ready(h, "kc_rk4_steps")) return rc;
  if (!time_host || n < 0 || !(dt > 0)) return kf_set_error("kc_rk4_steps: bad argument");
  if (h->cfg.jets_on && !actions_dev) return kf_set_error("kc_rk4_steps: actions required when jets_on");
  if (!h->have_spectral) return kf_set_error("kc_rk4_steps: no spectral state (call kc_to_spectral first)");
  const Ops& o = ops(*h);
  cudaStream_t s = (cudaStream_t)stream;
  if (h->cfg.jets_on)
    if (int rc = o.jets(*h, actions_dev, lo, hi, s)) return rc;
  if (int rc = o.gains(*h, time_host, n, dt, s)) return rc;
  return o.rk4_steps(*h, n, dt, s);
}

int kc_rhs(kc_handle h, const void* actions_dev, const double* time_host, double lo, double hi, void* out_dev,
           void* stream) {
  if (int rc = ready(h, "kc_rhs")) return rc;
  if (!time_host || !out_dev) return kf_set_error("kc_rhs: null argument");
  if (h->cfg.jets_on && !actions_dev) return kf_set_error("kc_rhs: actions required when jets_on");
  if (!h->have_spectral) return kf_set_error("kc_rhs: no spectral state (call kc_to_spectral first)");
  const Ops& o = ops(*h);
  cudaStream_t s = (cudaStream_t)stream;
  if (h->cfg.jets_on)
    if (int rc = o.jets(*h, actions_dev, lo, hi, s)) return rc;
  if (int rc = o.gains(*h, time_host, 1, 0.0, s)) return rc;
  return o.rhs(*h, out_dev, s);
}

int kc_get_spectral(kc_handle h, void* out_dev, void* stream) {
  if (!h || !out_dev) return kf_set_error("kc_get_spectral: null argument");
  if (!h->have_spectral) return kf_set_error("kc_get_spectral: no spectral state");
  KC_CUDA(cudaSetDevice(h->cfg.device));
  KC_CUDA(cudaMemcpyAsync(out_dev, h->S0, (size_t)3 * h->cfg.B * kc::V * 2 * h->w, cudaMemcpyDeviceToDevice,
                          (cudaStream_t)stream));
  return 0;
}

int kc_step(kc_handle h, const void* actions_dev, const double* time_host, int nsteps, double dt, double lo, double hi,
            void* obs_dev, void* reward_dev, void* stream) {
  if (int rc = ready(h, "kc_step")) return rc;
  if (!obs_dev || !reward_dev) return kf_set_error("kc_step: null output");
  if (!time_host || nsteps < 0 || !(dt > 0)) return kf_set_error("kc_step: bad argument");
  if (h->cfg.jets_on && !actions_dev) return kf_set_error("kc_step: actions required when jets_on");
  kc_ctx& c = *h;
  const Ops& o = ops(c);
  cudaStream_t s = (cudaStream_t)stream;
  // per-call inputs first (jet patterns of these actions, gains of these clocks), into fixed device buffers
  if (c.cfg.jets_on)
    if (int rc = o.jets(c, actions_dev, lo, hi, s)) return rc;
  if (int rc = o.gains(c, time_host, nsteps, dt, s)) return rc;
  auto body = [&](cudaStream_t q) -> int {
    if (int rc = o.to_spectral(c, q)) return rc;                                  // C:532
    if (int rc = o.rk4_steps(c, nsteps, dt, q)) return rc;                        // C:534-560
    return o.to_physical(c, q);                                                   // C:562-563
  };
  if (c.use_graph) {
    cudaStream_t q = c.own;   // the caller's stream may be the legacy default stream, which cannot be captured
    KC_CUDA(cudaEventRecord(c.ev_in, s));
    KC_CUDA(cudaStreamWaitEvent(q, c.ev_in, 0));
    if (!c.graph || c.graph_n != nsteps || c.graph_dt != dt) {
      if (c.graph) { cudaGraphExecDestroy(c.graph); c.graph = nullptr; }
      const long long l0 = c.launches;
      cudaGraph_t g = nullptr;
      KC_CUDA(cudaStreamBeginCapture(q, cudaStreamCaptureModeThreadLocal));
      const int rc = body(q);
      cudaError_t e = cudaStreamEndCapture(q, &g);
      if (rc) { if (g) cudaGraphDestroy(g); return rc; }
      if (e != cudaSuccess) return kf_set_error((std::string("kc_step: graph capture: ") + cudaGetErrorString(e)).c_str());
      e = cudaGraphInstantiate(&c.graph, g, 0);
      cudaGraphDestroy(g);
      if (e != cudaSuccess) { c.graph = nullptr; return kf_set_error((std::string("kc_step: graph instantiate: ") + cudaGetErrorString(e)).c_str()); }
      c.graph_n = nsteps; c.graph_dt = dt; c.graph_launches = c.launches - l0;
      c.launches = l0;   // captured, not launched yet
    }
    KC_CUDA(cudaGraphLaunch(c.graph, q));
    KC_CUDA(cudaEventRecord(c.ev_out, q));
    KC_CUDA(cudaStreamWaitEvent(s, c.ev_out, 0));
    c.launches += c.graph_launches;
    c.have_spectral = true;
  } else {
    if (int rc = body(s)) return rc;
  }
  return o.obs(c, obs_dev, reward_dev, s);                                        // C:566, 502, 515
}

long long kc_launch_count(kc_handle h) { return h ? h->launches : 0; }
unsigned long long kc_workspace_bytes(kc_handle h) { return h ? (unsigned long long)h->bytes : 0ull; }

}  // extern "C"
