// C-ABI of the slab-decomposed single-grid solver (include/kolmo_b200.h, kf_slab_*): per-rank workspace and
// launch sequencing.  The cross-rank all-to-all between the two layouts is issued by the host between
// kf_slab_col and kf_slab_row (torch.distributed / NCCL, hydrogym_b200/slab.py).  sm_100a only.
#include "../../include/kolmo_b200.h"

#include <cuda_runtime.h>

#include <cmath>
#include <string>
#include <cstdlib>
#include <vector>

#include "slab.cuh"

int kf_set_error(const char* msg);  // kolmo_b200.cu: sets the thread-local message, returns 1

namespace {

int fail(const std::string& msg) { return kf_set_error(msg.c_str()); }

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

struct SlabCtx;

struct SlabOps {
  size_t elem;  // bytes per complex
  int col_groups;  // columns per CTA of the ordinary column kernel
  cudaError_t (*prepare)();
  void (*fill_tw)(double* re_im);  // transposed twiddle table of the plan (fft.cuh), N complex
  void (*fill_tw2)(double* re_im);  // the same for the length-N/2 plan of the row kernel, N/2 complex
  void (*col)(const SlabCtx&, int stage, const void* adv_recv, void* fld_send, cudaStream_t);
  void (*row)(const SlabCtx&, const void* fld_recv, void* adv_send, cudaStream_t);
  void (*obs)(const SlabCtx&, const void* fld_recv, double* acc, void* obs, double obs_w, cudaStream_t);
  void (*import_)(const SlabCtx&, const void* user, cudaStream_t);
  void (*export_)(const SlabCtx&, void* user, cudaStream_t);
  void (*forcing)(const SlabCtx&, const void* action, cudaStream_t);
};

struct SlabCtx {
  kf_config cfg;
  const SlabOps* ops;
  int N, MH, rank, P, Q, X, lgQ;
  void *wk = nullptr, *hk = nullptr, *wnyq = nullptr, *hnyq = nullptr, *G = nullptr, *advk = nullptr, *fk = nullptr;
  void *tw = nullptr, *tw2 = nullptr, *twh = nullptr, *F = nullptr, *C = nullptr;
  size_t bytes = 0;
  long long launches = 0;
  bool have_constants = false;
  // peer-to-peer exchange: library-owned receive buffers and the peers' buffers mapped into this process
  void *own_fld = nullptr, *own_adv = nullptr;
  void* peer_fld[KF_MAX_PEERS] = {};
  void* peer_adv[KF_MAX_PEERS] = {};
  bool ipc_opened[KF_MAX_PEERS] = {};
  bool p2p = false;
  int sms = 148;
  int row_prefetch = 1;  // KF_SLAB_ROW_PREFETCH=0 switches the row kernel's L2 prefetch off (A/B measurements)
  // the one-CTA kernel of the packed column runs beside the main column kernel on its own stream
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  // peer-to-peer mode, P > 1, KF_SLAB_COL_CHUNKS = n > 1 (experiment, off by default): the column pass runs in n
  // pieces; the field transposes of piece i (remote stores over NVLink) run on `side2` while the column kernel of piece
  // i + 1 runs on the main stream.  Measured ms per RK step at 8192^2 fp32 with 1 / 2 / 4 pieces: 6.28 / 6.59 / 6.62
  // on 2 GPUs, 2.33 / 2.35 / 2.87 on 8 GPUs -- the transposes' blocks crowd out the 1024-thread column CTAs.
  static constexpr int kMaxColChunks = 8;
  int col_chunks = 1;
  // peer-to-peer mode with more than one rank (KF_SLAB_DIRECT=0/1 overrides): column kernels store field columns straight into the peers' buffers, the
  // receiver transposes them locally into fx_local before its row / observation kernels (see SlabArgs::direct)
  bool direct = false;
  void* fx_local = nullptr;
  cudaStream_t side2 = nullptr;
  cudaEvent_t ev_chunk[kMaxColChunks] = {}, ev_side2 = nullptr;
  // device-side cross-rank barrier (peer-to-peer mode, one process per GPU): flag words behind own_adv
  // ([0..P) = arrival epochs written by the peers, [KF_MAX_PEERS] = this rank's epoch counter)
  unsigned* flags = nullptr;
  unsigned* peer_flags[KF_MAX_PEERS] = {};
  bool local_peers = false;  // emulated ranks in one process: stream order is the barrier
  // one RK solver step (5 sub-stages) as a CUDA graph, replayed by kf_slab_rk_steps
  cudaGraphExec_t graph = nullptr;
};

// Cross-rank barrier as ONE small kernel per rank (replaces a 1-element NCCL all-reduce and can be captured in a CUDA
// graph): every rank bumps its epoch, writes it into slot `rank` of every peer's flag block over NVLink
// (st.release.sys) and spins until all P slots of its own block have reached the epoch (ld.acquire.sys).  The stream
// orders it after the kernels whose peer stores it publishes, and before the kernels that read what the peers stored.
// Ranks live on DIFFERENT GPUs (never launched for emulated ranks sharing a device), so every spinning kernel is
// resident while it waits.
struct BarrierArgs {
  unsigned* mine;
  unsigned* peers[KF_MAX_PEERS];
  int rank, P;
};

__global__ void slab_barrier_kernel(BarrierArgs a) {
  __shared__ unsigned s_epoch;
  if (threadIdx.x == 0) s_epoch = ++a.mine[KF_MAX_PEERS];
  __syncthreads();
  const unsigned e = s_epoch;
  __threadfence_system();
  if ((int)threadIdx.x < a.P) {
    unsigned* dst = a.peers[threadIdx.x] + a.rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst), "r"(e) : "memory");
    const unsigned* src = a.mine + threadIdx.x;
    unsigned v;
    do {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(src) : "memory");
      if ((int)(v - e) < 0) __nanosleep(64);
    } while ((int)(v - e) < 0);
  }
  __syncthreads();
}

// pieces the column pass is cut into (see SlabImpl::col): > 1 only with remote stores to several GPUs
int col_pieces(const SlabCtx& c, bool remote_stores) {
  const int ncb = c.Q / c.ops->col_groups;
  return (remote_stores && !c.direct && c.P > 1 && c.side2 && c.col_chunks > 1 && ncb % c.col_chunks == 0) ? c.col_chunks : 1;
}

template <typename T, class P, int GROUPS, int RGROUPS, class PH, int HGROUPS, int CMINB = 2, int TWS = 0, bool LEAN = false>
struct SlabImpl {
  static_assert(PH::L * 2 == P::L, "the row kernel's plan is half the grid size");
  using C = kf::Cx<T>;
  using C2 = kf::Cx<typename kf::PairOf<T>::type>;
  static constexpr int N = P::L, MH = N / 2;
  // LEAN (fp64 at 8192 points): split exchange + twiddles from global memory in the one-CTA packed-column kernel and
  // the observation kernel, whose whole-transform buffers would not fit 227 KB otherwise
  static size_t smem(int groups) {
    return LEAN ? sizeof(C) * groups * kf::slab_exchange_elems<N>()
                : sizeof(C2) * groups * kf::slab_exchange_elems<N>() + sizeof(C) * (size_t)N;
  }

  static size_t smem_scalar() { return kf::slab_scalar_smem_bytes<T, N, GROUPS, TWS>(); }
  static size_t smem_obs() { return sizeof(C) * (RGROUPS * kf::slab_exchange_elems<N>() + (LEAN ? 0 : (size_t)N)); }

  static cudaError_t prepare() {
    cudaError_t e;
#define KF_ATTR(k, bytes)                                                                              \
  e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes));              \
  if (e != cudaSuccess) return e;
    KF_ATTR((kf::slab_col_scalar_kernel<T, P, GROUPS, kf::kSlabStage, CMINB, TWS>), smem_scalar())
    KF_ATTR((kf::slab_col_scalar_kernel<T, P, GROUPS, kf::kSlabInit, CMINB, TWS>), smem_scalar())
    KF_ATTR((kf::slab_col_kernel<T, P, 1, kf::kSlabStage, true, 1, LEAN>), smem(1))
    KF_ATTR((kf::slab_col_kernel<T, P, 1, kf::kSlabInit, true, 1, LEAN>), smem(1))
    KF_ATTR((kf::slab_row_half_kernel<T, PH, HGROUPS, 2>), (kf::slab_half_smem_bytes<T, PH::L, HGROUPS>()))
    KF_ATTR((kf::slab_row_obs_kernel<T, P, RGROUPS, LEAN>), smem_obs())
#undef KF_ATTR
    return cudaSuccess;
  }

  static kf::SlabArgs<T> args(const SlabCtx& c) {
    kf::SlabArgs<T> a;
    a.wk = (C*)c.wk; a.hk = (C*)c.hk; a.wnyq = (C*)c.wnyq; a.hnyq = (C*)c.hnyq;
    a.G = (const C*)c.G; a.advk = (const C*)c.advk;
    a.fk = (C*)c.fk;
    a.fx = nullptr; a.fblk = 4 * (size_t)c.X * c.Q;
    for (int d = 0; d < KF_MAX_PEERS; ++d) a.adv_dst[d] = nullptr;
    a.tw = (const C*)c.tw; a.tw2 = (const C*)c.tw2; a.twh = (const C*)c.twh;
    a.beta = 0; a.gdt = 0; a.munu = 0;
    a.scale = (T)(1.0 / ((double)N * (double)N));
    a.dealias = c.cfg.dealias;
    a.q0 = c.rank * c.Q; a.x0 = c.rank * c.X;
    a.Q = c.Q; a.X = c.X; a.lgQ = c.lgQ;
    a.row_prefetch = c.row_prefetch;
    a.cb0 = 0; a.cb1 = c.Q / GROUPS;
    a.direct = 0; a.lgX = 0;
    while ((1 << a.lgX) < c.X) ++a.lgX;
    for (int d = 0; d < KF_MAX_PEERS; ++d) a.fld_dst[d] = nullptr;
    return a;
  }

  // the column / row kernels are persistent: at most one resident wave of CTAs
  template <class K> static int wave(const SlabCtx& c, K kernel, int threads, size_t smem_bytes, int blocks) {
    int occ = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, threads, smem_bytes) != cudaSuccess || occ < 1) occ = 1;
    return blocks < occ * c.sms ? blocks : occ * c.sms;
  }

  // column blocks [cb0, cb1) of the ordinary columns; first = true also runs the packed column 0 (rank 0) beside it
  template <int MODE>
  static void col_kernels(const SlabCtx& c, kf::SlabArgs<T> a, cudaStream_t s, int cb0, int cb1, bool first) {
    auto main_k = kf::slab_col_scalar_kernel<T, P, GROUPS, MODE, CMINB, TWS>;
    const int threads = P::TPF * GROUPS;
    if (first && c.rank == 0) {
      cudaEventRecord(c.ev_fork, s);
      cudaStreamWaitEvent(c.side, c.ev_fork, 0);
      kf::slab_col_kernel<T, P, 1, MODE, true, 1, LEAN><<<1, P::TPF, smem(1), c.side>>>(a);
      cudaEventRecord(c.ev_join, c.side);
    }
    a.cb0 = cb0; a.cb1 = cb1;
    main_k<<<wave(c, main_k, threads, smem_scalar(), cb1 - cb0), threads, smem_scalar(), s>>>(a);
    if (first && c.rank == 0) cudaStreamWaitEvent(s, c.ev_join, 0);
  }
  template <int MODE> static void col_kernels(const SlabCtx& c, const kf::SlabArgs<T>& a, cudaStream_t s) {
    col_kernels<MODE>(c, a, s, 0, c.Q / GROUPS, true);
  }

  template <typename V> static void transpose(const V* in, V* out, kf::TransposeArgs p, int nz, cudaStream_t s) {
    const dim3 grid((p.Cn + 31) / 32, (p.R + 31) / 32, nz);
    p.in_lg = p.out_lg = 0;  // rows per block are N, Q or X: powers of two (kf_slab_create checks nranks)
    while ((1 << p.in_lg) < p.in_rpb) ++p.in_lg;
    while ((1 << p.out_lg) < p.out_rpb) ++p.out_lg;
    kf::transpose_kernel<V><<<grid, 256, 0, s>>>(in, out, p);
  }
  // the same for logical input rows [r0, r1) only, for an input that is ONE block of plain rows and an output given by
  // out_tab (the field transposes): a launch over r1 - r0 rows with the input advanced by r0 rows and every output
  // block by r0 elements.  (A row-offset field in TransposeArgs instead made the kernel itself 36 % slower: 205 -> 279
  // us per 2C launch at 8192^2.)
  template <typename V>
  static void transpose_rows(const V* in, kf::TransposeArgs p, int nz, int r0, int r1, int nblocks, cudaStream_t s) {
    for (int d = 0; d < nblocks; ++d) p.out_tab[d] = (void*)((V*)p.out_tab[d] + r0);
    p.R = r1 - r0;
    transpose<V>(in + (size_t)r0 * p.in_ld, (V*)nullptr, p, nz, s);
  }

  // stage >= 0: received advection blocks -> K layout -> sub-stage `stage` -> fields -> send blocks;
  // stage < 0: fields of the stored omega_hat only
  static void col(const SlabCtx& c, int stage, const void* adv_recv, void* fld_send, cudaStream_t s) {
    kf::SlabArgs<T> a = args(c);
    const int Q = c.Q, X = c.X;
    // field planes [4][Q][N] -> blocks [4][X][Q] per destination rank d (rows x in [d*X, (d+1)*X), plane z at
    // + z*XQ): block d of the local send buffer, or slot `rank` of rank d's receive buffer written over NVLink
    kf::TransposeArgs pf{Q, N, Q, X, 0, (size_t)N, (size_t)Q * N, 4 * (size_t)X * Q, (size_t)Q, (size_t)X * Q, {}};
    for (int d = 0; d < c.P; ++d)
      pf.out_tab[d] = fld_send ? (void*)((C*)fld_send + (size_t)d * 4 * X * Q)
                               : (void*)((C*)c.peer_fld[d] + (size_t)c.rank * 4 * X * Q);
    if (stage >= 0) {
      // adv_recv [P][X][Q] is [N][Q] with plain rows -> advk [Q][N]
      kf::TransposeArgs p{N, Q, N, Q, 0, (size_t)Q, 0, 0, (size_t)N, 0, {}};
      transpose<C>((const C*)(adv_recv ? adv_recv : c.own_adv), (C*)c.advk, p, 1, s);
      a.beta = (T)kBeta[stage];
      a.gdt = (T)(kGamma[stage] * c.cfg.dt);                                        // base.py:55
      a.munu = (T)(0.5 * c.cfg.dt * (kAlpha[stage + 1] - kAlpha[stage]) * c.cfg.nu);  // base.py:54
    }
    if (!fld_send && c.direct) {  // fields leave from the column kernels themselves; no sender-side transposes
      a.direct = 1;
      for (int d = 0; d < c.P; ++d) a.fld_dst[d] = (C*)c.peer_fld[d] + (size_t)c.rank * 4 * X * Q;
      if (stage >= 0) col_kernels<kf::kSlabStage>(c, a, s);
      else col_kernels<kf::kSlabInit>(c, a, s);
      return;
    }
    const int ncb = Q / GROUPS;
    const int nch = col_pieces(c, !fld_send);
    if (nch == 1) {
      if (stage >= 0) col_kernels<kf::kSlabStage>(c, a, s);
      else col_kernels<kf::kSlabInit>(c, a, s);
      transpose<C>((const C*)c.fk, (C*)nullptr, pf, 4, s);
      return;
    }
    // Peer-to-peer exchange over several GPUs: the remote stores of the field transposes are bound by NVLink, the
    // column kernel sends nothing.  Pipeline them: piece i's transposes (stream side2) under piece i + 1's kernel.
    const int per = ncb / nch;
    for (int i = 0; i < nch; ++i) {
      if (stage >= 0) col_kernels<kf::kSlabStage>(c, a, s, i * per, (i + 1) * per, i == 0);
      else col_kernels<kf::kSlabInit>(c, a, s, i * per, (i + 1) * per, i == 0);
      cudaEventRecord(c.ev_chunk[i], s);
      cudaStreamWaitEvent(c.side2, c.ev_chunk[i], 0);
      transpose_rows<C>((const C*)c.fk, pf, 4, i * per * GROUPS, (i + 1) * per * GROUPS, c.P, c.side2);
    }
    cudaEventRecord(c.ev_side2, c.side2);
    cudaStreamWaitEvent(s, c.ev_side2, 0);
  }

  // direct mode: received K-layout pieces [P][4][Q][X] -> X-layout blocks [P][4][X][Q] in fx_local
  static const C* received_fields(const SlabCtx& c, const void* fld_recv, cudaStream_t s) {
    if (fld_recv || !c.direct) return (const C*)(fld_recv ? fld_recv : c.own_fld);
    kf::TransposeArgs p{c.Q, c.X, c.Q, c.X, 0, (size_t)c.X, (size_t)c.Q * c.X, 0, (size_t)c.Q, (size_t)c.X * c.Q, {}};
    transpose<C>((const C*)c.own_fld, (C*)c.fx_local, p, 4 * c.P, s);
    return (const C*)c.fx_local;
  }

  static void row(const SlabCtx& c, const void* fld_recv, void* adv_send, cudaStream_t s) {
    kf::SlabArgs<T> a = args(c);
    a.fx = received_fields(c, fld_recv, s);
    for (int d = 0; d < c.P; ++d)
      a.adv_dst[d] = adv_send ? (C*)adv_send + (size_t)d * c.X * c.Q : (C*)c.peer_adv[d] + (size_t)c.rank * c.X * c.Q;
    auto k = kf::slab_row_half_kernel<T, PH, HGROUPS, 2>;
    const int threads = PH::TPF * HGROUPS;
    const size_t sm = kf::slab_half_smem_bytes<T, PH::L, HGROUPS>();
    k<<<wave(c, k, threads, sm, c.X / HGROUPS), threads, sm, s>>>(a);
  }

  static void obs(const SlabCtx& c, const void* fld_recv, double* acc, void* obs_dev, double obs_w, cudaStream_t s) {
    kf::SlabArgs<T> a = args(c);
    a.fx = received_fields(c, fld_recv, s);
    const int st = N / c.cfg.obs_size;
    kf::slab_row_obs_kernel<T, P, RGROUPS, LEAN><<<c.X / RGROUPS, P::TPF * RGROUPS, smem_obs(), s>>>(
        a, acc, (T*)obs_dev, c.cfg.obs_size, st, st, (T)obs_w);
  }

  // user omega_hat [N][M/2+1] (full array on every rank): this rank's columns <-> wk [Q][N], ky = M/2 <-> wnyq
  static void import_(const SlabCtx& c, const void* user, cudaStream_t s) {
    const C* u = (const C*)user;
    kf::TransposeArgs p{N, c.Q, N, c.Q, 0, (size_t)(MH + 1), 0, 0, (size_t)N, 0, {}};
    transpose<C>(u + (size_t)c.rank * c.Q, (C*)c.wk, p, 1, s);
    if (c.rank == 0) kf::strided_copy_kernel<C><<<(N + 255) / 256, 256, 0, s>>>(u + MH, MH + 1, (C*)c.wnyq, 1, N);
  }
  static void export_(const SlabCtx& c, void* user, cudaStream_t s) {
    C* u = (C*)user;
    kf::TransposeArgs p{c.Q, N, c.Q, N, 0, (size_t)N, 0, 0, (size_t)(MH + 1), 0, {}};
    transpose<C>((const C*)c.wk, u + (size_t)c.rank * c.Q, p, 1, s);
    if (c.rank == 0) kf::strided_copy_kernel<C><<<(N + 255) / 256, 256, 0, s>>>((const C*)c.wnyq, 1, u + MH, MH + 1, N);
  }
  static void forcing(const SlabCtx& c, const void* action, cudaStream_t s) {
    kf::forcing_row_kernel<T><<<dim3((MH + 1 + 127) / 128, 1), 128, 0, s>>>(
        (C*)c.G, (const C*)c.F, (const C*)c.C, (const T*)action, (T)c.cfg.min_action, (T)c.cfg.max_action, MH + 1);
  }

  static SlabOps make() {
    SlabOps o;
    o.elem = sizeof(C);
    o.col_groups = GROUPS;
    o.prepare = prepare; o.fill_tw = kf::fill_twiddles<P>; o.fill_tw2 = kf::fill_twiddles<PH>; o.col = col; o.row = row; o.obs = obs;
    o.import_ = import_; o.export_ = export_; o.forcing = forcing;
    return o;
  }
};

const SlabOps* find_slab_ops(int dtype, int N) {
  using kf::Plan;
  // <dtype, column plan (length N), column groups, obs-row groups, row plan (length N/2), row groups
  //  [, column-kernel min CTAs per SM, twiddles of the column kernel in shared memory + per-group named barriers]>
  static const SlabOps f256 = SlabImpl<float, Plan<256, 16, 16, 16>, 8, 8, Plan<128, 8, 16, 8>, 16>::make();
  static const SlabOps f1024 = SlabImpl<float, Plan<1024, 64, 16, 4, 16>, 2, 2, Plan<512, 32, 8, 8, 8>, 4>::make();
  static const SlabOps f2048 = SlabImpl<float, Plan<2048, 128, 16, 8, 16>, 1, 1, Plan<1024, 64, 16, 4, 16>, 2>::make();
  static const SlabOps f4096 = SlabImpl<float, Plan<4096, 256, 16, 16, 16>, 1, 1, Plan<2048, 128, 16, 8, 16>, 1>::make();
#ifdef KF_SLAB_ROW_E8
  static const SlabOps f8192 =
      SlabImpl<float, Plan<8192, 512, 16, 16, 2, 16>, 1, 1, Plan<4096, 512, 8, 8, 8, 8>, 1>::make();
#else
#ifdef KF_SLAB_COL_2CTA
  // column kernel as two 512-thread CTAs per SM, twiddles through L1 (measured 815 us per sub-stage vs 790 below)
  static const SlabOps f8192 =
      SlabImpl<float, Plan<8192, 512, 16, 16, 2, 16>, 1, 1, Plan<4096, 256, 16, 16, 16>, 1>::make();
#else
  // column kernel: one 1024-thread CTA per SM, two groups on named barriers, twiddles in shared memory
  static const SlabOps f8192 =
      SlabImpl<float, Plan<8192, 512, 16, 16, 2, 16>, 2, 1, Plan<4096, 256, 16, 16, 16>, 1, 1, 1>::make();
#endif
#endif
  static const SlabOps d256 = SlabImpl<double, Plan<256, 16, 16, 16>, 8, 8, Plan<128, 8, 16, 8>, 16>::make();
  static const SlabOps d1024 = SlabImpl<double, Plan<1024, 64, 16, 4, 16>, 2, 2, Plan<512, 32, 8, 8, 8>, 4>::make();
  static const SlabOps d2048 = SlabImpl<double, Plan<2048, 128, 16, 8, 16>, 1, 1, Plan<1024, 64, 16, 4, 16>, 2>::make();
  static const SlabOps d4096 = SlabImpl<double, Plan<4096, 256, 16, 16, 16>, 1, 1, Plan<2048, 128, 16, 8, 16>, 1>::make();
  // fp64 at 8192 points: one transform per CTA everywhere (139 KB exchange buffers), twiddles of the column kernels
  // through L1, LEAN packed-column / observation kernels
  static const SlabOps d8192 =
      SlabImpl<double, Plan<8192, 512, 16, 16, 2, 16>, 1, 1, Plan<4096, 256, 16, 16, 16>, 1, 1, 0, true>::make();
  const bool f = dtype == KF_F32;
  switch (N) {
    case 256: return f ? &f256 : &d256;
    case 1024: return f ? &f1024 : &d1024;
    case 2048: return f ? &f2048 : &d2048;
    case 4096: return f ? &f4096 : &d4096;
    case 8192: return f ? &f8192 : &d8192;
    default: return nullptr;
  }
}

int dev_alloc(SlabCtx& c, void** p, size_t bytes) {
  KF_CUDA(cudaMalloc(p, bytes));
  KF_CUDA(cudaMemset(*p, 0, bytes));
  c.bytes += bytes;
  return 0;
}

void free_all(SlabCtx& c) {
  for (int d = 0; d < KF_MAX_PEERS; ++d) {
    if (c.ipc_opened[d]) {
      cudaIpcCloseMemHandle(c.peer_fld[d]);
      cudaIpcCloseMemHandle(c.peer_adv[d]);
    }
  }
  if (c.graph) cudaGraphExecDestroy(c.graph);
  if (c.side) cudaStreamDestroy(c.side);
  if (c.side2) cudaStreamDestroy(c.side2);
  if (c.ev_side2) cudaEventDestroy(c.ev_side2);
  for (cudaEvent_t e : c.ev_chunk)
    if (e) cudaEventDestroy(e);
  if (c.ev_fork) cudaEventDestroy(c.ev_fork);
  if (c.ev_join) cudaEventDestroy(c.ev_join);
  void* ptrs[] = {c.wk, c.hk, c.wnyq, c.hnyq, c.G, c.advk, c.fk, c.tw, c.tw2, c.twh, c.F, c.C, c.own_fld, c.own_adv,
                  c.fx_local};
  for (void* p : ptrs)
    if (p) cudaFree(p);
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

struct kf_slab_s {
  SlabCtx c;
};

extern "C" {

int kf_slab_create(kf_slab* out, const kf_config* cfg, int rank, int nranks) {
  if (!out || !cfg) return fail("kf_slab_create: null argument");
  *out = nullptr;
  if (cfg->N != cfg->M) return fail("kf_slab_create: only square grids are supported (N == M)");
  if (cfg->dtype != KF_F32 && cfg->dtype != KF_F64) return fail("kf_slab_create: dtype must be KF_F32 or KF_F64");
  const SlabOps* ops = find_slab_ops(cfg->dtype, cfg->N);
  if (!ops) return fail("kf_slab_create: grid size must be 256, 1024, 2048, 4096 or 8192");
  if (nranks < 1 || (nranks & (nranks - 1)) != 0) return fail("kf_slab_create: nranks must be a power of two");
  if (rank < 0 || rank >= nranks) return fail("kf_slab_create: rank out of range");
  if (cfg->N / 2 / nranks < 8) return fail("kf_slab_create: need at least 8 packed columns per rank");
  if (cfg->drag != 0.0) return fail("kf_slab_create: drag must be 0 (the reference has no drag term)");
  if (cfg->obs_size < 1 || cfg->N % cfg->obs_size != 0) return fail("kf_slab_create: obs_size must divide the grid size");
  if (!(cfg->dt > 0.0) || !(cfg->nu >= 0.0)) return fail("kf_slab_create: need dt > 0 and nu >= 0");
  int ndev = 0;
  KF_CUDA(cudaGetDeviceCount(&ndev));
  if (cfg->device < 0 || cfg->device >= ndev) return fail("kf_slab_create: no such CUDA device");
  DeviceGuard guard(cfg->device);
  cudaDeviceProp prop;
  KF_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major != 10) return fail("kf_slab_create: this library is built for sm_100a (B200) only");

  kf_slab h = new kf_slab_s();
  SlabCtx& c = h->c;
  c.cfg = *cfg;
  c.ops = ops;
  c.N = cfg->N;
  c.MH = cfg->N / 2;
  c.rank = rank;
  c.P = nranks;
  c.Q = c.MH / nranks;
  c.X = c.N / nranks;
  c.lgQ = 0;
  while ((1 << c.lgQ) < c.Q) ++c.lgQ;
  const size_t el = ops->elem, qn = (size_t)c.Q * c.N;
  c.sms = prop.multiProcessorCount;
  if (const char* e = getenv("KF_SLAB_ROW_PREFETCH")) c.row_prefetch = atoi(e) != 0;
  int rc = 0;
  {
    cudaError_t e = cudaStreamCreateWithFlags(&c.side, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c.ev_fork, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c.ev_join, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c.side2, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c.ev_side2, cudaEventDisableTiming);
    for (int i = 0; i < SlabCtx::kMaxColChunks && e == cudaSuccess; ++i)
      e = cudaEventCreateWithFlags(&c.ev_chunk[i], cudaEventDisableTiming);
    if (e != cudaSuccess) rc = fail(std::string("side stream: ") + cudaGetErrorString(e));
    if (const char* ev = getenv("KF_SLAB_COL_CHUNKS")) {
      const int v = atoi(ev);
      if (v >= 1 && v <= SlabCtx::kMaxColChunks) c.col_chunks = v;
    }
  }
  cudaError_t pe = ops->prepare();
  if (pe != cudaSuccess) rc = fail(std::string("kernel attribute setup: ") + cudaGetErrorString(pe));
  if (!rc) rc = dev_alloc(c, &c.wk, el * qn);
  if (!rc) rc = dev_alloc(c, &c.hk, el * qn);
  if (!rc) rc = dev_alloc(c, &c.wnyq, el * c.N);
  if (!rc) rc = dev_alloc(c, &c.hnyq, el * c.N);
  if (!rc) rc = dev_alloc(c, &c.G, el * (c.MH + 1));
  if (!rc) rc = dev_alloc(c, &c.advk, el * qn);
  if (!rc) rc = dev_alloc(c, &c.fk, el * 4 * qn);
  if (!rc) rc = dev_alloc(c, &c.tw, el * c.N);
  if (!rc) rc = dev_alloc(c, &c.F, el * (c.MH + 1));
  if (!rc) rc = dev_alloc(c, &c.C, el * 4 * (c.MH + 1));
  if (!rc) rc = dev_alloc(c, &c.tw2, el * c.MH);
  if (!rc) rc = dev_alloc(c, &c.twh, el * c.MH);
  if (!rc) {
    // tables are computed in fp64 and rounded once
    auto upload = [&](void* dst, const std::vector<double>& t64, size_t n) {
      if (cfg->dtype == KF_F32) {
        std::vector<float> t32(t64.begin(), t64.begin() + 2 * n);
        return cudaMemcpy(dst, t32.data(), el * n, cudaMemcpyHostToDevice);
      }
      return cudaMemcpy(dst, t64.data(), el * n, cudaMemcpyHostToDevice);
    };
    std::vector<double> t64(2 * (size_t)c.N);
    ops->fill_tw(t64.data());
    cudaError_t e = upload(c.tw, t64, c.N);
    if (e == cudaSuccess) {
      ops->fill_tw2(t64.data());
      e = upload(c.tw2, t64, c.MH);
    }
    if (e == cudaSuccess) {
      const double kPi = 3.14159265358979323846264338327950288;
      for (int k = 0; k < c.MH; ++k) {  // exp(-2 pi i k / N): pre / post twiddles of the half-length real transforms
        t64[2 * k] = std::cos(2.0 * kPi * k / c.N);
        t64[2 * k + 1] = -std::sin(2.0 * kPi * k / c.N);
      }
      e = upload(c.twh, t64, c.MH);
    }
    if (e != cudaSuccess) rc = fail(std::string("twiddle upload: ") + cudaGetErrorString(e));
  }
  if (rc) {
    free_all(c);
    delete h;
    return rc;
  }
  *out = h;
  return 0;
}

int kf_slab_destroy(kf_slab h) {
  if (!h) return 0;
  DeviceGuard guard(h->c.cfg.device);
  cudaDeviceSynchronize();
  free_all(h->c);
  delete h;
  return 0;
}

int kf_slab_set_constants(kf_slab h, const void* forcing_row, const void* control_rows) {
  if (!h || !forcing_row || !control_rows) return fail("kf_slab_set_constants: null argument");
  SlabCtx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  const size_t row = c.ops->elem * (c.MH + 1);
  KF_CUDA(cudaMemcpy(c.F, forcing_row, row, cudaMemcpyHostToDevice));
  KF_CUDA(cudaMemcpy(c.C, control_rows, 4 * row, cudaMemcpyHostToDevice));
  c.have_constants = true;
  return 0;
}

int kf_slab_set_action(kf_slab h, const void* action_dev, void* stream) {
  if (!h) return fail("kf_slab_set_action: null handle");
  SlabCtx& c = h->c;
  if (!c.have_constants) return fail("kf_slab_set_constants has not been called");
  DeviceGuard guard(c.cfg.device);
  c.ops->forcing(c, action_dev, (cudaStream_t)stream);
  c.launches += 1;
  KF_CUDA(cudaGetLastError());
  return 0;
}

int kf_slab_import(kf_slab h, const void* omega_hat_dev, void* stream) {
  if (!h || !omega_hat_dev) return fail("kf_slab_import: null argument");
  SlabCtx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  c.ops->import_(c, omega_hat_dev, (cudaStream_t)stream);
  c.launches += c.rank == 0 ? 2 : 1;
  KF_CUDA(cudaGetLastError());
  return 0;
}

int kf_slab_export(kf_slab h, void* omega_hat_dev, void* stream) {
  if (!h || !omega_hat_dev) return fail("kf_slab_export: null argument");
  SlabCtx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  c.ops->export_(c, omega_hat_dev, (cudaStream_t)stream);
  c.launches += c.rank == 0 ? 2 : 1;
  KF_CUDA(cudaGetLastError());
  return 0;
}

size_t kf_slab_exchange_bytes(kf_slab h, int which) {
  if (!h) return 0;
  const SlabCtx& c = h->c;
  const size_t tile = c.ops->elem * (size_t)c.N * c.Q;  // P blocks of [X][Q] complex
  return which == 0 ? 4 * tile : tile;
}

int kf_slab_col(kf_slab h, int stage, const void* adv_recv_dev, void* fld_send_dev, void* stream) {
  if (!h) return fail("kf_slab_col: null handle");
  if (!fld_send_dev && !h->c.p2p) return fail("kf_slab_col: no send buffer and no peer-to-peer exchange attached");
  if (stage > 4) return fail("kf_slab_col: stage must be 0..4, or negative for the fields of the stored state");
  if (stage >= 0 && !adv_recv_dev && !h->c.p2p)
    return fail("kf_slab_col: a sub-stage needs the received advection blocks");
  SlabCtx& c = h->c;
  if (stage >= 0 && !c.have_constants) return fail("kf_slab_set_constants has not been called");
  DeviceGuard guard(c.cfg.device);
  c.ops->col(c, stage, adv_recv_dev, fld_send_dev, (cudaStream_t)stream);
  c.launches += (stage >= 0 ? 3 : 2) + (c.rank == 0 ? 1 : 0) + 2 * (col_pieces(c, !fld_send_dev) - 1) -
                ((!fld_send_dev && c.direct) ? 1 : 0);
  KF_CUDA(cudaGetLastError());
  return 0;
}

int kf_slab_row(kf_slab h, const void* fld_recv_dev, void* adv_send_dev, void* stream) {
  if (!h) return fail("kf_slab_row: null handle");
  if ((!fld_recv_dev || !adv_send_dev) && !h->c.p2p)
    return fail("kf_slab_row: null buffer and no peer-to-peer exchange attached");
  SlabCtx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  c.ops->row(c, fld_recv_dev, adv_send_dev, (cudaStream_t)stream);
  c.launches += 1 + ((!fld_recv_dev && c.direct) ? 1 : 0);
  KF_CUDA(cudaGetLastError());
  return 0;
}

int kf_slab_obs(kf_slab h, const void* fld_recv_dev, double* acc4_dev, void* obs_dev, double obs_w, void* stream) {
  if (!h || !acc4_dev || (!fld_recv_dev && !h->c.p2p)) return fail("kf_slab_obs: null argument");
  SlabCtx& c = h->c;
  DeviceGuard guard(c.cfg.device);
  c.ops->obs(c, fld_recv_dev, acc4_dev, obs_dev, obs_w, (cudaStream_t)stream);
  c.launches += 1 + ((!fld_recv_dev && c.direct) ? 1 : 0);
  KF_CUDA(cudaGetLastError());
  return 0;
}

// -- peer-to-peer exchange -------------------------------------------------------------------------------

int kf_slab_p2p_init(kf_slab h, void* ipc_handles_out) {
  if (!h) return fail("kf_slab_p2p_init: null handle");
  SlabCtx& c = h->c;
  if (c.P > KF_MAX_PEERS) return fail("kf_slab_p2p_init: too many ranks");
  DeviceGuard guard(c.cfg.device);
  const size_t tile = c.ops->elem * (size_t)c.N * c.Q;
  if (!c.own_fld) {
    if (int rc = dev_alloc(c, &c.own_fld, 4 * tile)) return rc;
    if (int rc = dev_alloc(c, &c.own_adv, tile + 4096)) return rc;  // + the barrier's flag block (zeroed)
    c.flags = (unsigned*)((char*)c.own_adv + tile);
    c.direct = c.P > 1;  // measured at 8192^2 fp32: 2 GPUs 6.25 -> 5.27 ms per RK step; 1 rank 9.58 -> 9.80 (off there)
    if (const char* e = getenv("KF_SLAB_DIRECT")) c.direct = atoi(e) != 0;
    if (c.direct)
      if (int rc = dev_alloc(c, &c.fx_local, 4 * tile)) return rc;
  }
  if (ipc_handles_out) {
    cudaIpcMemHandle_t* out = (cudaIpcMemHandle_t*)ipc_handles_out;
    KF_CUDA(cudaIpcGetMemHandle(out + 0, c.own_fld));
    KF_CUDA(cudaIpcGetMemHandle(out + 1, c.own_adv));
  }
  return 0;
}

int kf_slab_p2p_attach(kf_slab h, const void* all_handles) {
  if (!h || !all_handles) return fail("kf_slab_p2p_attach: null argument");
  SlabCtx& c = h->c;
  if (!c.own_fld) return fail("kf_slab_p2p_attach: call kf_slab_p2p_init first");
  DeviceGuard guard(c.cfg.device);
  const cudaIpcMemHandle_t* hs = (const cudaIpcMemHandle_t*)all_handles;
  for (int d = 0; d < c.P; ++d) {
    if (d == c.rank) {
      c.peer_fld[d] = c.own_fld;
      c.peer_adv[d] = c.own_adv;
      continue;
    }
    KF_CUDA(cudaIpcOpenMemHandle(&c.peer_fld[d], hs[2 * d + 0], cudaIpcMemLazyEnablePeerAccess));
    KF_CUDA(cudaIpcOpenMemHandle(&c.peer_adv[d], hs[2 * d + 1], cudaIpcMemLazyEnablePeerAccess));
    c.ipc_opened[d] = true;
  }
  const size_t tile = c.ops->elem * (size_t)c.N * c.Q;
  for (int d = 0; d < c.P; ++d) c.peer_flags[d] = (unsigned*)((char*)c.peer_adv[d] + tile);
  c.p2p = true;
  c.local_peers = false;
  return 0;
}

int kf_slab_p2p_attach_local(kf_slab h, const kf_slab* peers) {
  if (!h || !peers) return fail("kf_slab_p2p_attach_local: null argument");
  SlabCtx& c = h->c;
  for (int d = 0; d < c.P; ++d) {
    if (!peers[d] || !peers[d]->c.own_fld) return fail("kf_slab_p2p_attach_local: peer without receive buffers");
    c.peer_fld[d] = peers[d]->c.own_fld;
    c.peer_adv[d] = peers[d]->c.own_adv;
  }
  c.p2p = true;
  c.local_peers = true;
  return 0;
}

namespace {

void enqueue_barrier(SlabCtx& c, cudaStream_t s) {
  if (c.P == 1 || c.local_peers) return;  // one rank, or emulated ranks on one stream: stream order suffices
  BarrierArgs a;
  a.mine = c.flags;
  for (int d = 0; d < KF_MAX_PEERS; ++d) a.peers[d] = d < c.P ? c.peer_flags[d] : nullptr;
  a.rank = c.rank;
  a.P = c.P;
  slab_barrier_kernel<<<1, 32, 0, s>>>(a);
  c.launches += 1;
}

// one RK solver step in peer-to-peer mode: 5 x (row kernel, barrier, column pass, barrier)
void enqueue_rk_step(SlabCtx& c, cudaStream_t s) {
  for (int stage = 0; stage < 5; ++stage) {
    c.ops->row(c, nullptr, nullptr, s);
    enqueue_barrier(c, s);  // every rank's advection blocks have landed
    c.ops->col(c, stage, nullptr, nullptr, s);
    enqueue_barrier(c, s);  // every rank's field blocks have landed
    c.launches += 4 + (c.rank == 0 ? 1 : 0) + 2 * (col_pieces(c, true) - 1);
  }
}

}  // namespace

int kf_slab_barrier(kf_slab h, void* stream) {
  if (!h) return fail("kf_slab_barrier: null handle");
  SlabCtx& c = h->c;
  if (!c.p2p) return fail("kf_slab_barrier: peer-to-peer exchange not attached");
  DeviceGuard guard(c.cfg.device);
  enqueue_barrier(c, (cudaStream_t)stream);
  KF_CUDA(cudaGetLastError());
  return 0;
}

int kf_slab_rk_steps(kf_slab h, int n_steps, void* stream) {
  if (!h) return fail("kf_slab_rk_steps: null handle");
  if (n_steps < 0) return fail("kf_slab_rk_steps: n_steps must be >= 0");
  SlabCtx& c = h->c;
  if (!c.p2p) return fail("kf_slab_rk_steps: needs the peer-to-peer exchange (kf_slab_p2p_init / attach)");
  if (c.local_peers && c.P > 1)
    return fail("kf_slab_rk_steps: emulated ranks in one process must be interleaved by the host (kf_slab_row / kf_slab_col)");
  if (!c.have_constants) return fail("kf_slab_set_constants has not been called");
  DeviceGuard guard(c.cfg.device);
  cudaStream_t s = (cudaStream_t)stream;
  if (!c.graph && n_steps >= 2) {
    // capture on a private stream (the caller's may be the legacy default stream, which cannot capture)
    cudaStream_t cs = nullptr;
    KF_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
    cudaGraph_t g = nullptr;
    const long long before = c.launches;
    cudaError_t ce = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
    if (ce == cudaSuccess) {
      enqueue_rk_step(c, cs);
      ce = cudaStreamEndCapture(cs, &g);
    }
    c.launches = before;
    cudaStreamDestroy(cs);
    if (ce != cudaSuccess) return fail(std::string("kf_slab_rk_steps: graph capture: ") + cudaGetErrorString(ce));
    ce = cudaGraphInstantiate(&c.graph, g, 0);
    cudaGraphDestroy(g);
    if (ce != cudaSuccess) return fail(std::string("kf_slab_rk_steps: cudaGraphInstantiate: ") + cudaGetErrorString(ce));
  }
  const long long per_step = 5LL * (4 + (c.rank == 0 ? 1 : 0) + (c.P > 1 ? 2 : 0) + 2 * (col_pieces(c, true) - 1));
  for (int i = 0; i < n_steps; ++i) {
    if (c.graph) {
      KF_CUDA(cudaGraphLaunch(c.graph, s));
      c.launches += per_step;
    } else {
      enqueue_rk_step(c, s);
    }
  }
  KF_CUDA(cudaGetLastError());
  return 0;
}

long long kf_slab_launch_count(kf_slab h) { return h ? h->c.launches : 0; }
size_t kf_slab_workspace_bytes(kf_slab h) { return h ? h->c.bytes : 0; }

}  // extern "C"
