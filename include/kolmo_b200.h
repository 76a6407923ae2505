/* kolmo_b200 -- C ABI of the B200-native Kolmogorov-flow environment step.
 *
 * Drop-in boundary for the hot path of hydrogym/jax (reference citations are
 * relative to the reference repository root):
 *   hydrogym/jax/envs/kolmogorov.py:397-420   KolmogorovFlow._rollout
 *   hydrogym/jax/solvers/base.py:32-59,82-116 RungeKuttaCrankNicolson.RK4_CN / solve
 *   hydrogym/jax/envs/kolmogorov.py:222-257   PseudoSpectralNavierStokes2D.nonlinear_terms
 *   hydrogym/jax/envs/kolmogorov.py:426-470   _trajectory_mean_obs / _avg_tke / _reward
 * The reference has no FFI for this path (it is pure JAX); these entry points are
 * what a ctypes binding inside KolmogorovFlow would call.
 *
 * Conventions: every function returns 0 on success, non-zero on error (message via
 * kf_last_error()); nothing throws across the ABI.  Complex numbers are interleaved
 * (re, im) exactly like numpy / torch complex64 / complex128.  Pointers named *_dev
 * are device pointers on the handle's device; `stream` is a cudaStream_t passed as
 * void* (NULL = legacy default stream).  Calls are asynchronous on `stream` unless
 * stated otherwise.  A handle is not thread-safe.
 */
#ifndef KOLMO_B200_H
#define KOLMO_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct kf_handle_s* kf_handle;

#define KF_F32 0
#define KF_F64 1

typedef struct kf_config {
  int N, M;          /* grid_size (kolmogorov.py:35); square power of two, 64..1024 */
  int B;             /* number of environments (VecEnv batch, env_core.py:609-613) */
  int dtype;         /* KF_F32 / KF_F64 */
  double nu;         /* 1/Re (kolmogorov.py:152-154) */
  double dt;         /* RK step (kolmogorov.py:317, fixed at construction, base.py:54-55) */
  int n_inner;       /* int(save_n // dt)        (base.py:100) */
  int n_outer;       /* int(end // dt) // n_inner (base.py:102-103) */
  int dealias;       /* 0 = reference behaviour (utils.py:72-86 is a no-op); 1 = intended 2/3 mask */
  double drag;       /* must be 0.0: the reference has no drag term */
  int device;        /* CUDA device ordinal */
  int obs_size;      /* probes per side (kolmogorov.py:30) */
  double min_action, max_action; /* clip range (kolmogorov.py:312-313, env_core.py:419-420) */
  int chunk;         /* environments advanced together (L2 blocking); 0 = auto */
  int use_graph;     /* 1 = replay RK steps from a CUDA graph */
} kf_config;

const char* kf_last_error(void);

/* Allocates the workspace (h buffers, x-space fields, twiddles) on cfg->device. */
int kf_create(kf_handle* out, const kf_config* cfg);
int kf_destroy(kf_handle h);

/* Host pointers, in the handle's dtype: forcing_row[M/2+1] complex is row kx=0 of
 * forcing_term() (kolmogorov.py:274-292); control_rows[4][M/2+1] complex are row kx=0
 * of control_term() for unit actions on k1..k4 (kolmogorov.py:259-272, 386-395).
 * Synchronous. */
int kf_set_constants(kf_handle h, const void* forcing_row, const void* control_rows);

/* _rollout: advance omega_hat_dev [B][N][M/2+1] complex IN PLACE by n_outer*n_inner RK
 * steps under actions_dev [B][4] real (NULL = no control, the reset path) and write the
 * n_outer snapshots to trajectory_dev [B][n_outer][N][M/2+1] (NULL = skip). */
int kf_rollout(kf_handle h, void* omega_hat_dev, const void* actions_dev, void* trajectory_dev, void* stream);

/* Same solver at RK-step granularity (tests / benchmarks): n_steps RK steps, no snapshots. */
int kf_rk_steps(kf_handle h, void* omega_hat_dev, const void* actions_dev, int n_steps, void* stream);

/* _trajectory_mean_obs + _avg_tke of trajectory_dev [B][n_snap][N][M/2+1]:
 * obs_dev [B][obs_size^2] real, mean_tke_dev [B] real, nonfinite_dev [B] int32 (each may be NULL). */
int kf_reward_obs(kf_handle h, const void* trajectory_dev, int n_snap, void* obs_dev, void* mean_tke_dev,
                  int* nonfinite_dev, void* stream);

/* kf_reward_obs without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * it).  Same arguments, same results as kf_reward_obs; the per-(environment, snapshot) partial sums live in a static
 * device array inside the library instead of a buffer that kf_reward_obs cudaMallocs and grows on demand.
 * Limit: B * n_snap <= KF_NOALLOC_MAX_ENV_SNAPSHOTS (16384; the array is 512 KiB, held by every GPU that loads the
 * library until the process exits); larger requests fail with a message (use kf_reward_obs).  The call itself calls no
 * cudaMalloc / cudaFree / allocating routine (the handle's workspace was allocated by kf_create).  ALL calls on a
 * device share the one array: it is cleared at the start of each call, and two calls that overlap on the GPU (different
 * streams or handles) race -- serialise them. */
#define KF_NOALLOC_MAX_ENV_SNAPSHOTS 16384
int kf_reward_obs_noalloc(kf_handle h, const void* trajectory_dev, int n_snap, void* obs_dev, void* mean_tke_dev,
                          int* nonfinite_dev, void* stream);

/* Physical-space fields of omega_hat_dev [B][N][M/2+1] (irfftn of the spectral fields of compute_velocity_fft
 * utils.py:53-69 and of the two vorticity gradients kolmogorov.py:240-247): out_dev [nfields][B][N][M] real,
 * fields in the order u, v, d(omega)/dx, d(omega)/dy (nfields = 1..4).  Scratches the workspace. */
int kf_physical_fields(kf_handle h, const void* omega_hat_dev, void* out_dev, int nfields, void* stream);

/* step_env's device work in one call: kf_rollout + kf_reward_obs, the reward/obs
 * reductions running on the snapshots while they are still on chip-side buffers. */
int kf_step(kf_handle h, void* omega_hat_dev, const void* actions_dev, void* trajectory_dev, void* obs_dev,
            void* mean_tke_dev, int* nonfinite_dev, void* stream);

/* kf_step with HOST buffers (pinned or pageable): copies omega_hat and actions in, runs
 * the step, copies omega_hat, obs, mean_tke, nonfinite out; synchronous. */
int kf_step_host(kf_handle h, void* omega_hat_host, const void* actions_host, void* obs_host,
                 void* mean_tke_host, int* nonfinite_host);

/* Solver surface the reference documents next to the env (docs/docs/api/jax/envs/kolmogorov.md):
 *
 * kf_set_control_field: the general control_field of RungeKuttaCrankNicolson.solve (base.py:82-116) --
 * control_hat_host [N][M/2+1] complex (host, handle's dtype) is control_term() of kolmogorov.py:259-272 for an
 * arbitrary (cfx(x,y), cfy(x,y)); it is added to the nonlinear term of every environment in every RK sub-stage
 * (base.py:53, kolmogorov.py:252-257) until cleared with NULL.  The env path does not need it (its control field is
 * x-independent: row kx = 0 only, carried by actions x control_rows).  Synchronous.
 *
 * kf_nonlinear_terms: PseudoSpectralNavierStokes2D.nonlinear_terms (kolmogorov.py:222-257) of omega_hat_dev
 * [B][N][M/2+1] -> out_dev (same shape): advection + forcing + control (actions_dev [B][4] or NULL, plus the dense
 * field if set).  Scratches the workspace. */
int kf_set_control_field(kf_handle h, const void* control_hat_host);
int kf_nonlinear_terms(kf_handle h, const void* omega_hat_dev, const void* actions_dev, void* out_dev, void* stream);

/* linear_terms (kolmogorov.py:209-211) and implicit_timestep (kolmogorov.py:213-220) of omega_hat_dev [B][N][N/2+1]
 * -> out_dev (may alias the input); no handle needed. */
int kf_linear_terms(int dtype, int N, int B, double nu, const void* omega_hat_dev, void* out_dev, void* stream);
int kf_implicit_timestep(int dtype, int N, int B, double nu, double time_step, const void* omega_hat_dev, void* out_dev,
                         void* stream);

/* compute_energy_dissipation (utils.py:145-172) and compute_divergence (utils.py:213-229) from the physical velocity
 * fields u_dev, v_dev [B][N][N] real (kf_physical_fields): dissipation_dev [B] double = sum 2 nu S_ij S_ij / (4 pi^2)
 * with the reference's index-unit finite differences (jnp.gradient: central inside, one-sided at the edges);
 * divergence_dev [B][N][N] real = du/dx + dv/dy.  Either output may be NULL. */
int kf_strain_diagnostics(int dtype, int N, int B, double nu, const void* u_dev, const void* v_dev,
                          double* dissipation_dev, void* divergence_dev, void* stream);

/* Measurement aid: average duration (ms, CUDA events on `stream`) of the column kernel and of the row kernel
 * measured inside `reps` repetitions of the real launch sequence (row, column, row, column, ...) on one lane's chunk
 * workspace.  Scratches the workspace; synchronous. */
int kf_profile_kernels(kf_handle h, int reps, float* ms_col, float* ms_row, void* stream);

/* Number of kernels this handle has launched (or replayed from graphs) so far. */
long long kf_launch_count(kf_handle h);
/* Bytes of device memory owned by the handle. */
size_t kf_workspace_bytes(kf_handle h);
/* Environments per chunk actually used. */
int kf_chunk(kf_handle h);
/* Stream lanes the batch is spread over (1 or 2; each lane owns a chunk-sized workspace). */
int kf_lanes(kf_handle h);

/* Radial kinetic-energy spectrum of omega_hat_dev [B][N][N/2+1] (dtype KF_F32 / KF_F64, no handle needed):
 * spectrum_dev [B][nbins] double, bin = round(|k|), per-mode energy as in compute_energy_mode
 * (hydrogym/jax/utils/utils.py:89-101) with velocities from compute_velocity_fft (utils.py:53-69), the
 * conjugate half of the spectrum counted through weights 2 on 0 < ky < N/2.  sum over bins = mean kinetic
 * energy per unit area (when nbins > N/sqrt(2)). */
int kf_energy_spectrum(int dtype, int N, int B, const void* omega_hat_dev, double* spectrum_dev, int nbins,
                       void* stream);

/* ---------------------------------------------------------------------------------------------
 * Slab-decomposed single grid (BASELINE.json configs[4]: one N x N flow over P GPUs, one process per
 * GPU).  The reference has no multi-device path; the arithmetic is the same RK sub-stage
 * (hydrogym/jax/solvers/base.py:50-57, hydrogym/jax/envs/kolmogorov.py:209-257).  Rank r owns, in
 * spectral space, Q = M/2/P packed ky columns of omega_hat and of the RK register (all kx), and in
 * x-space X = N/P rows (all ky).  One sub-stage is
 *     kf_slab_row(fld_recv -> adv_send);  all-to-all(adv_send -> adv_recv);
 *     kf_slab_col(stage, adv_recv -> fld_send);  all-to-all(fld_send -> fld_recv);
 * the host issues the two all-to-alls (NCCL through torch.distributed, equal splits: block d of a send
 * buffer goes to rank d, block s of a recv buffer comes from rank s).  Exchange buffers are caller-owned
 * device memory of kf_slab_exchange_bytes(h, 0) (fields) / (h, 1) (advection) bytes each.  With P = 1 the
 * all-to-all is the identity (send and recv may be the same buffer).
 * cfg->B, n_inner, n_outer, chunk and use_graph are ignored.  Grid sizes: 256, 1024, 2048, 4096, 8192
 * (fp32 and fp64).
 */
typedef struct kf_slab_s* kf_slab;

int kf_slab_create(kf_slab* out, const kf_config* cfg, int rank, int nranks);
int kf_slab_destroy(kf_slab h);
/* Same host arrays as kf_set_constants (every rank needs row kx = 0 of its own columns only, but takes all). */
int kf_slab_set_constants(kf_slab h, const void* forcing_row, const void* control_rows);
/* G = F + sum_m clip(a_m) C_m for action_dev[4] real (NULL = no control). */
int kf_slab_set_action(kf_slab h, const void* action_dev, void* stream);
/* omega_hat_dev: the FULL user-layout state [N][M/2+1] complex on this device; import reads this rank's
 * columns, export writes this rank's columns (and leaves the others untouched). */
int kf_slab_import(kf_slab h, const void* omega_hat_dev, void* stream);
int kf_slab_export(kf_slab h, void* omega_hat_dev, void* stream);
size_t kf_slab_exchange_bytes(kf_slab h, int which);
/* stage 0..4: RK sub-stage on this rank's columns; stage < 0: only the x-space fields of the stored
 * omega_hat (start of a run / observation), adv_recv_dev may then be NULL. */
int kf_slab_col(kf_slab h, int stage, const void* adv_recv_dev, void* fld_send_dev, void* stream);
int kf_slab_row(kf_slab h, const void* fld_recv_dev, void* adv_send_dev, void* stream);
/* Partial reward / observation sums of this rank's rows from the received fields of a snapshot:
 * acc4_dev[4] += (sum u, sum v, sum u^2, sum v^2) (fp64); obs_dev[obs_size^2] (real, may be NULL) +=
 * obs_w * |velocity| at the probes on this rank's rows.  The host sums both over the ranks
 * (compute_tke utils.py:175-194, _trajectory_mean_obs kolmogorov.py:426-446). */
int kf_slab_obs(kf_slab h, const void* fld_recv_dev, double* acc4_dev, void* obs_dev, double obs_w, void* stream);
/* Peer-to-peer exchange (instead of the two all-to-alls): the library owns one receive buffer per direction;
 * kf_slab_p2p_init allocates them and writes their two cudaIpcMemHandle_t (2 x 64 bytes) to ipc_handles_out (may
 * be NULL); after the host has gathered all ranks' handles (rank order, [P][128] bytes) kf_slab_p2p_attach maps the
 * peers' buffers over NVLink.  Then kf_slab_row / kf_slab_col / kf_slab_obs accept NULL exchange buffers: the row
 * kernel stores its advection blocks and the column pass's transposes store the field blocks DIRECTLY into the
 * owning rank's receive buffer, so the transfer overlaps the kernel that produces the data.  The host must
 * separate the two phases of a sub-stage with a cross-rank barrier on the stream (a 1-element NCCL all-reduce).
 * kf_slab_p2p_attach_local wires P handles living in ONE process (emulated ranks, tests). */
int kf_slab_p2p_init(kf_slab h, void* ipc_handles_out);
int kf_slab_p2p_attach(kf_slab h, const void* all_handles);
int kf_slab_p2p_attach_local(kf_slab h, const kf_slab* peers);
/* Peer-to-peer mode, one process per GPU.  kf_slab_barrier: the cross-rank barrier as one small kernel on `stream`
 * (epoch flags in the peer-mapped buffers, st.release.sys / ld.acquire.sys) -- no NCCL call, capturable.
 * kf_slab_rk_steps: n_steps RK solver steps, each 5 x (row kernel, barrier, column pass, barrier), replayed from ONE
 * CUDA graph per rank (the whole of RK4_CN.time_step_fn, hydrogym/jax/solvers/base.py:50-57, without a host call per
 * sub-stage).  Expects the x-space fields of the current state in the receive buffers (kf_slab_col(stage < 0) +
 * kf_slab_barrier) and leaves them there.  Not for emulated ranks sharing one device (the host interleaves those). */
int kf_slab_barrier(kf_slab h, void* stream);
int kf_slab_rk_steps(kf_slab h, int n_steps, void* stream);
long long kf_slab_launch_count(kf_slab h);
size_t kf_slab_workspace_bytes(kf_slab h);

#ifdef __cplusplus
}
#endif
#endif /* KOLMO_B200_H */
