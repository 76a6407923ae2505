/* C ABI of the 3-D channel-flow environment step, sm_100a kernels in libkolmo_b200.so.
 *
 * Reference (pure Python over JAX; it has no FFI, so these are the calls a ctypes stub behind
 * `ChannelFlowSpectralEnv.step_env` would bind -- INTEGRATION.md shows the stub):
 *   C = hydrogym/jax/envs/channel.py   S = hydrogym/jax/solvers/base.py   U = hydrogym/jax/utils/utils.py
 *
 * Grid 72 x 72 x 72 (C:420-422), Fourier in x and y, Chebyshev collocation in z (fastest axis).  All device arrays
 * are contiguous in the reference's own order: fields (B, 72, 72, 72) real, spectra (B, 72, 72, 72) complex
 * interleaved, actions (B, 24), observations (B, 128).  No torch types, no allocation by the caller beyond the
 * buffers it passes.  Every function returns 0 on success; otherwise kf_last_error() (kolmo_b200.h) has the text.
 */
#ifndef CHANNEL_B200_H
#define CHANNEL_B200_H
#ifdef __cplusplus
extern "C" {
#endif

#define KC_N 72          /* points per direction, C:420-422 */
#define KC_ACTIONS 24    /* wall jets, C:34 */
#define KC_OBS 128       /* 8 x 8 probes x (U, W), C:35-36 */
#define KC_NAB 37        /* |integer wavenumber| 0..36: pressure operators are shared by (+-kx, +-ky) */

typedef struct kc_ctx* kc_handle;

typedef struct {
  int B;                 /* environments advanced together */
  int dtype;             /* KF_F32 (0) or KF_F64 (1): the reference's env_config["dtype"], C:424-425 */
  int device;
  double nu;             /* C:419 */
  double lx, ly;         /* C:416-417 */
  double fx;             /* streamwise body force, C:541 */
  double target_bulk_u;  /* C:550 */
  int k_det;             /* observation plane, C:37 */
  int wss_plane;         /* plane of the one-sided wall-shear difference, C:566 (4) */
  int jets_on;           /* 1: jets written into planes 1..5 (action given); 0: `action=None` of C:237-239 */
} kc_config;

int kc_create(kc_handle* out, const kc_config* cfg);
int kc_destroy(kc_handle h);

/* Operator constants, host pointers, fp64 (converted to the working precision on upload).  They are CONSTANTS of the
 * solve, computed once by the host in fp64 exactly as the reference's constructor does (C:104-148, U:8-50):
 *   z[72]; dz, dzz [72][72] row-major (U:22-50); a_inv [37][37][72][72]: inverse pressure operator of the mode with
 *   |integer kx| = a, |integer ky| = b (C:128-148; k2 is even in both, so the four sign combinations share one);
 *   jets [24][72][72]: zero-mean Gaussian slit pattern of each actuator (C:194-217, linear in the amplitudes);
 *   jets_dx, jets_dy [24][72][72]: Re ifft2(i k fft2(jets)); jets_hat [24][72][72][2]: fft2(jets). */
int kc_set_operators(kc_handle h, const double* z, const double* dz, const double* dzz, const double* a_inv,
                     const double* jets, const double* jets_dx, const double* jets_dy, const double* jets_hat);

/* Physical state (U, V, W) of every environment, C:55-62 (ChannelEnvState). */
int kc_set_state(kc_handle h, const void* U_dev, const void* V_dev, const void* W_dev, void* stream);
int kc_get_state(kc_handle h, void* U_dev, void* V_dev, void* W_dev, void* stream);

/* `step_env`'s compute, C:486-519 + C:531-567: from the stored physical state, `nsteps` RK4 + projection +
 * mass-flux steps (S:119-185, C:243-372) with the jets of `actions_dev` (B, 24; clipped to [lo, hi] on the device),
 * step n of environment b running at t = time[b] + n (C:538); then the physical state is replaced, and
 * obs_dev (B, 128) = (U, W) probes of plane k_det (C:76-85), reward_dev (B) = -wall shear stress (C:88-92, 502).
 * `time_host` is a host array of B doubles. */
int kc_step(kc_handle h, const void* actions_dev, const double* time_host, int nsteps, double dt, double lo, double hi,
            void* obs_dev, void* reward_dev, void* stream);

/* Solver-surface pieces for parity tests and direct use (spectral state = fft2 of the physical state):
 *   kc_to_spectral: C:156-161 of the stored physical state into the internal spectral state;
 *   kc_rk4_steps:   `RungeKutta4.rk4_step` S:126-185 x n on the internal spectral state (actions as in kc_step;
 *                   needs kc_to_spectral or a previous kc_rk4_steps);
 *   kc_rhs:         `rhs` C:302-309 of the internal spectral state at time t into out_dev (3, B, 72, 72, 72) complex;
 *   kc_get_spectral: the internal spectral state, out_dev (3, B, 72, 72, 72) complex;
 *   kc_to_physical: C:163-168 of the internal spectral state into the stored physical state. */
int kc_to_spectral(kc_handle h, void* stream);
int kc_rk4_steps(kc_handle h, const void* actions_dev, const double* time_host, int n, double dt, double lo, double hi,
                 void* stream);
int kc_rhs(kc_handle h, const void* actions_dev, const double* time_host, double lo, double hi, void* out_dev,
           void* stream);
int kc_get_spectral(kc_handle h, void* out_dev, void* stream);
int kc_to_physical(kc_handle h, void* stream);

long long kc_launch_count(kc_handle h);      /* kernels launched so far */
unsigned long long kc_workspace_bytes(kc_handle h);

#ifdef __cplusplus
}
#endif
#endif
