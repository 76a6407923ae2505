"""NumPy model of the ALGORITHM the CUDA kernels implement (not of the reference).

It mirrors the kernels' data layout and tricks step by step -- integer
wavenumbers, x-space arrays with the ky=0 / ky=M/2 columns packed into one
complex column, two real rows per complex FFT, the mixed-radix Stockham index
formulas -- so that the maths can be checked against the oracle on the CPU,
where no GPU is available.  Used only by tests.
"""
import numpy as np

ALPHA = [0, 0.1496590219993, 0.3704009573644, 0.6222557631345, 0.9582821306748, 1]
BETA = [0, -0.4178904745, -1.192151694643, -1.697784692471, -1.514183444257]
GAMMA = [0.1496590219993, 0.3792103129999, 0.8229550293869, 0.6994504559488, 0.1530572479681]


def stockham_fft(x, radices, sign=-1):
    """Mixed-radix Stockham autosort FFT exactly as the kernels index it.

    Butterfly j of a radix-R pass with Ns = product of earlier radices reads
    in[j + r*L/R], multiplies by exp(sign*2i*pi*r*(j % Ns)/(Ns*R)), does a
    DFT-R and writes out[(j//Ns)*Ns*R + (j % Ns) + r*Ns].
    """
    x = np.asarray(x, dtype=complex)
    L = x.shape[-1]
    assert int(np.prod(radices)) == L
    Ns = 1
    for R in radices:
        out = np.empty_like(x)
        dft = np.exp(sign * 2j * np.pi * np.outer(np.arange(R), np.arange(R)) / R)
        for j in range(L // R):
            k = j % Ns
            v = np.stack([x[..., j + r * (L // R)] * np.exp(sign * 2j * np.pi * r * k / (Ns * R))
                          for r in range(R)], axis=-1)
            v = v @ dft.T
            base = (j // Ns) * Ns * R + k
            for r in range(R):
                out[..., base + r * Ns] = v[..., r]
        x = out
        Ns *= R
    return x


def wavenumbers(N, M):
    p = np.where(np.arange(N) < N // 2, np.arange(N), np.arange(N) - N)[:, None].astype(float)
    q = np.arange(M // 2 + 1)[None, :].astype(float)
    return p, q


def import_state(omega_hat):
    """user layout (N, Mh) -> internal (main (N, M/2), nyquist column (N,))."""
    return omega_hat[:, :-1].copy(), omega_hat[:, -1].copy()


def export_state(main, nyq):
    return np.concatenate([main, nyq[:, None]], axis=1)


def prologue(main, nyq):
    """omega_hat -> 4 x-space fields, each (N, M/2) with column 0 packed as
    Re(col ky=0) + i*Re(col ky=M/2); unnormalised inverse x-FFT, scale 1/(N*M)."""
    N, Mh1 = main.shape
    M = 2 * Mh1
    p, q = wavenumbers(N, M)
    w = export_state(main, nyq)
    k2 = p * p + q * q
    ik2 = np.where(k2 > 0, 1.0 / np.where(k2 > 0, k2, 1.0), 0.0)
    scale = 1.0 / (N * M)
    coefs = [1j * q * ik2 * scale, -1j * p * ik2 * scale, 1j * p * scale * np.ones_like(q),
             1j * q * scale * np.ones_like(p)]
    out = []
    for c in coefs:
        f = c * w
        a, b = f[:, 0], f[:, -1]
        mirror = (-np.arange(N)) % N
        a_h = 0.5 * (a + np.conj(a[mirror]))
        b_h = 0.5 * (b + np.conj(b[mirror]))
        g = f[:, :-1].copy()
        g[:, 0] = a_h + 1j * b_h
        out.append(np.fft.ifft(g, axis=0) * N)  # unnormalised inverse along x
    return out


def row_pass(fields):
    """4 packed x-space fields -> packed x-space advection rows (N, M/2)."""
    uh, vh, gxh, gyh = fields
    N, Mh1 = uh.shape
    M = 2 * Mh1

    def inv_pair(a, b):
        z = np.zeros((N, M), dtype=complex)
        z[:, 0] = a[:, 0].real + 1j * b[:, 0].real
        z[:, Mh1] = a[:, 0].imag + 1j * b[:, 0].imag
        z[:, 1:Mh1] = a[:, 1:] + 1j * b[:, 1:]
        z[:, Mh1 + 1:] = (np.conj(a[:, 1:]) + 1j * np.conj(b[:, 1:]))[:, ::-1]
        return np.fft.ifft(z, axis=1) * M

    z1 = inv_pair(uh, vh)
    z2 = inv_pair(gxh, gyh)
    adv = -(z2.real * z1.real + z2.imag * z1.imag)
    out = np.zeros((N, Mh1), dtype=complex)
    for x0 in range(0, N, 2):
        W = np.fft.fft(adv[x0] + 1j * adv[x0 + 1])
        Wm = np.conj(W[(-np.arange(M)) % M])
        A0 = 0.5 * (W + Wm)
        A1 = -0.5j * (W - Wm)
        out[x0, 1:] = A0[1:Mh1]
        out[x0 + 1, 1:] = A1[1:Mh1]
        out[x0, 0] = W[0].real + 1j * W[Mh1].real
        out[x0 + 1, 0] = W[0].imag + 1j * W[Mh1].imag
    return out


def column_forward(advx):
    """packed x-space advection -> spectral advection (main, nyq)."""
    N = advx.shape[0]
    Wc = np.fft.fft(advx, axis=0)
    Wm = np.conj(Wc[(-np.arange(N)) % N, 0])
    main = Wc.copy()
    main[:, 0] = 0.5 * (Wc[:, 0] + Wm)
    nyq = -0.5j * (Wc[:, 0] - Wm)
    return main, nyq


def rk_step(main, nyq, G, nu, dt):
    """One RK step in the kernels' formulation. G = forcing+control row (Mh,)."""
    N, Mh1 = main.shape
    M = 2 * Mh1
    p, q = wavenumbers(N, M)
    k2 = p * p + q * q
    h = np.zeros((N, Mh1 + 1), dtype=complex)
    w = export_state(main, nyq)
    for s in range(5):
        fields = prologue(w[:, :-1], w[:, -1])
        am, an = column_forward(row_pass(fields))
        nl = export_state(am, an)
        nl[0, :] += G
        h = nl + BETA[s] * h
        mu = 0.5 * dt * (ALPHA[s + 1] - ALPHA[s])
        w = (w + GAMMA[s] * dt * h - mu * nu * k2 * w) / (1.0 + mu * nu * k2)
    return w[:, :-1], w[:, -1]
