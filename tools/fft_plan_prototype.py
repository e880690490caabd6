"""numpy prototype of the fused kernel's 8192-point plan (index algebra check, no CUDA):
x[n] complex (L*w + i R*w), N = 2*M, M = 4096 = 16*16*16.
E/O = M-point DFTs of even/odd samples (three radix-16 passes, thread u in [0,256)),
then Z[k] = E[k] + W_N^k O[k], Z[k+M] = E[k] - W_N^k O[k], then the two-for-one split."""
import numpy as np

N = 8192; M = N // 2; R = 16; T = M // R  # 256 threads
rng = np.random.default_rng(0)
x = rng.standard_normal(N) + 1j * rng.standard_normal(N)

def dft16(a):  # a[..., 16] along last axis
    return np.fft.fft(a, axis=-1)

def fft_m(y):
    """M-point DFT with the 3-pass plan; y[n], n = n1*256 + n2*16 + n3; out k = k1 + 16 k2 + 256 k3."""
    u = np.arange(T)
    # pass 1: thread u = n2*16+n3 loads y[n1*256 + u], n1 = 0..15
    a = y[(np.arange(R)[None, :] * T) + u[:, None]]            # [u, n1]
    A = dft16(a)                                               # [u, k1]
    A = A * np.exp(-2j * np.pi * (u[:, None] * np.arange(R)[None, :]) / M)   # W_M^(u*k1)
    s1 = np.zeros(M, complex)
    s1[(np.arange(R)[None, :] * T) + u[:, None]] = A           # addr = k1*256 + u
    # pass 2: thread u' = k1*16 + n3 reads addr k1*256 + n2*16 + n3, n2 = 0..15
    k1 = u // 16; n3 = u % 16
    b = s1[k1[:, None] * T + np.arange(R)[None, :] * 16 + n3[:, None]]      # [u', n2]
    B = dft16(b)                                               # [u', k2]
    B = B * np.exp(-2j * np.pi * (n3[:, None] * np.arange(R)[None, :]) / 256)  # W_256^(n3*k2)
    s2 = np.zeros(M, complex)
    s2[n3[:, None] * T + np.arange(R)[None, :] * 16 + k1[:, None]] = B     # addr = n3*256 + k2*16 + k1
    # pass 3: thread u'' = k2*16 + k1 reads addr n3*256 + u'', n3 = 0..15
    c = s2[np.arange(R)[None, :] * T + u[:, None]]             # [u'', n3]
    Cc = dft16(c)                                              # [u'', k3]
    out = np.zeros(M, complex)
    out[u[:, None] + T * np.arange(R)[None, :]] = Cc           # k = u'' + 256*k3
    return out

E = fft_m(x[0::2]); O = fft_m(x[1::2])
assert np.allclose(E, np.fft.fft(x[0::2])) and np.allclose(O, np.fft.fft(x[1::2]))
k = np.arange(M)
WO = np.exp(-2j * np.pi * k / N) * O
Z = np.concatenate([E + WO, E - WO])
assert np.allclose(Z, np.fft.fft(x))
# split: thread owning bin j (1..M-1) needs Z[j] (own) and Z[N-j] = Zm[M-j] with Zm[m] = E[m] - W^m O[m]
Zp = E + WO; Zm = E - WO
j = np.arange(1, M)
zj = Zp[j]; znj = Zm[M - j]
A = 0.5 * (zj + np.conj(znj)); B = (zj - np.conj(znj)) / 2j
XL = np.fft.fft(x.real); XR = np.fft.fft(x.imag)
assert np.allclose(A, XL[j]) and np.allclose(B, XR[j])
# specials: DC = (Re Z[0], Im Z[0]); Nyquist packed in Z[M] = Zm[0]
assert np.allclose(Zp[0].real, XL[0].real) and np.allclose(Zp[0].imag, XR[0].real)
assert np.allclose(Zm[0].real, XL[M].real) and np.allclose(Zm[0].imag, XR[M].real)
print("plan OK")
