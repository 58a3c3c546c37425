"""Independent numpy restatement of Optimiser::FilterDynamicsMatrices (src/Optimiser/Optimiser.cpp:340-406), used to
cross-check the C++ oracle and the device kernel.  A is flat or shaped [T][col][row] (column-major matrices)."""
import numpy as np

DEFAULT_FIR = (0.1, 0.15, 0.5, 0.15, 0.1)   # include/Optimiser/Optimiser.h:220


def low_pass(x, a=0.25):
    """FilterIndValLowPass: y_t = (1-a) y_{t-1} + a (x_t + x_{t-1})/2, y_{-1} = x_{-1} = x_0."""
    y = np.empty_like(x)
    yp = xp = x[0]
    for t in range(len(x)):
        y[t] = ((1 - a) * yp) + a * ((x[t] + xp) / 2)
        xp, yp = x[t], y[t]
    return y


def fir(x, taps=DEFAULT_FIR):
    """FilterIndValFIRFilter: causal convolution, taps reaching before t = 0 dropped, accumulation in tap order."""
    y = np.zeros_like(x)
    for t in range(len(x)):
        acc = 0.0
        for j, c in enumerate(taps):
            if t - j >= 0:
                acc += x[t - j] * c
        y[t] = acc
    return y


def filter_dynamics(A, n, method, a=0.25, taps=DEFAULT_FIR):
    """Velocity rows (n..2n-1) of every column filtered along t; position rows untouched."""
    nx = 2 * n
    T = A.size // (nx * nx)
    out = A.reshape(T, nx, nx).copy()
    for row in range(n, nx):
        for col in range(nx):
            sig = out[:, col, row].copy()
            out[:, col, row] = low_pass(sig, a) if method == "low_pass" else fir(sig, taps)
    return out.reshape(A.shape)
