import numpy as np
from scipy.special import erf
from scipy.optimize import minimize
x = np.linspace(-8, 8, 40001)
ref = x * 0.5 * (1 + erf(x / np.sqrt(2)))
dref = 0.5 * (1 + erf(x / np.sqrt(2))) + x * np.exp(-x * x / 2) / np.sqrt(2 * np.pi)
def gelu_p(c, x):
    x2 = x * x
    p = c[0]
    for k in range(1, len(c)):
        p = p + c[k] * x2 ** k
    return 0.5 * x * (1 + np.tanh(x * p))
def cdf_p(c, x):
    x2 = x * x
    p = c[0]
    for k in range(1, len(c)):
        p = p + c[k] * x2 ** k
    return 0.5 * (1 + np.tanh(x * p))
def err(c):
    return np.max(np.abs(gelu_p(c, x) - ref))
std = [np.sqrt(2 / np.pi), np.sqrt(2 / np.pi) * 0.044715]
print("standard tanh-gelu max abs err", err(std))
for deg in (2, 3):
    c0 = std + [0.0] * (deg - 1)
    best = None
    for p_norm in (8, 32, 128):
        f = lambda c: np.sum(np.abs(gelu_p(c, x) - ref) ** p_norm) ** (1.0 / p_norm)
        r = minimize(f, c0, method="Nelder-Mead", options=dict(xatol=1e-12, fatol=1e-16, maxiter=20000))
        c0 = list(r.x)
    print(deg, [float("%.10g" % v) for v in c0], "max abs err gelu", err(c0),
          "max err cdf-based grad", np.max(np.abs(cdf_p(c0, x) + x * np.exp(-x * x / 2) / np.sqrt(2 * np.pi) - dref)))
