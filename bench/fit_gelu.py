"""Least-squares fit behind csrc/common.cuh::gauss_cdf: Phi(x) ~ 0.5 + 0.5 tanh(x (c0 + c1 x^2 + c2 x^4))."""
import numpy as np
from scipy.optimize import least_squares
from scipy.special import erf
x = np.linspace(1e-4, 6.0, 60001)
phi = 0.5 * (1 + erf(x / np.sqrt(2)))
model = lambda c, x: 0.5 * (1 + np.tanh(x * (c[0] + x * x * (c[1] + x * x * c[2]))))
c = least_squares(lambda c: (model(c, x) - phi) * (1 + x), [0.7978845608, 0.0356774, 0.0], xtol=1e-15, ftol=1e-15).x
print("coefficients", c.tolist())
print("max |Phi err|", np.abs(model(c, x) - phi).max(), " max |gelu err|", np.abs(x * (model(c, x) - phi)).max())
