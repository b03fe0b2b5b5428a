"""Initial conditions of the reference drivers."""
import numpy as np


def gaussian_hump(x, y, H0=10.0):
    """H=10, eta=exp(-10x^2-10y^2), u=v=0 (reference src/sw2d-simple/main.cpp:86-94). Returns H, h, hu, hv."""
    H = np.full_like(x, H0)
    eta = np.exp(-10 * (x * x) - 10 * (y * y))
    h = H + eta
    return H, h, np.zeros_like(x), np.zeros_like(x)
