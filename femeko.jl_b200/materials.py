"""Material tables -> mu(H), dmu/dH: host-side mirror of src/magneticProperties.jl (table-sized
work done once; the per-element spline evaluation runs inside the assembly kernel).

    DATA                    magneticProperties.jl:15-42
    gradient(x, y)          src/FEM.jl:66-92
    materialPermeability    magneticProperties.jl:157-188
    loadMaterial / addData  magneticProperties.jl:44-154  (reads HofM.dat, TofM.dat, M.dat)

The reference's data files (Materials/Gd_MFT, Pure_Iron_FEMM) are not redistributed here;
``loadMaterial`` reads them from any folder in the same format, and :func:`synthetic_gd_like` /
:func:`synthetic_iron_like` give analytic stand-ins with the same qualitative shape for the
benchmarks and tests.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import numpy as np

from .fem import Spline1D
from .mesh import mu0


@dataclass
class DATA:
    M: np.ndarray = None        # A/m
    HofM: np.ndarray = None     # A/m
    TofM: np.ndarray = None     # K
    B: np.ndarray = None        # T
    density: float = 0.0        # g/cm3
    mu: np.ndarray = None       # B/H
    dmu: np.ndarray = None      # d mu / dH
    nu: np.ndarray = None       # relative reluctance


def gradient(x, y):
    """Forward / central / backward differences (src/FEM.jl:66-92)."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    if x.size != y.size:
        raise ValueError("Input x and y must have the same length")
    if x.size < 2:
        raise ValueError("Input data must have more than 1 element")
    z = np.zeros(x.size)
    with np.errstate(divide="ignore", invalid="ignore"):
        z[0] = (y[1] - y[0]) / (x[1] - x[0])
        z[-1] = (y[-1] - y[-2]) / (x[-1] - x[-2])
        z[1:-1] = (y[2:] - y[:-2]) / (x[2:] - x[:-2])
    return z


def materialPermeability(data: DATA) -> None:
    """mu = B/H with non-finite -> max(mu); dmu = gradient(H, mu) with non-finite -> min(dmu);
    nu = mu0 H/B with non-finite -> 0 (magneticProperties.jl:157-188)."""
    H = np.asarray(data.HofM, dtype=np.float64)
    B = np.asarray(data.B, dtype=np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        mu = B / H
        idx = ~np.isfinite(mu)
        mu[idx] = 0.0
        mu[idx] = mu.max()
        dmu = gradient(H, mu)
        idx = ~np.isfinite(dmu)
        dmu[idx] = 0.0
        dmu[idx] = dmu.min()
        nu = mu0 * (H / B)
        nu[~np.isfinite(nu)] = 0.0
    data.mu, data.dmu, data.nu = mu, dmu, nu


def _readdlm(path):
    try:
        return np.loadtxt(path, ndmin=2)
    except ValueError:
        return np.loadtxt(path, delimiter=",", ndmin=2)


def loadMaterial(data: DATA, folder: str, name: str, key: str, density: float, T: float):
    """magneticProperties.jl:44-154 for the DATA-struct form: read M(H,T) in emu/g, Oe, K; convert
    to SI; interpolate M(H) at temperature T (bicubic Spline2D == a cubic Spline1D along T for each
    tabulated H, by uniqueness of the tensor-product interpolant); B = mu0 (H + M); permeability."""
    base = os.path.join(folder, name)
    data.HofM = _readdlm(os.path.join(base, "HofM.dat")).ravel()
    data.TofM = _readdlm(os.path.join(base, "TofM.dat")).ravel()
    M_emug = _readdlm(os.path.join(base, "M.dat"))
    if M_emug.shape != (data.HofM.size, data.TofM.size):
        raise ValueError("M.dat must be length(HofM) x length(TofM)")
    if np.isnan(M_emug).any() or np.isnan(data.TofM).any() or np.isnan(data.HofM).any():
        raise ValueError("Error in loadMaterial(). The data holds NaNs")
    data.density = density
    M = M_emug * density * 1e3            # emu/g -> A/m
    data.HofM = data.HofM * 1e-4 / mu0    # Oe    -> A/m
    data.M = np.array([float(Spline1D(data.TofM, M[i], bc="nearest")(T)) for i in range(data.HofM.size)])
    data.B = mu0 * (data.HofM + data.M)
    materialPermeability(data)
    return M_emug


def synthetic_gd_like(n: int = 62, Hmax: float = 4.0e6, chi0: float = 700.0, Ms: float = 1.6e6) -> DATA:
    """Analytic stand-in for the Gd @ 270 K table of NewtonRaphson.jl:41-47 (62 field points from 0):
    a saturating paramagnet M = Ms*L(3 chi0 H / Ms) (Langevin), mu_r from ~chi0 down to O(1)."""
    d = DATA()
    d.HofM = np.concatenate([[0.0], np.geomspace(Hmax / 2.0e5, Hmax, n - 1)])
    x = 3.0 * chi0 * d.HofM / Ms
    with np.errstate(divide="ignore", invalid="ignore"):
        L = 1.0 / np.tanh(x) - 1.0 / x
    L[0] = 0.0
    small = x < 1e-3
    L[small] = x[small] / 3.0
    d.M = Ms * L
    d.B = mu0 * (d.HofM + d.M)
    d.density = 7.9
    materialPermeability(d)
    return d


def synthetic_iron_like(n: int = 152, Hmax: float = 3.0e5, Bs: float = 2.1, mur: float = 4000.0) -> DATA:
    """Analytic soft-iron-like B(H) (Froehlich-Kennelly): B = mu0 H + H / (1/(mu0 (mur-1)) + H/Bs)."""
    d = DATA()
    d.HofM = np.concatenate([[0.0], np.geomspace(1.0, Hmax, n - 1)])
    d.B = mu0 * d.HofM + d.HofM / (1.0 / (mu0 * (mur - 1.0)) + d.HofM / Bs)
    d.density = 7.874
    materialPermeability(d)
    return d
