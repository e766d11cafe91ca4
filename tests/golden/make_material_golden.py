"""Generate tests/golden/gd_270K.npz: the mu(H), dmu/dH table NewtonRaphson.jl:41-63 builds for
Gd at T = 270 K (density 7.9 g/cm3), by running this repo's host-side mirror of
src/magneticProperties.jl:44-188 on the reference's own data files, and cross-checking every
spline piece against scipy's FITPACK wrapper (Dierckx.jl wraps the same Fortran; Spline2D =
RectBivariateSpline, Spline1D = splrep/splev).  The reference data lives only in
/root/reference, which is absent on the GPU box, so the 62-point derived table is committed as a
fixture.  Run in the dev container only:

    python tests/golden/make_material_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import femeko_jl_b200 as fk  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
MAT = "/root/reference/Magnetostatics/Materials"

if __name__ == "__main__":
    from scipy import interpolate

    T = 270.0
    d = fk.DATA()
    M_emug = fk.loadMaterial(d, MAT, "Gd_MFT", "Gd", 7.9, T)
    # independent path: bicubic Spline2D(H, T, M) evaluated at (H_i, T)  (magneticProperties.jl:61-69)
    H_oe = np.loadtxt(os.path.join(MAT, "Gd_MFT", "HofM.dat")).ravel()
    T_k = np.loadtxt(os.path.join(MAT, "Gd_MFT", "TofM.dat")).ravel()
    spl2 = interpolate.RectBivariateSpline(H_oe * 1e-4 / fk.mu0, T_k, M_emug * 7.9 * 1e3, kx=3, ky=3, s=0)
    M2 = spl2(d.HofM, T).ravel()
    assert np.allclose(M2, d.M, rtol=1e-9, atol=1e-9 * np.abs(d.M).max()), np.max(np.abs(M2 - d.M))
    spl = fk.Spline1D(d.HofM, d.mu, bc="error")
    tck = interpolate.splrep(d.HofM, d.mu, k=3, s=0)
    assert np.array_equal(spl.t, tck[0]) and np.allclose(spl.c[:len(d.mu)], tck[1][:len(d.mu)], rtol=1e-9)
    Hs = np.geomspace(d.HofM[1], d.HofM[-1], 257)
    np.savez_compressed(os.path.join(HERE, "gd_270K.npz"), T=T, density=7.9, HofM=d.HofM, M=d.M, B=d.B, mu=d.mu,
                        dmu=d.dmu, nu=d.nu, spl_t=tck[0], spl_c=tck[1], H_samples=Hs,
                        mu_samples=interpolate.splev(Hs, tck), M_spline2d=M2)
    print("gd_270K: points", d.HofM.size, "H range", d.HofM[0], d.HofM[-1], "mu_r range", d.mu.min() / fk.mu0,
          d.mu.max() / fk.mu0)
