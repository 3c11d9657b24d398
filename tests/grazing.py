"""Double-precision flagging of "grazing" rays (BASELINE.json north_star / SURVEY.md §8 D): rays
whose hit/miss, span topology or t hangs on float32 rounding of the reference's own formulas, which
the fast (FMA-contracting) arithmetic policy is allowed to round differently.  Flagged rays are
excluded from the fast-mode comparison and counted; the exact policy needs no exclusions (it is
bit-identical to the reference).

What float32 cannot pin down, per primitive (E = relative rounding error of the terms, ~2 ulp for an
untransformed ray, more after each TransformedObject matrix product):
  * sphere, disc = b^2 - a*c near its `<= eps` gate: hit/miss may flip (narrow band, topology);
  * sphere hit at shallow incidence, k = disc/b^2 small: dt/t ~ E / (2 sqrt(k)) exceeds 1e-5;
  * origin (almost) on a surface: c = |oc|^2 - r^2 (or the plane's numerator) cancels and the near
    root carries ulp(|oc|^2)/|c| relative error;
  * ray (almost) parallel to a plane.
The last three make that primitive's boundaries "uncertain": they matter when the primitive is the
one hit, when its boundary sits next to another boundary (ordering), or near the t >= eps gate.
"""
import numpy as np

EPS = 1e-3
EPS2 = float(np.float32(1e-3) * np.float32(1e-3))
ULP = 6e-8


def flag_grazing(desc, origins, dirs, tol=1e-5):
    """Returns (topology flags [n], uncertain-boundary flags [n, n_prims])."""
    o = np.asarray(origins, np.float64).reshape(-1, 3)
    d = np.asarray(dirs, np.float64).reshape(-1, 3)
    n = len(o)
    flagged = np.zeros(n, bool)
    uncertain = np.zeros((n, desc.n_prims), bool)
    ts, tprim = [], []
    for i in range(desc.n_prims):
        p = desc.prims[i]
        lo, ld = o, d
        nchain = 0
        if p.chain >= 0:
            c = desc.chains[p.chain]
            nchain = c.count
            for k in range(c.count):
                m = np.array(list(desc.xforms[desc.chain_xf[c.first + k]].m), np.float64).reshape(3, 4)
                lo = lo @ m[:, :3].T + m[:, 3]
                ld = ld @ m[:, :3].T
        E = 4 * ULP * (1 + 2 * nchain)
        cond = 2 * E / tol  # |c| / |oc|^2 below this: the near root cannot be held to `tol`
        g = np.array(list(p.p), np.float64)
        if p.type == 0:
            oc = lo - g[:3]
            a = (ld * ld).sum(1)
            b = (oc * ld).sum(1)
            c = (oc * oc).sum(1) - g[3]
            disc = b * b - a * c
            big = np.maximum(b * b, np.abs(a * c)) + 1e-300
            flagged |= np.abs(disc - EPS) <= 16 * E * big  # the sign of disc - eps itself is in doubt
            kmin = (E / (2 * tol)) ** 2 * 4  # E / (2 sqrt(k)) <= tol, with a 2x margin
            uncertain[:, i] = (disc > EPS) & (disc <= kmin * big)
            uncertain[:, i] |= np.abs(c) <= cond * np.maximum((oc * oc).sum(1), g[3])
            s = np.sqrt(np.maximum(disc, 0))
            hit = disc > EPS
            for t in ((-b - s) / a, (-b + s) / a):
                ts.append(np.where(hit, t, np.nan))
                tprim.append(i)
        else:
            div = (ld * g[:3]).sum(1)
            num = -g[3] - (lo * g[:3]).sum(1)
            terms = np.abs(ld * g[:3]).sum(1) + 1e-300
            flagged |= np.abs(np.abs(div) - EPS2) <= 16 * E * terms  # the parallel gate itself
            uncertain[:, i] = np.abs(div) <= (E / tol) * terms  # shallow incidence: div cancels
            uncertain[:, i] |= np.abs(num) <= cond * np.maximum(np.abs(g[3]), np.abs(lo * g[:3]).sum(1))
            with np.errstate(divide="ignore", invalid="ignore"):
                t = num / div
            ts.append(np.where(np.abs(t) < 1e19, t, np.nan))
            tprim.append(i)
    T = np.stack(ts, 1)
    U = np.where(uncertain[:, tprim], 1e-3, 4e-6)  # relative uncertainty of each boundary
    with np.errstate(invalid="ignore"):
        flagged |= (np.abs(T - EPS) <= U * EPS).any(1)  # a boundary sitting on the t >= eps gate
    order = np.argsort(T, 1)  # NaNs last
    S = np.take_along_axis(T, order, 1)
    SU = np.take_along_axis(U, order, 1)
    with np.errstate(invalid="ignore"):
        close = np.abs(S[:, 1:] - S[:, :-1]) <= (SU[:, 1:] + SU[:, :-1]) * np.maximum(np.abs(S[:, 1:]), np.abs(S[:, :-1]))
    flagged |= np.nan_to_num(close, nan=False).astype(bool).any(1)
    return flagged, uncertain
