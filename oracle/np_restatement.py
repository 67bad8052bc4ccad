"""oracle/np_restatement.py -- TEST INFRASTRUCTURE, never imported by the product.

A second, independent restatement of the reference's hot-path kernels in numpy, written directly from the
Fortran sources (file:line cited per function), vectorised over everything the reference's loop nest leaves
independent.  Its only job is to pin the C++ oracle port (oracle/cz_f90_port.cpp): two restatements written
separately, in different languages and with different loop structures, must agree bit for bit on random input
(tests/test_oracle.py).  numpy's elementwise float32 / float64 arithmetic is IEEE, unfused and evaluated in the
order written, which is what the reference's un-contracted CPU build does per cell.  The Fortran cannot be
compiled in this image (no Fortran compiler); since round 2 the port is ALSO pinned to the reference's own Fortran
executed through a mechanical translation (oracle/f90_to_c.py, tests/test_f90c_pin.py) -- which
caught the one error both restatements shared (the doubled residual of the _maf line routines, read off the same
misunderstood #ifdef).

Arrays follow tests/oracle_util.py: C order [j][i][k] with g = 2 guard layers, i.e. Fortran p(k,i,j).
`idx` = (ist, ied, jst, jed, kst, ked) in Fortran indices (1-based inner numbering)."""
import numpy as np

G = 2


def _rng(idx):
    ist, ied, jst, jed, kst, ked = idx
    o = G - 1                      # Fortran index f -> C index f + g - 1
    return (slice(jst + o, jed + o + 1), slice(ist + o, ied + o + 1), slice(kst + o, ked + o + 1))


def _shift(sl, dj=0, di=0, dk=0):
    return (slice(sl[0].start + dj, sl[0].stop + dj), slice(sl[1].start + di, sl[1].stop + di),
            slice(sl[2].start + dk, sl[2].stop + dk))


def _stencil_sum(p, sl, cf):
    """ss = c1 p(i+1) + c2 p(i-1) + c3 p(j+1) + c4 p(j-1) + c5 p(k+1) + c6 p(k-1), left to right
    (cz_solver.f90:335-340, 470-475; cz_blas.f90:622-627)."""
    c = [p.dtype.type(v) for v in cf]
    ss = c[0] * p[_shift(sl, di=1)]
    ss = ss + c[1] * p[_shift(sl, di=-1)]
    ss = ss + c[2] * p[_shift(sl, dj=1)]
    ss = ss + c[3] * p[_shift(sl, dj=-1)]
    ss = ss + c[4] * p[_shift(sl, dk=1)]
    ss = ss + c[5] * p[_shift(sl, dk=-1)]
    return ss


def jacobi(p, b, idx, cf, omg):
    """cz_solver.f90:284-387: every inner cell from the OLD field, then copied back.  Returns sum dp^2 (double)."""
    t = p.dtype.type
    sl = _rng(idx)
    ss = _stencil_sum(p, sl, cf)
    pp = p[sl]
    dp = ((ss - b[sl]) / t(cf[6]) - pp) * t(omg)          # :341
    res = float(np.sum((dp * dp).astype(np.float64)))
    p[sl] = pp + dp                                        # :342-343, :369-375
    return res


def psor2sma_core(p, b, idx, cf, ofst, color, omg):
    """cz_solver.f90:404-493: cells with mod(i+j+k+kp, 2) == mod(kst, 2)... i.e. k = kst + mod(i+j+kp,2), step 2.
    Cells of one colour only read cells of the other colour, so the sweep vectorises."""
    t = p.dtype.type
    ist, ied, jst, jed, kst, ked = idx
    sl = _rng(idx)
    kp = ofst + color
    jj, ii, kk = np.meshgrid(np.arange(jst, jed + 1), np.arange(ist, ied + 1), np.arange(kst, ked + 1), indexing="ij")
    sel = ((kk - kst) % 2) == ((ii + jj + kp) % 2)        # :466  do k = kst+mod(i+j+kp,2), ked, 2
    ss = _stencil_sum(p, sl, cf)
    pp = p[sl]
    dp = ((ss - b[sl]) / t(cf[6]) - pp) * t(omg)          # :476
    dp = np.where(sel, dp, t(0))
    res = float(np.sum((dp[sel] * dp[sel]).astype(np.float64)))
    p[sl] = np.where(sel, pp + dp, pp)
    return res


def blas_calc_ax(ap, p, idx, cf):
    """cz_blas.f90:579-644."""
    sl = _rng(idx)
    ap[sl] = _stencil_sum(p, sl, cf) - p.dtype.type(cf[6]) * p[sl]


def blas_calc_rk(r, p, b, idx, cf):
    """cz_blas.f90:658-723."""
    sl = _rng(idx)
    r[sl] = b[sl] - (_stencil_sum(p, sl, cf) - p.dtype.type(cf[6]) * p[sl])


def _line_source(x, msk, rhs, idx, jc, ic):
    """d(k) of the lines (jc, ic) (C indices), cz_solver.f90:569-579: [line][k] array, plus the mask rows."""
    t = x.dtype.type
    kst, ked = idx[4], idx[5]
    o = G - 1
    n = ked - kst + 1
    r = t(1.0) / t(6.0)
    ks = slice(kst + o, ked + o + 1)
    mk = msk[jc, ic, ks]
    d = ((x[jc - 1, ic, ks] + x[jc + 1, ic, ks] + x[jc, ic - 1, ks] + x[jc, ic + 1, ks] - rhs[jc, ic, ks]) * r) * mk
    d[:, 0] = (d[:, 0] + x[jc, ic, kst + o - 1] * r) * mk[:, 0]
    d[:, n - 1] = (d[:, n - 1] + x[jc, ic, ked + o + 1] * r) * mk[:, n - 1]
    return d, mk, ks


def _pcr_stages(dsrc, nstages, P):
    """The PCR stages (cz_solver.f90:584-608 / 1346-1363) on work arrays padded by P zeros per side (an index the
    reference clamps to kst-1 / ked+1, or reads from its zeroed pads, finds a zero here).  a and c start from the
    per-line constants (:556-564) and are the same for every line, so they stay 1-D.  Returns padded a, c, d."""
    t = dsrc.dtype.type
    L, n = dsrc.shape
    r = t(1.0) / t(6.0)
    a = np.zeros(n + 2 * P, dtype=dsrc.dtype); c = np.zeros(n + 2 * P, dtype=dsrc.dtype)
    a[P + 1:P + n] = -r                                    # a(kst) = 0, a(kst+1..ked) = -r
    c[P:P + n - 1] = -r                                    # c(kst..ked-1) = -r, c(ked) = 0
    d = np.zeros((L, n + 2 * P), dtype=dsrc.dtype)
    d[:, P:P + n] = dsrc
    k = np.arange(P, P + n)
    for p in range(1, nstages + 1):
        s = 2 ** (p - 1)
        ap = a[k]; cp = c[k]
        e = t(1.0) / (t(1.0) - ap * c[k - s] - cp * a[k + s])
        a1 = -e * ap * a[k - s]
        c1 = -e * cp * c[k + s]
        d1 = e * (d[:, k] - ap * d[:, k - s] - cp * d[:, k + s])
        a[k] = a1; c[k] = c1; d[:, k] = d1
    return a, c, d


def _solve_2x2(dsrc, pn):
    """pn-1 stages, then the 2x2 systems {k, k+s}, s = 2^(pn-1) (cz_solver.f90:612-630)."""
    t = dsrc.dtype.type
    n = dsrc.shape[1]
    s = 2 ** (pn - 1)
    P = 2 * s
    a, c, d = _pcr_stages(dsrc, pn - 1, P)
    k = np.arange(P, P + s)
    cc1 = c[k]; aa2 = a[k + s]
    f1 = d[:, k]; f2 = d[:, k + s]
    jj = t(1.0) / (t(1.0) - aa2 * cc1)
    d1 = np.zeros_like(d)
    d1[:, k] = (f1 - cc1 * f2) * jj
    d1[:, k + s] = (f2 - aa2 * f1) * jj
    return d1[:, P:P + n]


def _solve_4x4(dsrc, pn):
    """pn-2 stages, then the 4x4 systems {k, k+s, k+2s, k+3s}, s = 2^(pn-2), by Cramer's rule
    (cz_solver.f90:788-844, 1177-1225, 1391-1438)."""
    t = dsrc.dtype.type
    n = dsrc.shape[1]
    sq = 2 ** (pn - 2)
    P = 4 * sq
    a, c, d = _pcr_stages(dsrc, pn - 2, P)
    k = np.arange(P, P + sq)
    kl, km, kr = k + sq, k + 2 * sq, k + 3 * sq
    cc1, cc2, cc3 = c[k], c[kl], c[km]
    aa2, aa3, aa4 = a[kl], a[km], a[kr]
    dd1, dd2, dd3, dd4 = d[:, k], d[:, kl], d[:, km], d[:, kr]
    one = t(1.0)
    inv = one / (one - aa4 * cc3 - aa3 * cc2 - aa2 * cc1 * (one - cc3 * aa4))
    det1 = -(cc3 * (aa4 * dd1 + cc1 * cc2 * dd4 - aa4 * cc1 * dd2)) + dd1 + cc1 * cc2 * dd3 - aa3 * cc2 * dd1 - cc1 * dd2
    det2 = dd2 + cc2 * cc3 * dd4 - aa4 * cc3 * dd2 - cc2 * dd3 - aa2 * (dd1 - aa4 * cc3 * dd1)
    det3 = dd3 - cc3 * dd4 - aa3 * dd2 - aa2 * (cc1 * dd3 - cc1 * cc3 * dd4 - aa3 * dd1)
    det4 = dd4 + aa3 * aa4 * dd2 - aa4 * dd3 - aa3 * cc2 * dd4 - aa2 * (cc1 * dd4 + aa3 * aa4 * dd1 - aa4 * cc1 * dd3)
    d1 = np.zeros_like(d)
    d1[:, k] = det1 * inv; d1[:, kl] = det2 * inv; d1[:, km] = det3 * inv; d1[:, kr] = det4 * inv
    return d1[:, P:P + n]


def _relax_lines(x, dst, msk, rhs, idx, pn, omg, jc, ic, four):
    """Source term, line solve and relaxation (cz_solver.f90:642-648) of the lines (jc, ic): reads x, writes dst."""
    t = x.dtype.type
    d, mk, ks = _line_source(x, msk, rhs, idx, jc, ic)
    d1 = _solve_4x4(d, pn) if four else _solve_2x2(d, pn)
    pp = x[jc, ic, ks]
    dp = (d1 - pp) * t(omg) * mk
    dst[jc, ic, ks] = pp + dp
    return float(np.sum((dp * dp).astype(np.float64)))


def _lines(idx, color=None):
    ist, ied, jst, jed = idx[:4]
    o = G - 1
    J, I = np.meshgrid(np.arange(jst, jed + 1), np.arange(ist, ied + 1), indexing="ij")
    m = np.ones_like(J, dtype=bool) if color is None else ((I + J) % 2) == color   # :549 mod(i+j,2) /= color -> cycle
    return J[m] + o, I[m] + o


def pcr_rb(x, msk, rhs, idx, pn, color, omg):
    """cz_solver.f90:497-662: the lines of one colour only read lines of the other colour -> all at once, in place."""
    jc, ic = _lines(idx, color)
    return _relax_lines(x, x, msk, rhs, idx, pn, omg, jc, ic, False)


def pcr_rb_esa(x, msk, rhs, idx, pn, color, omg):
    """cz_solver.f90:1261-1469: as pcr_rb with pn-2 stages and the 4x4 last stage."""
    jc, ic = _lines(idx, color)
    return _relax_lines(x, x, msk, rhs, idx, pn, omg, jc, ic, True)


def pcr_j_esa(x, msk, rhs, idx, pn, omg):
    """cz_solver.f90:1473-1676: every line from the OLD field (source term to `src`, result to `wrk`, copy back)."""
    jc, ic = _lines(idx)
    wrk = x.copy()
    res = _relax_lines(x, wrk, msk, rhs, idx, pn, omg, jc, ic, False)
    x[...] = wrk
    return res


def _lexicographic(x, msk, rhs, idx, pn, omg, four):
    """pcr :666-872 (4x4), pcr_esa :1050-1257 (4x4, padded), pcr_eda :883-1046 (2x2): the lines one after the
    other in (j, i) order, in place (Gauss-Seidel between lines; the reference's 1-thread order)."""
    jc, ic = _lines(idx)
    res = 0.0
    for q in range(len(jc)):
        res += _relax_lines(x, x, msk, rhs, idx, pn, omg, jc[q:q + 1], ic[q:q + 1], four)
    return res


def pcr(x, msk, rhs, idx, pn, omg):
    return _lexicographic(x, msk, rhs, idx, pn, omg, True)


def pcr_esa(x, msk, rhs, idx, pn, omg):
    return _lexicographic(x, msk, rhs, idx, pn, omg, True)


def pcr_eda(x, msk, rhs, idx, pn, omg):
    return _lexicographic(x, msk, rhs, idx, pn, omg, False)


def psor(p, b, idx, cf, omg):
    """cz_solver.f90:207-269: lexicographic SOR, cell by cell in (j, i, k) order (scalar loop: small grids only)."""
    t = p.dtype.type
    ist, ied, jst, jed, kst, ked = idx
    o = G - 1
    c = [t(v) for v in cf]
    om = t(omg)
    res = 0.0
    for j in range(jst + o, jed + o + 1):
        for i in range(ist + o, ied + o + 1):
            for k in range(kst + o, ked + o + 1):
                pp = p[j, i, k]
                ss = c[0] * p[j, i + 1, k]
                ss = ss + c[1] * p[j, i - 1, k]
                ss = ss + c[2] * p[j + 1, i, k]
                ss = ss + c[3] * p[j - 1, i, k]
                ss = ss + c[4] * p[j, i, k + 1]
                ss = ss + c[5] * p[j, i, k - 1]
                dp = ((ss - b[j, i, k]) / c[6] - pp) * om
                p[j, i, k] = pp + dp
                res += float(dp * dp)
    return res


def blas_triad(z, x, y, a, idx):
    """cz_blas.f90:255-307  z = a*x + y on the inner cells."""
    sl = _rng(idx)
    z[sl] = z.dtype.type(a) * x[sl] + y[sl]


def blas_bicg_1(p, r, q, beta, omg, idx):
    """cz_blas.f90:452-504  p = r + beta*(p - omg*q)."""
    t = p.dtype.type
    sl = _rng(idx)
    p[sl] = r[sl] + t(beta) * (p[sl] - t(omg) * q[sl])


def blas_bicg_2(z, x, y, a, b, idx):
    """cz_blas.f90:517-569  z = a*x + b*y + z."""
    t = z.dtype.type
    sl = _rng(idx)
    z[sl] = t(a) * x[sl] + t(b) * y[sl] + z[sl]


# ======================================================================================================
# Stretched-grid ("MAF") twins, src/cz_f90/cz_maf.f90 and cz_blas.f90:738-1039.  Coordinate arrays are the
# Fortran X(-1:sz(1)+2) etc.: Fortran index m -> C offset m + 1.
# ======================================================================================================
def _maf_metrics(X, Y, Z, idx):
    """C1, C2, C3, C7, C8, C9 of every inner cell as [j][i][k] arrays, cz_maf.f90:70-99 statement by statement;
    the products involve factors from all three axes, so every cell rounds on its own."""
    t = X.dtype.type
    ist, ied, jst, jed, kst, ked = idx
    i = np.arange(ist, ied + 1)[None, :, None]
    j = np.arange(jst, jed + 1)[:, None, None]
    k = np.arange(kst, ked + 1)[None, None, :]
    half, two, one = t(0.5), t(2.0), t(1.0)
    XG = half * (X[i + 2] - X[i]); YE = half * (Y[j + 2] - Y[j]); ZT = half * (Z[k + 2] - Z[k])
    XGG = X[i + 2] - two * X[i + 1] + X[i]
    YEE = Y[j + 2] - two * Y[j + 1] + Y[j]
    ZTT = Z[k + 2] - two * Z[k + 1] + Z[k]
    YJA = XG * YE * ZT
    YJAI = one / YJA
    GX = YE * ZT * YJAI; EY = XG * ZT * YJAI; TZ = XG * YE * YJAI
    C1 = GX * GX; C2 = EY * EY; C3 = TZ * TZ
    C7 = -XGG * C1 * GX; C8 = -YEE * C2 * EY; C9 = -ZTT * C3 * TZ
    return C1, C2, C3, C7, C8, C9


def _maf_offdiag(p, sl, m):
    """(C1+.5C7) p(i+1) + (C1-.5C7) p(i-1) + (C2+.5C8) p(j+1) + (C2-.5C8) p(j-1) + (C3+.5C9) p(k+1) + (C3-.5C9) p(k-1),
    left to right (cz_maf.f90:103-108)."""
    C1, C2, C3, C7, C8, C9 = m
    h = p.dtype.type(0.5)
    s = (C1 + h * C7) * p[_shift(sl, di=1)]
    s = s + (C1 - h * C7) * p[_shift(sl, di=-1)]
    s = s + (C2 + h * C8) * p[_shift(sl, dj=1)]
    s = s + (C2 - h * C8) * p[_shift(sl, dj=-1)]
    s = s + (C3 + h * C9) * p[_shift(sl, dk=1)]
    s = s + (C3 - h * C9) * p[_shift(sl, dk=-1)]
    return s


def _maf_dp(p, b, sl, m, omg):
    t = p.dtype.type
    dd = t(2.0) * (m[0] + m[1] + m[2])                     # :102
    rp = _maf_offdiag(p, sl, m) + b[sl]                    # :103-109, the source term is ADDED
    return (rp / dd - p[sl]) * t(omg)                      # :110


def jacobi_maf(p, b, idx, X, Y, Z, omg):
    """cz_maf.f90:131-282."""
    sl = _rng(idx)
    dp = _maf_dp(p, b, sl, _maf_metrics(X, Y, Z, idx), omg)
    p[sl] = p[sl] + dp
    return float(np.sum((dp * dp).astype(np.float64)))


def psor2sma_core_maf(p, b, idx, X, Y, Z, ofst, color, omg):
    """cz_maf.f90:301-438."""
    ist, ied, jst, jed, kst, ked = idx
    sl = _rng(idx)
    kp = ofst + color
    jj, ii, kk = np.meshgrid(np.arange(jst, jed + 1), np.arange(ist, ied + 1), np.arange(kst, ked + 1), indexing="ij")
    sel = ((kk - kst) % 2) == ((ii + jj + kp) % 2)
    dp = _maf_dp(p, b, sl, _maf_metrics(X, Y, Z, idx), omg)
    p[sl] = np.where(sel, p[sl] + dp, p[sl])
    return float(np.sum((dp[sel] * dp[sel]).astype(np.float64)))


def psor_maf(p, b, idx, X, Y, Z, omg):
    """cz_maf.f90:23-114: lexicographic, cell by cell in (j, i, k) order (scalar loop: small grids only)."""
    t = p.dtype.type
    ist, ied, jst, jed, kst, ked = idx
    o = G - 1
    C1, C2, C3, C7, C8, C9 = _maf_metrics(X, Y, Z, idx)
    h, two, om = t(0.5), t(2.0), t(omg)
    res = 0.0
    for j in range(jst, jed + 1):
        for i in range(ist, ied + 1):
            for k in range(kst, ked + 1):
                q = (j - jst, i - ist, k - kst)
                c1, c2, c3, c7, c8, c9 = C1[q], C2[q], C3[q], C7[q], C8[q], C9[q]
                J, I, K = j + o, i + o, k + o
                pp = p[J, I, K]
                dd = two * (c1 + c2 + c3)
                rp = (c1 + h * c7) * p[J, I + 1, K]
                rp = rp + (c1 - h * c7) * p[J, I - 1, K]
                rp = rp + (c2 + h * c8) * p[J + 1, I, K]
                rp = rp + (c2 - h * c8) * p[J - 1, I, K]
                rp = rp + (c3 + h * c9) * p[J, I, K + 1]
                rp = rp + (c3 - h * c9) * p[J, I, K - 1]
                rp = rp + b[J, I, K]
                dp = (rp / dd - pp) * om
                p[J, I, K] = pp + dp
                res += float(dp * dp)
    return res


def calc_ax_maf(ap, p, idx, X, Y, Z, pvt):
    """cz_blas.f90:845-935."""
    t = p.dtype.type
    sl = _rng(idx)
    m = _maf_metrics(X, Y, Z, idx)
    ap[sl] = (_maf_offdiag(p, sl, m) - t(2.0) * (m[0] + m[1] + m[2]) * p[sl]) * pvt[sl]


def calc_rk_maf(r, p, b, idx, X, Y, Z, pvt):
    """cz_blas.f90:738-829: b + 2(C1+C2+C3) p - ... , each neighbour product subtracted in turn."""
    t = p.dtype.type
    sl = _rng(idx)
    C1, C2, C3, C7, C8, C9 = _maf_metrics(X, Y, Z, idx)
    h = t(0.5)
    s = b[sl] + t(2.0) * (C1 + C2 + C3) * p[sl]
    s = s - (C1 + h * C7) * p[_shift(sl, di=1)]
    s = s - (C1 - h * C7) * p[_shift(sl, di=-1)]
    s = s - (C2 + h * C8) * p[_shift(sl, dj=1)]
    s = s - (C2 - h * C8) * p[_shift(sl, dj=-1)]
    s = s - (C3 + h * C9) * p[_shift(sl, dk=1)]
    s = s - (C3 - h * C9) * p[_shift(sl, dk=-1)]
    r[sl] = s * pvt[sl]


def search_pivot(pvt, idx, X, Y, Z):
    """cz_blas.f90:947-1039: reciprocal of the largest |coefficient| of the row."""
    t = X.dtype.type
    C1, C2, C3, C7, C8, C9 = _maf_metrics(X, Y, Z, idx)
    h = t(0.5)
    cand = [np.abs(C1 + h * C7), np.abs(C1 - h * C7), np.abs(C2 + h * C8), np.abs(C2 - h * C8),
            np.abs(C3 + h * C9), np.abs(C3 - h * C9), np.abs(t(2.0) * (C1 + C2 + C3))]
    ss = cand[0]
    for c in cand[1:]:
        ss = np.maximum(ss, c)
    pvt[_rng(idx)] = t(1.0) / ss


def _relax_lines_maf(x, msk, rhs, idx, pn, omg, jc, ic, XX, YY, ZZ):
    """One batch of MAF lines (jc, ic) (C indices), cz_maf.f90:522-647: line-dependent coefficients a, c are carried
    through the PCR stages next to d; pn-1 stages + 2x2 in all five MAF line routines."""
    t = x.dtype.type
    kst, ked = idx[4], idx[5]
    o = G - 1
    n = ked - kst + 1
    L = len(jc)
    half, two, one = t(0.5), t(2.0), t(1.0)
    fi = ic - o; fj = jc - o                               # Fortran i, j of each line
    GX = two / (XX[fi + 2] - XX[fi]); EY = two / (YY[fj + 2] - YY[fj])
    C1 = GX * GX; C2 = EY * EY
    C7 = -(XX[fi + 2] - two * XX[fi + 1] + XX[fi]) * C1 * GX
    C8 = -(YY[fj + 2] - two * YY[fj + 1] + YY[fj]) * C2 * EY
    dd1 = (C1 + half * C7)[:, None]; dd2 = (C1 - half * C7)[:, None]
    cc1 = (C2 + half * C8)[:, None]; cc2 = (C2 - half * C8)[:, None]
    fk = np.arange(kst, ked + 1)
    f1 = ZZ[fk + 2]; f2 = ZZ[fk]
    TZ = two / (f1 - f2)
    ZTT = f1 - two * ZZ[fk + 1] + f2
    f3 = TZ * TZ
    aw = np.broadcast_to(f3[None, :], (L, n))              # C3
    cw = np.broadcast_to((-ZTT * f3 * TZ)[None, :], (L, n))   # C9
    dw = half / ((C1 + C2)[:, None] + f3[None, :])         # 1/R7
    s = 2 ** (pn - 1)
    P = 2 * s
    a = np.zeros((L, n + 2 * P), dtype=x.dtype); c = np.zeros_like(a); d = np.zeros_like(a)
    av = -(aw - half * cw) * dw; cv = -(aw + half * cw) * dw
    a[:, P:P + n] = av; c[:, P:P + n] = cv
    a[:, P] = 0; c[:, P + n - 1] = 0                       # :545, :557
    ks = slice(kst + o, ked + o + 1)
    mk = msk[jc, ic, ks]
    dk = (dd1 * x[jc, ic + 1, ks] + dd2 * x[jc, ic - 1, ks] + cc1 * x[jc + 1, ic, ks] + cc2 * x[jc - 1, ic, ks]
          - rhs[jc, ic, ks]) * dw * mk                     # :560-568, the source term is SUBTRACTED here
    dk[:, 0] = (dk[:, 0] + (aw[:, 0] - half * cw[:, 0]) * dw[:, 0] * x[jc, ic, kst + o - 1]) * mk[:, 0]
    dk[:, n - 1] = (dk[:, n - 1] + (aw[:, n - 1] + half * cw[:, n - 1]) * dw[:, n - 1] * x[jc, ic, ked + o + 1]) * mk[:, n - 1]
    d[:, P:P + n] = dk
    k = np.arange(P, P + n)
    for p in range(1, pn):                                 # :580-601
        q = 2 ** (p - 1)
        ap = a[:, k]; cp = c[:, k]
        e = one / (one - ap * c[:, k - q] - cp * a[:, k + q])
        a1 = -e * ap * a[:, k - q]
        c1 = -e * cp * c[:, k + q]
        d1 = e * (d[:, k] - ap * d[:, k - q] - cp * d[:, k + q])
        a[:, k] = a1; c[:, k] = c1; d[:, k] = d1
    k = np.arange(P, P + s)                                # :605-620
    cc = c[:, k]; aa = a[:, k + s]; g1 = d[:, k]; g2 = d[:, k + s]
    jj = one / (one - aa * cc)
    sol = np.zeros_like(d)
    sol[:, k] = (g1 - cc * g2) * jj
    sol[:, k + s] = (g2 - aa * g1) * jj
    pp = x[jc, ic, ks]
    dp = (sol[:, P:P + n] - pp) * t(omg) * mk              # :630-637
    x[jc, ic, ks] = pp + dp
    return float(np.sum((dp * dp).astype(np.float64)))   # :645 (`tmp(k)` at :643 is the other branch of #ifdef _SVR)


def pcr_rb_maf(x, msk, rhs, idx, pn, color, omg, XX, YY, ZZ):
    """cz_maf.f90:442-668 (and pcr_rb_esa_maf :1339-1560, same arithmetic on padded arrays)."""
    jc, ic = _lines(idx, color)
    return _relax_lines_maf(x, msk, rhs, idx, pn, omg, jc, ic, XX, YY, ZZ)


def pcr_maf(x, msk, rhs, idx, pn, omg, XX, YY, ZZ):
    """cz_maf.f90:672-892 (pcr_eda_maf :896-1113, pcr_esa_maf :1117-1335: same arithmetic): lines in (j, i) order."""
    jc, ic = _lines(idx)
    res = 0.0
    for q in range(len(jc)):
        res += _relax_lines_maf(x, msk, rhs, idx, pn, omg, jc[q:q + 1], ic[q:q + 1], XX, YY, ZZ)
    return res
