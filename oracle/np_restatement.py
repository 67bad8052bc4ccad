"""oracle/np_restatement.py -- TEST INFRASTRUCTURE, never imported by the product.

A second, independent restatement of the reference's hot-path kernels in numpy, written directly from the
Fortran sources (file:line cited per function), vectorised over everything the reference's loop nest leaves
independent.  Its only job is to pin the C++ oracle port (oracle/cz_f90_port.cpp): two restatements written
separately, in different languages and with different loop structures, must agree bit for bit on random input
(tests/test_oracle.py).  numpy's elementwise float32 / float64 arithmetic is IEEE, unfused and evaluated in the
order written, which is what the reference's un-contracted CPU build does per cell.  The Fortran cannot be
compiled in this image (no Fortran compiler), so this is the strongest check of the port available here;
parity with reference OUTPUTS stays unpinned, see DESIGN.md section 2.

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


def pcr_rb(x, msk, rhs, idx, pn, color, omg):
    """cz_solver.f90:497-662, all lines of one colour at once ([line][k] arrays; a and c are re-initialised
    per line to the same constants, :556-564, so they stay 1-D).  Work arrays carry one zero slot on each
    side: the reference clamps k-s / k+s to kst-1 / ked+1, slots it never writes (:588-589)."""
    t = x.dtype.type
    ist, ied, jst, jed, kst, ked = idx
    o = G - 1
    n = ked - kst + 1
    r = t(1.0) / t(6.0)
    J, I = np.meshgrid(np.arange(jst, jed + 1), np.arange(ist, ied + 1), indexing="ij")
    m = ((I + J) % 2) == color                             # :549  if(mod(i+j,2) /= color) cycle
    jc, ic = J[m] + o, I[m] + o                            # C indices of the selected lines
    ks = slice(kst + o, ked + o + 1)
    a = np.zeros(n + 2, dtype=x.dtype); c = np.zeros(n + 2, dtype=x.dtype)
    a[2:n + 1] = -r                                        # a(kst) = 0, a(kst+1..ked) = -r
    c[1:n] = -r                                            # c(kst..ked-1) = -r, c(ked) = 0
    mk = msk[jc, ic, ks]
    d = np.zeros((len(jc), n + 2), dtype=x.dtype)
    d[:, 1:n + 1] = ((x[jc - 1, ic, ks] + x[jc + 1, ic, ks] + x[jc, ic - 1, ks] + x[jc, ic + 1, ks]
                      - rhs[jc, ic, ks]) * r) * mk         # :569-575
    d[:, 1] = (d[:, 1] + x[jc, ic, kst + o - 1] * r) * mk[:, 0]          # :578
    d[:, n] = (d[:, n] + x[jc, ic, ked + o + 1] * r) * mk[:, n - 1]      # :579
    k = np.arange(1, n + 1)
    for p in range(1, pn):                                 # :584-608
        s = 2 ** (p - 1)
        kl = np.maximum(k - s, 0); kr = np.minimum(k + s, n + 1)
        ap = a[k]; cp = c[k]
        e = t(1.0) / (t(1.0) - ap * c[kl] - cp * a[kr])
        a1 = -e * ap * a[kl]
        c1 = -e * cp * c[kr]
        d1 = e * (d[:, k] - ap * d[:, kl] - cp * d[:, kr])
        a[k] = a1; c[k] = c1; d[:, k] = d1
    s = 2 ** (pn - 1)                                      # :612-630
    d1 = np.zeros_like(d)
    k = np.arange(1, s + 1)
    kr = np.minimum(k + s, n + 1)
    cc1 = c[k]; aa2 = a[kr]
    f1 = d[:, k]; f2 = d[:, kr]
    jj = t(1.0) / (t(1.0) - aa2 * cc1)
    dd1 = (f1 - cc1 * f2) * jj
    dd2 = (f2 - aa2 * f1) * jj
    # the reference stores d1(k) and d1(kr) pair by pair in ascending k; a clamped kr = ked+1 is a guard slot that
    # is rewritten by later pairs and never read, so the pairwise order does not matter
    d1[:, k] = dd1
    d1[:, kr] = dd2
    pp = x[jc, ic, ks]                                     # :642-648
    dp = (d1[:, 1:n + 1] - pp) * t(omg) * mk
    x[jc, ic, ks] = pp + dp
    return float(np.sum((dp * dp).astype(np.float64)))


def pcr_rb_esa(x, msk, rhs, idx, pn, color, omg):
    """cz_solver.f90:1261-1469: pn-2 PCR stages on zero-padded work arrays (no index clamps), then the 4x4
    systems {k, k+s, k+2s, k+3s}, s = 2^(pn-2), by Cramer's rule (:1391-1438).  The pads are wide enough for
    every index the last stage forms (the reference pads by s and relies on zeroed work arrays, cz_Poisson.cpp:655-661)."""
    t = x.dtype.type
    ist, ied, jst, jed, kst, ked = idx
    o = G - 1
    n = ked - kst + 1
    r = t(1.0) / t(6.0)
    J, I = np.meshgrid(np.arange(jst, jed + 1), np.arange(ist, ied + 1), indexing="ij")
    m = ((I + J) % 2) == color
    jc, ic = J[m] + o, I[m] + o
    ks = slice(kst + o, ked + o + 1)
    sq = 2 ** (pn - 2)
    P = 4 * sq                                             # pad; cell k (0-based) lives at index P + k
    a = np.zeros(n + 2 * P, dtype=x.dtype); c = np.zeros(n + 2 * P, dtype=x.dtype)
    a[P + 1:P + n] = -r                                    # :1311-1319
    c[P:P + n - 1] = -r                                    # :1320-1328
    mk = msk[jc, ic, ks]
    d = np.zeros((len(jc), n + 2 * P), dtype=x.dtype)
    d[:, P:P + n] = ((x[jc - 1, ic, ks] + x[jc + 1, ic, ks] + x[jc, ic - 1, ks] + x[jc, ic + 1, ks]
                      - rhs[jc, ic, ks]) * r) * mk
    d[:, P] = (d[:, P] + x[jc, ic, kst + o - 1] * r) * mk[:, 0]
    d[:, P + n - 1] = (d[:, P + n - 1] + x[jc, ic, ked + o + 1] * r) * mk[:, n - 1]
    k = np.arange(P, P + n)
    for p in range(1, pn - 1):                             # :1346-1363
        s = 2 ** (p - 1)
        ap = a[k]; cp = c[k]
        e = t(1.0) / (t(1.0) - ap * c[k - s] - cp * a[k + s])
        a1 = -e * ap * a[k - s]
        c1 = -e * cp * c[k + s]
        d1 = e * (d[:, k] - ap * d[:, k - s] - cp * d[:, k + s])
        a[k] = a1; c[k] = c1; d[:, k] = d1
    k = np.arange(P, P + sq)                               # :1365-1438
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
    pp = x[jc, ic, ks]
    dp = (d1[:, P:P + n] - pp) * t(omg) * mk
    x[jc, ic, ks] = pp + dp
    return float(np.sum((dp * dp).astype(np.float64)))
